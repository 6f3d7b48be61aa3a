"""GPU parity: the CUDA phasing scorer against the oracle — integer bit-exact (north_star)."""
import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = N.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("key", ["test_var", "toy_varblock"])
def test_golden_fixtures(ctx, golden, golden_files, key):
    l = O.Links.load(golden_files[key])
    a, b, w = l.export()
    n = l.size()
    ctx.set_phase_links(n, a, b, w)
    genes = np.array([c["gene"] for c in golden[key]["cases"]], np.uint8)
    want = np.array([c["score"] for c in golden[key]["cases"]])
    oi, of = ctx.phase_score_batch(genes)
    assert of.tolist() == want.tolist()
    assert oi.tolist() == [int(v) for v in want]


def test_all_bitstrings_of_the_reference_fixture(ctx, golden_files):
    l = O.Links.load(golden_files["test_var"])
    a, b, w = l.export()
    ctx.set_phase_links(11, a, b, w)
    genes = np.array([[m >> i & 1 for i in range(11)] for m in range(1 << 11)], np.uint8)
    ref = l.phase_score_batch(genes)                 # the reference's loop
    oi, of = ctx.phase_score_batch(genes)
    assert of.tolist() == ref.tolist()
    assert of.min() == -679.0                        # the global optimum


@pytest.mark.parametrize("n,P,seed,far", [(50, 10, 1, 0), (1000, 128, 2, 500), (4097, 33, 3, 100),
                                          (25000, 64, 5, 5000),
                                          (4100, 130, 7, 300),    # 5 groups: a block of 4 + a block with 3 absent
                                          (2048, 200, 8, 0), (1001, 97, 9, 50)])
def test_synthetic_blocks_bit_exact(ctx, n, P, seed, far):
    a, b, w, truth = S.phase_links(n, seed, far_pairs=far)
    genes = S.phase_population(n, P, seed + 100, truth)
    _, exact = O.phase_score_links_batch(n, a, b, w, genes, nthreads=4)
    ctx.set_phase_links(n, a, b, w)
    oi, of = ctx.phase_score_batch(genes)
    assert oi.tolist() == exact.tolist()
    assert of.tolist() == exact.astype(np.float64).tolist()
    if n <= 4097:                                    # the reference-order loop (map look-ups)
        table = O.Links.from_records(n, a, b, w)
        ref = table.phase_score_batch(genes[:16], nthreads=4)
        assert of[:16].tolist() == ref.tolist()
    order = N.rank(of)
    assert order[0] in (0, 1)                        # the truth and its complement tie for best
    assert of[0] == of[1]


def test_large_weights_take_the_64_bit_chunk_sums(ctx):
    """|w| up to 2^30 - 1 (the setter's limit): a chunk of 2048 records no longer fits 32 bits, the
    kernel's wide variant sums it in 64; small weights take the 32-bit variant.  Both exact."""
    n, P = 3000, 70
    a, b, w, truth = S.phase_links(n, 11, far_pairs=50)
    genes = S.phase_population(n, P, 111, truth)
    rng = np.random.default_rng(11)
    for scale in (1, (1 << 19) - 1, (1 << 30) - 1):
        ws = np.sign(w) * rng.integers(scale // 2 + 1, scale + 1, len(w)).astype(np.float64)
        _, exact = O.phase_score_links_batch(n, a, b, ws, genes, nthreads=4)
        ctx.set_phase_links(n, a, b, ws)
        oi, of = ctx.phase_score_batch(genes)
        assert oi.tolist() == exact.tolist()


def test_edge_cases(ctx):
    ctx.set_phase_links(5, np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0))
    oi, of = ctx.phase_score_batch(np.ones((3, 5), np.uint8))
    assert oi.tolist() == [0, 0, 0]
    # head quirk: pair (0,1) counts once, pair (1,2) twice
    ctx.set_phase_links(3, [0, 2], [1, 1], [3.0, 5.0])
    oi, _ = ctx.phase_score_batch(np.array([[0, 0, 0], [0, 1, 0], [1, 1, 0], [2, 2, 7]], np.uint8))
    assert oi.tolist() == [-13, 13, 7, -13]          # any non-zero byte is `true`
    with pytest.raises(N.LxyError):
        ctx.set_phase_links(3, [0], [3], [1.0])      # id out of range
    with pytest.raises(N.LxyError):
        ctx.set_phase_links(3, [0], [1], [0.5])      # phasing weights are integer counts


def test_full_size_properties(ctx):
    """Config 5 at N=300k, P=1024: complement invariance and linearity in the link set."""
    n, P = 300_000, 1024
    a, b, w, truth = S.phase_links(n, 5)
    genes = S.phase_population(n, P, 105, truth)
    ctx.set_phase_links(n, a, b, w)
    oi, _ = ctx.phase_score_batch(genes)
    oc, _ = ctx.phase_score_batch(1 - genes)
    assert oi.tolist() == oc.tolist()                # s(i,j) is invariant under complement
    h = len(a) // 2
    ctx.set_phase_links(n, a[:h], b[:h], w[:h])
    o1, _ = ctx.phase_score_batch(genes)
    ctx.set_phase_links(n, a[h:], b[h:], w[h:])
    o2, _ = ctx.phase_score_batch(genes)
    assert (o1 + o2).tolist() == oi.tolist()
    _, exact = O.phase_score_links_batch(n, a, b, w, genes[:4], nthreads=4)
    assert oi[:4].tolist() == exact.tolist()
