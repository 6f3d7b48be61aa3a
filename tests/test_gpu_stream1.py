"""GPU parity of the streaming-regime kernel (k_scaff_stream1, batches of <= 8 orderings: the
reference's literal boundary is ONE ordering per score() call, scaff/genome.go:153-159).

Checked against the reference-order oracle (scaff.go:246-310 restated), against the exact integer
sums of the link-centric oracle, and against the batch kernel (LXY_FLAG_NO_STREAM1) bit for bit.
"""
import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu
REL = 1e-9


@pytest.fixture(scope="module")
def ctxs():
    s, b = N.Context(0, path=N.PATH_STREAM), N.Context(0, path=N.PATH_STREAM, flags=N.FLAG_NO_STREAM1)
    yield s, b
    s.close()
    b.close()


@pytest.mark.parametrize("n,L,P,seed", [
    (11, 55, 1, 1), (300, 40000, 1, 2), (300, 40000, 2, 3), (777, 90000, 3, 4), (1500, 200000, 4, 5),
    (2049, 300000, 5, 6), (5000, 400000, 8, 7), (256, 4097, 7, 8), (257, 511, 1, 9),
])
def test_stream1_unit_weights(ctxs, n, L, P, seed):
    cs, cb = ctxs
    a, b, w = S.scaff_links(n, L, seed)
    perms = S.scaff_population(n, P, seed + 100)
    table = O.Links.from_records(n, a, b, w)
    ref = table.scaff_score_batch(perms, nthreads=4)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=4)
    cs.set_links(n, a, b, None)
    cb.set_links(n, a, b, None)
    got, x100 = cs.scaff_score_batch(perms, want_x100=True)
    assert cs.timing()["launches"] == 3                 # head + stream1 + finish: the new path ran
    gb, xb = cb.scaff_score_batch(perms, want_x100=True)
    assert x100.tolist() == ox100.tolist()
    assert x100.tolist() == xb.tolist() and got.tolist() == gb.tolist()
    np.testing.assert_allclose(got, ref, rtol=REL, atol=0)


@pytest.mark.parametrize("P", [1, 2, 6])
def test_stream1_integer_and_real_weights(ctxs, P):
    cs, cb = ctxs
    n, L = 900, 120000
    a, b, _ = S.scaff_links(n, L, 21)
    rng = np.random.default_rng(21)
    perms = S.scaff_population(n, P, 121)
    wi = rng.integers(-50, 2000, L).astype(np.float64)
    _, ox100 = O.scaff_score_links_batch(n, a, b, wi, perms, nthreads=4)
    cs.set_links(n, a, b, wi)
    got, x100 = cs.scaff_score_batch(perms, want_x100=True)
    assert x100.tolist() == ox100.tolist()
    ref = O.Links.from_records(n, a, b, wi).scaff_score_batch(perms, nthreads=4)
    np.testing.assert_allclose(got, ref, rtol=REL, atol=1e-9)
    wr = rng.gamma(2.0, 1.7, L)
    ref = O.Links.from_records(n, a, b, wr).scaff_score_batch(perms, nthreads=4)
    cs.set_links(n, a, b, wr)
    got, x100 = cs.scaff_score_batch(perms, want_x100=True)
    assert (x100 == N.INT64_MIN).all()
    np.testing.assert_allclose(got, ref, rtol=REL, atol=0)


def test_stream1_u16_table_limit(ctxs):
    """n = 65536 is the largest placement table the u16 format holds (pos <= 65535); 65537 falls
    back to the batch kernel.  Both agree with the oracle's exact integers."""
    cs, _ = ctxs
    for n, launches in ((65536, 3), (65537, None)):
        L = 300000
        a, b, w = S.scaff_links(n, L, 31)
        perms = S.scaff_population(n, 1, 131)
        _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=4)
        cs.set_links(n, a, b, None)
        _, x100 = cs.scaff_score_batch(perms, want_x100=True)
        if launches:
            assert cs.timing()["launches"] == launches
        else:
            assert cs.timing()["launches"] != 3
        assert x100.tolist() == ox100.tolist()


def test_stream1_rejects_bad_orderings(ctxs):
    cs, _ = ctxs
    n = 64
    a, b, _ = S.scaff_links(n, 3000, 41)
    cs.set_links(n, a, b, None)
    perms = S.scaff_population(n, 2, 141)
    bad = perms.copy()
    bad[1, 3] = bad[1, 4]
    with pytest.raises(N.LxyError) as e:
        cs.scaff_score_batch(bad)
    assert e.value.status == -4
    bad = perms.copy()
    bad[0, 0] = -1
    with pytest.raises(N.LxyError):
        cs.scaff_score_batch(bad)
    assert np.isfinite(cs.scaff_score_batch(perms)).all()


@pytest.mark.parametrize("P", [1, 2, 8])
def test_stream1_duplicate_far_apart_is_always_rejected(ctxs, P):
    """A contig id repeated at ordering positions that lie in different warps and different
    validation slices (every CTA builds its own table with unordered stores, so which duplicate wins
    the slot differs from CTA to CTA): the check goes by contig id — the id the duplicate displaced
    was never written — and must fire on every one of many repeats, for both streaming kernels."""
    cs, _ = ctxs
    n = 5000
    a, b, _ = S.scaff_links(n, 400_000, 43)          # enough records for one CTA per 512: many CTAs
    cs.set_links(n, a, b, None)
    rng = np.random.default_rng(143)
    perms = S.scaff_population(n, P, 143)
    assert np.isfinite(cs.scaff_score_batch(perms)).all()
    for rep in range(40):
        bad = perms.copy()
        row = int(rng.integers(0, P))
        i1 = int(rng.integers(0, n // 4))
        i2 = int(rng.integers(3 * n // 4, n))
        bad[row, i2] = bad[row, i1]
        with pytest.raises(N.LxyError) as e:
            cs.scaff_score_batch(bad)
        assert e.value.status == -4, rep


def test_stream1_full_size_single_genome_score(ctxs):
    """A lone Score() call at BASELINE config-3 size (5,000 contigs / 200 M records): the streaming
    kernel and the batch kernel return the same exact integer, and the hidden true ordering beats
    a random one."""
    torch = pytest.importorskip("torch")
    import bench
    cs, cb = ctxs
    n, L = 5000, 200_000_000
    a, b, order = bench.gen_links(n, L, 3, torch.device("cuda", 0))
    cs.set_links_dev(n, L, a.data_ptr(), b.data_ptr(), None)
    cb.set_links_dev(n, L, a.data_ptr(), b.data_ptr(), None)
    del a, b
    true = order.cpu().numpy().astype(np.int32)[None, :]
    rnd = np.random.default_rng(5).permutation(n).astype(np.int32)[None, :]
    for g in (true, rnd):
        s, xs = cs.scaff_score_batch(g, want_x100=True)
        _, xb = cb.scaff_score_batch(g, want_x100=True)
        assert xs.tolist() == xb.tolist()
    assert cs.scaff_score_batch(true)[0] < cs.scaff_score_batch(rnd)[0]


@pytest.mark.parametrize("n,L,P,passes", [(300, 3_000_000, 9, 2), (300, 3_000_000, 16, 2), (700, 6_000_000, 13, 2),
                                          (700, 20_000_000, 24, 3), (5000, 20_000_000, 17, 3)])
def test_small_populations_take_several_passes(ctxs, n, L, P, passes):
    """Populations above 8 run as passes of <= 8 orderings while that beats the batch kernel (the
    reference's own GA holds 40 orderings, scaff.go:94): same integers as the batch kernel and the
    oracle, the error flags survive the passes, a repeated call (graph replay) returns the same."""
    cs, cb = ctxs
    a, b, w = S.scaff_links(n, L, 31)
    perms = S.scaff_population(n, P, 131)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=8)
    cs.set_links(n, a, b, None)
    cb.set_links(n, a, b, None)
    got, x100 = cs.scaff_score_batch(perms, want_x100=True)
    assert cs.timing()["launches"] == 3 * passes
    gb, xb = cb.scaff_score_batch(perms, want_x100=True)
    assert x100.tolist() == ox100.tolist()
    assert x100.tolist() == xb.tolist() and got.tolist() == gb.tolist()
    for _ in range(3):                                      # the second identical call is captured, then replayed
        g2, x2 = cs.scaff_score_batch(perms, want_x100=True)
        assert x2.tolist() == ox100.tolist() and g2.tolist() == got.tolist()
    bad = perms.copy()
    bad[P - 1, 5] = bad[P - 1, 6]                           # a repeated id in the LAST pass
    with pytest.raises(N.LxyError):
        cs.scaff_score_batch(bad)
    bad = perms.copy()
    bad[0, 5] = bad[0, 6]                                   # ... and in the FIRST one (flags are cleared once)
    with pytest.raises(N.LxyError):
        cs.scaff_score_batch(bad)
    assert cs.scaff_score_batch(perms).tolist() == got.tolist()


@pytest.mark.parametrize("n,streamed", [(65515, False), (37000, True), (32768, True), (32769, True), (18000, True)])
def test_two_orderings_per_table_word_extreme_differences(ctxs, n, streamed):
    """2..8 orderings with unit weights interleave two placement tables in one word and take differences
    modulo 2^16 (scaff_stream1p.cuh; exact up to n = 65,515; the tables of two orderings fit the shared
    memory up to n ~ 38,000, the batch kernel runs beyond).  n = 32,769 holds the ordering index 32,768,
    whose negation modulo 2^16 is itself (a compiler-sensitive spot of the permutation check).  Identity and reversed orderings
    put the records (0, n-1), (0, n-2), (1, n-1) at the extreme differences, neighbours at the weighted ones."""
    cs, cb = ctxs
    rng = np.random.default_rng(n)
    a = np.concatenate([[0, 0, 1, 5, 7], rng.integers(0, n, 200_000)]).astype(np.uint32)
    b = np.concatenate([[n - 1, n - 2, n - 1, 6, 27], rng.integers(0, n, 200_000)]).astype(np.uint32)
    keep = a != b
    a, b = a[keep], b[keep]
    near = rng.integers(0, n - 21, 100_000).astype(np.uint32)          # pairs at the six weighted steps
    step = rng.choice([1, 2, 3, 5, 11, 20], 100_000).astype(np.uint32)
    a, b = np.concatenate([a, near]), np.concatenate([b, near + step])
    w = np.ones(len(a))
    ident = np.arange(n, dtype=np.int32)
    perms = np.stack([ident, ident[::-1], np.roll(ident, 1), rng.permutation(n).astype(np.int32)])
    for P in (2, 3, 4):
        _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms[:P], nthreads=8)
        cs.set_links(n, a, b, None)
        cb.set_links(n, a, b, None)
        got, x100 = cs.scaff_score_batch(perms[:P], want_x100=True)
        if streamed and (P == 2 or n <= 18000):
            assert cs.timing()["launches"] == 3
        gb, xb = cb.scaff_score_batch(perms[:P], want_x100=True)
        assert x100.tolist() == ox100.tolist()
        assert x100.tolist() == xb.tolist() and got.tolist() == gb.tolist()
