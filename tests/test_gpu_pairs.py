"""GPU parity of the pair-table path (cfg.path = LXY_PATH_PAIRS): the reference's own algorithm —
aggregate the records per pair (Links.Add), then 6n table gathers per candidate — against the
reference-order oracle and against the streaming path (identical exact integers)."""
import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu
REL = 1e-9


@pytest.fixture(scope="module")
def ctxs():
    p, s = N.Context(0, path=N.PATH_PAIRS), N.Context(0, path=N.PATH_STREAM)
    yield p, s
    p.close(); s.close()


@pytest.mark.parametrize("key", ["test_ctg", "toy_ctg"])
def test_golden_fixtures(ctxs, golden, golden_files, key):
    ctx, _ = ctxs
    l = O.Links.load(golden_files[key])
    a, b, w = l.export()
    ctx.set_links(l.size(), a, b, w)
    genes = np.array([c["gene"] for c in golden[key]["cases"]], np.int32)
    want = np.array([c["score"] for c in golden[key]["cases"]])
    np.testing.assert_allclose(ctx.scaff_score_batch(genes), want, rtol=REL, atol=0)


@pytest.mark.parametrize("n,L,P,seed", [(40, 3000, 24, 1), (500, 50000, 128, 3), (700, 200000, 130, 4),
                                        (257, 10000, 1, 6), (3000, 300000, 96, 8), (20, 500, 7, 9),
                                        (21, 500, 7, 10), (5, 60, 4, 11)])
def test_unit_records_vs_reference_and_streaming(ctxs, n, L, P, seed):
    pc, sc = ctxs
    a, b, w = S.scaff_links(n, L, seed)
    table = O.Links.from_records(n, a, b, w)
    perms = S.scaff_population(n, P, seed + 100)
    ref = table.scaff_score_batch(perms, nthreads=4)
    pc.set_links(n, a, b, None)
    sc.set_links(n, a, b, None)
    got, x100 = pc.scaff_score_batch(perms, want_x100=True)
    sgot, sx100 = sc.scaff_score_batch(perms, want_x100=True)
    assert x100.tolist() == sx100.tolist() and got.tolist() == sgot.tolist()
    np.testing.assert_allclose(got, ref, rtol=REL, atol=0)


def test_weights_int_real_and_errors(ctxs):
    pc, sc = ctxs
    n, L, P = 600, 80000, 70
    a, b, _ = S.scaff_links(n, L, 21)
    rng = np.random.default_rng(21)
    perms = S.scaff_population(n, P, 121)
    wi = rng.integers(-50, 2000, L).astype(np.float64)
    _, ox = O.scaff_score_links_batch(n, a, b, wi, perms, nthreads=4)
    pc.set_links(n, a, b, wi)
    _, x100 = pc.scaff_score_batch(perms, want_x100=True)
    assert x100.tolist() == ox.tolist()
    wr = rng.gamma(2.0, 1.7, L)
    ref = O.Links.from_records(n, a, b, wr).scaff_score_batch(perms, nthreads=4)
    pc.set_links(n, a, b, wr)
    np.testing.assert_allclose(pc.scaff_score_batch(perms), ref, rtol=REL, atol=0)
    bad = perms.copy()
    bad[3, 10] = bad[3, 11]
    with pytest.raises(N.LxyError) as e:
        pc.scaff_score_batch(bad)
    assert e.value.status == -4
    bad = perms.copy()
    bad[0, 0] = -1
    with pytest.raises(N.LxyError):
        pc.scaff_score_batch(bad)
    assert np.isfinite(pc.scaff_score_batch(perms)).all()


def test_large_table_n20000(ctxs):
    pc, sc = ctxs
    n, L, P = 20000, 600000, 24                          # 1.6 GB dense int32 table
    a, b, w = S.scaff_links(n, L, 61)
    perms = S.scaff_population(n, P, 161)
    _, ox = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=4)
    pc.set_links(n, a, b, None)
    _, x100 = pc.scaff_score_batch(perms, want_x100=True)
    assert x100.tolist() == ox.tolist()
