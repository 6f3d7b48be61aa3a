"""World-size > 1 on real devices: one process, one context per GPU, one host thread per context
(the model the reference's single Go process can host, scaff.go:88-94: a goroutine per device), the
NCCL communicator created by lxy_comm_init_all.  Skipped with fewer than two devices.

Checked: population-sharded and link-sharded scores on every rank equal the single-GPU result and the
oracle's exact integers; an invalid row is reported on every rank; the sharded GA loop walks the same
trajectory as the local one.
"""
import threading

import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu


def _world():
    return min(N.device_count(), 4)


def _run(ctxs, fn):
    """fn(rank, ctx) on one thread per context (a sharded call returns when its collective is done)."""
    out, err = [None] * len(ctxs), [None] * len(ctxs)

    def work(r):
        try:
            out[r] = fn(r, ctxs[r])
        except Exception as e:      # noqa: BLE001 - re-raised below on the main thread
            err[r] = e

    ts = [threading.Thread(target=work, args=(r,)) for r in range(len(ctxs))]
    for t in ts:
        t.start()
    for t in ts:
        t.join(300)
    return out, err


@pytest.fixture(scope="module")
def ranks():
    W = _world()
    if W < 2:
        pytest.skip("needs two CUDA devices")
    ctxs = [N.Context(d, path=N.PATH_STREAM) for d in range(W)]
    N.comm_init_all(ctxs)
    assert [c.comm_info() for c in ctxs] == [(W, r) for r in range(W)]
    yield ctxs
    for c in ctxs:
        c.close()


def test_population_sharded_equals_single_gpu_and_oracle(ranks):
    W = len(ranks)
    n, L, P = 1200, 500_000, 203                          # 203 does not divide by the world size
    a, b, w = S.scaff_links(n, L, 21)
    perms = S.scaff_population(n, P, 121)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=8)
    with N.Context(0, path=N.PATH_STREAM) as one:
        one.set_links(n, a, b, None)
        ref, _ = one.scaff_score_batch(perms, want_x100=True)

    def score(r, c):
        c.set_links(n, a, b, None)                        # links replicated
        return c.scaff_score_sharded(N.SHARD_POPULATION, perms, want_x100=True)

    out, err = _run(ranks, score)
    assert err == [None] * W, err
    for got, x100 in out:
        assert x100.tolist() == ox100.tolist() and got.tolist() == ref.tolist()

    bad = perms.copy()
    bad[P - 1, 3] = bad[P - 1, 4]                         # only the last rank scores this row

    def score_bad(r, c):
        return c.scaff_score_sharded(N.SHARD_POPULATION, bad)

    _, err = _run(ranks, score_bad)
    assert all(isinstance(e, N.LxyError) and e.status == -4 for e in err), err   # reported on EVERY rank


@pytest.mark.parametrize("weights", ["unit", "real"])
def test_link_sharded_equals_single_gpu_and_oracle(ranks, weights):
    W = len(ranks)
    n, L, P = 1200, 600_001, 130
    a, b, w = S.scaff_links(n, L, 22)
    if weights == "real":
        w = np.random.default_rng(23).gamma(2.0, 1.5, L)
    perms = S.scaff_population(n, P, 122)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=8)
    ref = O.Links.from_records(n, a, b, w).scaff_score_batch(perms, nthreads=8)
    cut = [L * r // W for r in range(W + 1)]

    def score(r, c):
        sl = slice(cut[r], cut[r + 1])                    # any split of the records is valid
        c.set_links(n, a[sl], b[sl], None if weights == "unit" else w[sl])
        return c.scaff_score_sharded(N.SHARD_LINKS, perms, want_x100=True)

    out, err = _run(ranks, score)
    assert err == [None] * W, err
    for got, x100 in out:
        if weights == "unit":
            assert x100.tolist() == ox100.tolist()        # all-reduce of exact integers: bit-identical
            assert got.tolist() == (-(ox100 / 100.0)).tolist()
        else:
            assert (x100 == N.INT64_MIN).all()
            np.testing.assert_allclose(got, ref, rtol=1e-9, atol=0)
    assert all(o[0].tolist() == out[0][0].tolist() for o in out)      # the same bits on every rank


@pytest.mark.parametrize("mode", [N.SHARD_POPULATION, N.SHARD_LINKS])
def test_sharded_ga_walks_the_local_trajectory(ranks, mode):
    W = len(ranks)
    n, L = 500, 200_000
    a, b, _ = S.scaff_links(n, L, 24)
    with N.Context(0, path=N.PATH_STREAM) as one:
        one.set_links(n, a, b, None)
        with N.GA(one, N.GA_SCAFF, popsize=96, seed=5) as ga:
            ga.init()
            ga.optimize(5)
            want = (ga.population()[0].tolist(), ga.population()[1].tolist())
    cut = [L * r // W for r in range(W + 1)]

    def run(r, c):
        if mode == N.SHARD_LINKS:
            c.set_links(n, a[cut[r]:cut[r + 1]], b[cut[r]:cut[r + 1]], None)
        else:
            c.set_links(n, a, b, None)
        with N.GA(c, N.GA_SCAFF, popsize=96, seed=5) as ga:
            ga.set_sharding(mode)
            ga.init()
            ga.optimize(5)
            return ga.population()[0].tolist(), ga.population()[1].tolist()

    out, err = _run(ranks, run)
    assert err == [None] * W, err
    assert all(o == want for o in out)
