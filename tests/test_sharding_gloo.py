"""world_size-2 gloo tests of the multi-GPU composition (lxy_b200/sharding.py) on the CPU.  The
per-rank scorer is the oracle here (no GPU in this tier); what is tested is the split and the
combine: sharded results must equal the unsharded ones — bit for bit for the integer path."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _oracle as O
import _synth as S
from lxy_b200 import sharding as SH


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, L, P = 120, 9001, 37                      # ragged: neither splits evenly
        a, b, w = S.scaff_links(n, L, 5)
        perms = S.scaff_population(n, P, 105)
        full, fx = O.scaff_score_links_batch(n, a, b, w, perms)

        # population-sharded: links replicated, candidates split, all-gather of scores
        lo, hi = SH.shard_range(P, world, rank)
        mine, _ = O.scaff_score_links_batch(n, a, b, w, perms[lo:hi])
        got = SH.gather_scores(torch.from_numpy(mine), P)
        assert got.numpy().tolist() == full.tolist()

        # link-sharded: records split, whole population, exact integer all-reduce
        lo, hi = SH.shard_range(L, world, rank)
        _, px = O.scaff_score_links_batch(n, a[lo:hi], b[lo:hi], w[lo:hi], perms)
        tot = SH.combine_link_partials_exact(torch.from_numpy(px))
        assert tot.numpy().tolist() == fx.tolist()
        np.testing.assert_array_equal(SH.fitness_from_x100(tot).numpy(), -(fx / 100.0))

        # link-sharded with real weights: fp64 partials, rank-ordered sum
        wr = np.random.default_rng(1).gamma(2.0, 1.5, L)
        fullr, _ = O.scaff_score_links_batch(n, a, b, wr, perms)
        pr, _ = O.scaff_score_links_batch(n, a[lo:hi], b[lo:hi], wr[lo:hi], perms)
        totr = SH.combine_link_partials_f64(torch.from_numpy(pr))
        np.testing.assert_allclose(totr.numpy(), fullr, rtol=1e-12)

        # GA generation with sharded scoring: same plan and children on every rank (oracle operators),
        # each rank scores its slice, all-gather of the fitness vector, identical survivors everywhere
        import _ga_oracle as G
        Pg = 24
        pop = perms[:Pg]
        s_pop, _ = O.scaff_score_links_batch(n, a, b, w, pop)
        order0 = np.argsort(s_pop, kind="stable")
        pop, s_pop = pop[order0], s_pop[order0]
        ch = G.children(9, 0, pop, 0.4, 0.6, 0.7, 5, 7)
        C = len(ch)
        lo, cnt, width = SH.ga_child_slice(C, world, rank)
        assert sum(SH.ga_child_slice(C, world, r)[1] for r in range(world)) == C
        buf = torch.full((world * width + 3,), float("nan"), dtype=torch.float64)
        if cnt:
            buf[lo:lo + cnt] = torch.from_numpy(O.scaff_score_links_batch(n, a, b, w, ch[lo:lo + cnt])[0])
        SH.ga_gather_child_scores(buf, C)
        s_ch, _ = O.scaff_score_links_batch(n, a, b, w, ch)
        assert buf[:C].numpy().tolist() == s_ch.tolist()
        survivors = G.select(s_pop, buf[:C].numpy())
        assert survivors.tolist() == G.select(s_pop, s_ch).tolist()

        # elite broadcast
        elite = torch.from_numpy(perms[:3].copy()) if rank == 0 else torch.zeros(3, n, dtype=torch.int32)
        SH.broadcast_elite(elite, src=0)
        assert elite.numpy().tolist() == perms[:3].tolist()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_sharded_equals_unsharded_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(results) == [(0, "ok"), (1, "ok")], results


def test_shard_range_partitions():
    for total in (0, 1, 7, 128, 1000):
        for world in (1, 2, 3, 8):
            spans = [SH.shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == SH.shard_counts(total, world)


def test_python_slicing_mirrors_the_library():
    """lxy_b200/sharding.py is the CPU-testable mirror of the split comm.cu uses on the GPUs
    (lxy_comm_slice): same balanced contiguous ranges."""
    import lxy_b200._native as N
    from lxy_b200 import sharding as SH
    for total, world in [(1024, 8), (1028, 8), (5, 8), (0, 3), (1000, 7), (129, 2)]:
        for r in range(world):
            lo, hi = SH.shard_range(total, world, r)
            assert (lo, hi - lo) == N.comm_slice(total, world, r)
