#!/usr/bin/env python
"""Small fixed workloads for ncu: `python tools/prof_workload.py {r|g|pairs|phase|s1|r3} [steps]`.
cfg2-sized (500 contigs, 5 M records, population 128); phase: 100k blocks, population 256."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lxy_b200._native as N
from lxy_b200 import synth as S

mode = sys.argv[1] if len(sys.argv) > 1 else "r"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
n, L, P = 500, 5_000_000, 128
if mode == "phase":
    pn, pP = 300_000, 1024       # the bench's phasing shape (BASELINE config 5)
    a, b, w, truth = S.phase_links(pn, 5)
    genes = S.phase_population(pn, pP, 105, truth)
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_phase_links(pn, a, b, w)
        for _ in range(steps):
            ctx.phase_score_batch(genes)
        print(ctx.timing())
    sys.exit(0)
if mode in ("r3", "r4"):   # r3: the bench default (config 3, population 1024, batch kernel); r4: config-4 contigs
    import torch
    import bench
    n, L, P = (5000, 200_000_000, int(os.environ.get("R3_P", "1024"))) if mode == "r3" else (50000, 250_000_000, 1024)   # r4: one GPU's shard of config 4
    ta, tb, order = bench.gen_links(n, L, 3 if mode == "r3" else 4, torch.device("cuda", 0))
    perms = torch.from_numpy(bench.gen_population(n, P, 103, order.cpu().numpy())).cuda()
    out = torch.zeros(P, dtype=torch.float64, device="cuda")
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), None)
        for _ in range(steps):
            ctx.scaff_score_batch_dev(perms.data_ptr(), P, out.data_ptr(), None)
        print(ctx.timing())
    sys.exit(0)
if mode == "g3":      # G-profile at config-3 size (bench.py's g_profile inputs): k_scaffg2r_score
    import torch
    import bench
    n, L, P = 5000, int(float(os.environ.get("G3_L", "2e8"))), 1024
    dev = torch.device("cuda", 0)
    ta, tb, order = bench.gen_links(n, L, 3, dev)
    g = torch.Generator(device=dev)
    g.manual_seed(3 + 17)
    lens_t = torch.exp(torch.randn(n, generator=g, device=dev, dtype=torch.float64) * 0.6
                       + float(np.log(min(1_000_000, 1_600_000_000 // n)))).clamp_(1000).long()
    xa = (torch.rand(L, generator=g, device=dev, dtype=torch.float64) * lens_t[ta.long()]).int()
    xb = (torch.rand(L, generator=g, device=dev, dtype=torch.float64) * lens_t[tb.long()]).int()
    perms = torch.from_numpy(bench.gen_population(n, P, 103, order.cpu().numpy())).cuda()
    oris = torch.from_numpy((np.random.default_rng(134).random((P, n)) < 0.5).astype(np.uint8)).cuda()
    out = torch.zeros(P, dtype=torch.float64, device="cuda")
    unit = int(lens_t.double().median().item())
    with N.Context(0) as ctx:
        ctx.set_links_g_dev(lens_t.cpu().numpy().astype(np.uint32), unit, L, ta.data_ptr(), xa.data_ptr(),
                            tb.data_ptr(), xb.data_ptr(), None)
        for _ in range(steps):
            ctx.scaffg_score_batch_dev(perms.data_ptr(), oris.data_ptr(), P, out.data_ptr(), None)
        print(ctx.timing())
    sys.exit(0)
if mode == "ga":      # GA generations at config-3 shape on the pair-table path: the operators dominate
    import torch
    import bench
    n, L, P = 5000, 20_000_000, 1024
    ta, tb, order = bench.gen_links(n, L, 3, torch.device("cuda", 0))
    torch.cuda.synchronize()
    with N.Context(0) as ctx:
        ctx.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), None)
        with N.GA(ctx, N.GA_SCAFF, popsize=P, seed=1234) as ga:
            ga.init()
            ga.optimize(steps)
            print(ga.stats())
    sys.exit(0)
if mode == "s1":      # streaming regime at config-3 size: one ordering against 200 M records
    import torch
    import bench
    import os
    n, L = 5000, int(float(os.environ.get("S1_L", "2e8")))
    ta, tb, order = bench.gen_links(n, L, 3, torch.device("cuda", 0))
    sP = int(os.environ.get("S1_P", "1"))      # S1_P=8: one pass of the packed kernel (k_scaff_stream1p)
    rng5 = np.random.default_rng(5)
    perm = torch.from_numpy(np.stack([rng5.permutation(n) for _ in range(sP)]).astype(np.int32)).cuda()
    out = torch.zeros(max(8, sP), dtype=torch.float64, device="cuda")
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), None)
        for _ in range(steps):
            ctx.scaff_score_batch_dev(perm.data_ptr(), sP, out.data_ptr(), None)
        print(ctx.timing())
    sys.exit(0)
a, b, w = S.scaff_links(n, L, 2)
perms = S.scaff_population(n, P, 102)
if mode == "g":
    lens = S.contig_lengths(n, 2)
    rng = np.random.default_rng(31)
    xa = (rng.random(L) * lens[a]).astype(np.uint32)
    xb = (rng.random(L) * lens[b]).astype(np.uint32)
    oris = S.orientations(n, P, 2)
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links_g(lens, int(np.median(lens)), a, xa, b, xb, None)
        for _ in range(steps):
            ctx.scaffg_score_batch(perms, oris)
        print(ctx.timing())
else:
    with N.Context(0, path=N.PATH_PAIRS if mode == "pairs" else N.PATH_STREAM) as ctx:
        ctx.set_links(n, a, b, None)
        for _ in range(steps):
            ctx.scaff_score_batch(perms)
        print(ctx.timing())
