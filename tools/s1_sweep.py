#!/usr/bin/env python
"""Kernel time of k_scaff_stream1 (one ordering) against the number of records: separates the fixed
cost (table build, tile search, launch) from the per-byte streaming rate.  `python tools/s1_sweep.py`"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import lxy_b200._native as N

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
dev = torch.device("cuda", 0)
Lmax = 400_000_000
ta, tb, order = bench.gen_links(n, Lmax, 3, dev)
perm = torch.from_numpy(np.random.default_rng(5).permutation(n).astype(np.int32)).cuda()
out = torch.zeros(8, dtype=torch.float64, device="cuda")
with N.Context(0, path=N.PATH_STREAM, flags=int(os.environ.get("S1_FLAGS", "0"))) as ctx:
    Ls = [int(float(x)) for x in os.environ.get("S1_LS", "1e6,1e7,5e7,1e8,2e8,4e8").split(",")]
    for L in Ls:
        ctx.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), None)
        ts = []
        for _ in range(8):
            ctx.scaff_score_batch_dev(perm.data_ptr(), 1, out.data_ptr(), None)
            ts.append(ctx.timing()["ms_score"])
        t = float(np.median(ts[2:]))
        print(f"L={L:>11d}  k_scaff_stream1 {t*1e3:8.1f} us   {2*L/t/1e6:8.1f} GB/s")
