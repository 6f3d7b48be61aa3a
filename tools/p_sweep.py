#!/usr/bin/env python
"""Scoring time against the population size on config-3's link store (5,000 contigs / 200 M records):
where the streaming-regime kernel (P <= 8 per pass) hands over to the batch kernel.
`python tools/p_sweep.py`  (S1_FLAGS=2 forces the batch kernel everywhere)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import lxy_b200._native as N

n, L = 5000, int(float(os.environ.get("PS_L", "2e8")))
ta, tb, order = bench.gen_links(n, L, 3, torch.device("cuda", 0))
rng = np.random.default_rng(5)
Ps = [int(x) for x in os.environ.get("PS_P", "1,2,4,8,9,12,16,24,32,48,64,128").split(",")]
perms = torch.from_numpy(np.stack([rng.permutation(n) for _ in range(max(Ps))]).astype(np.int32)).cuda()
out = torch.zeros(max(Ps), dtype=torch.float64, device="cuda")
with N.Context(0, path=N.PATH_STREAM, flags=int(os.environ.get("S1_FLAGS", "0"))) as ctx:
    ctx.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), None)
    for P in Ps:
        ts = []
        for _ in range(7):
            ctx.scaff_score_batch_dev(perms.data_ptr(), P, out.data_ptr(), None)
            t = ctx.timing()
            ts.append(t["ms_total"])
        t = float(np.median(ts[2:]))
        print(f"P={P:4d}  {t*1e3:9.1f} us   {L*P/t/1e9:8.2f} e12 evals/s   launches {ctx.timing()['launches']}")
