#!/usr/bin/env python
"""Scoring time against the population size: where the streaming-regime kernels (passes of <= 8
orderings) hand over to the batch kernel, and what the dispatch of scaff_score_enqueue picks.
For every (n, L, weights, P) the call is timed twice: as dispatched (cost model) and with the batch
kernel forced (LXY_FLAG_NO_STREAM1); one JSON line per point.

    python tools/p_sweep.py            # PS_N=500,5000,20000  PS_L=5e6,5e7,2e8  PS_W=unit,int  PS_P=...
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import lxy_b200._native as N

Ns = [int(x) for x in os.environ.get("PS_N", "500,5000,20000").split(",")]
Ls = [int(float(x)) for x in os.environ.get("PS_L", "5e6,5e7,2e8").split(",")]
Ws = os.environ.get("PS_W", "unit,int").split(",")
Ps = [int(x) for x in os.environ.get("PS_P", "1,2,4,8,9,12,16,24,32,40,48,64,96,128,129,136,160,200,264").split(",")]
dev = torch.device("cuda", 0)
rng = np.random.default_rng(5)
for n in Ns:
    perms = torch.from_numpy(np.stack([rng.permutation(n) for _ in range(max(Ps))]).astype(np.int32)).cuda()
    out = torch.zeros(max(Ps), dtype=torch.float64, device="cuda")
    for L in Ls:
        ta, tb, order = bench.gen_links(n, L, 3, dev)
        for wm in Ws:
            w = None
            if wm == "int":
                w = torch.randint(1, 9, (L,), device=dev).double()
            torch.cuda.synchronize()
            with N.Context(0, path=N.PATH_STREAM) as ctx, N.Context(0, path=N.PATH_STREAM, flags=N.FLAG_NO_STREAM1) as cb:
                for c in (ctx, cb):
                    c.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), w.data_ptr() if w is not None else None)
                for P in Ps:
                    res = {}
                    for name, c in (("dispatched", ctx), ("batch", cb)):
                        ts = []
                        for _ in range(6):
                            c.scaff_score_batch_dev(perms.data_ptr(), P, out.data_ptr(), None)
                            ts.append(c.timing()["ms_total"])
                        res[name] = float(np.median(ts[2:]))
                        res[name + "_launches"] = c.timing()["launches"]
                    print(json.dumps({"n": n, "L": L, "weights": wm, "P": P, "ms_dispatched": round(res["dispatched"], 4),
                                      "ms_batch_forced": round(res["batch"], 4), "launches": res["dispatched_launches"],
                                      "dispatch_vs_batch": round(res["dispatched"] / res["batch"], 3)}), flush=True)
            del w
        del ta, tb
