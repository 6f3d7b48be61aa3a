#!/usr/bin/env python
"""Per-CTA time stamps of k_scaff_stream1 (library built with EXTRA=-DLXY_S1_TRACE, LXY_B200_LIB pointing at it):
where the microseconds that do not scale with the record stream go.  `python tools/s1_trace.py [L] [P]`."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import lxy_b200._native as N

L = int(float(sys.argv[1])) if len(sys.argv) > 1 else 200_000_000
sP = int(sys.argv[2]) if len(sys.argv) > 2 else 1
n = 5000
ta, tb, order = bench.gen_links(n, L, 3, torch.device("cuda", 0))
rng5 = np.random.default_rng(5)
perm = torch.from_numpy(np.stack([rng5.permutation(n) for _ in range(sP)]).astype(np.int32)).cuda()
out = torch.zeros(8, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
with N.Context(0, path=N.PATH_STREAM, flags=int(os.environ.get("S1_FLAGS", "0"))) as ctx:
    ctx.set_links_dev(n, L, ta.data_ptr(), tb.data_ptr(), None)
    for _ in range(6):
        ctx.scaff_score_batch_dev(perm.data_ptr(), sP, out.data_ptr(), None)
    tm = ctx.timing()
    grid = tm["grid"]
    print("L", L, "P", sP, "ms_score", tm["ms_score"], "ms_total", tm["ms_total"], "grid", grid)
    fn = getattr(N.lib(), "lxy_debug_s1_trace", None)
    if fn is not None:
        buf = np.zeros(1024 * 8 + 256, np.uint64)
        rc = fn(buf.ctypes.data_as(C.c_void_p), 1024 * 8 + 256)
        wst = buf[1024 * 8:].astype(np.int64).reshape(8, 32)
        full = buf[:grid * 8].reshape(grid, 8).astype(np.int64)
        t, smid = full[:, :4], full[:, 4]
        t0 = t[:, 0].min()
        r = (t - t0) / 1000.0
        def q(x):
            return "min %.2f p10 %.2f med %.2f p90 %.2f max %.2f" % (x.min(), np.percentile(x, 10), np.median(x), np.percentile(x, 90), x.max())
        print("rc", rc, "us since the first CTA started:")
        print("  start      ", q(r[:, 0]))
        print("  table done ", q(r[:, 1]))
        print("  stream done", q(r[:, 2]))
        print("  end        ", q(r[:, 3]))
        print("  prologue   ", q(r[:, 1] - r[:, 0]))
        if full[:, 5].max() > 0:
            e = (full[:, 5:8] - t0) / 1000.0
            print("    loads requested + ring filled", q(e[:, 0] - r[:, 0]), "\n    luts + tile stored", q(e[:, 1] - r[:, 0]), "\n    tables scattered", q(e[:, 2] - r[:, 0]))
        print("  stream     ", q(r[:, 2] - r[:, 1]))
        print("  epilogue   ", q(r[:, 3] - r[:, 2]))
        st = r[:, 2] - r[:, 1]
        print("  stream time by blockIdx octile:", " ".join("%.1f" % st[i * grid // 8:(i + 1) * grid // 8].mean() for i in range(8)))
        order_sm = np.argsort(smid)
        print("  stream time by smid octile:    ", " ".join("%.1f" % st[order_sm][i * grid // 8:(i + 1) * grid // 8].mean() for i in range(8)))
        per_sm = {}
        for b in range(grid):
            per_sm.setdefault(int(smid[b]), []).append(st[b])
        pairs = np.array([v for v in per_sm.values() if len(v) == 2])
        print("  SMs with two CTAs:", len(pairs), " |difference of the two CTAs of one SM| mean %.2f" % np.abs(pairs[:, 0] - pairs[:, 1]).mean(),
              " std of SM means %.2f" % pairs.mean(axis=1).std())
        print("  smid -> stream us:", " ".join("%d:%.0f" % (k, np.mean(v)) for k, v in sorted(per_sm.items())))
        for b in range(2):
            print("  CTA", b, "warps done at (us):", " ".join("%.1f" % ((x - t0) / 1000.0) for x in wst[b] if x > 0))
        worst = np.argsort(-st)[:4]
        for b in worst:
            print("  slowest CTA %d on SM %d: start %.1f table %.1f stream done %.1f end %.1f" % (b, smid[b], r[b, 0], r[b, 1], r[b, 2], r[b, 3]))
