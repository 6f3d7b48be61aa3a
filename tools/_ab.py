import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, lxy_b200._native as N
n, L = 5000, 200_000_000
dev = torch.device("cuda", 0)
a, b, order = bench.gen_links(n, L, 3, dev)
perms_all = bench.gen_population(n, 1024, 103, order.cpu().numpy())
with N.Context(0) as ctx:
    ctx.set_links_dev(n, L, a.data_ptr(), b.data_ptr(), None)
    for P in (1024, 512, 256, 128):
        perms = torch.from_numpy(perms_all[:P].copy()).cuda()
        out = torch.zeros(P, dtype=torch.float64, device=dev)
        ts = []
        for _ in range(7):
            ctx.scaff_score_batch_dev(perms.data_ptr(), P, out.data_ptr(), None); ts.append(ctx.timing()["ms_score"])
        print(os.environ.get("LXY_B200_LIB", "HEAD"), "P=%d k_scaff_score %.3f ms" % (P, np.median(ts[2:])))
