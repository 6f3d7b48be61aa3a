#!/usr/bin/env python
"""bench.py — GA fitness link-evaluations/sec of the LXY scaffolding scorer on N B200s.

    python bench.py --gpus N --steps K --warmup W [--workload cfg3|cfg2] [--impl reference]

A step = one generation's fitness pass: score every candidate ordering of the population against
every Hi-C link record.  Workload (BASELINE.json configs[2], the one the metric's "at 1/2/4/8 B200"
is quoted on): synthetic 5,000 contigs / 200M link records, population 1024, population-sharded
(strong scaling: the population is split across ranks, links replicated).  One evaluation = one link
record scored for one candidate, so a step is P x L evaluations.

  value  device-resident: the population is already in HBM; timed with CUDA events on the stream
         the kernels run on; max over ranks.
  e2e    through the C ABI with HOST buffers (lxy_scaff_score_batch): host->device copy of the
         population and device->host copy of the scores inside the timed region; wall clock, max
         over ranks.
  --impl reference   the reference's own CPU algorithm (the oracle's line-for-line restatement of
         scaff.score + Links.Get; the Go toolchain is absent) on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (contigs, link records, population, link seed, population seed)
    "cfg2": (500, 5_000_000, 128, 2, 102),
    "cfg3": (5_000, 200_000_000, 1024, 3, 103),
    "tiny": (300, 200_000, 64, 9, 109),
    # link-sharded (BASELINE.json configs[3]): records split across ranks, NCCL all-reduce of partials
    "cfg4": (50_000, 2_000_000_000, 1024, 4, 104),
    "cfg4s": (50_000, 64_000_000, 256, 4, 104),
    "cfg4m": (50_000, 500_000_000, 1024, 4, 104),
}
LINK_SHARDED = {"cfg4", "cfg4s", "cfg4m"}
METRIC = "ga_fitness_link_evals_per_sec"
UNIT = "link-evals/s"


# ------------------------------------------------------------------------------------------------
# synthetic inputs (torch is plumbing here: device RNG + memory)
# ------------------------------------------------------------------------------------------------
def gen_links(n, L, seed, device, stream_seed=None):
    """contig pairs with P(a,b) ~ 1/|i-j| in a hidden true order plus 10 % uniform
    noise pairs; one record per read pair (weight 1).  Returns int32 tensors a, b, order."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    order = torch.randperm(n, generator=g, device=device, dtype=torch.int64)
    if stream_seed is not None:      # same hidden order on every rank, a different record stream
        g.manual_seed(stream_seed)
    d = torch.arange(1, n, device=device, dtype=torch.float64)
    pmf = (n - d) / d
    cdf = torch.cumsum(pmf / pmf.sum(), 0)
    a = torch.empty(L, dtype=torch.int32, device=device)
    b = torch.empty(L, dtype=torch.int32, device=device)
    chunk = 1 << 24
    for lo in range(0, L, chunk):
        m = min(chunk, L - lo)
        u = torch.rand(m, generator=g, device=device, dtype=torch.float64)
        dist = (torch.searchsorted(cdf, u, right=True) + 1).clamp_(1, n - 1)
        i = (torch.rand(m, generator=g, device=device, dtype=torch.float64) * (n - dist)).long()
        x, y = order[i], order[i + dist]
        noise = torch.rand(m, generator=g, device=device) < 0.1
        nx = torch.randint(0, n, (m,), generator=g, device=device)
        ny = (nx + torch.randint(1, n, (m,), generator=g, device=device)) % n
        x = torch.where(noise, nx, x)
        y = torch.where(noise, ny, y)
        flip = torch.rand(m, generator=g, device=device) < 0.5
        a[lo:lo + m] = torch.where(flip, y, x).int()
        b[lo:lo + m] = torch.where(flip, x, y).int()
    return a, b, order


def gen_population(n, P, seed, order):
    """The true order, its reverse, k-swap neighbours of it, random permutations (int32 [P][n])."""
    rng = np.random.default_rng(seed)
    order = np.asarray(order, np.int32)
    pop = np.empty((P, n), np.int32)
    for p in range(P):
        if p == 0:
            g = order.copy()
        elif p == 1:
            g = order[::-1].copy()
        elif p % 2 == 0:
            g = order.copy()
            for _ in range(int(rng.integers(1, max(2, n // 8)))):
                x, y = rng.integers(0, n, 2)
                g[x], g[y] = g[y], g[x]
        else:
            g = rng.permutation(n).astype(np.int32)
        pop[p] = g
    return pop


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (NVML)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
               0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
               0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


class _DevArray:
    """A device allocation owned by liblxyb200, exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr, count, typestr):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2}


def measure_ga(ctx, N, torch, dist, dev, P, n, L_eval, rank, world, generations):
    """GA generations/sec (BASELINE.json's third metric): the device-resident loop of lxy_ga_* with
    the reference's Scaffold parameters (PBreed .2, PMutate .6, tournament(0.7, 5), shift + switch +
    invert; scaff.go:72-94) at the workload's population size.  N > 1: every rank runs the same
    deterministic plan and makes the same children; rank r scores its slice of them and the fitness
    vector is all-gathered (the only per-generation traffic between GPUs)."""
    from lxy_b200 import sharding as SH
    ga = N.GA(ctx, N.GA_SCAFF, popsize=P, seed=1234)
    ga.init()
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)

    def generation():
        if world == 1:
            ga.optimize(1)
            return
        C, _, sc_ptr, _ = ga.step_begin()
        lo, cnt, q = SH.ga_child_slice(C, world, rank)
        ga.score_children(lo, cnt)
        ctx.sync()
        sc = torch.as_tensor(_DevArray(sc_ptr, world * q, "<f8"), device=dev)   # capacity 3P + 64 >= world*q
        SH.ga_gather_child_scores(sc, C)
        torch.cuda.current_stream().synchronize()
        ga.step_end()

    for _ in range(3):
        generation()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    first = ga.best()[1]
    t0 = time.perf_counter()
    children = 0
    parts = {"ms_operators": 0.0, "ms_score": 0.0, "ms_select": 0.0}
    for _ in range(generations):
        generation()
        st = ga.stats()
        children += st["last_children"]
        for k in parts:
            parts[k] += st[k]
    ctx.sync()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if dist is not None:
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt[0])
    best = ga.best()[1]
    ga.close()
    return {"what": "device-resident GA loop (selection, PMX, mutators, scoring of the children, sort/"
                    "truncate on the GPU; only the child count crosses PCIe per generation); wall clock "
                    "around the generation calls, max over ranks",
            "generations_per_sec": generations / dt, "ms_per_generation": 1e3 * dt / generations,
            "population": P, "children_per_generation": children / generations,
            "link_evals_per_sec": children * float(L_eval) / dt,
            "device_ms_per_generation": {k: v / generations for k, v in parts.items()},
            "best_fitness_before_after": [first, best]}


def pinned(arr):
    """A page-locked copy of a host array (the e2e legs copy their inputs from pinned host memory:
    the library hands page-locked caller memory straight to the copy engine)."""
    import torch
    return torch.from_numpy(np.ascontiguousarray(arr)).pin_memory().numpy()


def measured_peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# CPU legs (the oracle is the checker / baseline only)
# ------------------------------------------------------------------------------------------------
def cpu_link_streaming_baseline(n, a_host, b_host, perms, seconds=10.0):
    """The oracle's link-centric port (same algorithm as the GPU kernels) on all host cores, on a
    bounded sample: the first records x the first candidates, sized for ~`seconds` of work."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    cores = os.cpu_count() or 1
    Ls, Ps = min(len(a_host), 2_000_000), min(len(perms), 2 * cores)
    w = np.ones(Ls, np.float64)
    t0 = time.perf_counter()
    O.scaff_score_links_batch(n, a_host[:Ls], b_host[:Ls], w, perms[:Ps], nthreads=cores)
    t1 = time.perf_counter() - t0
    reps = max(1, int(seconds / max(t1, 1e-3)) - 1)
    reps = min(reps, 20)
    t0 = time.perf_counter()
    for _ in range(reps):
        O.scaff_score_links_batch(n, a_host[:Ls], b_host[:Ls], w, perms[:Ps], nthreads=cores)
    dt = (time.perf_counter() - t0) / reps
    return {"value": Ls * Ps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle link-streaming port (C++ restatement; Go toolchain absent): first "
                      f"{Ls} records x first {Ps} candidates, {reps} reps, all {cores} host threads"}


def run_reference(args, name, n, L, P):
    """--impl reference: the reference's algorithm (scaff.score's 12N-90 Links.Get look-ups per
    candidate against the map-of-maps table, scaff.go:246-310 / links.go:190-221) as restated by
    the oracle, all host threads.  Its cost does not depend on the number of records, so
    link-evals/s = candidates/s x L (equivalent evaluations)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import _oracle as O
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    seed, pseed = WORKLOADS[name][3], WORKLOADS[name][4]
    a, b, order = gen_links(n, L, seed, dev)
    # Links.Add per record == add the per-pair counts once (links.go:149-177); aggregate with torch
    # so that building the table stays bounded (the table is built once, outside the timed steps).
    lo, hi = torch.minimum(a, b).long(), torch.maximum(a, b).long()
    keys, counts = torch.unique(lo * n + hi, return_counts=True)
    ua = (keys // n).cpu().numpy().astype(np.uint32)
    ub = (keys % n).cpu().numpy().astype(np.uint32)
    uw = counts.cpu().numpy().astype(np.float64)
    del a, b, lo, hi, keys, counts
    table = O.Links.from_records(n, ua, ub, uw)
    perms = gen_population(n, P, pseed, order.cpu().numpy())
    Ps = min(P, max(cores, 16))                       # bounded sample of the population per step
    for _ in range(args.warmup):
        table.scaff_score_batch(perms[:Ps], nthreads=cores)
    t0 = time.perf_counter()
    for s in range(args.steps):
        table.scaff_score_batch(perms[(s * Ps) % (P - Ps + 1):][:Ps], nthreads=cores)
    dt = time.perf_counter() - t0
    cand_per_s = Ps * args.steps / dt
    value = cand_per_s * L
    # one GA generation the reference's way: its own O(n^2) PMX / mutators for ~P new individuals
    # (oracle/ga_oracle.cpp, one thread per go-galib's breeding loop) + scoring them on all cores
    import _ga_oracle as G
    t0 = time.perf_counter()
    ch = G.children(1234, 0, perms, 0.2, 0.6, 0.7, 5, 7)
    t_ops = time.perf_counter() - t0
    ga_ref = {"generations_per_sec": 1.0 / (t_ops + len(ch) / cand_per_s),
              "operators_s": t_ops, "children": int(len(ch)),
              "what": "reference-style generation on the host: genome operators as restated from "
                      "scaff/genome.go (PMX O(n^2)) on one thread + scaff.score of the children on "
                      f"{cores} threads"}
    sample = (f"{Ps} of {P} candidates per step against the full aggregated table "
              f"({len(ua)} pairs from {L} records), {cores} host threads; "
              f"value = candidates/s x L (the reference's cost is 12N-90 look-ups per candidate, "
              f"independent of L)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "candidates_per_sec": cand_per_s, "ga": ga_ref,
        "config": {"workload": f"{name}: {n} contigs / {L} links / population {P}"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the pair-table / G-profile / phasing side measurements")
    args = ap.parse_args()
    n, L, P, seed, pseed = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference(args, args.workload, n, L, P)

    import torch
    import lxy_b200._native as N
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (lxy_b200 has no CPU fallback)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    # ---- inputs: links replicated on every rank, the population sharded ----
    link_mode = args.workload in LINK_SHARDED
    if link_mode:       # records sharded, the whole population on every rank
        L_loc, P_loc = L // world, P
        a, b, order = gen_links(n, L_loc, seed, dev, stream_seed=seed * 1000 + rank)
        perms_all = gen_population(n, P, pseed, order.cpu().numpy())
        perms = perms_all
    else:               # links replicated, the population sharded
        L_loc, P_loc = L, P // world
        a, b, order = gen_links(n, L, seed, dev)
        perms_all = gen_population(n, P, pseed, order.cpu().numpy())
        perms = np.ascontiguousarray(perms_all[rank * P_loc:(rank + 1) * P_loc])
    perms = pinned(perms)
    torch.cuda.synchronize()
    ctx = N.Context(local)
    t0 = time.perf_counter()
    ctx.set_links_dev(n, L_loc, a.data_ptr(), b.data_ptr(), None)
    t_build = time.perf_counter() - t0
    a_host = a[:2_000_000].cpu().numpy().astype(np.uint32) if rank == 0 else None
    b_host = b[:2_000_000].cpu().numpy().astype(np.uint32) if rank == 0 else None

    perm_dev = torch.from_numpy(perms).to(dev)
    out_dev = torch.zeros(P_loc, dtype=torch.float64, device=dev)
    x100_dev = torch.zeros(P_loc, dtype=torch.int64, device=dev)
    gathered = torch.zeros(P, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def bracket():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def time_path(cx, dev_call, host_call, steps, sample_clocks=False):
        """K device-resident steps (CUDA events on the context's stream) and K host-buffer steps
        (wall clock); every step ends with the all-gather of the score vector when N > 1."""
        stream = torch.cuda.ExternalStream(cx.stream(), device=dev)

        def step_device():
            dev_call()
            if dist is not None and link_mode:      # partial exact sums -> all-reduce, then -x/100
                dist.all_reduce(x100_dev, op=dist.ReduceOp.SUM)
                torch.div(x100_dev.double(), -100.0, out=out_dev)
            elif dist is not None:                  # only the score vector moves between GPUs
                dist.all_gather_into_tensor(gathered, out_dev)

        for _ in range(max(args.warmup, 3)):
            step_device()
        sampler = ClockSampler(local) if sample_clocks else None
        bracket()
        if sampler:
            sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        tms = {"ms_score": [], "ms_layout": [], "ms_finish": []}
        launches = 0
        for _ in range(steps):
            step_device()
            t = cx.timing()
            for k in tms:
                tms[k].append(t[k])
            launches += t["launches"]
        if dist is not None:
            torch.cuda.current_stream().synchronize()
        ev1.record(stream)
        bracket()
        clk = sampler.stop() if sampler else None
        ms_d = ev0.elapsed_time(ev1)
        last = cx.timing()
        for _ in range(2):
            host_call()
        bracket()
        t0 = time.perf_counter()
        for _ in range(steps):
            host_scores = host_call()
            if dist is not None and link_mode:
                x100_dev.copy_(torch.from_numpy(host_scores[1]))
                dist.all_reduce(x100_dev, op=dist.ReduceOp.SUM)
                final = (x100_dev.double() / -100.0).cpu()
            elif dist is not None:
                out_dev.copy_(torch.from_numpy(host_scores))
                dist.all_gather_into_tensor(gathered, out_dev)
                torch.cuda.synchronize()
        bracket()
        ms_h = 1e3 * (time.perf_counter() - t0)
        if dist is not None:
            tt = torch.tensor([ms_d, ms_h], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms_d, ms_h = float(tt[0]), float(tt[1])
        return {"ms_dev": ms_d, "ms_e2e": ms_h, "launches": launches, "clocks": clk, "last": last,
                "kernel_ms": {k: float(np.mean(v)) for k, v in tms.items()}}

    # ---- the headline path: streaming kernel, R-profile ----
    main = time_path(ctx, lambda: ctx.scaff_score_batch_dev(perm_dev.data_ptr(), P_loc, out_dev.data_ptr(),
                                                            x100_dev.data_ptr()),
                     lambda: ctx.scaff_score_batch(perms, want_x100=link_mode), args.steps,
                     sample_clocks=True)
    ms_dev, ms_e2e, launches, clocks, timing = (main["ms_dev"], main["ms_e2e"], main["launches"],
                                                main["clocks"], main["last"])
    ms_score, ms_layout, ms_finish = ([main["kernel_ms"]["ms_score"]], [main["kernel_ms"]["ms_layout"]],
                                      [main["kernel_ms"]["ms_finish"]])
    stream_x100 = x100_dev.clone()
    # the same host-buffer call with PAGEABLE caller memory (what a cgo caller passes): the library
    # stages it through its pinned buffer with a few host threads
    perms_pageable = np.array(perms, copy=True)
    ctx.scaff_score_batch(perms_pageable, want_x100=link_mode)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ctx.scaff_score_batch(perms_pageable, want_x100=link_mode)
    ms_e2e_pageable = 1e3 * (time.perf_counter() - t0) / args.steps
    del perms_pageable

    extras = {}
    if not args.no_extras and not link_mode and world == 1:
        # ---- streaming regime: a lone Score() call (the reference's literal boundary hands ONE
        # ordering to score(), scaff/genome.go:153-159).  No cross-candidate reuse, so this pass is
        # HBM-bound: k_scaff_stream1 reads every 2-byte record once and scores it once. ----
        peak_s, peak_src_s = measured_peak_hbm()
        sweep = {}
        for ps in (1, 2, 4, 8):
            pdev = perm_dev[:ps].contiguous()
            call = lambda: ctx.scaff_score_batch_dev(pdev.data_ptr(), ps, out_dev.data_ptr(), x100_dev.data_ptr())
            for _ in range(3):
                call()
            stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = max(args.steps, 10)
            kms = []
            e0.record(stream)
            for _ in range(reps):
                call()
                kms.append(ctx.timing()["ms_score"])
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            k = float(np.mean(kms))
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.scaff_score_batch(perms[:ps])
            ms_h = 1e3 * (time.perf_counter() - t0) / reps
            alg = 2.0 * L_loc + 4.0 * n * ps            # every record once + the orderings once
            sweep[f"P={ps}"] = {
                "value": ps * float(L_loc) / (ms * 1e-3), "unit": UNIT, "ms_per_call": ms,
                "e2e": {"value": ps * float(L_loc) / (ms_h * 1e-3), "unit": UNIT, "ms_per_call": ms_h},
                "k_scaff_stream1_ms": k,
                "roofline": {"bound": "hbm", "achieved": alg / (k * 1e-3) / 1e9, "peak": peak_s,
                             "unit": "GB/s", "frac": alg / (k * 1e-3) / 1e9 / peak_s,
                             "algorithmic_bytes_per_launch": alg, "peak_source": peak_src_s}}
        extras["streaming_regime"] = {
            "what": "lone Score() calls / tiny batches on the same resident link store: lane = record, "
                    "the whole placement table in shared memory (k_scaff_stream1); P=1 is the "
                    "HBM-bound case north_star's '% of HBM roofline' refers to",
            "sweep": sweep}
    if not args.no_extras and not link_mode:
        # ---- pair-table path (same inputs, same results; cost independent of L) ----
        pctx = N.Context(local, path=N.PATH_PAIRS)
        t0 = time.perf_counter()
        pctx.set_links_dev(n, L, a.data_ptr(), b.data_ptr(), None)
        t_pbuild = time.perf_counter() - t0
        pr = time_path(pctx, lambda: pctx.scaff_score_batch_dev(perm_dev.data_ptr(), P_loc, out_dev.data_ptr(),
                                                                x100_dev.data_ptr()),
                       lambda: pctx.scaff_score_batch(perms), args.steps)
        same = bool((x100_dev == stream_x100).all().item())
        ga_pairs = measure_ga(pctx, N, torch, dist, dev, P, n, L, rank, world, max(args.steps, 5))
        pctx.close()
        extras["pair_table_path"] = {
            "what": "LXY_PATH_PAIRS: records aggregated per pair at set_links (Links.Add), 6n table "
                    "gathers per candidate - the reference's algorithm on the GPU; value = "
                    "candidates/s x L (the convention of the reference arm)",
            "value": float(P) * L * args.steps / (pr["ms_dev"] * 1e-3), "unit": UNIT,
            "candidates_per_sec": float(P) * args.steps / (pr["ms_dev"] * 1e-3),
            "ms_per_step": pr["ms_dev"] / args.steps,
            "e2e": {"value": float(P) * L * args.steps / (pr["ms_e2e"] * 1e-3), "unit": UNIT,
                    "ms_per_step": pr["ms_e2e"] / args.steps},
            "kernel_ms": pr["kernel_ms"], "table_bytes": 4 * n * n, "table_build_s": round(t_pbuild, 3),
            "gather_sector_GBps": 6.0 * P_loc * n * 32 / (pr["kernel_ms"]["ms_score"] * 1e-3) / 1e9,
            "identical_x100_to_streaming_path": same, "ga": ga_pairs}

        # ---- G-profile (north_star's generalisation): lengths, orientations, in-contig offsets ----
        g = torch.Generator(device=dev)
        g.manual_seed(seed + 17)
        lens_t = torch.exp(torch.randn(n, generator=g, device=dev, dtype=torch.float64) * 0.6
                           + float(np.log(min(1_000_000, 1_600_000_000 // n)))).clamp_(1000).long()
        xa = (torch.rand(L, generator=g, device=dev, dtype=torch.float64) * lens_t[a.long()]).int()
        xb = (torch.rand(L, generator=g, device=dev, dtype=torch.float64) * lens_t[b.long()]).int()
        ori_host = pinned((np.random.default_rng(pseed + 31).random((P_loc, n)) < 0.5).astype(np.uint8))
        ori_dev = torch.from_numpy(ori_host).to(dev)
        unit = int(lens_t.double().median().item())
        gctx = N.Context(local)
        gctx.set_links_g_dev(lens_t.cpu().numpy().astype(np.uint32), unit, L, a.data_ptr(), xa.data_ptr(),
                             b.data_ptr(), xb.data_ptr(), None)
        del xa, xb
        gr = time_path(gctx, lambda: gctx.scaffg_score_batch_dev(perm_dev.data_ptr(), ori_dev.data_ptr(), P_loc,
                                                                 out_dev.data_ptr(), x100_dev.data_ptr()),
                       lambda: gctx.scaffg_score_batch(perms, ori_host), args.steps)
        gctx.close()
        extras["g_profile"] = {
            "what": "contig lengths (log-normal), per-candidate orientations, per-record in-contig "
                    "offsets; bucket = genomic distance / unit (unit = median contig length)",
            "value": float(P) * L * args.steps / (gr["ms_dev"] * 1e-3), "unit": UNIT,
            "ms_per_step": gr["ms_dev"] / args.steps,
            "e2e": {"value": float(P) * L * args.steps / (gr["ms_e2e"] * 1e-3), "unit": UNIT,
                    "ms_per_step": gr["ms_e2e"] / args.steps},
            "kernel_ms": gr["kernel_ms"], "record_bytes": 10,
            "total_length_bp": int(lens_t.sum().item()), "unit_bp": unit,
            "coordinate_mode": "signed 32-bit origins (total < 2^31)" if int(lens_t.sum().item()) < 0x7FFFFFF0
                               else "unsigned 32-bit origins + orientation bits"}
        # ---- phasing (BASELINE.json configs[4]): simvarblocks-style blocks, integer-exact ----
        if world == 1:
            from lxy_b200 import synth as SY
            pn, pP = 300_000, 1024
            pa, pb, pw, truth = SY.phase_links(pn, 5)
            genes = pinned(SY.phase_population(pn, pP, 105, truth))
            hctx = N.Context(local)
            hctx.set_phase_links(pn, pa, pb, pw)
            genes_dev = torch.from_numpy(genes).to(dev)
            oi = torch.zeros(pP, dtype=torch.int64, device=dev)
            of = torch.zeros(pP, dtype=torch.float64, device=dev)
            torch.cuda.synchronize()
            for _ in range(3):
                hctx.phase_score_batch_dev(genes_dev.data_ptr(), pP, oi.data_ptr(), of.data_ptr())
            hstream = torch.cuda.ExternalStream(hctx.stream(), device=dev)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(hstream)
            for _ in range(args.steps):
                hctx.phase_score_batch_dev(genes_dev.data_ptr(), pP, oi.data_ptr(), of.data_ptr())
            e1.record(hstream)
            torch.cuda.synchronize()
            pms = e0.elapsed_time(e1) / args.steps
            pt = hctx.timing()
            hctx.phase_score_batch(genes)               # warm-up: sizes the library's pinned staging buffer
            t0 = time.perf_counter()
            for _ in range(args.steps):
                hctx.phase_score_batch(genes)
            pms_h = 1e3 * (time.perf_counter() - t0) / args.steps
            genes_pageable = np.array(genes, copy=True)
            hctx.phase_score_batch(genes_pageable)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                hctx.phase_score_batch(genes_pageable)
            pms_hp = 1e3 * (time.perf_counter() - t0) / args.steps
            del genes_pageable
            hctx.close()
            extras["phasing"] = {
                "what": f"phase.score over {pn} simulated variant blocks, {len(pa)} block-pair records, "
                        f"population {pP}; int64 exact",
                "value": float(pP) * len(pa) / (pms * 1e-3), "unit": UNIT, "ms_per_step": pms,
                "e2e": {"value": float(pP) * len(pa) / (pms_h * 1e-3), "unit": UNIT, "ms_per_step": pms_h,
                        "h2d_bytes_per_step": int(genes.nbytes), "d2h_bytes_per_step": 16 * pP,
                        "host_memory": "pinned", "pageable_ms_per_step": pms_hp},
                "kernel_ms": {"k_phase_score": pt["ms_score"], "k_phase_bits": pt["ms_layout"]}}
            del genes_dev
    if not args.no_extras and not link_mode:
        extras["ga"] = measure_ga(ctx, N, torch, dist, dev, P, n, L, rank, world, max(args.steps, 5))
    del a, b
    torch.cuda.empty_cache()

    L_total = L_loc * world if link_mode else L      # records scored per candidate, whole job
    evals_per_step = float(P) * float(L_total)
    value = evals_per_step * args.steps / (ms_dev * 1e-3)
    e2e_value = evals_per_step * args.steps / (ms_e2e * 1e-3)

    if rank == 0:
        peak, peak_src = measured_peak_hbm()
        # algorithmic bytes of one k_scaff_score launch: every record once (2 B),
        # both placement tables of every local candidate once, the per-CTA partial sums once
        wide = n > 65515                              # 32-bit placement tables (scaff_kernels.cu)
        cmax = 64 if wide else 128
        cand = cmax if P_loc > cmax // 2 else (cmax // 2 if P_loc > cmax // 4 else cmax // 4)
        groups = (P_loc + cand - 1) // cand
        es = 4 if wide else 2
        alg_bytes = 2.0 * L_loc + es * 2.0 * groups * cand * n + 8.0 * timing["grid"] * 2 * cand
        # shared-memory wavefronts per record-warp: row B + decay look-ups + row A / 8 (per packet)
        evals_per_cycle = (64 / 3.375) if wide else (128 / 4.375)
        cfg_index = {"cfg2": 1, "cfg3": 2, "cfg4": 3}.get(args.workload)
        k_ms = float(np.mean(ms_score))
        achieved = alg_bytes / (k_ms * 1e-3) / 1e9
        traffic = None
        try:                                          # ncu-measured DRAM bytes of the same launch shape
            with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
                ent = json.load(f).get(args.workload)
            if ent and ent["population_per_gpu"] == P_loc and not link_mode:
                traffic = ent["bytes"]
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int64 (exact integer accumulation of 100*T*w; fitness = -x/100 in f64)",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {n} contigs / {L_total} link records / "
                                   f"population {P}"
                                   + (f" (BASELINE.json configs[{cfg_index}])" if cfg_index else ""),
                       "parallelism": (f"link-sharded x{world} (NCCL all-reduce of exact int64 partials), "
                                       "population replicated") if link_mode
                                      else f"population-sharded x{world}, links replicated",
                       "population_per_gpu": P_loc, "records_per_gpu": L_loc,
                       "l2_policy": f"inputs larger than L2 ({2 * L_loc / 1e6:.0f} MB of records per pass)"
                                    if 2 * L_loc > 130e6 else "records fit L2; no flush (steady-state GA loop)",
                       "link_store_build_s": round(t_build, 3)},
            "candidates_per_sec": P * args.steps / (ms_dev * 1e-3),
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(perms.nbytes), "d2h_bytes_per_step": int(16 * P_loc),
                    "host_memory": "pinned (page-locked caller buffers go straight to the copy engine)",
                    "pageable_ms_per_step": ms_e2e_pageable},
            "gpu_launches": int(launches),
            "kernel_ms": {"k_scaff_score": k_ms, "layout+check": float(np.mean(ms_layout)),
                          "head+finish": float(np.mean(ms_finish))},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "kernel": "k_scaff_score", "algorithmic_bytes_per_launch": alg_bytes,
                         "smem_wavefront": {
                             "what": "the roof that binds: per record-warp (128 evaluations; 64 with "
                                     "32-bit tables) 2 wavefronts for row B + 2 (1) decay-table look-ups "
                                     "(pair table: one 32-bit look-up per two candidates) + 2/8 for "
                                     "row A (gathered once per 8-record packet) + 1/8 for the packet's "
                                     "own 16 bytes (four lanes load a word each, REDUX.OR spreads them "
                                     "through uniform registers) = 4.375 (3.375) wavefronts of the L1/shared data pipe, "
                                     "one wavefront/cycle/SM",
                             "evals_per_cycle_per_sm": evals_per_cycle,
                             "peak_evals_per_s": evals_per_cycle * 148 * (clocks["sm_mhz"] or 1965.0) * 1e6,
                             "achieved_evals_per_s": float(P_loc) * L_loc / (k_ms * 1e-3),
                             "frac": float(P_loc) * L_loc / (k_ms * 1e-3)
                                     / (evals_per_cycle * 148 * (clocks["sm_mhz"] or 1965.0) * 1e6)},
                         "note": "each record byte is reused by every local candidate "
                                 f"({P_loc} evaluations per 2-byte record), so the kernel sits on "
                                 "the shared-memory-gather/issue roof, not the HBM roof; see "
                                 "DESIGN.md §4 and profiles/"},
            "clocks": clocks,
        }
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_link_streaming_baseline(n, a_host, b_host, perms_all)
        print(json.dumps(line))
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
