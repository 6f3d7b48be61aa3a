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


def workload_string(name):
    """The one description both arms print (the driver compares the two `config.workload` strings)."""
    n, L, P = WORKLOADS[name][:3]
    idx = {"cfg2": 1, "cfg3": 2, "cfg4": 3}.get(name)
    return (f"{name}: {n} contigs / {L} link records / population {P}"
            + (f" (BASELINE.json configs[{idx}])" if idx is not None else ""))


def kernel_source_sha():
    """sha256 of the sources the scaffolding scoring kernels are built from: an ncu capture quoted in the
    bench line (roofline.traffic) names the build it was taken from, and is dropped when they differ."""
    import hashlib
    h = hashlib.sha256()
    for f in ("scaff_kernels.cu", "scaff_stream1.cuh", "scaff_stream1p.cuh", "device_utils.cuh", "common.cuh"):
        with open(os.path.join(ROOT, "lxy_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


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


def link_owners(a, b, n, world):
    """torch mirror of lxy_comm_link_owners (comm.cu): contiguous tile ranges of the sorted store."""
    import torch
    nblk = (n + 255) // 256
    ntiles = nblk * (nblk + 1) // 2
    lo, hi = torch.minimum(a, b).long() >> 8, torch.maximum(a, b).long() >> 8
    t = lo * nblk - lo * (lo - 1) // 2 + (hi - lo)
    return ((t + 1) * world - 1) // ntiles


def gen_links_tile_shard(n, L, seed, device, world, rank):
    """The same L records as the `world` per-rank streams of gen_links(..., stream_seed=seed * 1000 + r)
    together, but rank `rank` keeps the ones lxy_comm_link_owners assigns to it (whole tiles per rank)."""
    import torch
    keep_a, keep_b, order = [], [], None
    for r in range(world):
        a, b, order = gen_links(n, L // world, seed, device, stream_seed=seed * 1000 + r)
        mine = link_owners(a, b, n, world) == rank
        keep_a.append(a[mine])
        keep_b.append(b[mine])
        del a, b, mine
    return torch.cat(keep_a), torch.cat(keep_b), order


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


def measure_ga(ctx, N, torch, dist, dev, P, n, L_eval, rank, world, generations, shard_mode=None):
    """GA generations/sec (BASELINE.json's third metric): the device-resident loop of lxy_ga_* with
    the reference's Scaffold parameters (PBreed .2, PMutate .6, tournament(0.7, 5), shift + switch +
    invert; scaff.go:72-94) at the workload's population size.  N > 1: lxy_ga_set_sharding - every
    rank runs the same deterministic plan and makes the same children, the library scores them sharded
    over the context's communicator (NCCL on the context's stream: the score vector is the only
    per-generation traffic between GPUs)."""
    ga = N.GA(ctx, N.GA_SCAFF, popsize=P, seed=1234)
    if world > 1:
        ga.set_sharding(N.SHARD_POPULATION if shard_mode is None else shard_mode)
    ga.init()
    for _ in range(3):
        ga.optimize(1)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    first = ga.best()[1]
    t0 = time.perf_counter()
    children = 0
    parts = {"ms_operators": 0.0, "ms_score": 0.0, "ms_select": 0.0}
    for _ in range(generations):
        ga.optimize(1)
        st = ga.stats()
        children += st["last_children"]
        for k in parts:
            parts[k] += st[k]
    ctx.sync()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    best_gene, best = ga.best()
    check = None
    if dist is not None:
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt[0])
        # every rank must hold the same best individual after the sharded generations
        h = torch.tensor([int(np.int64(np.sum(best_gene.astype(np.int64) * np.arange(1, n + 1)) % (1 << 61))),
                          int(round(best * 100))], dtype=torch.int64, device=dev)
        lo, hi = h.clone(), h.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        check = "ok" if bool((lo == hi).all().item()) else "FAILED: ranks disagree on the best individual"
    ga.close()
    out = {"what": "device-resident GA loop (selection, PMX, mutators, scoring of the children, sort/"
                   "truncate on the GPU; only the child count crosses PCIe per generation); wall clock "
                   "around lxy_ga_optimize(1) calls, max over ranks",
           "generations_per_sec": generations / dt, "ms_per_generation": 1e3 * dt / generations,
           "population": P, "children_per_generation": children / generations,
           "link_evals_per_sec": children * float(L_eval) / dt,
           "device_ms_per_generation": {k: v / generations for k, v in parts.items()},
           "best_fitness_before_after": [first, best]}
    if check is not None:
        out["multi_gpu_check"] = check
    return out


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
    the oracle, all host threads, the WHOLE population per step.  Its cost does not depend on the
    number of records, so link-evals/s = candidates/s x L (equivalent evaluations)."""
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
    # bounded sample: the whole population per step unless that would take minutes
    t0 = time.perf_counter()
    table.scaff_score_batch(perms[:cores], nthreads=cores)
    per_cand = (time.perf_counter() - t0) / cores
    budget = 120.0 / max(1, args.steps + args.warmup)
    Ps = P if per_cand * P <= budget else max(cores, int(budget / per_cand) // cores * cores)
    Ps = min(P, Ps)
    for _ in range(args.warmup):
        table.scaff_score_batch(perms[:Ps], nthreads=cores)
    t0 = time.perf_counter()
    for s in range(args.steps):
        table.scaff_score_batch(perms[:Ps], nthreads=cores)
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
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps * (P / Ps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "candidates_per_sec": cand_per_s, "ga": ga_ref,
        "config": {"workload": workload_string(name)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
# one scoring workload on the ranks of this job
# ------------------------------------------------------------------------------------------------
class Job:
    """torch / distributed plumbing shared by the measurements."""

    def __init__(self, torch, N, dist, dev, rank, world, local):
        self.torch, self.N, self.dist, self.dev = torch, N, dist, dev
        self.rank, self.world, self.local = rank, world, local

    def bracket(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, *vals):
        if self.dist is None:
            return vals
        tt = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
        return tuple(float(v) for v in tt)

    def init_comm(self, ctx):
        """The NCCL communicator of the library's context: rank 0's unique id travels through
        torch.distributed once; from then on every collective of the scoring path is issued by
        liblxyb200 on the context's own stream."""
        if self.world == 1:
            return
        torch = self.torch
        idt = torch.zeros(self.N.COMM_ID_BYTES, dtype=torch.uint8, device=self.dev)
        if self.rank == 0:
            idt.copy_(torch.frombuffer(bytearray(self.N.comm_unique_id()), dtype=torch.uint8))
        self.dist.broadcast(idt, 0)
        ctx.comm_init_rank(self.world, self.rank, bytes(idt.cpu().numpy().tobytes()))


def time_path(job, cx, dev_call, host_call, steps, warmup, sample_clocks=False):
    """K device-resident steps (CUDA events on the context's stream) and K host-buffer steps (wall
    clock); at N > 1 every step includes the library's collective.  Both are the max over ranks."""
    torch = job.torch
    stream = torch.cuda.ExternalStream(cx.stream(), device=job.dev)
    for _ in range(max(warmup, 3)):
        dev_call()
    sampler = ClockSampler(job.local) if sample_clocks else None
    job.bracket()
    if sampler:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    tms = {"ms_score": [], "ms_layout": [], "ms_finish": []}
    # The calls queue back to back: the library's own event times are read once, after the K steps (reading
    # them per step waits for the step).  A context with LXY_FLAG_DEFER_ERRORS does not wait for its
    # validation flags either; cx.sync() below reports them (and raises if any step saw a bad row).
    for _ in range(steps):
        dev_call()
    ev1.record(stream)
    cx.sync()
    t = cx.timing()
    for k in tms:
        tms[k].append(t[k])
    launches = t["launches"] * steps
    job.bracket()
    clk = sampler.stop() if sampler else None
    ms_d = ev0.elapsed_time(ev1)
    last = cx.timing()
    for _ in range(2):
        host_call()
    job.bracket()
    t0 = time.perf_counter()
    for _ in range(steps):
        host_call()
    job.bracket()
    ms_h = 1e3 * (time.perf_counter() - t0)
    ms_d, ms_h = job.max_over_ranks(ms_d, ms_h)
    by_rank = None
    if job.dist is not None:   # every rank's own kernel time: the step waits for the slowest (collective per step)
        mine = torch.tensor([float(np.mean(tms["ms_score"]))], dtype=torch.float64, device=job.dev)
        allr = [torch.zeros_like(mine) for _ in range(job.world)]
        job.dist.all_gather(allr, mine)
        by_rank = [round(float(t.item()), 4) for t in allr]
    return {"ms_dev": ms_d, "ms_e2e": ms_h, "launches": launches, "clocks": clk, "last": last,
            "kernel_ms": {k: float(np.mean(v)) for k, v in tms.items()}, "ms_score_by_rank": by_rank}


def check_multi_gpu(job, name, ctx, link_mode, n, L_loc, P, a, b, perms_all, perm_dev, x100_dev, seed):
    L_all = WORKLOADS[name][1]
    """Outside the timed region: is the sharded result the single-GPU result, and the oracle's?
    (1) every rank holds the same [P] vector of exact integers; (2) rank 0 scores the whole population
    against ALL records on its own GPU (link-sharded: shard by shard, the other ranks' record streams
    regenerated from their seeds) and compares all P integers; (3) two sampled candidates against the
    oracle over all records (link-sharded: every rank runs the oracle over its own shard on the host,
    the partial sums are added)."""
    torch, dist, N, dev = job.torch, job.dist, job.N, job.dev
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    mode = N.SHARD_LINKS if link_mode else N.SHARD_POPULATION
    fails = []
    ctx.scaff_score_sharded_dev(mode, perm_dev.data_ptr(), P, torch.zeros(P, dtype=torch.float64, device=dev).data_ptr(),
                                x100_dev.data_ptr())
    ctx.sync()
    got = x100_dev.clone()
    lo, hi = got.clone(), got.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if not bool((lo == hi).all().item()):
        fails.append("ranks hold different score vectors")
    # (2) single GPU, all records
    if job.rank == 0:
        one = torch.zeros(P, dtype=torch.int64, device=dev)
        tmp_o = torch.zeros(P, dtype=torch.float64, device=dev)
        tmp_x = torch.zeros(P, dtype=torch.int64, device=dev)
        if not link_mode:
            ctx.scaff_score_batch_dev(perm_dev.data_ptr(), P, tmp_o.data_ptr(), one.data_ptr())
            ctx.sync()
        else:
            with N.Context(job.local, path=N.PATH_STREAM) as sc:
                for r in range(job.world):   # the `world` record streams the shards were cut from, one at a time
                    ra, rb, _ = gen_links(n, L_all // job.world, seed, dev, stream_seed=seed * 1000 + r)
                    torch.cuda.synchronize()   # the library's stream does not wait for torch's
                    sc.set_links_dev(n, int(ra.numel()), ra.data_ptr(), rb.data_ptr(), None)
                    sc.scaff_score_batch_dev(perm_dev.data_ptr(), P, tmp_o.data_ptr(), tmp_x.data_ptr())
                    sc.sync()
                    one += tmp_x
                    del ra, rb
        if not bool((one == got).all().item()):
            bad = torch.nonzero(one != got).flatten()[:3].tolist()
            fails.append(f"{int((one != got).sum().item())} of {P} sharded scores differ from the single-GPU result "
                         f"(first: {[(i, int(one[i]), int(got[i])) for i in bad]})")
    # (3) the oracle on two sampled candidates, over all records
    pick = np.array([1, P // 2 + 1], np.int64)
    cores = max(1, (os.cpu_count() or 1) // (job.world if link_mode else 1))
    if link_mode or job.rank == 0:
        ah, bh = a.cpu().numpy().astype(np.uint32), b.cpu().numpy().astype(np.uint32)
        _, ox = O.scaff_score_links_batch(n, ah, bh, np.ones(len(ah)), perms_all[pick], nthreads=min(cores, 2))
        del ah, bh
        ox = torch.from_numpy(ox).to(dev)
    else:
        ox = torch.zeros(2, dtype=torch.int64, device=dev)
    if link_mode:
        dist.all_reduce(ox, op=dist.ReduceOp.SUM)
    if (link_mode or job.rank == 0) and ox.tolist() != got[torch.from_numpy(pick).to(dev)].tolist():
        fails.append("sampled candidates differ from the oracle over all records")
    flag = torch.tensor([len(fails)], dtype=torch.int64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.SUM)
    if job.rank == 0:
        return "ok" if not fails and int(flag.item()) == 0 else "FAILED: " + ("; ".join(fails) or "on another rank")
    return None


def run_scoring(job, name, steps, warmup, sample_clocks, check):
    """Inputs, link store, communicator and the two timed legs of one named workload.
    Population-sharded (links replicated) unless the workload is a link-sharded one."""
    torch, N, dev, rank, world = job.torch, job.N, job.dev, job.rank, job.world
    n, L, P, seed, pseed = WORKLOADS[name]
    link_mode = name in LINK_SHARDED
    if link_mode:       # records sharded by tile range (lxy_comm_link_owners), the whole population on every rank
        if world > 1:
            a, b, order = gen_links_tile_shard(n, L, seed, dev, world, rank)
        else:
            a, b, order = gen_links(n, L, seed, dev, stream_seed=seed * 1000)
        L_loc = int(a.numel())
    else:               # links replicated, the population sharded by the library
        L_loc = L
        a, b, order = gen_links(n, L, seed, dev)
    perms_all = gen_population(n, P, pseed, order.cpu().numpy())
    perms = pinned(perms_all)
    torch.cuda.synchronize()
    # the raw record stream: what `value` is defined on.  LXY_FLAG_DEFER_ERRORS: the device-pointer calls of the
    # timed region run their on-device validation every step and hand its verdict to the one lxy_ctx_sync()
    # after the K steps; the host-buffer calls of the e2e leg wait for results and errors as always
    ctx = N.Context(job.local, path=N.PATH_STREAM, flags=N.FLAG_DEFER_ERRORS)
    t0 = time.perf_counter()
    ctx.set_links_dev(n, L_loc, a.data_ptr(), b.data_ptr(), None)
    t_build = time.perf_counter() - t0
    job.init_comm(ctx)
    perm_dev = torch.from_numpy(perms).to(dev)
    out_dev = torch.zeros(P, dtype=torch.float64, device=dev)
    x100_dev = torch.zeros(P, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    mode = N.SHARD_LINKS if link_mode else N.SHARD_POPULATION
    perms16 = pinned(perms_all.astype(np.uint16)) if (world == 1 and n <= 65536) else None
    if world == 1:
        dev_call = lambda: ctx.scaff_score_batch_dev(perm_dev.data_ptr(), P, out_dev.data_ptr(), x100_dev.data_ptr())
        # the call the cgo shim makes (go/scaff/gpu.go): 16-bit genes while the contig ids fit them
        host_call = (lambda: ctx.scaff_score_batch_u16(perms16, want_x100=True)) if perms16 is not None else \
                    (lambda: ctx.scaff_score_batch(perms, want_x100=True))
    else:
        dev_call = lambda: ctx.scaff_score_sharded_dev(mode, perm_dev.data_ptr(), P, out_dev.data_ptr(), x100_dev.data_ptr())
        host_call = lambda: ctx.scaff_score_sharded(mode, perms, want_x100=True)
    main = time_path(job, ctx, dev_call, host_call, steps, warmup, sample_clocks=sample_clocks)
    res = {"name": name, "n": n, "L": L, "L_loc": L_loc, "P": P, "seed": seed, "pseed": pseed,
           "link_mode": link_mode, "ctx": ctx, "a": a, "b": b, "perms_all": perms_all, "perms": perms,
           "perm_dev": perm_dev, "out_dev": out_dev, "x100_dev": x100_dev, "t_build": t_build, "main": main,
           "h2d_bytes": int(perms16.nbytes if perms16 is not None else perms.nbytes),
           "host_entry": "lxy_scaff_score_batch_u16 (16-bit genes, as the cgo shim passes them)" if perms16 is not None
                         else ("lxy_scaff_score_sharded" if world > 1 else "lxy_scaff_score_batch"),
           "P_loc": P if link_mode else N.comm_slice(P, world, rank)[1]}
    L_total = (L // world) * world if link_mode else L
    evals = float(P) * float(L_total)
    res["L_total"] = L_total
    res["value"] = evals * steps / (main["ms_dev"] * 1e-3)
    res["e2e_value"] = evals * steps / (main["ms_e2e"] * 1e-3)
    res["stream_x100"] = x100_dev.clone()
    if world > 1 and check:
        res["multi_gpu_check"] = check_multi_gpu(job, name, ctx, link_mode, n, L_loc, P, a, b, perms_all,
                                                 perm_dev, x100_dev, seed)
    return res


def parallelism_string(res, world):
    if res["link_mode"]:
        return (f"link-sharded x{world} (ncclAllReduce of exact int64 partials, issued by liblxyb200 on the "
                "context's stream), population replicated")
    return (f"population-sharded x{world} (ncclAllGather of the score slices, issued by liblxyb200 on the "
            "context's stream), links replicated")


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
                    help="skip the side measurements (streaming sweep, best exact path, G-profile, phasing, GA, cfg4)")
    ap.add_argument("--cfg4", default="auto", choices=["auto", "on", "off"],
                    help="also run BASELINE configs[3] link-sharded (auto: at 8 GPUs)")
    ap.add_argument("--cfg4-workload", default="cfg4", choices=sorted(LINK_SHARDED))
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
    job = Job(torch, N, dist, dev, rank, world, local)

    # ---- the headline path: the raw record stream, R-profile ----
    R = run_scoring(job, args.workload, args.steps, args.warmup, sample_clocks=True, check=True)
    ctx, a, b, perms, perms_all = R["ctx"], R["a"], R["b"], R["perms"], R["perms_all"]
    perm_dev, out_dev, x100_dev = R["perm_dev"], R["out_dev"], R["x100_dev"]
    link_mode, L_loc, P_loc = R["link_mode"], R["L_loc"], R["P_loc"]
    main_t = R["main"]
    ms_dev, ms_e2e, launches, clocks, timing = (main_t["ms_dev"], main_t["ms_e2e"], main_t["launches"],
                                                main_t["clocks"], main_t["last"])
    stream_x100 = R["stream_x100"]
    a_host = a[:2_000_000].cpu().numpy().astype(np.uint32) if rank == 0 else None
    b_host = b[:2_000_000].cpu().numpy().astype(np.uint32) if rank == 0 else None
    # the same host-buffer call with PAGEABLE caller memory (what a cgo caller passes): the library
    # stages it through its pinned buffer with a few host threads
    ms_e2e_pageable = None
    if world == 1:
        narrow = n <= 65536                      # the same entry point as the e2e leg, caller memory pageable
        perms_pageable = np.array(perms.astype(np.uint16) if narrow else perms, copy=True)
        call_pageable = ctx.scaff_score_batch_u16 if narrow else ctx.scaff_score_batch
        call_pageable(perms_pageable)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            call_pageable(perms_pageable)
        ms_e2e_pageable = 1e3 * (time.perf_counter() - t0) / args.steps
        del perms_pageable

    extras = {}
    peak, peak_src = measured_peak_hbm()
    streaming = None
    if not args.no_extras and not link_mode and world == 1:
        # ---- streaming regime: a lone Score() call (the reference's literal boundary hands ONE
        # ordering to score(), scaff/genome.go:153-159).  No cross-candidate reuse, so this pass is
        # HBM-bound: k_scaff_stream1 reads every 2-byte record once and scores it once. ----
        sweep, streaming = {}, {}
        nv_ctx = N.Context(local, path=N.PATH_STREAM, flags=N.FLAG_NO_VALIDATE)   # no error-flag round trip
        nv_ctx.set_links_dev(n, L, a.data_ptr(), b.data_ptr(), None)
        for ps in (1, 2, 4, 8):
            pdev = perm_dev[:ps].contiguous()
            call = lambda: ctx.scaff_score_batch_dev(pdev.data_ptr(), ps, out_dev.data_ptr(), x100_dev.data_ptr())
            for _ in range(3):
                call()
            stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = max(args.steps, 10)
            kms, tms_call = [], []
            e0.record(stream)
            for _ in range(reps):
                call()
                tt = ctx.timing()
                kms.append(tt["ms_score"])
                tms_call.append(tt["ms_total"])
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            k = float(np.mean(kms))
            # the same calls without validation: nothing comes back to the host, the calls queue up
            ncall = lambda: nv_ctx.scaff_score_batch_dev(pdev.data_ptr(), ps, out_dev.data_ptr(), x100_dev.data_ptr())
            for _ in range(3):
                ncall()
            nstream = torch.cuda.ExternalStream(nv_ctx.stream(), device=dev)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(nstream)
            for _ in range(reps):
                ncall()
            e1.record(nstream)
            torch.cuda.synchronize()
            ms_nv = e0.elapsed_time(e1) / reps
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.scaff_score_batch(perms[:ps])
            ms_h = 1e3 * (time.perf_counter() - t0) / reps
            alg = 2.0 * L_loc + 4.0 * n * ps            # every record once + the orderings once
            frac = alg / (k * 1e-3) / 1e9 / peak
            sweep[f"P={ps}"] = {
                "value": ps * float(L_loc) / (ms * 1e-3), "unit": UNIT, "ms_per_call": ms,
                "ms_per_call_no_validate": ms_nv,
                "e2e": {"value": ps * float(L_loc) / (ms_h * 1e-3), "unit": UNIT, "ms_per_call": ms_h},
                "kernel_ms": k,
                "roofline": {"bound": "hbm", "achieved": alg / (k * 1e-3) / 1e9, "peak": peak,
                             "unit": "GB/s", "frac": frac,
                             "algorithmic_bytes_per_launch": alg, "peak_source": peak_src}}
            if ps <= 4:
                streaming[f"P{ps}"] = {"kernel": "k_scaff_stream1" if ps == 1 else "k_scaff_stream1p",
                                       "kernel_ms": k, "bytes": alg, "achieved_GBps": alg / (k * 1e-3) / 1e9,
                                       "frac": frac, "call_ms_device": ms, "call_ms_device_no_validate": ms_nv,
                                       # first to last CUDA event of one call, by the library's own events
                                       # (call_ms_device also holds the host's turn-around between calls)
                                       "call_ms_library_events": float(np.mean(tms_call))}
        nv_ctx.close()
        extras["streaming_regime"] = {
            "what": "lone Score() calls / tiny batches on the same resident link store: lane = record, "
                    "the whole placement table in shared memory (k_scaff_stream1 / k_scaff_stream1p); "
                    "P=1 is the HBM-bound case north_star's '% of HBM roofline' refers to",
            "sweep": sweep}
    if not args.no_extras and not link_mode and world == 1:
        # ---- what a default-constructed context does with the same inputs (LXY_PATH_AUTO) ----
        pctx = N.Context(local)
        t0 = time.perf_counter()
        pctx.set_links_dev(n, L, a.data_ptr(), b.data_ptr(), None)
        t_pbuild = time.perf_counter() - t0
        chosen, stream_records = pctx.path()
        pr = time_path(job, pctx, lambda: pctx.scaff_score_batch_dev(perm_dev.data_ptr(), P, out_dev.data_ptr(),
                                                                     x100_dev.data_ptr()),
                       lambda: pctx.scaff_score_batch(perms), args.steps, args.warmup)
        same = bool((x100_dev == stream_x100).all().item())
        e2e_u16 = None
        if n <= 65536:      # the same call with 16-bit genes: half the bytes over PCIe
            perms16 = pinned(perms.astype(np.uint16))
            for _ in range(2):
                pctx.scaff_score_batch_u16(perms16)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                _, x16 = pctx.scaff_score_batch_u16(perms16, want_x100=True)
            ms16 = 1e3 * (time.perf_counter() - t0) / args.steps
            e2e_u16 = {"entry_point": "lxy_scaff_score_batch_u16", "ms_per_step": ms16,
                       "candidates_per_sec": float(P) / (ms16 * 1e-3), "h2d_bytes_per_step": int(perms16.nbytes),
                       "identical_x100": bool((x16 == stream_x100.cpu().numpy()).all())}
        ga_auto = measure_ga(pctx, N, torch, dist, dev, P, n, L, rank, world, max(args.steps, 5))
        pctx.close()
        extras["best_exact_path"] = {
            "what": "the default context (LXY_PATH_AUTO) on the same inputs: the path chosen at set_links from "
                    "a cost model (dense pair table = the reference's own 6n gathers per candidate when 4n^2 "
                    "bytes fit and the link set is dense; else the stream with the records of a pair summed, "
                    "Links.Add); value = candidates/s x L (the convention of the reference arm)",
            "path": {N.PATH_STREAM: "stream", N.PATH_PAIRS: "pairs"}[chosen],
            "stream_records_after_summing": int(stream_records),
            "value": float(P) * L * args.steps / (pr["ms_dev"] * 1e-3), "unit": UNIT,
            "candidates_per_sec": float(P) * args.steps / (pr["ms_dev"] * 1e-3),
            "ms_per_step": pr["ms_dev"] / args.steps,
            "e2e": {"value": float(P) * L * args.steps / (pr["ms_e2e"] * 1e-3), "unit": UNIT,
                    "candidates_per_sec": float(P) * args.steps / (pr["ms_e2e"] * 1e-3),
                    "ms_per_step": pr["ms_e2e"] / args.steps},
            "e2e_u16": e2e_u16,
            "kernel_ms": pr["kernel_ms"], "table_bytes": 4 * n * n, "set_links_s": round(t_pbuild, 3),
            "identical_x100": same, "ga": ga_auto}

        # ---- G-profile (north_star's generalisation): lengths, orientations, in-contig offsets ----
        g = torch.Generator(device=dev)
        g.manual_seed(seed + 17)
        lens_t = torch.exp(torch.randn(n, generator=g, device=dev, dtype=torch.float64) * 0.6
                           + float(np.log(min(1_000_000, 1_600_000_000 // n)))).clamp_(1000).long()
        xa = (torch.rand(L, generator=g, device=dev, dtype=torch.float64) * lens_t[a.long()]).int()
        xb = (torch.rand(L, generator=g, device=dev, dtype=torch.float64) * lens_t[b.long()]).int()
        ori_host = pinned((np.random.default_rng(pseed + 31).random((P, n)) < 0.5).astype(np.uint8))
        ori_dev = torch.from_numpy(ori_host).to(dev)
        unit = int(lens_t.double().median().item())
        gctx = N.Context(local)
        gctx.set_links_g_dev(lens_t.cpu().numpy().astype(np.uint32), unit, L, a.data_ptr(), xa.data_ptr(),
                             b.data_ptr(), xb.data_ptr(), None)
        del xa, xb
        gr = time_path(job, gctx, lambda: gctx.scaffg_score_batch_dev(perm_dev.data_ptr(), ori_dev.data_ptr(), P,
                                                                      out_dev.data_ptr(), x100_dev.data_ptr()),
                       lambda: gctx.scaffg_score_batch(perms, ori_host), args.steps, args.warmup)
        g_grid = gr["last"]["grid"]
        gctx.close()
        g_ms = gr["kernel_ms"]["ms_score"]
        g_alg = 10.0 * L + 4.0 * P * n + 16.0 * ((P + 127) // 128) * n
        extras["g_profile"] = {
            "what": "contig lengths (log-normal), per-candidate orientations, per-record in-contig "
                    "offsets; bucket = genomic distance / unit (unit = median contig length)",
            "value": float(P) * L * args.steps / (gr["ms_dev"] * 1e-3), "unit": UNIT,
            "ms_per_step": gr["ms_dev"] / args.steps,
            "e2e": {"value": float(P) * L * args.steps / (gr["ms_e2e"] * 1e-3), "unit": UNIT,
                    "ms_per_step": gr["ms_e2e"] / args.steps},
            "kernel_ms": gr["kernel_ms"], "record_bytes": 10, "grid": g_grid,
            "roofline": {"bound": "hbm", "kernel": "k_scaffg_score", "achieved": g_alg / (g_ms * 1e-3) / 1e9,
                         "peak": peak, "unit": "GB/s", "frac": g_alg / (g_ms * 1e-3) / 1e9 / peak,
                         "algorithmic_bytes_per_launch": g_alg, "peak_source": peak_src,
                         "achieved_evals_per_s": float(P) * L / (g_ms * 1e-3)},
            "total_length_bp": int(lens_t.sum().item()), "unit_bp": unit}
        # ---- phasing (BASELINE.json configs[4]): simvarblocks-style blocks, integer-exact ----
        from lxy_b200 import synth as SY
        pn, pP = 300_000, 1024
        pa, pb, pw, truth = SY.phase_links(pn, 5)
        genes = pinned(SY.phase_population(pn, pP, 105, truth))
        words = pinned(N.pack_bits(genes))
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

        def wall(fn):
            fn()                                        # warm-up: sizes the library's pinned staging buffer
            t0 = time.perf_counter()
            for _ in range(args.steps):
                r = fn()
            return 1e3 * (time.perf_counter() - t0) / args.steps, r

        pms_h, by = wall(lambda: hctx.phase_score_batch(genes))
        pms_bits, bi = wall(lambda: hctx.phase_score_batch_bits(words))
        pt_bits = hctx.timing()
        genes_pageable = np.array(genes, copy=True)
        pms_hp, _ = wall(lambda: hctx.phase_score_batch(genes_pageable))
        del genes_pageable
        hctx.close()
        extras["phasing"] = {
            "what": f"phase.score over {pn} simulated variant blocks, {len(pa)} block-pair records, "
                    f"population {pP}; int64 exact",
            "value": float(pP) * len(pa) / (pms * 1e-3), "unit": UNIT, "ms_per_step": pms,
            "e2e": {"value": float(pP) * len(pa) / (pms_bits * 1e-3), "unit": UNIT, "ms_per_step": pms_bits,
                    "entry_point": "lxy_phase_score_batch_bits (row-packed genes, 1 bit per gene)",
                    "h2d_bytes_per_step": int(words.nbytes), "d2h_bytes_per_step": 16 * pP,
                    "host_memory": "pinned", "identical_to_byte_entry_point": bool((by[0] == bi[0]).all())},
            "e2e_bytes": {"value": float(pP) * len(pa) / (pms_h * 1e-3), "unit": UNIT, "ms_per_step": pms_h,
                          "entry_point": "lxy_phase_score_batch (one byte per gene)",
                          "h2d_bytes_per_step": int(genes.nbytes), "pageable_ms_per_step": pms_hp},
            "kernel_ms": {"k_phase_score": pt["ms_score"], "k_phase_bits4": pt["ms_layout"],
                          "k_phase_bits_packed": pt_bits["ms_layout"]}}
        del genes_dev
    if not args.no_extras and not link_mode:
        extras["ga"] = measure_ga(ctx, N, torch, dist, dev, P, n, L, rank, world, max(args.steps, 5))
    del a, b
    R["a"] = R["b"] = None
    torch.cuda.empty_cache()

    # ---- BASELINE configs[3], link-sharded, next to the headline (the 8-GPU record carries it) ----
    want_cfg4 = (args.cfg4 == "on" or (args.cfg4 == "auto" and world == 8)) and not link_mode and not args.no_extras
    if want_cfg4:
        ctx.close()
        torch.cuda.empty_cache()
        C4 = run_scoring(job, args.cfg4_workload, args.steps, args.warmup, sample_clocks=False, check=True)
        c4 = {"what": "BASELINE.json configs[3]: the records split across the ranks (each rank generates and "
                      "holds its own share), every rank scores the whole population against its share, "
                      "lxy_scaff_score_sharded* combines the exact int64 partial sums with ncclAllReduce; "
                      "e2e: each rank uploads 1/N of the population, ncclAllGather over NVLink",
              "config": {"workload": workload_string(args.cfg4_workload), "parallelism": parallelism_string(C4, world),
                         "records_per_gpu": C4["L_loc"], "population_per_gpu": C4["P"]},
              "value": C4["value"], "unit": UNIT, "ms_per_step": C4["main"]["ms_dev"] / args.steps,
              "e2e": {"value": C4["e2e_value"], "unit": UNIT, "ms_per_step": C4["main"]["ms_e2e"] / args.steps,
                      "h2d_bytes_per_step": int(C4["perms"].nbytes), "d2h_bytes_per_step": int(16 * C4["P"]) * world},
              "kernel_ms": C4["main"]["kernel_ms"], "multi_gpu_check": C4.get("multi_gpu_check")}
        C4["ctx"].close()
        n4, L4, P4 = WORKLOADS[args.cfg4_workload][:3]
        del C4
        torch.cuda.empty_cache()
        if world > 1:
            # its own N = 1: rank 0 alone, all records resident on one GPU (the other ranks wait)
            if rank == 0:
                one = Job(torch, N, None, dev, 0, 1, local)
                S1 = run_scoring(one, args.cfg4_workload, max(2, args.steps // 2), 1, sample_clocks=False, check=False)
                c4["n1"] = {"value": S1["value"], "e2e": S1["e2e_value"],
                            "ms_per_step": S1["main"]["ms_dev"] / max(2, args.steps // 2)}
                c4["efficiency_vs_own_n1"] = {"device": c4["value"] / (world * S1["value"]),
                                              "e2e": c4["e2e"]["value"] / (world * S1["e2e_value"])}
                S1["ctx"].close()
                del S1
            dist.barrier()
        extras["cfg4_link_sharded"] = c4

    value, e2e_value = R["value"], R["e2e_value"]
    if rank == 0:
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
        k_ms = main_t["kernel_ms"]["ms_score"]
        achieved = alg_bytes / (k_ms * 1e-3) / 1e9
        sha = kernel_source_sha()
        traffic, traffic_source = None, None
        try:                                          # ncu-measured DRAM bytes of the same launch shape
            with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
                ent = json.load(f).get(args.workload)
            if ent and ent["population_per_gpu"] == P_loc and not link_mode:
                if ent.get("kernel_source_sha") == sha:
                    traffic = ent["bytes"]
                    traffic_source = {"capture": ent["source"], "kernel_source_sha": sha}
                else:
                    traffic_source = {"capture": ent["source"], "kernel_source_sha": ent.get("kernel_source_sha"),
                                      "stale": f"the kernels were rebuilt from other sources ({sha}); not quoted"}
        except Exception:
            pass
        sm_hz = (clocks["sm_mhz"] or 1965.0) * 1e6
        ceiling = evals_per_cycle * 148 * sm_hz
        achieved_evals = float(P_loc) * L_loc / (k_ms * 1e-3)
        roofline = {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_source,
            "peak_source": peak_src, "kernel": "k_scaff_score", "algorithmic_bytes_per_launch": alg_bytes,
            "binding": {
                "roof": "l1_shared_wavefronts",
                "what": "per record-warp (128 evaluations; 64 with 32-bit tables) 2 wavefronts for row B + 2 "
                        "(1) decay-table look-ups + 2/8 for row A (once per 8-record packet) + 1/8 for the "
                        "packet's own 16 bytes = 4.375 (3.375) wavefronts of the L1/shared data pipe, one "
                        "wavefront/cycle/SM; ceiling = evals/cycle/SM x 148 SMs x the SM clock sampled in the run",
                "evals_per_cycle_per_sm": evals_per_cycle, "unit": "link-evals/s",
                "ceiling": ceiling, "achieved": achieved_evals, "frac": achieved_evals / ceiling},
            "streaming": streaming,
            "note": "each record byte is reused by every local candidate "
                    f"({P_loc} evaluations per 2-byte record), so k_scaff_score sits on the shared-memory "
                    "wavefront roof (`binding`), not the HBM roof; the HBM-bound regime north_star's '>= 70 % "
                    "of HBM roofline' refers to is `streaming` (P1 / P2 / P4 orderings per pass over the same "
                    "records: kernel_ms, algorithmic bytes, fraction of the measured peak)"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int64 (exact integer accumulation of 100*T*w; fitness = -x/100 in f64)",
            "data": "synthetic",
            "config": {"workload": workload_string(args.workload),
                       "parallelism": parallelism_string(R, world),
                       "population_per_gpu": P_loc, "records_per_gpu": L_loc,
                       "l2_policy": f"inputs larger than L2 ({2 * L_loc / 1e6:.0f} MB of records per pass)"
                                    if 2 * L_loc > 130e6 else "records fit L2; no flush (steady-state GA loop)",
                       "link_store_build_s": round(R["t_build"], 3),
                       "validation": "every step runs the on-device permutation check; `value`: its verdict is read "
                                     "by one lxy_ctx_sync() after the K steps (LXY_FLAG_DEFER_ERRORS); `e2e`: per call"},
            "candidates_per_sec": P * args.steps / (ms_dev * 1e-3),
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "candidates_per_sec": P * args.steps / (ms_e2e * 1e-3),
                    "h2d_bytes_per_step": R["h2d_bytes"], "d2h_bytes_per_step": int(16 * P) * world,
                    "entry_point": R["host_entry"],
                    "host_memory": "pinned (page-locked caller buffers go straight to the copy engine)",
                    "pageable_ms_per_step": ms_e2e_pageable},
            "gpu_launches": int(launches),
            "kernel_ms": {"k_scaff_score": k_ms, "layout+check": main_t["kernel_ms"]["ms_layout"],
                          "head+finish": main_t["kernel_ms"]["ms_finish"]},
            "k_scaff_score_ms_by_rank": main_t.get("ms_score_by_rank"),
            "roofline": roofline,
            "clocks": clocks,
        }
        if "multi_gpu_check" in R:
            line["multi_gpu_check"] = R["multi_gpu_check"]
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_link_streaming_baseline(n, a_host, b_host, perms_all)
        print(json.dumps(line))
    if not want_cfg4:
        ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
