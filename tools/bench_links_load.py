"""Side measurement for SURVEY.md §8(f) rank 2: LoadLinks (links.go:427-491) on a synthetic `.links` text
of cfg2's shape (500 contigs, one row per read pair) — the host loader (lxy_links_load, hash maps on
one core) against the device loader (lxy_links_load_gpu / lxy_ctx_load_links).  Prints one JSON line.

    python tools/bench_links_load.py [--lines 5000000] [--names 500] [--reps 3]
    python tools/bench_links_load.py --sam [--lines 4000000]    # ScaffoldLinksFromSam (alignment.go:624-676)
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lxy_b200._native as N  # noqa: E402


def write_links(path, n_names, n_lines, seed):
    """Fixed-width rows `ctgAAAAA ctgBBBBB VVVV.000000\\n`, built as one byte matrix."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, n_names, n_lines)
    b = (a + 1 + rng.geometric(0.02, n_lines)) % n_names
    v = rng.integers(1, 9999, n_lines)
    W = 8 + 1 + 8 + 1 + 11 + 1
    m = np.full((n_lines, W), ord(" "), np.uint8)

    def digits(col, x, width):
        for k in range(width):
            m[:, col + width - 1 - k] = ord("0") + (x // 10 ** k) % 10

    for col, x in ((0, a), (9, b)):
        m[:, col:col + 3] = np.frombuffer(b"ctg", np.uint8)
        digits(col + 3, x, 5)
    digits(18, v, 4)
    m[:, 22] = ord(".")
    m[:, 23:29] = ord("0")
    m[:, 29] = ord("\n")
    with open(path, "wb") as f:
        f.write(b"# synthetic links, cfg2 shape\n")
        f.write(m.tobytes())
    return os.path.getsize(path)


def write_sam(path, n_names, n_lines, seed):
    """Name-sorted SAM of read pairs (two lines per read id), fixed-width rows built as one byte matrix:
    `rdNNNNNNNN\t99\tctgAAAAA\tPPPPPPPP\t60\t100M\t=\t0\t0\t*\t*\n`."""
    rng = np.random.default_rng(seed)
    n_pairs = n_lines // 2
    a = rng.integers(0, n_names, n_pairs)
    b = np.where(rng.random(n_pairs) < 0.6, a, (a + rng.geometric(0.02, n_pairs)) % n_names)
    ctg = np.stack([a, b], 1).reshape(-1)
    rid = np.repeat(np.arange(n_pairs), 2)
    pos = rng.integers(1, 99_999_999, 2 * n_pairs)
    tail = b"\t60\t100M\t=\t0\t0\t*\t*\n"
    W = 10 + 1 + 2 + 1 + 8 + 1 + 8 + len(tail)
    m = np.full((2 * n_pairs, W), ord("\t"), np.uint8)

    def digits(col, x, width):
        for k in range(width):
            m[:, col + width - 1 - k] = ord("0") + (x // 10 ** k) % 10

    m[:, 0:2] = np.frombuffer(b"rd", np.uint8)
    digits(2, rid, 8)
    m[:, 11:13] = np.frombuffer(b"99", np.uint8)
    m[:, 14:17] = np.frombuffer(b"ctg", np.uint8)
    digits(17, ctg, 5)
    digits(23, pos, 8)
    m[:, 31:] = np.frombuffer(tail, np.uint8)
    with open(path, "wb") as f:
        f.write(b"@HD\tVN:1.0\tSO:queryname\n")
        f.write(m.tobytes())
    return os.path.getsize(path)


def main_sam(args):
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "bench.sam")
        size = write_sam(p, args.names, args.lines, 3)
        host_s = []
        for _ in range(max(1, args.reps - 1)):
            t0 = time.perf_counter()
            h = N.NativeLinks.from_sam(p)
            host_s.append(time.perf_counter() - t0)
        gpu_s, stats = [], None
        for _ in range(args.reps + 1):           # the first call also creates the CUDA context
            t0 = time.perf_counter()
            g = N.NativeLinks.from_sam(p, device=args.device)
            gpu_s.append(time.perf_counter() - t0)
            stats = g.load_stats
        ha, hb, hw = h.records()
        ga, gb, gw = g.records()
        same = (h.Size() == g.Size() and h.Decode(range(h.Size())) == g.Decode(range(g.Size()))
                and np.array_equal(ha, ga) and np.array_equal(hb, gb) and np.array_equal(hw, gw))
    host, gpu = min(host_s), min(gpu_s[1:])
    print(json.dumps({
        "what": "ScaffoldLinksFromSam: SAM text -> contig link table (groups of two hits per read id, +1 per pair)",
        "file_bytes": size, "lines": args.lines, "names": stats["names"], "pairs": stats["pairs"],
        "read_pairs_added": stats["records"], "identical_tables": bool(same),
        "host_loop": {"s": host, "MB_per_s": size / host / 1e6, "cores": 1},
        "device_pipeline": {"s": gpu, "MB_per_s": size / gpu / 1e6, "stats_last_call": stats,
                            "device_part_GB_per_s": size / (stats["ms_device"] * 1e-3) / 1e9},
        "speedup_wall": host / gpu,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sam", action="store_true")
    ap.add_argument("--lines", type=int, default=5_000_000)
    ap.add_argument("--names", type=int, default=500)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args()
    if args.sam:
        return main_sam(args)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "bench.links")
        size = write_links(p, args.names, args.lines, 2)
        host_s = []
        for _ in range(max(1, args.reps - 1)):
            t0 = time.perf_counter()
            h = N.NativeLinks(p)
            host_s.append(time.perf_counter() - t0)
        gpu_s, stats = [], None
        for _ in range(args.reps + 1):           # the first call also creates the CUDA context
            t0 = time.perf_counter()
            g = N.NativeLinks(p, device=args.device)
            gpu_s.append(time.perf_counter() - t0)
            stats = g.load_stats
        ctx_s = []
        with N.Context(args.device) as c:
            for _ in range(args.reps):
                t0 = time.perf_counter()
                N.NativeLinks(p, ctx=c)
                ctx_s.append(time.perf_counter() - t0)
        ha, hb, hw = h.records()
        ga, gb, gw = g.records()
        same = (h.Size() == g.Size() and h.Decode(range(h.Size())) == g.Decode(range(g.Size()))
                and np.array_equal(ha, ga) and np.array_equal(hb, gb)
                and np.array_equal(hw.view(np.uint64), gw.view(np.uint64)))
    host, gpu, ctx = min(host_s), min(gpu_s[1:]), min(ctx_s)
    print(json.dumps({
        "what": "LoadLinks: .links text -> link table (ids by first appearance, last row of a pair wins)",
        "file_bytes": size, "lines": args.lines, "names": stats["names"], "pairs": stats["pairs"],
        "identical_tables": bool(same),
        "host_loader": {"s": host, "MB_per_s": size / host / 1e6, "cores": 1},
        "device_loader": {"s": gpu, "MB_per_s": size / gpu / 1e6, "stats_last_call": stats,
                          "device_part_GB_per_s": size / (stats["ms_device"] * 1e-3) / 1e9,
                          "text_passes": 3},
        "device_loader_into_link_store": {"s": ctx, "what": "lxy_ctx_load_links: parse + build of the scoring store"},
        "speedup_wall": host / gpu,
    }))


if __name__ == "__main__":
    main()
