#!/usr/bin/env python
"""Randomised parity sweep of the scaffolding scorers against the oracle: random contig counts, record
counts (including tiny ones: fewer packets than warps), weight modes, population sizes and flags, so that
the record rings see partial slices, empty warps, one-record tiles and shares that end inside a packet.
`python tools/stress_scaff.py [cases] [seed]` -> prints one line per failure and a summary; exit 1 on failure."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O
import lxy_b200._native as N
from lxy_b200 import synth as S

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
fails = 0
for case in range(cases):
    n = int(rng.choice([2, 3, 11, 40, 255, 256, 257, 300, 700, 2049, 5000, 20000]))
    L = int(rng.choice([1, 7, 8, 9, 63, 64, 65, 500, 4095, 4096, 16383, 16385, 70000, 300000, 1500000]))
    P = int(rng.choice([1, 2, 3, 4, 5, 8, 9, 16, 33, 64, 127, 128, 129, 200, 260]))
    wm = str(rng.choice(["unit", "int", "real"]))
    flags = int(rng.choice([0, 0, N.FLAG_NO_STREAM1, N.FLAG_NO_GRAPH, N.FLAG_DEFER_ERRORS]))
    if n < 3:
        L = min(L, 500)
    a, b, w = S.scaff_links(n, L, int(rng.integers(1, 1 << 30)))
    if len(a) == 0:
        continue
    if wm == "int":
        w = rng.integers(-5, 50, len(a)).astype(np.float64)
    elif wm == "real":
        w = rng.gamma(2.0, 1.5, len(a))
    perms = S.scaff_population(n, P, int(rng.integers(1, 1 << 30)))
    ref = O.Links.from_records(n, a, b, w).scaff_score_batch(perms, nthreads=8)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=8)
    try:
        with N.Context(0, path=N.PATH_STREAM, flags=flags) as c:
            c.set_links(n, a, b, None if wm == "unit" else w)
            for rep in range(2):       # the second call replays the captured graph
                got, x100 = c.scaff_score_batch(perms, want_x100=True)
                if wm != "real":
                    ok = x100.tolist() == ox100.tolist()
                else:
                    ok = np.allclose(got, ref, rtol=1e-9, atol=1e-9)
                if not ok:
                    raise AssertionError("scores differ (call %d)" % rep)
    except Exception as e:   # noqa: BLE001
        fails += 1
        print("FAIL case", case, dict(n=n, L=len(a), P=P, w=wm, flags=flags), repr(e)[:200], flush=True)
# ---- G-profile (lengths, orientations, in-contig offsets): unit weights take k_scaffg2r_score ----
gcases = cases // 3
for case in range(gcases):
    n = int(rng.choice([2, 5, 60, 255, 257, 700, 3000]))
    L = int(rng.choice([1, 8, 9, 127, 129, 2000, 4097, 50000, 300000]))
    P = int(rng.choice([1, 3, 33, 64, 127, 128, 129, 200]))
    unit = int(rng.choice([2, 3, 1000, 13972, 13973, 250000, 777777, 1 << 20]))
    wm = str(rng.choice(["unit", "unit", "int"]))
    seed = int(rng.integers(1, 1 << 30))
    lens = S.contig_lengths(n, seed)
    a, xa, b, xb, w = S.scaffg_links(n, L, seed, lens)
    if len(a) == 0:
        continue
    if wm == "int":
        w = rng.integers(1, 9, len(a)).astype(np.float64)
    perms = S.scaff_population(n, P, seed + 1)
    oris = S.orientations(n, P, seed) if rng.random() < 0.7 else None
    want, ox100 = O.scaffg_score_links_batch(lens, unit, a, xa, b, xb, w, perms, oris, nthreads=8)
    try:
        with N.Context(0) as c:
            c.set_links_g(lens, unit, a, xa, b, xb, None if wm == "unit" else w)
            for rep in range(2):
                got, x100 = c.scaffg_score_batch(perms, oris, want_x100=True)
                if x100.tolist() != ox100.tolist():
                    raise AssertionError("G scores differ (call %d)" % rep)
    except Exception as e:   # noqa: BLE001
        fails += 1
        print("FAIL G case", case, dict(n=n, L=len(a), P=P, unit=unit, w=wm), repr(e)[:200], flush=True)
cases += gcases
print("stress_scaff: %d cases, %d failures" % (cases, fails))
sys.exit(1 if fails else 0)
