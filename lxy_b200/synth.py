"""Seeded synthetic inputs of the BASELINE.json shapes (numpy only; no device code).

Scaffolding links follow SURVEY.md §8(d): contig pairs are drawn with P(a,b) ∝ 1/|i-j| in a
hidden true order plus 10 % uniform noise pairs, one record per read pair with weight 1 — the
records `ScaffoldLinksFromSam` would `Add(..., 1)` (/root/reference/src/sequtil/alignment.go:
656-660).  Phasing links follow `TabulateVariantBlockLinks` (src/sequtil/links.go:372-420):
signed counts (#in-phase − #out-of-phase read pairs) per block pair.
"""
import numpy as np


def true_order(n, seed):
    return np.random.default_rng(seed ^ 0x5EED).permutation(n).astype(np.int32)


def scaff_links(n, L, seed, noise=0.1, order=None):
    """-> (a, b, w): uint32, uint32, float64 arrays of length L (raw records, w = 1)."""
    rng = np.random.default_rng(seed)
    order = true_order(n, seed) if order is None else np.asarray(order, np.int32)
    d = np.arange(1, n)
    pmf = (n - d) / d.astype(np.float64)          # (#pairs at distance d) * 1/d
    cdf = np.cumsum(pmf / pmf.sum())
    n_noise = int(L * noise)
    n_sig = L - n_noise
    dist = (np.searchsorted(cdf, rng.random(n_sig), side="right") + 1).clip(1, n - 1)
    i = (rng.random(n_sig) * (n - dist)).astype(np.int64)
    a = order[i]
    b = order[i + dist]
    na = rng.integers(0, n, n_noise)
    nb = (na + rng.integers(1, n, n_noise)) % n   # never equal to na
    a = np.concatenate([a, na]).astype(np.uint32)
    b = np.concatenate([b, nb]).astype(np.uint32)
    flip = rng.random(L) < 0.5
    a2 = np.where(flip, b, a)
    b2 = np.where(flip, a, b)
    shuffle = rng.permutation(L)
    return a2[shuffle], b2[shuffle], np.ones(L, np.float64)


def scaff_population(n, P, seed, order=None):
    """-> int32 [P][n]: the true order, its reverse, k-swap neighbours, random permutations."""
    rng = np.random.default_rng(seed)
    order = true_order(n, seed - 100) if order is None else np.asarray(order, np.int32)
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


def contig_lengths(n, seed, median=1_000_000, sigma=0.6, lo=1000):
    rng = np.random.default_rng(seed + 17)
    ln = rng.lognormal(np.log(median), sigma, n)
    return np.maximum(ln, lo).astype(np.uint32)


def scaffg_links(n, L, seed, lens, noise=0.1, order=None):
    """G-profile records: (a, xa, b, xb, w) with in-contig offsets xa < len[a], xb < len[b]."""
    a, b, w = scaff_links(n, L, seed, noise, order)
    rng = np.random.default_rng(seed + 29)
    xa = (rng.random(L) * lens[a]).astype(np.uint32)
    xb = (rng.random(L) * lens[b]).astype(np.uint32)
    return a, xa, b, xb, w


def orientations(n, P, seed):
    return (np.random.default_rng(seed + 31).random((P, n)) < 0.5).astype(np.uint8)


def phase_links(n, seed, lam=12.0, far_pairs=0):
    """simvarblocks-style block links -> (a, b, w, truth): ids in genomic order, signed counts."""
    rng = np.random.default_rng(seed)
    truth = (rng.random(n) < 0.5).astype(np.uint8)
    A, B, W = [], [], []
    for step in range(1, 31):
        if n - step <= 0:
            break
        i = np.arange(0, n - step, dtype=np.int64)
        cnt = rng.poisson(lam / step, len(i))
        agree_true = truth[i] == truth[i + step]
        good = rng.binomial(cnt, 0.95)
        w = np.where(agree_true, good - (cnt - good), (cnt - good) - good)
        keep = cnt > 0
        A.append(i[keep]); B.append((i + step)[keep]); W.append(w[keep])
    if far_pairs:
        fa = rng.integers(0, max(1, n - 40), far_pairs)
        fb = np.minimum(fa + rng.integers(31, 40, far_pairs), n - 1)
        ok = fb - fa > 30
        A.append(fa[ok]); B.append(fb[ok]); W.append(rng.integers(-5, 6, int(ok.sum())))
    a = np.concatenate(A).astype(np.uint32)
    b = np.concatenate(B).astype(np.uint32)
    w = np.concatenate(W).astype(np.float64)
    flip = rng.random(len(a)) < 0.5
    a2, b2 = np.where(flip, b, a), np.where(flip, a, b)
    sh = rng.permutation(len(a))
    return a2[sh], b2[sh], w[sh], truth


def phase_population(n, P, seed, truth=None):
    rng = np.random.default_rng(seed)
    pop = (rng.random((P, n)) < 0.5).astype(np.uint8)
    if truth is not None and P > 0:
        pop[0] = truth
        if P > 1:
            pop[1] = 1 - truth
    return pop
