"""GPU parity of the G-profile scorer (lengths, orientations, in-contig offsets) against the oracle.
The G-profile is this repository's generalisation (BASELINE.json north_star); the oracle defines
it, and it must collapse to the reference scorer for len == 1, x == 0, unit == 1."""
import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu
REL = 1e-9


@pytest.fixture(scope="module")
def ctx():
    c = N.Context(0)
    yield c
    c.close()


def test_collapses_to_the_reference_scorer(ctx):
    n, L, P = 300, 40000, 70
    a, b, w = S.scaff_links(n, L, 7)
    perms = S.scaff_population(n, P, 107)
    oris = S.orientations(n, P, 7)
    zeros, ones = np.zeros(L, np.uint32), np.ones(n, np.uint32)
    table = O.Links.from_records(n, a, b, w)
    ref = table.scaff_score_batch(perms, nthreads=4)          # the reference's own loop
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=4)
    ctx.set_links_g(ones, 1, a, zeros, b, zeros, None)
    got, x100 = ctx.scaffg_score_batch(perms, oris, want_x100=True)
    assert x100.tolist() == ox100.tolist()
    np.testing.assert_allclose(got, ref, rtol=REL, atol=0)
    start, pos = ctx.scaffg_layout(perms[:3])
    for p in range(3):
        assert start[p].tolist() == list(range(n))              # unit lengths: offset = index
        assert pos[p].tolist() == np.argsort(perms[p]).tolist()


@pytest.mark.parametrize("n,L,P,seed,unit,with_ori", [
    (60, 5000, 10, 1, 1_000_000, True), (500, 80000, 128, 2, 1_000_000, True),
    (700, 100000, 65, 3, 250_000, False), (1500, 150000, 40, 4, 1 << 20, True),
    (257, 20000, 64, 5, 777_777, True), (3000, 200000, 33, 6, 3, True),
])
def test_synthetic_vs_oracle(ctx, n, L, P, seed, unit, with_ori):
    lens = S.contig_lengths(n, seed)
    a, xa, b, xb, w = S.scaffg_links(n, L, seed, lens)
    perms = S.scaff_population(n, P, seed + 100)
    oris = S.orientations(n, P, seed) if with_ori else None
    want, ox100 = O.scaffg_score_links_batch(lens, unit, a, xa, b, xb, w, perms, oris, nthreads=4)
    ctx.set_links_g(lens, unit, a, xa, b, xb, None)
    got, x100 = ctx.scaffg_score_batch(perms, oris, want_x100=True)
    assert x100.tolist() == ox100.tolist()                       # exact integers
    np.testing.assert_allclose(got, want, rtol=REL, atol=0)


def test_layout_offsets_and_distances_bit_exact(ctx):
    n, L = 400, 30000
    lens = S.contig_lengths(n, 11)
    a, xa, b, xb, w = S.scaffg_links(n, L, 11, lens)
    perms = S.scaff_population(n, 4, 111)
    oris = S.orientations(n, 4, 11)
    ctx.set_links_g(lens, 1_000_000, a, xa, b, xb, None)
    start, pos = ctx.scaffg_layout(perms)
    for p in range(4):
        ostart, opos = O.scaffg_layout(lens, perms[p])
        assert start[p].tolist() == ostart.tolist() and pos[p].tolist() == opos.tolist()
        _, _, odist = O.scaffg_score_links(lens, 1_000_000, a, xa, b, xb, w, perms[p], oris[p],
                                           want_dist=True)
        got = ctx.scaffg_link_distances(a, xa, b, xb, perms[p], oris[p])
        assert got.tolist() == odist.tolist()


def test_weights_int_and_real(ctx):
    n, L, P = 350, 50000, 48
    lens = S.contig_lengths(n, 21)
    a, xa, b, xb, _ = S.scaffg_links(n, L, 21, lens)
    perms = S.scaff_population(n, P, 121)
    oris = S.orientations(n, P, 21)
    rng = np.random.default_rng(21)
    wi = rng.integers(1, 50, L).astype(np.float64)
    want, ox = O.scaffg_score_links_batch(lens, 500_000, a, xa, b, xb, wi, perms, oris, nthreads=4)
    ctx.set_links_g(lens, 500_000, a, xa, b, xb, wi)
    got, x100 = ctx.scaffg_score_batch(perms, oris, want_x100=True)
    assert x100.tolist() == ox.tolist()
    wr = rng.gamma(2.0, 1.5, L)
    want, _ = O.scaffg_score_links_batch(lens, 500_000, a, xa, b, xb, wr, perms, oris, nthreads=4)
    ctx.set_links_g(lens, 500_000, a, xa, b, xb, wr)
    got = ctx.scaffg_score_batch(perms, oris)
    np.testing.assert_allclose(got, want, rtol=REL, atol=0)


def test_hand_case_and_errors(ctx):
    lens = np.array([10, 5, 7], np.uint32)
    ctx.set_links_g(lens, 4, [0], [2], [1], [4], [3.0])
    perm = np.array([[2, 0, 1]], np.int32)
    ori = np.array([[0, 1, 0]], np.uint8)            # flip the contig at ordering index 1 (c0)
    got, x100 = ctx.scaffg_score_batch(perm, ori, want_x100=True)
    assert got.tolist() == [-6.0] and x100.tolist() == [600]     # d = 7 -> bucket 1, both pos > 0
    assert ctx.scaffg_link_distances([0], [2], [1], [4], perm[0], ori[0]).tolist() == [7]
    with pytest.raises(N.LxyError):
        ctx.set_links_g(np.array([5, 0], np.uint32), 1, [0], [0], [1], [0], None)   # zero length
    with pytest.raises(N.LxyError):
        ctx.set_links_g(np.full(3, 2_000_000_000, np.uint32), 1, [0], [0], [1], [0], None)  # > 2^32


@pytest.mark.parametrize("unit", [2, 3, 7, 4096, 13_971, 13_973, 318_574, 999_983, 1 << 20, 50_000_001, 190_000_000])
def test_one_multiply_division_is_exact_at_the_bucket_edges(ctx, unit):
    """k_scaffg2_score divides the clamped distance by `unit` with one multiply-high (scaffg2.cuh,
    g2_magic).  Distances of exactly k * unit - 1, k * unit and k * unit + 1 for every k up to 23 — and
    far beyond the clamp — must land in the oracle's bucket, for units on either side of the magic's
    regime changes (22 unit^2 < 2^32 up to 13,972; the fall-back to k_scaffg_score above ~1.9e8)."""
    n = 40 if unit < 10_000_000 else 4
    span = 30 * unit + 10
    lens = np.full(n, min(span, (2**31 - 1000) // n), np.uint32)
    lens[0] = min(span, 2**31 - 1000 - int(lens[1:].sum(dtype=np.int64)))
    rng = np.random.default_rng(unit % 1000)
    a, xa, b, xb = [], [], [], []
    for k in list(range(0, 24)) + [30, 100, 1000]:
        for delta in (-1, 0, 1):
            d = k * unit + delta
            # the last base of contig 0 and base d - 1 of contig 1, adjacent and forward in the identity
            # ordering: exactly d apart
            if 1 <= d <= int(lens[1]):
                a.append(0); xa.append(int(lens[0]) - 1); b.append(1); xb.append(d - 1)
    a, xa, b, xb = (np.array(v, np.uint32) for v in (a, xa, b, xb))
    assert len(a) > (20 if unit < 10_000_000 else 1)      # the 2^31 bp coordinate range bounds k for huge units
    perms = np.stack([np.arange(n, dtype=np.int32)] + [rng.permutation(n).astype(np.int32) for _ in range(140)])
    oris = (rng.random(perms.shape) < 0.5).astype(np.uint8)
    oris[0] = 0
    w = np.ones(len(a))
    want, ox = O.scaffg_score_links_batch(lens, unit, a, xa, b, xb, w, perms, oris, nthreads=4)
    _, _, dist = O.scaffg_score_links(lens, unit, a, xa, b, xb, w, perms[0], oris[0], want_dist=True)
    assert {int(d) % unit for d in dist} >= ({0, 1, unit - 1} if unit < 10_000_000 else {1}) or unit == 2     # the edges are really hit
    ctx.set_links_g(lens, unit, a, xa, b, xb, None)
    got, x100 = ctx.scaffg_score_batch(perms, oris, want_x100=True)
    assert x100.tolist() == ox.tolist()


@pytest.mark.parametrize("n,L,P,unit", [(900, 300_000, 129, 250_000), (2100, 500_000, 300, 1 << 19), (300, 100_000, 1024, 123_457)])
def test_128_candidate_groups_vs_oracle(ctx, n, L, P, unit):
    """Populations above one group, group boundaries inside the population, both orientations, the
    first contig of many candidates among the records (the `i-k > 0` multiplicity path)."""
    lens = S.contig_lengths(n, 31)
    a, xa, b, xb, w = S.scaffg_links(n, L, 31, lens)
    perms = S.scaff_population(n, P, 131)
    oris = S.orientations(n, P, 31)
    want, ox100 = O.scaffg_score_links_batch(lens, unit, a, xa, b, xb, w, perms, oris, nthreads=8)
    ctx.set_links_g(lens, unit, a, xa, b, xb, None)
    got, x100 = ctx.scaffg_score_batch(perms, oris, want_x100=True)
    assert x100.tolist() == ox100.tolist()
    np.testing.assert_allclose(got, want, rtol=REL, atol=0)
    bad = perms.copy()
    bad[P - 1, 0] = bad[P - 1, 1]
    with pytest.raises(N.LxyError):
        ctx.scaffg_score_batch(bad, oris)
