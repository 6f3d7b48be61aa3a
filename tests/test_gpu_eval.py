"""GPU parity of the quality metrics (lxy_eval_scaffolding / lxy_eval_phasing) against the
reference's own loops restated in the oracle (scaff.go:329-384 O(n^3), phase.go:211-263 O(n^2)).
The counts are integers, the results quotients of exactly representable doubles: compared exactly."""
import numpy as np
import pytest

import _ga_oracle as G
import lxy_b200._native as N

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,P,seed", [(3, 4, 1), (4, 6, 2), (11, 40, 3), (64, 16, 4), (150, 8, 5)])
def test_eval_scaffolding_matches_reference_loop(n, P, seed):
    rng = np.random.default_rng(seed)
    key = rng.permutation(n).astype(np.int32)
    perms = np.stack([rng.permutation(n) for _ in range(P)]).astype(np.int32)
    perms[0] = key                                   # perfect: (1, 1)
    perms[1] = key[::-1]                             # reversed key is as good (scaff.go:378-382)
    if P > 2:
        perms[2] = key
        perms[2, :2] = perms[2, 1::-1]               # one adjacent swap
    with N.Context(0) as ctx:
        got = N.eval_scaffolding(ctx, perms, key)
    want = np.stack([G.eval_scaffolding(p, key) for p in perms])
    assert got.tolist() == want.tolist()
    assert got[0].tolist() == [1.0, 1.0] and got[1].tolist() == [1.0, 1.0]


def test_eval_scaffolding_degenerate_and_errors():
    with N.Context(0) as ctx:
        two = N.eval_scaffolding(ctx, np.array([[0, 1]], np.int32), np.array([0, 1], np.int32))
        assert np.isnan(two).all()                   # 0/0 in the reference as well (no triple)
        assert np.isnan(G.eval_scaffolding([0, 1], [0, 1])).all()
        with pytest.raises(N.LxyError):              # an ordering that is no permutation
            N.eval_scaffolding(ctx, np.array([[0, 0, 5]], np.int32), np.array([0, 1, 2], np.int32))
        big = np.arange(5000, dtype=np.int32)
        rng = np.random.default_rng(7)
        perms = np.stack([big, big[::-1], rng.permutation(5000).astype(np.int32)])
        q = N.eval_scaffolding(ctx, perms, big)      # config-3 size: infeasible for the O(n^3) loop
        assert q[0].tolist() == [1.0, 1.0] and q[1].tolist() == [1.0, 1.0]
        assert 0.15 < q[2, 0] < 0.19                 # a random order: about 1/6 of the triples


@pytest.mark.parametrize("n,P,seed", [(1, 2, 1), (2, 3, 2), (4, 5, 3), (11, 10, 4), (300, 16, 5)])
def test_eval_phasing_matches_reference_loop(n, P, seed):
    rng = np.random.default_rng(seed)
    key = rng.integers(0, 2, n).astype(np.uint8)
    genes = rng.integers(0, 2, (P, n)).astype(np.uint8)
    genes[0] = key
    genes[1] = 1 - key                               # the complement phases identically
    with N.Context(0) as ctx:
        got = N.eval_phasing(ctx, genes, key)
    want = np.stack([G.eval_phasing(g, key) for g in genes])
    assert np.array_equal(got, want, equal_nan=True)
    assert got[0, 0] == 1.0 and got[1, 0] == 1.0


@pytest.mark.parametrize("n,P,seed,kind", [(1, 2, 1, "seq"), (2, 3, 2, "seq"), (7, 5, 3, "gaps"), (40, 10, 4, "seq"),
                                           (300, 16, 5, "gaps"), (300, 4, 6, "dups"), (64, 3, 7, "zero"),
                                           (257, 6, 8, "shuffled")])
def test_eval_phasing_dev_matches_reference_loop(n, P, seed, kind):   # phase.go:266-325
    rng = np.random.default_rng(seed)
    key = rng.integers(0, 2, n).astype(np.uint8)
    genes = rng.integers(0, 2, (P, n)).astype(np.uint8)
    genes[0] = key
    genes[1] = 1 - key
    if kind == "seq":
        index = np.arange(n)
    elif kind == "gaps":                             # blocks missing from the key leave holes
        index = np.cumsum(rng.integers(1, 6, n))
    elif kind == "dups":                             # repeated indices (several chromosomes numbered alike)
        index = rng.integers(0, 40, n)
    elif kind == "zero":                             # names whose suffix does not parse: Atoi -> 0 for all
        index = np.zeros(n, np.int64)
    else:                                            # map iteration order is arbitrary in the reference
        index = rng.permutation(n) - 100
    index = index.astype(np.int32)
    with N.Context(0) as ctx:
        got = N.eval_phasing_dev(ctx, genes, key, index)
    want = np.stack([G.eval_phasing_dev(g, key, index) for g in genes])
    assert got.tolist() == want.tolist()
    assert got[0].tolist() == [1.0, 1.0, 1.0] and got[1].tolist() == [1.0, 1.0, 1.0]


def test_eval_phasing_dev_large_and_host_signature():
    from lxy_b200 import host as H
    n = 300_000                                      # config-5 size: 9e10 pairs in the reference's loop
    rng = np.random.default_rng(9)
    key = rng.integers(0, 2, n).astype(np.uint8)
    flip = (rng.random(n) < 0.25).astype(np.uint8)
    genes = np.stack([key, 1 - key, key ^ flip])
    with N.Context(0) as ctx:
        q = N.eval_phasing_dev(ctx, genes, key, np.arange(n, dtype=np.int32))
    assert q[0].tolist() == [1.0, 1.0, 1.0] and q[1].tolist() == [1.0, 1.0, 1.0]
    c1 = int(flip.sum())
    assert q[2, 0] == (float(c1) ** 2 + float(n - c1) ** 2) / float(n) ** 2
    # other blocks agree with probability 0.75^2 + 0.25^2 = 0.625, a block always agrees with itself
    assert abs(q[2, 1] - (1 + 4 * 0.625) / 5) < 0.005 and abs(q[2, 2] - (1 + 18 * 0.625) / 19) < 0.005
    # the reference's map signature (names `chr_<index>`), blocks absent from the key skipped
    names = [f"blk_{i}" for i in range(50)]
    phasing = {nm: bool(rng.integers(0, 2)) for nm in names}
    keym = {nm: bool(rng.integers(0, 2)) for nm in names[:45]}
    keym["other_999"] = True
    got = H.EvalPhasingDev(phasing, keym)
    common = [nm for nm in phasing if nm in keym]
    want = G.eval_phasing_dev([phasing[nm] for nm in common], [keym[nm] for nm in common],
                              [int(nm.split("_")[1]) for nm in common])
    assert list(got) == want.tolist()
