"""CPU tests of the GA-operator oracle (oracle/ga_oracle.cpp): the reference's genome methods
restated (scaff/genome.go:37-97,125-128,161-209; phase/genome.go:37-59,95-131) against the
independent Python mirror in lxy_b200/host.py, plus the properties of the generation plan
(include/lxy_ga_plan.h)."""
import numpy as np
import pytest

import _ga_oracle as G
from lxy_b200 import host as H


def test_pmx_hand_case():
    # a = 0..7, b reversed, window [2,4]: worked by hand from genome.go:47-91
    a = np.arange(8, dtype=np.int32)
    b = a[::-1].copy()
    ca, cb = G.pmx(a, b, 2, 4)
    assert sorted(ca.tolist()) == list(range(8)) and sorted(cb.tolist()) == list(range(8))
    assert ca[2:5].tolist() == b[2:5].tolist() and cb[2:5].tolist() == a[2:5].tolist()
    # window values 5,4,3 entered ca; 5 sat at index 5 of a (outside) -> repaired from cb's duplicates
    assert ca.tolist() == [0, 1, 5, 4, 3, 2, 6, 7]
    assert cb.tolist() == [7, 6, 2, 3, 4, 5, 1, 0]


@pytest.mark.parametrize("n,seed", [(2, 1), (3, 2), (11, 3), (64, 4), (257, 5), (1000, 6)])
def test_pmx_matches_python_mirror_and_is_a_permutation(n, seed):
    rng = np.random.default_rng(seed)
    for _ in range(40):
        a, b = rng.permutation(n).astype(np.int32), rng.permutation(n).astype(np.int32)
        p1, p2 = sorted(rng.integers(0, n, 2).tolist())
        ca, cb = G.pmx(a, b, p1, p2)
        ga, gb = H.GAOrderedIntGenome(a, None), H.GAOrderedIntGenome(b, None)
        ha, hb = ga.Crossover(gb, p1, p2)
        assert ca.tolist() == ha.Gene and cb.tolist() == hb.Gene
        assert sorted(ca.tolist()) == list(range(n)) and sorted(cb.tolist()) == list(range(n))
        assert ca[p1:p2 + 1].tolist() == b[p1:p2 + 1].tolist()


def test_pmx_identical_parents_and_full_window():
    a = np.random.default_rng(9).permutation(50).astype(np.int32)
    ca, cb = G.pmx(a, a, 10, 30)
    assert ca.tolist() == a.tolist() and cb.tolist() == a.tolist()
    b = np.random.default_rng(10).permutation(50).astype(np.int32)
    ca, cb = G.pmx(a, b, 0, 49)          # whole genome exchanged
    assert ca.tolist() == b.tolist() and cb.tolist() == a.tolist()


@pytest.mark.parametrize("seed", range(5))
def test_mutators_match_genome_methods(seed):
    rng = np.random.default_rng(seed)
    n = 37
    g = rng.permutation(n).astype(np.int32)
    p1, p2 = sorted(rng.integers(0, n, 2).tolist())
    h = H.GAOrderedIntGenome(g, None)
    h.Switch(p1, p2)                                             # genome.go:125-128
    assert G.mutate(G.CHILD_SWITCH, g, p1, p2).tolist() == h.Gene
    h = H.GAOrderedIntGenome(g, None)
    h.Invert(p1, p2)                                             # genome.go:161-177
    assert G.mutate(G.CHILD_INVERT, g, p1, p2).tolist() == h.Gene
    h, src = H.GAOrderedIntGenome(g, None), H.GAOrderedIntGenome(g, None)
    h.Splice(src, p2, p1, 1)                                     # go-galib GAShiftMutator: p2 -> p1 ...
    h.Splice(src, p1, p1 + 1, p2 - p1)                           # ... [p1, p2) one to the right
    assert G.mutate(G.CHILD_SHIFT, g, p1, p2).tolist() == h.Gene
    bits = rng.integers(0, 2, n).astype(np.uint8)
    hb = H.GAFixedBitstringGenome(bits, None)
    hb.Invert(p1, p2)                                            # phase/genome.go:95-109: [p1, p2)
    assert G.mutate(G.CHILD_INVERT, bits, p1, p2, bits=True).tolist() == [int(x) for x in hb.Gene]
    hb = H.GAFixedBitstringGenome(bits, None)
    hb.Switch(p1, p2)
    assert G.mutate(G.CHILD_SWITCH, bits, p1, p2, bits=True).tolist() == [int(x) for x in hb.Gene]


def test_plan_is_deterministic_ordered_and_has_the_right_rates():
    P, n = 2000, 300
    a = G.plan(7, 3, P, n, 0.2, 0.6, 0.7, 5, 7)
    b = G.plan(7, 3, P, n, 0.2, 0.6, 0.7, 5, 7)
    assert a.tobytes() == b.tobytes()
    assert G.plan(8, 3, P, n, 0.2, 0.6, 0.7, 5, 7).tobytes() != a.tobytes()
    assert (np.diff(a["slot"]) >= 0).all()                       # go-galib's append order
    assert (a["p1"] <= a["p2"]).all() and (a["p2"] < n).all() and (a["p1"] >= 0).all()
    ia = np.flatnonzero(a["kind"] == G.CHILD_BREED_A)
    assert (a["kind"][ia + 1] == G.CHILD_BREED_B).all()           # siblings adjacent
    assert abs(len(ia) / P - 0.2) < 0.03
    nm = (a["kind"] >= G.CHILD_SHIFT).sum()
    assert abs(nm / P - 0.6) < 0.04
    for k in (G.CHILD_SHIFT, G.CHILD_SWITCH, G.CHILD_INVERT):     # GAMultiMutator: uniform choice
        assert abs((a["kind"] == k).sum() / nm - 1 / 3) < 0.05
    # tournament(0.7, 5) prefers good ranks: the mean parent rank is far below the uniform P/2
    assert a["pa"][ia].mean() < 0.3 * P
    only_invert = G.plan(7, 3, P, n, 0.0, 1.0, 0.7, 5, 4)
    assert len(only_invert) == P and (only_invert["kind"] == G.CHILD_INVERT).all()
    assert (only_invert["pa"] == np.arange(P)).all()              # Mutate(pop[p])


def test_children_and_select():
    rng = np.random.default_rng(3)
    P, n = 30, 25
    pop = np.stack([rng.permutation(n) for _ in range(P)]).astype(np.int32)
    ch = G.children(5, 0, pop, 0.5, 0.5, 0.7, 5, 7)
    pl = G.plan(5, 0, P, n, 0.5, 0.5, 0.7, 5, 7)
    assert len(ch) == len(pl)
    for c, k in zip(ch, pl):
        assert sorted(c.tolist()) == list(range(n))
    s_pop = np.sort(rng.random(P))
    s_ch = rng.random(len(ch))
    s_ch[0] = s_pop[3]                                             # a tie: the survivor keeps its place
    order = G.select(s_pop, s_ch)
    allk = np.concatenate([s_pop, s_ch])
    assert order.tolist() == np.argsort(allk, kind="stable")[:P].tolist()


def test_eval_phasing_dev_oracle_against_a_second_restatement():
    """EvalPhasingDev (phase.go:266-325): the C++ oracle against an independent restatement over the
    reference's own data shape (maps keyed by `name_<index>`, names missing from the key skipped)."""
    import numpy as np
    import _ga_oracle as G
    rng = np.random.default_rng(21)
    for n, extra in ((1, 0), (5, 2), (37, 5)):
        phasing = {f"b_{3 * i}": bool(rng.integers(0, 2)) for i in range(n + extra)}
        key = {nm: bool(rng.integers(0, 2)) for nm in list(phasing)[:n]}
        m = c = m3 = c3 = m1 = c1 = 0.0
        for i in phasing:                                           # :277-278
            for j in phasing:
                d = abs(int(i.split("_")[1]) - int(j.split("_")[1]))   # :279-285
                if i not in key or j not in key:                    # :289-293
                    continue
                agree = (phasing[i] == phasing[j]) == (key[i] == key[j])   # :301-314
                c += 1; c3 += d <= 2; c1 += d <= 9
                if agree:
                    m += 1; m3 += d <= 2; m1 += d <= 9
        common = [nm for nm in phasing if nm in key]
        got = G.eval_phasing_dev([phasing[nm] for nm in common], [key[nm] for nm in common],
                                 [int(nm.split("_")[1]) for nm in common])
        assert got.tolist() == [m / c, m3 / c3, m1 / c1]


def test_parallel_bracket_matching_rule_equals_the_two_stack_replay():
    """k_ga_children pairs PMX's hits without walking them one by one: with S the prefix sum of the
    events (+1: a ca-side hit, -1: a cb-side hit, 0: none), an event that moves S towards 0 pops the
    first event after the last earlier slot u with S[u] == S[t].  This is the rule, stated in Python,
    against the reference's stack discipline (scaff/genome.go:60-88) on random event strings."""
    import random

    def replay(ev):
        depth, st, pairs = 0, [], set()
        for t, d in enumerate(ev):
            if d == 1:
                if depth < 0:
                    pairs.add((t, st.pop()))
                else:
                    st.append(t)
                depth += 1
            elif d == -1:
                if depth > 0:
                    pairs.add((st.pop(), t))
                else:
                    st.append(t)
                depth -= 1
        return pairs

    def matched(ev):
        S, s = [], 0
        for d in ev:
            s += d
            S.append(s)
        pairs = set()
        for t, d in enumerate(ev):
            if d == 0:
                continue
            cur, prev = S[t], S[t] - d
            if abs(cur) >= abs(prev):
                continue
            u = t - 1
            while u >= 0 and S[u] != cur:
                u -= 1
            p = u + 1
            while ev[p] == 0:
                p += 1
            pairs.add((p, t) if ev[p] == 1 else (t, p))
        return pairs

    rng = random.Random(7)
    for _ in range(2000):
        ev = [rng.choice([0, 0, 1, -1]) for _ in range(rng.randint(1, 80))]
        assert replay(ev) == matched(ev)
