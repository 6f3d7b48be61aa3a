"""GPU parity of the device-resident GA generation (lxy_b200/csrc/ga_kernels.cu) against the CPU
oracle of the reference's genome operators (oracle/ga_oracle.cpp: scaff/genome.go:47-91 PMX in its
O(n^2) form, Splice/Switch/Invert, phase/genome.go bit operators) and of the survivor selection,
with the children scored by the already-verified GPU scorers and re-checked against the scoring
oracle.  Bit-exact throughout (genes are integers; fitness compared exactly as produced)."""
import numpy as np
import pytest

import _ga_oracle as G
import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu
PRM = dict(p_breed=0.35, p_mutate=0.6, tour_p=0.7, tour_k=5, mutators=7)


def _check_generation(ga, seed, gen, bits=False):
    """One generation through the split API, every stage compared with the oracle."""
    pop, s_pop = ga.population()
    C, _, _, _ = ga.step_begin()
    got, plan = ga.children(C)
    want_plan = G.plan(seed, gen, ga.popsize, ga.n, PRM["p_breed"], PRM["p_mutate"], PRM["tour_p"],
                       PRM["tour_k"], PRM["mutators"])
    assert plan.tobytes() == want_plan.tobytes()
    want = G.children(seed, gen, pop, PRM["p_breed"], PRM["p_mutate"], PRM["tour_p"], PRM["tour_k"],
                      PRM["mutators"], bits=bits)
    assert got.shape == want.shape and (got == want).all()
    ga.score_children(0, C)
    ga.step_end()
    npop, ns = ga.population()
    return pop, s_pop, got, npop, ns


@pytest.mark.parametrize("n,L,P,seed", [(40, 3000, 12, 1), (300, 40000, 40, 2), (1000, 100000, 130, 3),
                                         (2500, 200000, 64, 4)])
def test_scaff_generation_matches_oracle(n, L, P, seed):
    a, b, w = S.scaff_links(n, L, seed)
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links(n, a, b, None)
        with N.GA(ctx, N.GA_SCAFF, popsize=P, seed=seed, **PRM) as ga:
            ga.init()
            pop0, s0 = ga.population()
            assert all(sorted(r.tolist()) == list(range(n)) for r in pop0)      # Randomize keeps permutations
            assert (np.diff(s0) >= 0).all()                                     # sorted best-first
            assert s0.tolist() == ctx.scaff_score_batch(pop0).tolist()
            for gen in range(3):
                pop, s_pop, ch, npop, ns = _check_generation(ga, seed, gen)
                s_ch = ctx.scaff_score_batch(ch) if len(ch) else np.zeros(0)
                order = G.select(s_pop, s_ch)
                allg = np.concatenate([pop, ch])
                assert (npop == allg[order]).all()
                assert ns.tolist() == np.concatenate([s_pop, s_ch])[order].tolist()
            # the survivors' memoised scores are the reference-order oracle's (scaff.go:246-310)
            table = O.Links.from_records(n, a, b, w)
            np.testing.assert_allclose(ns, table.scaff_score_batch(npop, nthreads=4), rtol=1e-9, atol=0)
            st = ga.stats()
            assert st["generations"] == 3 and st["scored"] >= P


def test_scaff_optimize_equals_split_steps_and_improves():
    n, L, P, seed = 400, 60000, 40, 11
    a, b, _ = S.scaff_links(n, L, seed)
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links(n, a, b, None)
        kw = dict(popsize=P, seed=seed, p_breed=0.2, p_mutate=0.6, tour_p=0.7, tour_k=5, mutators=7)
        with N.GA(ctx, N.GA_SCAFF, **kw) as g1, N.GA(ctx, N.GA_SCAFF, **kw) as g2:
            g1.init()
            g2.init()
            first = g1.best()[1]
            g1.optimize(25)
            for _ in range(25):
                C, _, _, _ = g2.step_begin()
                g2.score_children(0, C // 2)          # scored in two pieces, as a sharded caller would
                g2.score_children(C // 2, C - C // 2)
                g2.step_end()
            p1, s1 = g1.population()
            p2, s2 = g2.population()
            assert (p1 == p2).all() and s1.tolist() == s2.tolist()
            assert s1[0] <= first and (np.diff(s1) >= 0).all()
            assert s1.tolist() == ctx.scaff_score_batch(p1).tolist()
            gene, sc = g1.best()
            assert gene.tolist() == p1[0].tolist() and sc == s1[0]


def test_scaff_ga_recovers_a_small_hidden_order():
    """The reference's own e2e test idea (scaff_test.go:10-30: run Scaffold, expect a perfect
    ordering) on the reference's fixture shape: 11 contigs, Toeplitz weights 10 - (j - i - 1)."""
    n = 11
    ia, ib = np.triu_indices(n, 1)
    w = (10 - (ib - ia - 1)).astype(np.float64)
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links(n, ia.astype(np.uint32), ib.astype(np.uint32), w)
        with N.GA(ctx, N.GA_SCAFF, popsize=40, seed=5) as ga:      # the reference's parameters
            ga.init()
            ga.optimize(300)
            gene, sc = ga.best()
    ident = ctx_score = None
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_links(n, ia.astype(np.uint32), ib.astype(np.uint32), w)
        ident = ctx.scaff_score_batch(np.arange(n, dtype=np.int32)[None, :])[0]
    assert sc <= ident + 1e-9                                       # at least as good as the true order


@pytest.mark.parametrize("n,P,seed", [(300, 10, 1), (2000, 64, 2)])
def test_phase_generation_matches_oracle(n, P, seed):
    pa, pb, pw, truth = S.phase_links(n, seed)
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        ctx.set_phase_links(n, pa, pb, pw)
        with N.GA(ctx, N.GA_PHASE, popsize=P, seed=seed, **PRM) as ga:
            ga.init()
            pop0, s0 = ga.population()
            assert set(np.unique(pop0).tolist()) <= {0, 1} and (np.diff(s0) >= 0).all()
            for gen in range(3):
                pop, s_pop, ch, npop, ns = _check_generation(ga, seed, gen, bits=True)
                s_ch = ctx.phase_score_batch(ch)[1] if len(ch) else np.zeros(0)
                order = G.select(s_pop, s_ch)
                assert (npop == np.concatenate([pop, ch])[order]).all()
                assert ns.tolist() == np.concatenate([s_pop, s_ch])[order].tolist()
            _, exact = O.phase_score_links_batch(n, pa, pb, pw, npop, nthreads=2)
            assert ns.tolist() == [float(x) for x in exact]


def test_ga_error_paths():
    with N.Context(0, path=N.PATH_STREAM) as ctx:
        with pytest.raises(N.LxyError):                 # links not set
            N.GA(ctx, N.GA_SCAFF)
        a, b, _ = S.scaff_links(30, 500, 1)
        ctx.set_links(30, a, b, None)
        with pytest.raises(N.LxyError):
            N.GA(ctx, N.GA_SCAFF, popsize=1)
        with N.GA(ctx, N.GA_SCAFF, popsize=8) as ga:
            with pytest.raises(N.LxyError):             # optimise before init
                ga.optimize(1)
            bad = np.tile(np.arange(30, dtype=np.int32), (8, 1))
            bad[3, 4] = bad[3, 5]
            with pytest.raises(N.LxyError) as e:
                ga.set_population(bad)
            assert e.value.status == -4
            ga.init()
            with pytest.raises(N.LxyError):             # step_end without step_begin
                ga.step_end()
