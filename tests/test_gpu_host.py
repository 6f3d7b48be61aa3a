"""The reference-facing host interface on the GPU: genomes scored through ScaffScorer / PhaseScorer
(the `sfunc` of the reference) and the per-generation ScoreBatch pre-pass, against the oracle run
on the reference's own fixture files."""
import random

import numpy as np
import pytest

import _oracle as O
from lxy_b200 import host as H

pytestmark = pytest.mark.gpu


def test_scaffold_fixture_through_the_reference_interface(golden, golden_files):
    links = H.LoadLinks(golden_files["test_ctg"])            # scaff/_testdata/test.ctg.links
    ref = O.Links.load(golden_files["test_ctg"])
    sc = H.ScaffScorer(links)
    genome = H.NewOrderedIntGenome(links.IntIDs(), sc.score)  # scaff.go:89-92
    genome.data = links
    assert genome.Score() == pytest.approx(-319.2999999999998, rel=1e-12)
    rng = random.Random(7)
    pop = []
    for _ in range(40):                                        # gao.Init(40, genome), scaff.go:94
        g = genome.Copy()
        g.Randomize(rng)
        pop.append(g)
    assert sc.ScoreBatch(pop) == 40 and all(g.hasscore for g in pop)
    before = sc.scores
    for g in pop:
        assert g.Score() == pytest.approx(ref.scaff_score(g.Gene), rel=1e-9)
    assert sc.scores == before                                 # memo hits: no further device calls
    child, _ = pop[0].Crossover(pop[1], 2, 6)
    assert not child.hasscore and sc.ScoreBatch(pop + [child]) == 1
    assert child.Score() == pytest.approx(ref.scaff_score(child.Gene), rel=1e-9)
    best = H.Rank(pop)[0]
    assert best.Score() == min(g.Score() for g in pop)
    bad = H.NewOrderedIntGenome([0] * links.Size(), sc.score)  # not a permutation
    assert bad.Score() == float("inf")                         # worst fitness, never elite
    sc.close()


def test_phasing_fixture_through_the_reference_interface(golden_files):
    links = H.LoadLinks(golden_files["test_var"])             # phase/_testdata/test.var.links
    ref = O.Links.load(golden_files["test_var"])
    sc = H.PhaseScorer(links)
    g = H.NewFixedBitstringGenome([False] * links.Size(), sc.score)   # phase.go:59
    assert g.Score() == -265.0
    truth = [c == "A" for c in "AABAAAAAABA"]
    t = H.NewFixedBitstringGenome(truth, sc.score)
    assert t.Score() == -679.0
    rng = random.Random(3)
    pop = []
    for _ in range(10):                                        # gao.Init(10, genome), phase.go:63
        c = g.Copy()
        c.Randomize(rng)
        pop.append(c)
    sc.ScoreBatch(pop)
    for c in pop:
        assert c.Score() == ref.phase_score([int(b) for b in c.Gene])
    sc.close()


def test_scaffold_workflow_on_the_reference_fixture(golden_files, tmp_path):
    """scaff.Scaffold end to end (the reference's own e2e idea, scaff_test.go:10-30) on its fixture
    test.ctg.links: 11 contigs chr2_0..chr2_10 with Toeplitz weights — the GA must find an ordering at
    least as good as the true one, and the output file holds one contig name per line."""
    import lxy_b200.host as H
    links = H.LoadLinks(golden_files["test_ctg"])
    out = tmp_path / "x.scaff.txt"
    scaffolding, inter = H.Scaffold(links, str(out), 400, 0.01, seed=3)
    assert sorted(scaffolding) == sorted(links.StringIDs())
    assert H.ReadScaffolding(str(out)) == scaffolding
    scorer = H.ScaffScorer(links)
    true = H.NewOrderedIntGenome(list(range(11)), scorer.score)
    found = H.NewOrderedIntGenome([links.idKey[x] for x in scaffolding], scorer.score)
    assert found.Score() <= true.Score() + 1e-9
    assert all(i.score >= found.Score() - 1e-9 for i in inter)      # best-so-far never gets worse
    scorer.close()
    assert H.Scaffold.scores > 400                                   # children were scored


def test_phase_workflow_on_the_reference_fixture(golden_files, tmp_path):
    """phase.Phase on test.var.links (11 variants, true phase A A B A A A A A A B A in its header):
    the optimum (-679, brute-forced in SURVEY.md §4) is the key phasing or its complement."""
    import lxy_b200.host as H
    links = H.LoadLinks(golden_files["test_var"])
    out = tmp_path / "x.phase.txt"
    phasing = H.Phase(links, str(out), 300, seed=4)
    bits = [phasing[links.idKeyRev[i]] for i in range(11)]
    key = [c == "A" for c in "AABAAAAAABA"]
    assert bits == key or bits == [not x for x in key]
    assert len(out.read_text().splitlines()) == 11


def test_native_links_attach_and_file_workflows(golden, golden_files, tmp_path):
    """The library's own LoadLinks -> resident store -> scores (golden values of the reference
    fixtures), and the file-to-file workflows lxy_scaffold_file / lxy_phase_file."""
    import lxy_b200._native as N
    nl = N.NativeLinks(golden_files["test_ctg"])
    with N.Context(0) as ctx:
        nl.attach(ctx)
        genes = np.array([c["gene"] for c in golden["test_ctg"]["cases"]], np.int32)
        want = np.array([c["score"] for c in golden["test_ctg"]["cases"]])
        np.testing.assert_allclose(ctx.scaff_score_batch(genes), want, rtol=1e-9, atol=0)
        ident = ctx.scaff_score_batch(np.arange(11, dtype=np.int32)[None, :])[0]
    out = tmp_path / "s.txt"
    best = N.scaffold_file(golden_files["test_ctg"], str(out), 300)
    names = out.read_text().splitlines()
    assert sorted(names) == sorted(nl.Decode(range(11))) and best <= ident + 1e-9
    pout = tmp_path / "p.txt"
    pbest = N.phase_file(golden_files["test_var"], str(pout), 300)
    assert pbest == -679.0                                           # the global optimum
    rows = [r.split("\t") for r in pout.read_text().splitlines()]
    assert len(rows) == 11 and {r[1] for r in rows} <= {"A", "B"}
    with pytest.raises(N.LxyError):
        N.scaffold_file(str(tmp_path / "nope.links"), str(out), 10)


def test_concurrent_score_calls_from_seven_threads():
    """go-galib's GAParallel enters Score() from 7 goroutines at once (scaff.go:88): the C ABI takes
    the context mutex per call, so concurrent single-genome calls (and batch calls mixed in) from 7
    OS threads must return exactly what the same calls return one after the other."""
    import threading

    import _synth as S
    import lxy_b200._native as N
    n, L = 700, 150000
    a, b, w = S.scaff_links(n, L, 77)
    perms = S.scaff_population(n, 7 * 12, 177)
    with N.Context(0) as ctx:
        ctx.set_links(n, a, b, None)
        want, want_x = ctx.scaff_score_batch(perms, want_x100=True)
        got = np.zeros(len(perms))
        gotx = np.zeros(len(perms), np.int64)
        errors = []

        def worker(t):
            try:
                for k in range(12):
                    p = t * 12 + k
                    if k % 4 == 3:                       # a small batch now and then
                        s, x = ctx.scaff_score_batch(perms[p - 1:p + 1], want_x100=True)
                        got[p], gotx[p] = s[1], x[1]
                    else:                                # the reference's literal call: one genome
                        s, x = ctx.scaff_score_batch(perms[p:p + 1], want_x100=True)
                        got[p], gotx[p] = s[0], x[0]
            except Exception as e:                       # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=worker, args=(t,)) for t in range(7)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        assert not errors
        assert gotx.tolist() == want_x.tolist()
        assert got.tolist() == want.tolist()
