"""The reference-facing host interface on the GPU: genomes scored through ScaffScorer / PhaseScorer
(the `sfunc` of the reference) and the per-generation ScoreBatch pre-pass, against the oracle run
on the reference's own fixture files."""
import random

import numpy as np
import pytest

import _ga_oracle as G
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
    assert H.readPhasing(str(out)) == links.DecodePhasing([1 if b else 0 for b in bits])   # phase.go:119-167 round trip


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
    # writePhasing's format (phase.go:151-167): `name 1` / `name 0`; round trip through readPhasing
    rows = [r.split(" ") for r in pout.read_text().splitlines()]
    assert len(rows) == 11 and {r[1] for r in rows} <= {"0", "1"}
    back = H.readPhasing(str(pout))
    nv = N.NativeLinks(golden_files["test_var"])
    assert list(back) == nv.Decode(range(11))
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


def test_infer_workflows_file_to_file(tmp_path, golden_files):
    """`lxy scaff infer` / `lxy phase infer` (scaff/cmd.go:135-202, phase/cmd.go:114-165) without the
    plots: links parsed on the device, GA, quality metrics, the reference's output files."""
    import numpy as np
    from lxy_b200 import host as H
    # a chain of 12 contigs: adjacent pairs strongly linked, so the GA recovers the order (or its reverse)
    names = [f"ctg{i}" for i in range(12)]
    lp = tmp_path / "chain.links"
    with open(lp, "w") as f:
        for i in range(12):
            for j in range(i + 1, 12):
                f.write(f"{names[i]} {names[j]} {1000 // (j - i) ** 2}\n")
    kp = tmp_path / "key.txt"
    kp.write_text("".join(nm + "\n" for nm in names))
    dev = H.LoadLinksDevice(str(lp))
    ref = H.LoadLinks(str(lp))
    assert dev.idKey == ref.idKey and dev.data == ref.data
    prefix = str(tmp_path / "run")
    scaffolding, score, nscore = H.ScaffoldInfer(str(lp), prefix, str(kp), iterations=300, optReportFreq=0.05, seed=3)
    assert sorted(scaffolding) == sorted(names)
    assert H.ReadScaffolding(prefix + ".scaff.txt") == scaffolding
    assert open(prefix + ".qscore.txt").read() == "score=%f;score_neighbor=%f\n" % (score, nscore)
    want = G.eval_scaffolding([names.index(nm) for nm in scaffolding], list(range(12)))
    assert [score, nscore] == want.tolist()
    assert score > 0.9
    rows = [r.split(" ") for r in open(prefix + ".opt.txt").read().splitlines()]
    assert len(rows) >= 3 and all(len(r) == 3 for r in rows)
    assert [int(r[0]) for r in rows] == sorted(int(r[0]) for r in rows)
    assert float(rows[-1][1]) <= float(rows[0][1])                    # the best fitness never gets worse
    # phasing: the reference fixture, key = the known phasing of SURVEY.md §4 in readPhasing's format
    links = H.LoadLinks(golden_files["test_var"])
    truth = "AABAAAAAABA"
    pk = tmp_path / "phase.key"
    pk.write_text("".join(f"{links.idKeyRev[i]} {1 if c == 'A' else 0}\n" for i, c in enumerate(truth)))
    if all("_" in nm for nm in links.idKey):
        phasing, scores = H.PhaseInfer(golden_files["test_var"], str(tmp_path / "ph"), str(pk), iterations=300,
                                       subruntag="t", seed=4)
        assert set(phasing) == set(links.idKey)
        text = open(str(tmp_path / "ph") + ".t.qscore.txt").read()
        assert text == "score=%f;300kbp_score=%f;1mbp_score=%f\n" % scores
        key = H.readPhasing(str(pk))
        common = [nm for nm in phasing if nm in key]
        want = G.eval_phasing_dev([phasing[nm] for nm in common], [key[nm] for nm in common],
                                  [H._atoi_or_zero(nm.split("_")[1]) for nm in common])
        assert list(scores) == want.tolist()
        assert scores[0] == 1.0                                        # 11 blocks: the GA finds the phasing
