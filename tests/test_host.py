"""Host-side mirror of the reference interface (lxy_b200/host.py): link table semantics against the
oracle's restatement and the reference's own link-table tests; genome operators and the Score()
memo contract.  CPU only (no fitness is computed here)."""
import io
import random

import numpy as np
import pytest

import _oracle as O
from lxy_b200 import host as H


def test_loadlinks_matches_oracle_on_reference_fixtures(golden, golden_files):
    for key in ("test_ctg", "toy_ctg", "test_var", "toy_varblock"):
        mine, ref = H.LoadLinks(golden_files[key]), O.Links.load(golden_files[key])
        assert mine.Size() == ref.size()
        assert [mine.idKeyRev[i] for i in range(mine.Size())] == golden[key]["names"]
        for i in range(mine.Size()):
            for j in range(mine.Size()):
                assert mine.Get(i, j)[0] == ref.get(i, j)[0]
        a, b, w = mine.records()
        ra, rb, rw = ref.export()
        assert a.tolist() == ra.tolist() and b.tolist() == rb.tolist() and w.tolist() == rw.tolist()


def test_reference_link_table_tests_restated():
    # links_test.go:41-56 (toy table), :84-100 (Set + Add), :103-117 (Get), :130-139, :146-159
    l = H.Links()
    l.addKey("a"); l.addKey("b")
    assert l.IntIDs() == [0, 1] and l.Size() == 2
    l.Set(0, 1, 0.1)
    assert l.Get(0, 1) == (0.1, None) and l.Get(1, 0)[0] == 0.1
    l.Add(1, 0, 0.1)
    assert l.Get(0, 1)[0] == 0.2
    assert l.Decode([0, 1]) == ["a", "b"]
    val, err = l.Get(0, 7)
    assert val == -1.0 and err is not None                       # links.go:193-200
    with pytest.raises(H.LinksError):
        l.Set(0, 7, 1.0)
    with pytest.raises(H.LinksError):
        l.Decode([5])
    assert l.DecodePhasing([True, False]) == {"a": True, "b": False}


def test_loadlinks_quirks(tmp_path):
    p = tmp_path / "x.links"
    p.write_text("# header a:0\nc1 c2 3\nc2 c1 5\nc3 c1 oops\nc4  c1 2\n")
    l = H.LoadLinks(str(p))
    assert l.idKey == {"c1": 0, "c2": 1, "c3": 2, "c4": 3, "": 4}   # double space -> empty name
    assert l.Get(0, 1)[0] == 5.0                                 # last row wins, (a,b) == (b,a)
    assert l.Get(0, 2)[0] == 0.0                                 # ParseFloat error -> 0
    assert l.Get(3, 4)[0] == 0.0                                 # "c4", "" , value token "c1" -> 0
    (tmp_path / "e.links").write_text("# only a comment\n")
    with pytest.raises(H.LinksError):
        H.LoadLinks(str(tmp_path / "e.links"))
    (tmp_path / "b.links").write_text("a b 1\n\nc d 2\n")
    with pytest.raises(H.LinksError):
        H.LoadLinks(str(tmp_path / "b.links"))
    out = io.StringIO()
    l.Write(out)
    assert out.getvalue().splitlines()[0].startswith("# c1:0 c2:1")
    assert "c1 c2 5.000000" in out.getvalue()


def _pmx_reference(a, b, p1, p2):
    """scaff/genome.go:37-91 transliterated naively (linear pmxmap scans) as the test oracle."""
    ca, cb = list(a), list(b)
    ca[p1:p2 + 1] = b[p1:p2 + 1]
    cb[p1:p2 + 1] = a[p1:p2 + 1]

    def pmxmap(g, v):
        for i, c in enumerate(g):
            if c == v and (i < p1 or i > p2):
                return i, True
        return 0, False

    amap, bmap = [], []
    for i in range(p1, p2 + 1):
        ma, found = pmxmap(ca, ca[i])
        if found:
            amap.append(ma)
            if bmap:
                i1, i2 = amap.pop(), bmap.pop()
                ca[i1], cb[i2] = cb[i2], ca[i1]
        mb, found = pmxmap(cb, cb[i])
        if found:
            bmap.append(mb)
            if amap:
                i1, i2 = amap.pop(), bmap.pop()
                ca[i1], cb[i2] = cb[i2], ca[i1]
    return ca, cb


def test_pmx_crossover_equals_the_reference_algorithm():
    rng = random.Random(3)
    for n in (2, 3, 8, 31, 200):
        for _ in range(60):
            a, b = list(range(n)), list(range(n))
            rng.shuffle(a); rng.shuffle(b)
            p1, p2 = sorted((rng.randrange(n), rng.randrange(n)))
            ga, gb = H.NewOrderedIntGenome(a, None), H.NewOrderedIntGenome(b, None)
            ca, cb = ga.Crossover(gb, p1, p2)
            ra, rb = _pmx_reference(a, b, p1, p2)
            assert ca.Gene == ra and cb.Gene == rb
            assert ca.Valid() and cb.Valid() and not ca.hasscore and not cb.hasscore
            assert ga.Gene == a and gb.Gene == b                  # parents untouched


def test_memo_contract_and_operators():
    calls = []

    def sfunc(g):
        calls.append(tuple(g.Gene))
        return float(sum(i * v for i, v in enumerate(g.Gene)))

    g = H.NewOrderedIntGenome([3, 1, 0, 2], sfunc)
    assert g.Score() == g.Score() and len(calls) == 1             # genome.go:153-159
    c = g.Copy()
    assert c.hasscore and c.Score() == g.Score() and len(calls) == 1   # Copy carries the memo
    c.Switch(0, 3)
    assert not c.hasscore and c.Gene == [2, 1, 0, 3]
    c.Invert(3, 0)
    assert c.Gene == [3, 0, 1, 2]
    c.Splice(g, 0, 2, 2)
    assert c.Gene == [3, 0, 3, 1] and not c.Valid()
    g.Randomize(random.Random(1))
    assert sorted(g.Gene) == [0, 1, 2, 3] and not g.hasscore
    assert g.Len() == 4 and g.String().startswith("[")

    b = H.NewFixedBitstringGenome([0, 1, 1, 0, 1], lambda x: 1.0)
    assert b.Score() == 1.0 and b.hasscore
    b.Invert(4, 1)
    assert b.Gene == [False, False, False, True, True] and not b.hasscore
    x, y = b.Crossover(H.NewFixedBitstringGenome([1] * 5, None), 1, 2)
    assert x.Gene == [False, True, True, True, True] and y.Gene == [True, False, False, True, True]
    assert b.Valid() and b.String() == "[false false false true true]"


def test_phasing_file_format_round_trip(tmp_path):
    """writePhasing emits `name 1` / `name 0`, space separated (phase.go:151-167), which is what
    readPhasing accepts (phase.go:119-149); a TAB / A|B file is rejected like the reference does."""
    ph = {"chr1_0": True, "chr1_1": False, "chr1_2": True}
    p = tmp_path / "x.phasing.txt"
    H.writePhasing(ph, str(p))
    assert p.read_text() == "chr1_0 1\nchr1_1 0\nchr1_2 1\n"
    assert H.readPhasing(str(p)) == ph
    bad = tmp_path / "bad.txt"
    bad.write_text("chr1_0\tA\n")
    with pytest.raises(ValueError):
        H.readPhasing(str(bad))
