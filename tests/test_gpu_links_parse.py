"""LoadLinks on the device (lxy_b200/csrc/links_parse.cu: lxy_links_load_gpu / lxy_ctx_load_links)
against the host loader (lxy_links_load, itself checked against the reference's links_test.go in
tests/test_native_links.py) and against the oracle's restatement of LoadLinks
(/root/reference/src/sequtil/links.go:427-491): names in id order, records in order of first
appearance, every value bit for bit, the same error returns."""
import os

import numpy as np
import pytest

import _oracle as O
import lxy_b200._native as N

pytestmark = pytest.mark.gpu


def _same_table(g, h):
    assert g.Size() == h.Size()
    assert g.Decode(range(g.Size())) == h.Decode(range(h.Size()))
    ga, gb, gw = g.records()
    ha, hb, hw = h.records()
    assert ga.tolist() == ha.tolist() and gb.tolist() == hb.tolist()
    assert gw.view(np.uint64).tolist() == hw.view(np.uint64).tolist()      # bit for bit, NaN included


def _sorted_table(l):
    a, b, w = l.records()
    return sorted(zip(a.tolist(), b.tolist(), w.tolist()))


@pytest.mark.parametrize("key", ["test_ctg", "toy_ctg", "test_var", "toy_varblock"])
def test_reference_fixtures(golden_files, key):
    g = N.NativeLinks(golden_files[key], device=0)
    _same_table(g, N.NativeLinks(golden_files[key]))
    o = O.Links.load(golden_files[key])
    assert g.Decode(range(g.Size())) == [o.name(i) for i in range(o.size())]
    oa, ob, ow = o.export()
    assert _sorted_table(g) == sorted(zip(oa.tolist(), ob.tolist(), ow.tolist()))
    assert g.load_stats["slow_values"] == 0 and g.load_stats["attempts"] == 1


def test_toy_table_of_the_reference_test(golden_files):  # links_test.go:41-56
    g = N.NativeLinks(golden_files["toy_ctg"], device=0)
    assert g.Size() == 3 and g.Decode([0, 1, 2]) == ["chr1", "chr2", "chr3"]
    assert _sorted_table(g) == [(0, 1, 2.0), (1, 2, 1.0)]
    assert g.Get(0, 1) == (2.0, None) and g.Get(2, 1)[0] == 1.0 and g.Get(0, 2)[0] == 0.0
    g.Add(0, 1, 0.5)                                       # the lazily built pair index takes updates
    assert g.Get(1, 0)[0] == 2.5


QUIRKS = ("# c:0 a:1 header written by Write is ignored\n"
          "c a 5\n"
          "a b 2.5\n"
          "b a 7\n"            # same pair reversed: the last row wins (Set, links.go:477)
          "a c oops\n"         # ParseFloat error ignored -> 0 (links.go:473), overwrites 5
          "d  x 3\n"           # two spaces: the second name is the empty string
          "e f 1e2 trailing fields are ignored\n"
          "#x y 1\n"
          "g g 4\n"            # a self pair can be stored (never queried by the scorers)
          "h i \n"             # empty value token: syntax error -> 0
          "c a 6\r\n"          # one trailing CR is dropped by the line scanner
          "j k -0\n")


@pytest.mark.parametrize("tail", ["", "l m 9", "l m 9\r"])     # with and without a final newline
def test_loader_quirks(tmp_path, tail):
    p = tmp_path / "q.links"
    p.write_bytes((QUIRKS + tail).encode())
    g = N.NativeLinks(str(p), device=0)
    h = N.NativeLinks(str(p))
    _same_table(g, h)
    o = O.Links.load(str(p))
    assert g.Decode(range(g.Size())) == [o.name(i) for i in range(o.size())]
    assert g.Get(g.ID("a"), g.ID("b"))[0] == 7.0 and g.Get(g.ID("c"), g.ID("a"))[0] == 6.0
    assert g.Get(g.ID("e"), g.ID("f"))[0] == 100.0
    assert np.signbit(g.Get(g.ID("j"), g.ID("k"))[0])


VALUES = ["0", "1", "12.000000", "0.100000", "3.14159", ".5", "5.", "+3.25", "-17.125", "1.5E-3", "2e22",
          "1e-22", "1234567890123.45", "9007199254740991", "0.000001", "00012.50", "1e0", "7e+2",
          # outside the exactly rounded fast path: resolved by the host loader's conversion
          "9007199254740993", "1e23", "1e-23", "123456789012345678901234567890", "0.30000000000000004441",
          "2.2250738585072014e-308", "1e400", "-1e400", "inf", "-Infinity", "nan", "1e", "e5", ".", "-",
          "1..2", "1.2.3", "0x10", "1_000", "12abc", "1e5x", "--1", "4.9e-324"]


def test_value_tokens_bit_for_bit(tmp_path):
    p = tmp_path / "v.links"
    p.write_text("".join(f"n{i} m{i} {v}\n" for i, v in enumerate(VALUES)))
    g = N.NativeLinks(str(p), device=0)
    _same_table(g, N.NativeLinks(str(p)))
    _, _, w = g.records()
    assert w[:18].tolist() == [float(v) for v in VALUES[:18]]           # correctly rounded, as Python parses
    assert g.load_stats["slow_values"] == len(VALUES) - 18


def test_errors_match_the_host_loader(tmp_path):
    def both(text):
        p = tmp_path / "e.links"
        p.write_bytes(text)
        with pytest.raises(N.LxyError) as eh:
            N.NativeLinks(str(p))
        with pytest.raises(N.LxyError) as eg:
            N.NativeLinks(str(p), device=0)
        assert eg.value.status == eh.value.status
        assert str(eg.value) == str(eh.value)
        return eg.value
    with pytest.raises(N.LxyError):
        N.NativeLinks(str(tmp_path / "missing.links"), device=0)
    assert both(b"# only a comment\n").status == -3                      # links.go:485-487
    assert both(b"").status == -3
    assert ":2:" in str(both(b"a b 1\n\nc d 2\n"))                      # empty line: line[0] panics
    assert both(b"a b 1\r\n\r\n").status == -1
    assert ":1:" in str(both(b"a b\n"))                                 # arr[2] out of range
    e = both(b"a b 1\nc d\n\ne f\n")                                    # the first bad line decides
    assert ":2:" in str(e) and "three fields" in str(e)


def _synthetic(path, n_names, n_lines, seed, comments=True):
    rng = np.random.default_rng(seed)
    names = [f"ctg{i:06d}_{'x' * int(rng.integers(0, 12))}" for i in range(n_names)]
    a = rng.integers(0, n_names, n_lines)
    b = rng.integers(0, min(n_names, 64), n_lines)                       # many repeated pairs
    v = rng.integers(0, 5000, n_lines) / 8.0
    with open(path, "w") as f:
        f.write("# " + " ".join(f"{nm}:{i}" for i, nm in enumerate(names[:50])) + "\n")
        for k in range(n_lines):
            if comments and k % 977 == 5:
                f.write("# a comment in the middle\n")
            f.write(f"{names[a[k]]} {names[b[k]]} {v[k]:f}\n")


def test_synthetic_file_against_the_host_loader(tmp_path):
    p = str(tmp_path / "s.links")
    _synthetic(p, 3000, 200_000, 11)
    g = N.NativeLinks(p, device=0)
    _same_table(g, N.NativeLinks(p))
    st = g.load_stats
    assert st["bytes"] == os.path.getsize(p) and st["records"] == 200_000
    assert st["names"] == g.Size() and st["pairs"] == g.npairs()
    assert st["slow_values"] == 0 and st["attempts"] == 1


def test_name_table_growth(tmp_path):
    p = str(tmp_path / "many.links")
    with open(p, "w") as f:                                              # 120,000 distinct names > 2^15
        for k in range(60_000):
            f.write(f"L{k} R{k} {k % 7}\n")
    g = N.NativeLinks(p, device=0)
    _same_table(g, N.NativeLinks(p))
    assert g.Size() == 120_000 and g.npairs() == 60_000


def test_hash_collisions_are_detected_and_reseeded(tmp_path, monkeypatch):
    p = str(tmp_path / "c.links")
    _synthetic(p, 500, 5_000, 12, comments=False)
    monkeypatch.setenv("LXY_PARSE_HASH_BITS", "5")                       # 32 hash values for 500 names
    g = N.NativeLinks(p, device=0)
    monkeypatch.delenv("LXY_PARSE_HASH_BITS")
    assert g.load_stats["attempts"] == 2
    _same_table(g, N.NativeLinks(p))


def test_context_store_from_the_device_loader(tmp_path, golden_files):
    p = str(tmp_path / "s.links")
    _synthetic(p, 400, 50_000, 13)
    rng = np.random.default_rng(5)
    with N.Context(0) as cg, N.Context(0) as ch:
        g = N.NativeLinks(p, ctx=cg)                                     # records never leave the device
        h = N.NativeLinks(p)
        h.attach(ch)
        _same_table(g, h)
        perms = np.stack([rng.permutation(g.Size()) for _ in range(33)]).astype(np.int32)
        sg, sh = cg.scaff_score_batch(perms), ch.scaff_score_batch(perms)
        np.testing.assert_allclose(sg, sh, rtol=1e-13)
        o = O.Links.load(p)
        assert sg[0] == pytest.approx(o.scaff_score(perms[0].tolist()), rel=1e-9)
    with N.Context(0) as cg, N.Context(0) as ch:                         # phasing store
        g = N.NativeLinks(golden_files["test_var"], ctx=cg, phase=True)
        h = N.NativeLinks(golden_files["test_var"])
        h.attach(ch, phase=True)
        genes = rng.integers(0, 2, (17, g.Size())).astype(np.uint8)
        assert cg.phase_score_batch(genes)[0].tolist() == ch.phase_score_batch(genes)[0].tolist()
        zero = np.zeros((1, g.Size()), np.uint8)
        assert cg.phase_score_batch(zero)[0].tolist() == [-265]
