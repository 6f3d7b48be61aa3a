"""The library's own link table (lxy_b200/csrc/links_io.cu, `lxy_links_*`) against the reference's
tests for sequtil.Links (/root/reference/src/sequtil/links_test.go, cited per test) and against the
oracle's loader on the reference's fixture files and on generated files with the loader's quirks
(comments, duplicate rows, reversed pairs, malformed numbers, extra fields).  Host-only: runs
without a GPU; the device side of lxy_ctx_set_links_from is in tests/test_gpu_host.py."""
import numpy as np
import pytest

import _oracle as O
import lxy_b200._native as N


def _table(l):
    a, b, w = l.records()
    return sorted(zip(a.tolist(), b.tolist(), w.tolist()))


def test_add_key_and_ids():  # links_test.go:10-24, 26-39
    l = N.NativeLinks()
    assert l.ID("hi") == 0 and l.ID("hello") == 1 and l.ID("hi") == 0
    assert l.Decode([0, 1]) == ["hi", "hello"]          # links_test.go:146-159
    assert l.Size() == 2                                 # links_test.go:130-139
    assert l.ID("nobody", create=False) == -1


def test_load_links_toy(golden_files):  # links_test.go:41-56
    l = N.NativeLinks(golden_files["toy_ctg"])
    assert l.Size() == 3 and l.Decode([0, 1, 2]) == ["chr1", "chr2", "chr3"]
    assert _table(l) == [(0, 1, 2.0), (1, 2, 1.0)]


def test_set_add_get():  # links_test.go:72-82, 84-100, 103-117
    l = N.NativeLinks()
    l.ID("hi"), l.ID("hello")
    l.Set(0, 1, 0.1)
    assert l.Get(0, 1) == (0.1, None) and l.Get(1, 0)[0] == 0.1
    l.Add(1, 0, 0.1)
    assert l.Get(0, 1)[0] == 0.2
    assert l.Get(0, 0) == (0.0, None)                    # absent pair -> 0 (links.go:211-215)
    v, err = l.Get(0, 7)                                 # unknown id -> -1 + error (links.go:193-200)
    assert v == -1.0 and err is not None
    with pytest.raises(N.LxyError):
        l.Set(0, 7, 1.0)


@pytest.mark.parametrize("key", ["test_ctg", "toy_ctg", "test_var", "toy_varblock"])
def test_load_matches_oracle_on_reference_fixtures(golden_files, key):
    l = N.NativeLinks(golden_files[key])
    o = O.Links.load(golden_files[key])
    assert l.Size() == o.size()
    assert l.Decode(range(l.Size())) == [o.name(i) for i in range(o.size())]
    oa, ob, ow = o.export()
    assert _table(l) == sorted(zip(oa.tolist(), ob.tolist(), ow.tolist()))


def test_loader_quirks(tmp_path):
    p = tmp_path / "q.links"
    p.write_text("# c:0 a:1 header written by Write is ignored\n"
                 "c a 5\n"
                 "a b 2.5\n"
                 "b a 7\n"            # same pair reversed: the last row wins (Set, links.go:477)
                 "a c oops\n"         # ParseFloat error ignored -> 0 (links.go:473), overwrites 5
                 "d  x 3\n"           # two spaces: the second name is the empty string
                 "e f 1e2 trailing fields are ignored\n"
                 "#x y 1\n")
    l = N.NativeLinks(str(p))
    o = O.Links.load(str(p))
    names = l.Decode(range(l.Size()))
    assert names == [o.name(i) for i in range(o.size())] == ["c", "a", "b", "d", "", "e", "f"]
    oa, ob, ow = o.export()
    assert _table(l) == sorted(zip(oa.tolist(), ob.tolist(), ow.tolist()))
    assert l.Get(l.ID("a"), l.ID("b"))[0] == 7.0 and l.Get(0, 1)[0] == 0.0
    assert l.Get(l.ID("e"), l.ID("f"))[0] == 100.0


def test_loader_errors(tmp_path):
    with pytest.raises(N.LxyError):
        N.NativeLinks(str(tmp_path / "missing.links"))
    p = tmp_path / "empty.links"
    p.write_text("# only a comment\n")
    with pytest.raises(N.LxyError) as e:                 # links.go:485-487
        N.NativeLinks(str(p))
    assert e.value.status == -3
    p.write_text("a b 1\n\nc d 2\n")                      # empty line: the reference panics (line[0])
    with pytest.raises(N.LxyError):
        N.NativeLinks(str(p))
    p.write_text("a b\n")                                 # arr[2] out of range in the reference
    with pytest.raises(N.LxyError):
        N.NativeLinks(str(p))


def test_write_round_trip(tmp_path):  # Write (links.go:242-264) then LoadLinks
    rng = np.random.default_rng(1)
    l = N.NativeLinks()
    for i in range(30):
        l.ID(f"ctg{i}")
    for _ in range(200):
        a, b = rng.integers(0, 30, 2)
        l.Add(int(a), int(b), int(rng.integers(1, 9)))
    p = tmp_path / "w.links"
    l.Write(str(p))
    text = p.read_text().splitlines()
    assert text[0].startswith("# ctg0:0 ctg1:1")
    assert all(len(t.split(" ")) == 3 and "." in t.split(" ")[2] for t in text[1:])   # %f
    back = N.NativeLinks(str(p))
    # ids are re-assigned by first appearance in the rows; compare by name
    def named(x):
        a, b, w = x.records()
        n = x.Decode(range(x.Size()))
        return sorted((tuple(sorted((n[i], n[j]))), v) for i, j, v in zip(a.tolist(), b.tolist(), w.tolist()))
    assert named(back) == named(l)
