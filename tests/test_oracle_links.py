"""Pins the oracle's link table against the reference's own tests.

Each test restates one test of /root/reference/src/sequtil/links_test.go (cited per test).
"""
import pytest

import _oracle as O


def test_add_key():  # links_test.go:10-24
    l = O.Links()
    l.add_key("hi")
    assert l.has_name("hi")
    assert l.name(l.id("hi")) == "hi"


def test_int_ids_are_dense_first_appearance():  # links_test.go:26-39
    l = O.Links()
    l.add_key("hi")
    l.add_key("hello")
    assert [l.id("hi"), l.id("hello")] == [0, 1]


def test_load_links_toy(golden_files):  # links_test.go:41-56
    loaded = O.Links.load(golden_files["toy_ctg"])
    key = O.Links()
    key.set(key.id("chr1"), key.id("chr2"), 2)
    key.set(key.id("chr2"), key.id("chr3"), 1)
    assert loaded.size() == key.size() == 3
    assert [loaded.name(i) for i in range(3)] == [key.name(i) for i in range(3)]
    la, lb, lw = loaded.export()
    ka, kb, kw = key.export()
    assert la.tolist() == ka.tolist() and lb.tolist() == kb.tolist() and lw.tolist() == kw.tolist()
    assert list(zip(la.tolist(), lb.tolist(), lw.tolist())) == [(0, 1, 2.0), (1, 2, 1.0)]


def test_set_returns_nil():  # links_test.go:72-82
    l = O.Links()
    l.add_key("hi")
    l.add_key("hello")
    assert l.set(0, 1, 0.01) == 0


def test_set_then_add():  # links_test.go:84-100
    l = O.Links()
    l.add_key("a")
    l.add_key("b")
    l.set(l.id("a"), l.id("b"), 0.1)
    l.add(l.id("a"), l.id("b"), 0.1)
    val, rc = l.get(l.id("a"), l.id("b"))
    assert rc == 0 and val == 0.2


def test_get():  # links_test.go:103-117
    l = O.Links()
    l.add_key("a")
    l.add_key("b")
    l.set(l.id("a"), l.id("b"), 0.1)
    assert l.get(l.id("a"), l.id("b")) == (0.1, 0)


def test_size():  # links_test.go:130-139
    l = O.Links()
    l.add_key("hi")
    l.add_key("hello")
    assert l.size() == 2


def test_decode():  # links_test.go:146-159
    l = O.Links()
    l.add_key("a")
    l.add_key("b")
    assert l.decode([0, 1]) == ["a", "b"]
    with pytest.raises(KeyError):
        l.decode([0, 7])


# ---- behaviours of links.go the reference's tests do not cover but score() depends on --------
def test_get_semantics():
    l = O.Links()
    l.add_key("a"); l.add_key("b"); l.add_key("c")
    l.set(1, 0, 3.0)                       # stored under (0,1): links.go:130-132
    assert l.get(0, 1) == (3.0, 0) and l.get(1, 0) == (3.0, 0)
    assert l.get(0, 2) == (0.0, 0)         # known ids, absent pair: links.go:211-215
    assert l.get(0, 9) == (-1.0, -1)       # unknown id: links.go:193-200
    assert l.set(0, 9, 1.0) == -1 and l.add(9, 0, 1.0) == -1


def test_load_last_duplicate_wins_and_comments(tmp_path):  # links.go:450,477
    p = tmp_path / "d.links"
    p.write_text("# x:0 y:1\nx y 1\ny x 5\nx z notanumber\n")
    l = O.Links.load(str(p))
    assert l.size() == 3
    assert l.get(0, 1)[0] == 5.0           # (y,x) == (x,y), last row wins
    assert l.get(0, 2)[0] == 0.0           # ParseFloat error ignored -> 0 (links.go:473)


def test_load_missing_and_empty(tmp_path):  # links.go:435-438,485
    with pytest.raises(OSError):
        O.Links.load(str(tmp_path / "nope"))
    p = tmp_path / "e.links"
    p.write_text("# only a header\n")
    with pytest.raises(ValueError):
        O.Links.load(str(p))
