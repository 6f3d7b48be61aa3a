"""ScaffoldLinksFromSam on the device (lxy_b200/csrc/links_parse.cu, SAM mode: lxy_links_from_sam_gpu /
lxy_scaffold_links_from_sam / lxy_ctx_load_sam) against the library's host loop (lxy_links_from_sam,
checked against the oracle and the golden in tests/test_sam_links.py) and against the oracle's
restatement of /root/reference/src/sequtil/alignment.go:624-676: names in id order, records in order
of first appearance, every count exact, the same error returns."""
import json
import os

import numpy as np
import pytest

import _oracle as O
import _sam_cases as S
import lxy_b200._native as N
import lxy_b200.host as H

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "sam_fixtures.json")))


def _same_table(g, h):
    assert g.Size() == h.Size()
    assert g.Decode(range(g.Size())) == h.Decode(range(h.Size()))
    ga, gb, gw = g.records()
    ha, hb, hw = h.records()
    assert ga.tolist() == ha.tolist() and gb.tolist() == hb.tolist()
    assert gw.tolist() == hw.tolist()


def _against_oracle(g, path):
    o = O.Links.from_sam(path)
    assert g.Decode(range(g.Size())) == [o.name(i) for i in range(o.size())]
    a, b, w = g.records()
    oa, ob, ow = o.export()
    assert sorted(zip(a.tolist(), b.tolist(), w.tolist())) == sorted(zip(oa.tolist(), ob.tolist(), ow.tolist()))


@pytest.mark.parametrize("key", ["toy_sam", "toy_sam_paired"])
def test_reference_fixture(key, tmp_path):
    g0 = GOLD[key]
    p = tmp_path / "x.sam"
    p.write_bytes(g0["text"].encode())
    g = N.NativeLinks.from_sam(str(p), device=0)
    assert g.Decode(range(g.Size())) == g0["names"]
    a, b, w = g.records()
    assert sorted(zip(a.tolist(), b.tolist(), w.tolist())) == [tuple(r) for r in g0["records"]]
    _same_table(g, N.NativeLinks.from_sam(str(p)))


@pytest.mark.parametrize("name,text,status", S.edge_cases(), ids=[c[0] for c in S.edge_cases()])
def test_edge_cases_match_the_host_loop(name, text, status, tmp_path):
    p = tmp_path / (name + ".sam")
    p.write_bytes(text.encode())
    if status == "arg":
        with pytest.raises(N.LxyError) as eh:
            N.NativeLinks.from_sam(str(p))
        with pytest.raises(N.LxyError) as eg:
            N.NativeLinks.from_sam(str(p), device=0)
        assert eg.value.status == eh.value.status == -1
        assert str(eg.value) == str(eh.value)                  # same line number, same wording
    else:
        g = N.NativeLinks.from_sam(str(p), device=0)
        _same_table(g, N.NativeLinks.from_sam(str(p)))
        _against_oracle(g, str(p))


def test_empty_and_missing_files(tmp_path):
    p = tmp_path / "empty.sam"
    p.write_bytes(b"")
    g = N.NativeLinks.from_sam(str(p), device=0)
    assert g.Size() == 0 and g.npairs() == 0
    with pytest.raises(N.LxyError):
        N.NativeLinks.from_sam(str(tmp_path / "absent.sam"), device=0)


@pytest.mark.parametrize("kw", [dict(), dict(crlf=True), dict(interleave_headers=True, final_newline=False),
                                dict(header=False, long_names=True)])
def test_random_texts(kw, tmp_path):
    p = str(tmp_path / "r.sam")
    with open(p, "wb") as f:
        f.write(S.random_sam(21, 20_000, 300, **kw).encode())
    g = N.NativeLinks.from_sam(p, device=0)
    _same_table(g, N.NativeLinks.from_sam(p))
    _against_oracle(g, p)
    st = g.load_stats
    assert st["bytes"] == os.path.getsize(p) and st["names"] == g.Size() and st["pairs"] == g.npairs()
    a, b, w = g.records()
    assert st["records"] == int(w.sum()) and st["attempts"] == 1


def test_many_contigs_grow_the_name_table(tmp_path):
    p = str(tmp_path / "many.sam")
    with open(p, "w") as f:                                     # 100,000 distinct contig names > 2^15
        for k in range(50_001):
            f.write(f"q{k}\t0\tL{k}\t1\nq{k}\t16\tR{k}\t9\n")
    g = N.NativeLinks.from_sam(p, device=0)
    _same_table(g, N.NativeLinks.from_sam(p))
    assert g.Size() == 100_000 and g.npairs() == 50_000         # the last pair is never added


def test_hash_collisions_are_detected_and_reseeded(tmp_path, monkeypatch):
    p = str(tmp_path / "c.sam")
    with open(p, "wb") as f:
        f.write(S.random_sam(22, 4_000, 500).encode())
    monkeypatch.setenv("LXY_PARSE_HASH_BITS", "5")              # 32 hash values for 500 names
    g = N.NativeLinks.from_sam(p, device=0)
    monkeypatch.delenv("LXY_PARSE_HASH_BITS")
    assert g.load_stats["attempts"] == 2
    _same_table(g, N.NativeLinks.from_sam(p))


def test_scaffold_links_from_sam_file_to_file(tmp_path):
    """ScaffoldLinksFromSam(samPath, outPath) then LoadLinks(outPath): the reference's two steps."""
    p = str(tmp_path / "r.sam")
    with open(p, "wb") as f:
        f.write(S.random_sam(23, 5_000, 60).encode())
    out = str(tmp_path / "r.ctg.links")
    st = H.ScaffoldLinksFromSam(p, out)
    assert st["pairs"] > 100
    back = N.NativeLinks(out, device=0)
    _same_table(back, N.NativeLinks.from_sam(p))
    o = O.Links.load(out)                                       # the oracle's LoadLinks reads it too
    oa, ob, ow = o.export()
    a, b, w = back.records()
    assert sorted(zip(a.tolist(), b.tolist(), w.tolist())) == sorted(zip(oa.tolist(), ob.tolist(), ow.tolist()))
    with pytest.raises(H.LinksError):
        H.ScaffoldLinksFromSam(str(tmp_path / "absent.sam"), out)


def test_context_store_straight_from_sam(tmp_path):
    """SAM -> link store without the .links file: scores equal those of the host table attached."""
    p = str(tmp_path / "s.sam")
    with open(p, "wb") as f:
        f.write(S.random_sam(24, 30_000, 200).encode())
    rng = np.random.default_rng(6)
    with N.Context(0) as cg, N.Context(0) as ch:
        g = N.NativeLinks.from_sam(p, ctx=cg)
        h = N.NativeLinks.from_sam(p)
        h.attach(ch)
        _same_table(g, h)
        perms = np.stack([rng.permutation(g.Size()) for _ in range(33)]).astype(np.int32)
        sg, sh = cg.scaff_score_batch(perms), ch.scaff_score_batch(perms)
        assert sg.tolist() == sh.tolist()
        o = O.Links.from_sam(p)
        assert sg[0] == pytest.approx(o.scaff_score(perms[0].tolist()), rel=1e-9)
    q = tmp_path / "none.sam"
    q.write_bytes(b"@HD\tVN:1.0\n")
    with N.Context(0) as c, pytest.raises(N.LxyError):
        N.NativeLinks.from_sam(str(q), ctx=c)


def test_text_staged_through_the_pinned_ring(tmp_path, monkeypatch):
    """Large texts reach the device through a page-locked ring filled by four threads (links_parse.cu,
    h2d_through_ring); the threshold is lowered so that a 21 MB text (5 pieces + a partial one) takes it."""
    p = str(tmp_path / "big.sam")
    with open(p, "wb") as f:
        f.write(S.random_sam(31, 230_000, 700).encode())
    assert os.path.getsize(p) > 5 * (4 << 20)
    monkeypatch.setenv("LXY_PARSE_RING_MIN_BYTES", "1")
    g = N.NativeLinks.from_sam(p, device=0)
    monkeypatch.setenv("LXY_PARSE_NO_RING", "1")
    g2 = N.NativeLinks.from_sam(p, device=0)
    monkeypatch.delenv("LXY_PARSE_NO_RING")
    _same_table(g, g2)
    _same_table(g, N.NativeLinks.from_sam(p))
    out = str(tmp_path / "big.links")                           # the .links loader shares the staging
    g.Write(out)
    with open(out, "ab") as f:
        f.write((b"# padding\n" * 10 + b"x y 1.5\n") * 200_000)
    assert os.path.getsize(out) > (4 << 20)
    a = N.NativeLinks(out, device=0)
    monkeypatch.delenv("LXY_PARSE_RING_MIN_BYTES")
    b = N.NativeLinks(out)
    assert a.Decode(range(a.Size())) == b.Decode(range(b.Size()))
    assert [x.tolist() for x in a.records()] == [x.tolist() for x in b.records()]
