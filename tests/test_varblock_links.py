"""TabulateVariantBlockLinks (/root/reference/src/sequtil/links.go:372-420), the accumulation step of
BlockLinksFromSam (alignment.go:461-555): the oracle's restatement against the golden vectors of the
independent Python restatement (tests/golden/gen_varblock_golden.py), and the library's host loop
(lxy_links_tabulate_varblocks, device < 0) against both.  Tables compare BY NAME: the reference assigns
ids in Go map iteration order, which is random from run to run.

Parity unpinned by the reference's own tests (TestBlockLinksFromSam is commented out; see the header of
the generator for what its fixture does and does not exercise).
"""
import json
import os

import numpy as np
import pytest

import _oracle as O
import lxy_b200._native as N

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "varblock_fixtures.json")) as f:
    CASES = json.load(f)["cases"]


def _arrays(c):
    return (np.array(c["off1"], np.int64), np.array(c["blk1"], np.int32).reshape(-1, 4),
            np.array(c["off2"], np.int64), np.array(c["blk2"], np.int32).reshape(-1, 4))


def _want(c):
    return {(a, b): float(v) for a, b, v in c["table"]}


@pytest.mark.parametrize("c", CASES, ids=[c["name"] for c in CASES])
def test_oracle_matches_the_golden_vectors(c):
    t = O.Links()
    ct, bal = t.tabulate_varblocks(c["chrom"], *_arrays(c))
    assert (ct, bal) == (c["ct"], c["balance"])
    assert t.by_name() == _want(c)


@pytest.mark.parametrize("c", CASES, ids=[c["name"] for c in CASES])
def test_host_loop_matches_the_golden_vectors_and_the_oracle(c):
    nl = N.NativeLinks()
    ct, bal = nl.TabulateVariantBlockLinks(c["chrom"], *_arrays(c))
    assert (ct, bal) == (c["ct"], c["balance"])
    assert nl.by_name() == _want(c)
    assert nl.Size() == len({n for k in _want(c) for n in k})          # only linked blocks receive ids (links.go:402,409)


def test_the_toy_fixture_case_reads_as_the_reference_documents_it():
    """toy.sam / toy.vcf (read ids without the .1/.2 suffix, BLOCK:bN read as N): pair 1 reads block 1
    ref and block 2 alt (out of phase, -1), pair 3's second mate shows neither allele at position 24
    (ambiguous: nothing), pairs 2 and 4 read blocks 3 and 4 alt/alt and ref... in phase (+2)."""
    c = next(c for c in CASES if c["name"] == "toy_chr1")
    assert _want(c) == {("chr1_1", "chr1_2"): -1.0, ("chr1_3", "chr1_4"): 2.0}


def test_accumulates_over_calls_and_chromosomes():
    """BlockLinksFromSam adds read pair after read pair into one table (alignment.go:524-535)."""
    nl, t = N.NativeLinks(), O.Links()
    for c in CASES:
        arr = _arrays(c)
        half = (len(arr[0]) - 1) // 2
        for lo, hi in ((0, half), (half, len(arr[0]) - 1)):                 # two batches per case
            o1, o2 = arr[0][lo:hi + 1], arr[2][lo:hi + 1]
            b1, b2 = arr[1][o1[0]:o1[-1]], arr[3][o2[0]:o2[-1]]
            a = nl.TabulateVariantBlockLinks(c["chrom"], o1 - o1[0], b1, o2 - o2[0], b2)
            b = t.tabulate_varblocks(c["chrom"], o1 - o1[0], b1, o2 - o2[0], b2)
            assert a == b
    assert nl.by_name() == t.by_name()
    # the table is an ordinary Links object: Write / LoadLinks round trip, phase store
    assert nl.Size() == t.size()


def test_bad_arguments():
    nl = N.NativeLinks()
    with pytest.raises(N.LxyError):
        nl.TabulateVariantBlockLinks("c", [0, 2, 1], np.zeros((2, 4), np.int32), [0, 0, 0], np.zeros((0, 4), np.int32))
    assert nl.TabulateVariantBlockLinks("c", [0], np.zeros((0, 4), np.int32), [0], np.zeros((0, 4), np.int32)) == (0, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES, ids=[c["name"] for c in CASES])
def test_device_tabulation_equals_the_host_loop(c):
    host, dev = N.NativeLinks(), N.NativeLinks()
    assert host.TabulateVariantBlockLinks(c["chrom"], *_arrays(c)) == dev.TabulateVariantBlockLinks(c["chrom"], *_arrays(c), device=0)
    assert dev.by_name() == _want(c)
    assert dev.Decode(range(dev.Size())) == host.Decode(range(host.Size()))     # same ids, same record order
    assert [x.tolist() for x in dev.records()] == [x.tolist() for x in host.records()]


@pytest.mark.gpu
def test_device_tabulation_at_scale_feeds_the_phasing_scorer():
    """300,000 read pairs over 20,000 blocks: device == host == oracle (by name), and the table goes
    straight into the phasing store (the path BlockLinksFromSam -> LoadLinks -> phase.score)."""
    rng = np.random.default_rng(17)
    n_pairs, n_blocks = 300_000, 20_000
    c1, c2 = rng.integers(0, 4, n_pairs), rng.integers(0, 4, n_pairs)
    off1, off2 = np.concatenate([[0], np.cumsum(c1)]), np.concatenate([[0], np.cumsum(c2)])

    def entries(counts, shift):
        rows = np.zeros((int(counts.sum()), 4), np.int32)
        base = np.repeat(rng.integers(0, n_blocks - 40, len(counts)), counts)
        rows[:, 0] = base + shift + 7 * (np.arange(len(base)) - np.repeat(np.concatenate([[0], np.cumsum(counts)])[:-1], counts))
        rows[:, 1] = rng.integers(0, 3, len(base))
        rows[:, 2] = rng.integers(0, 3, len(base))
        rows[:, 3] = rng.random(len(base)) < 0.1
        return rows
    blk1, blk2 = entries(c1, 9), entries(c2, 0)
    host, dev, ora = N.NativeLinks(), N.NativeLinks(), O.Links()
    r = host.TabulateVariantBlockLinks("chr1", off1, blk1, off2, blk2)
    assert r == dev.TabulateVariantBlockLinks("chr1", off1, blk1, off2, blk2, device=0)
    assert r == ora.tabulate_varblocks("chr1", off1, blk1, off2, blk2) and r[0] > 100_000
    assert [x.tolist() for x in dev.records()] == [x.tolist() for x in host.records()]
    assert dev.by_name() == ora.by_name()
    with N.Context(0) as ctx:
        dev.attach(ctx, phase=True)
        genes = (rng.random((8, dev.Size())) < 0.5).astype(np.uint8)
        got, _ = ctx.phase_score_batch(genes)
        a, b, w = dev.records()
        _, want = O.phase_score_links_batch(dev.Size(), a, b, w, genes, nthreads=2)
        assert got.tolist() == want.tolist()
