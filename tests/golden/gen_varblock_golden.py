#!/usr/bin/env python
"""Golden vectors for TabulateVariantBlockLinks (/root/reference/src/sequtil/links.go:372-420) and the
two functions that feed it in BlockLinksFromSam (GenomePositions alignment.go:192-220, GetVariantBlocks
alignment.go:265-323) — an independent Python restatement of the Go source text, written without
looking at oracle/lxy_oracle.cpp.  Run here (reads the reference's fixtures), commit the JSON:

    python tests/golden/gen_varblock_golden.py > tests/golden/varblock_fixtures.json

PARITY UNPINNED by the reference's own tests: TestBlockLinksFromSam is commented out
(alignment_test.go:282-299), toy.sam's mates carry different read ids (no group of two ever forms) and
its VCF names blocks "b1".."b5", which strconv.Atoi turns into 0 (alignment.go:289) — on the fixture as
it stands the current code emits nothing, while toy.varblock.links holds the output of an older
implementation.  Case "toy" below therefore uses the fixture's reads and variants with the read-id
suffixes dropped and BLOCK:bN read as N, which is the smallest change that makes the fixture exercise
the function; the other cases are synthetic.
"""
import json
import os
import random
import sys

REF = "/root/reference/src/sequtil/_testdata"


def genome_positions(pos, cigar, seq):            # alignment.go:192-220 (M / S / H only)
    hits, offset, first_h = {}, 0, True
    num = ""
    for ch in cigar:
        if ch.isdigit():
            num += ch
            continue
        v, num = int(num), ""
        if ch == "M":
            for _ in range(v):
                hits[pos + offset] = seq[offset]
                offset += 1
        elif ch == "H":
            if not first_h:
                break
            first_h = False
        elif ch == "S":
            offset += v
        else:
            raise ValueError("Unsupported code in CIGAR.")
    return hits


def get_variant_blocks(chrom, gp, variants):      # alignment.go:265-323
    ret = {}
    if chrom not in variants:
        return {}
    for pos, v in list(gp.items()):
        if pos not in variants[chrom]:
            continue
        ref, alt, block = variants[chrom][pos]
        if block is None:
            continue
        e = ret.setdefault(block, [0, 0, 0])
        if v == ref:
            e[0] += 1
        elif v == alt:
            e[1] += 1
        else:
            e[2] += 1
    return ret


def tabulate(chrom, fgp1, fgp2, table):           # links.go:372-420; table: {(name, name) sorted: value}
    ct = balance = 0
    for k1 in fgp1:
        for k2 in fgp2:
            if k1 > k2:
                if fgp1[k1][2] == 0 and fgp2[k2][2] == 0:
                    one = fgp1[k1][0] > fgp1[k1][1] and fgp2[k2][0] > fgp2[k2][1]
                    two = fgp1[k1][0] < fgp1[k1][1] and fgp2[k2][0] < fgp2[k2][1]
                    key = tuple(sorted((f"{chrom}_{k1}", f"{chrom}_{k2}")))
                    d = 1 if (one or two) else -1
                    table[key] = table.get(key, 0) + d
                    balance += d
                    ct += 1
    return ct, balance


def csr(maps):
    off, rows = [0], []
    for m in maps:
        for k, e in m.items():
            rows.append([k] + list(e))
        off.append(len(rows))
    return off, rows


def case(name, chrom, pairs):
    table, ct, bal = {}, 0, 0
    for m1, m2 in pairs:
        c, b = tabulate(chrom, m1, m2, table)
        ct += c
        bal += b
    off1, blk1 = csr([p[0] for p in pairs])
    off2, blk2 = csr([p[1] for p in pairs])
    return {"name": name, "chrom": chrom, "off1": off1, "blk1": blk1, "off2": off2, "blk2": blk2, "ct": ct,
            "balance": bal, "table": [[a, b, v] for (a, b), v in sorted(table.items())]}


def toy_case():
    variants = {}
    with open(os.path.join(REF, "toy.vcf")) as f:
        for line in f:
            if line.startswith("#") or not line.strip():
                continue
            a = line.rstrip("\n").split("\t")
            info = dict(kv.split(":") for kv in a[7].split(";"))
            blk = info.get("BLOCK")
            variants.setdefault(a[0], {})[int(a[1])] = (a[3], a[4], int(blk[1:]) if blk else None)
    reads = {}
    with open(os.path.join(REF, "toy.sam")) as f:
        for line in f:
            if line.startswith("@") or not line.strip():
                continue
            a = line.rstrip("\n").split("\t")
            reads.setdefault(a[0].rsplit(".", 1)[0], []).append((a[2], int(a[3]), a[5], a[9]))
    by_chrom = {}
    for rid, hits in reads.items():
        if len(hits) == 2 and hits[0][0] == hits[1][0]:
            chrom = hits[0][0]
            m = [get_variant_blocks(chrom, genome_positions(h[1], h[2], h[3]), variants) for h in hits]
            by_chrom.setdefault(chrom, []).append((m[0], m[1]))
            by_chrom[chrom].append((m[1], m[0]))   # and with the mates exchanged: k1 > k2 needs the later block first
    return [case(f"toy_{c}", c, pairs) for c, pairs in sorted(by_chrom.items())]


def random_case(name, seed, n_pairs, n_blocks, neg=False):
    rng = random.Random(seed)
    pairs = []
    for _ in range(n_pairs):
        ms = []
        for _m in range(2):
            ks = rng.sample(range(-3 if neg else 0, n_blocks), rng.randint(0, min(4, n_blocks)))
            ms.append({k: [rng.randint(0, 3), rng.randint(0, 3), 0 if rng.random() < 0.8 else rng.randint(1, 2)] for k in ks})
        pairs.append((ms[0], ms[1]))
    return case(name, "chr7" if not neg else "ctg-1", pairs)


if __name__ == "__main__":
    cases = toy_case()
    cases.append(case("empty_mates", "chr1", [({}, {}), ({3: [1, 0, 0]}, {}), ({}, {1: [0, 2, 0]})]))
    cases.append(case("ties_are_out_of_phase", "chr1", [({5: [1, 1, 0]}, {2: [1, 1, 0]}), ({5: [2, 0, 0]}, {2: [1, 1, 0]}),
                                                         ({5: [0, 0, 0]}, {2: [0, 0, 0]})]))
    cases.append(case("cancels_to_zero", "chr2", [({9: [1, 0, 0]}, {4: [1, 0, 0]}), ({9: [1, 0, 0]}, {4: [0, 1, 0]})]))
    cases.append(random_case("random_small", 1, 40, 12))
    cases.append(random_case("random_dense", 2, 400, 6))
    cases.append(random_case("random_negative_block_numbers", 3, 60, 5, neg=True))
    json.dump({"_what": "TabulateVariantBlockLinks golden vectors (tests/golden/gen_varblock_golden.py)", "cases": cases},
              sys.stdout, indent=1)
