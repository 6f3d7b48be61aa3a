#!/usr/bin/env python
"""Generate tests/golden/fixtures.json from the reference's own fixture files.

Run in the build container (needs /root/reference):  python tests/golden/gen_golden.py

This is an INDEPENDENT second restatement (pure Python dict-of-dicts, plain float arithmetic =
IEEE double, no FMA) of
  * LoadLinks / Set / Get      /root/reference/src/sequtil/links.go:113-145,190-221,427-491
  * scaff.score                /root/reference/src/lxy/scaff/scaff.go:246-310
  * phase.score                /root/reference/src/lxy/phase/phase.go:171-207
written separately from oracle/lxy_oracle.cpp so that a transcription slip in one is caught by
the other.  The reference asserts no score values in its tests, so these goldens
are "what the reference source yields on the reference's fixtures", not reference-asserted.
The link-table goldens (toy.ctg.links -> chr1-chr2=2, chr2-chr3=1) ARE reference-asserted
(links_test.go:41-56).
"""
import itertools
import json
import os
import random
import struct

REF = "/root/reference"
FIXTURES = {
    "test_ctg": "src/lxy/scaff/_testdata/test.ctg.links",
    "toy_ctg": "src/sequtil/_testdata/toy.ctg.links",
    "test_var": "src/lxy/phase/_testdata/test.var.links",
    "toy_varblock": "src/sequtil/_testdata/toy.varblock.links",
}


class PyLinks:
    def __init__(self):
        self.id_key, self.id_key_rev, self.maxid, self.data = {}, {}, 0, {}

    def add_key(self, k):
        self.id_key[k] = self.maxid
        self.id_key_rev[self.maxid] = k
        self.maxid += 1

    def set(self, a, b, v):
        if a not in self.id_key_rev or b not in self.id_key_rev:
            return False
        if a > b:
            a, b = b, a
        self.data.setdefault(a, {})[b] = v
        return True

    def get(self, a, b):
        if a not in self.id_key_rev or b not in self.id_key_rev:
            return -1.0
        if a > b:
            a, b = b, a
        return self.data.get(a, {}).get(b, 0.0)

    def size(self):
        return len(self.id_key)


def go_parse_float(tok):
    try:
        return float(tok)
    except ValueError:
        return 0.0


def load_links(path):
    l = PyLinks()
    with open(path) as f:
        for line in f.read().split("\n"):
            if line == "":
                continue  # only the final newline produces this in the fixtures
            if line[0] == "#":
                continue
            arr = line.split(" ")
            for name in arr[:2]:
                if name not in l.id_key:
                    l.add_key(name)
            l.set(l.id_key[arr[0]], l.id_key[arr[1]], go_parse_float(arr[2]))
    return l


def scaff_score(l, gene):
    total = 0.0
    size = l.size()
    for i, c in enumerate(gene):
        for k, w in ((1, None), (2, 0.5), (3, 0.33), (5, 0.2), (11, 0.1), (20, 0.05)):
            if i + k < size:
                v = l.get(c, gene[i + k])
                total += v if w is None else w * v
            if i - k > 0:
                v = l.get(c, gene[i - k])
                total += v if w is None else w * v
    return -total


PHASE_STEPS = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 15, 20, 25, 30]


def phase_score(l, gene):
    total = 0.0
    size = l.size()
    for i, c in enumerate(gene):
        for v in PHASE_STEPS:
            if i + v < size:
                val = l.get(i, i + v)
                total = total - val if c != gene[i + v] else total + val
            if i - v > 0:
                val = l.get(i, i - v)
                total = total - val if c != gene[i - v] else total + val
    return -total


def hexf(x):
    """Bit-exact float64 as hex so JSON round-trips cannot blur it."""
    return struct.pack(">d", x).hex()


def table_records(l):
    return sorted([a, b, v] for a, inner in l.data.items() for b, v in inner.items())


def main():
    out = {"_generated_by": "tests/golden/gen_golden.py", "_reference_commit": "1406eafa"}
    rng = random.Random(20261104)
    for key, rel in FIXTURES.items():
        path = os.path.join(REF, rel)
        l = load_links(path)
        n = l.size()
        entry = {
            "source": rel,
            "text": open(path).read(),
            "names": [l.id_key_rev[i] for i in range(n)],
            "records": table_records(l),
            "cases": [],
        }
        if key in ("test_ctg", "toy_ctg"):
            genes = []
            if n <= 6:
                genes = [list(p) for p in itertools.permutations(range(n))]
            else:
                genes = [list(range(n)), list(range(n - 1, -1, -1)),
                         [3, 7, 0, 10, 5, 1, 8, 2, 9, 4, 6]]
                for _ in range(64):
                    g = list(range(n))
                    rng.shuffle(g)
                    genes.append(g)
            for g in genes:
                s = scaff_score(l, g)
                entry["cases"].append({"gene": g, "score": s, "score_hex": hexf(s)})
            entry["kind"] = "scaff"
        else:
            genes = []
            if key == "test_var":
                keyline = open(path).read().split("\n")[1].lstrip("#").split()
                true = [c == "A" for c in keyline]
                genes = [true, [not b for b in true], [False] * n, [True] * n]
                for _ in range(64):
                    genes.append([rng.random() < 0.5 for _ in range(n)])
                best = min(phase_score(l, [bool(m >> i & 1) for i in range(n)])
                           for m in range(1 << n))
                entry["global_optimum"] = best
            else:
                genes = [[bool(m >> i & 1) for i in range(n)] for m in range(1 << n)]
            for g in genes:
                s = phase_score(l, g)
                entry["cases"].append({"gene": [int(b) for b in g], "score": s,
                                       "score_hex": hexf(s)})
            entry["kind"] = "phase"
        out[key] = entry
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fixtures.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", dst, {k: len(v["cases"]) for k, v in out.items() if isinstance(v, dict)})


if __name__ == "__main__":
    main()
