#!/usr/bin/env python
"""Generate tests/golden/sam_fixtures.json from the reference's SAM fixture.

Run in the build container (needs /root/reference):  python tests/golden/gen_sam_golden.py

An INDEPENDENT restatement (plain Python lists and dicts) of ScaffoldLinksFromSam,
/root/reference/src/sequtil/alignment.go:624-676, written separately from oracle/lxy_oracle.cpp's
links_from_sam so that a transcription slip in one is caught by the other.  The reference's own test
of this function (alignment_test.go:319-333) is commented out and its key file toy.ctg.links
(chr1-chr2 = 2, chr2-chr3 = 1) is NOT what the function yields on toy.sam: the mates there carry
different read ids (readid-1.1 / readid-1.2), so every group holds one hit and the table stays empty
("toy_sam").  With the mate suffix removed ("toy_sam_paired", derived here from the same file) the
loop still differs from the key: it also counts the four same-contig pairs (chr1-chr1 = 4) and never
adds the last read pair (chr2-chr3), because a group is only added when the next read id arrives.
The goldens below are therefore "what the reference source yields on the reference's fixture" -
parity pinned by source text, not by a reference test.
"""
import json
import os
import re

REF = "/root/reference"
SAM = "src/sequtil/_testdata/toy.sam"


def scaffold_links_from_sam(text):
    id_key, data = {}, {}
    current, hits = "", []

    def ident(name):
        if name not in id_key:
            id_key[name] = len(id_key)
        return id_key[name]

    for line in text.split("\n")[:-1] if text.endswith("\n") else text.split("\n"):
        if line.endswith("\r"):
            line = line[:-1]
        if line[0] != "@":                      # IndexError on an empty line = the Go panic
            arr = line.split("\t")
            if arr[0] != current:
                if len(hits) == 2:
                    a, b = ident(hits[0]), ident(hits[1])
                    if a > b:
                        a, b = b, a
                    data[(a, b)] = data.get((a, b), 0.0) + 1.0
                current, hits = arr[0], []
            hits.append(arr[2])
    names = [None] * len(id_key)
    for k, v in id_key.items():
        names[v] = k
    return names, sorted([a, b, w] for (a, b), w in data.items())


def main():
    text = open(os.path.join(REF, SAM)).read()
    names, records = scaffold_links_from_sam(text)
    header = "@HD\tVN:1.0\tSO:queryname\n@SQ\tSN:chr1\tLN:1000\n"
    paired = header + re.sub(r"^(readid-\d+)\.[12]\t", r"\1\t", text, flags=re.M)
    pnames, precords = scaffold_links_from_sam(paired)
    out = {
        "_generated_by": "tests/golden/gen_sam_golden.py",
        "toy_sam": {"source": SAM, "text": text, "names": names, "records": records},
        "toy_sam_paired": {"source": SAM + " (mate suffix .1/.2 removed, two header lines added)",
                           "text": paired, "names": pnames, "records": precords},
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sam_fixtures.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print(path, names, records, pnames, precords)


if __name__ == "__main__":
    main()
