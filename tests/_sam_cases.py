"""SAM texts exercising ScaffoldLinksFromSam's loop (alignment.go:624-676): shared by the host-loader
tests (tests/test_sam_links.py, CPU) and the device-loader tests (tests/test_gpu_sam.py)."""
import numpy as np


def _line(qname, rname, pos=1, extra=True):
    f = [qname, "16", rname, str(pos)]
    if extra:
        f += ["60", "3M", "*", "0", "0", "ATG", "III"]
    return "\t".join(f)


def random_sam(seed, n_reads, n_contigs, header=True, crlf=False, interleave_headers=False,
               final_newline=True, long_names=False):
    """Name-sorted SAM text: groups of 1 (single-end / unpaired), 2 (a read pair), 3 or 4 (a pair
    with supplementary hits) hits per read id, contigs drawn with a skew so pairs repeat."""
    rng = np.random.default_rng(seed)
    names = [("contig_%d" % i) + ("_" + "x" * int(rng.integers(1, 40)) if long_names else "")
             for i in range(n_contigs)]
    out = []
    if header:
        out.append("@HD\tVN:1.0\tSO:queryname")
        out += ["@SQ\tSN:%s\tLN:100000" % n for n in names[: min(n_contigs, 50)]]
    sizes = rng.choice([1, 2, 2, 2, 2, 3, 4], size=n_reads)
    for r in range(n_reads):
        for _ in range(int(sizes[r])):
            c = int(min(n_contigs - 1, abs(rng.normal(0, n_contigs / 3))))
            if rng.random() < 0.5:
                c = int(rng.integers(0, n_contigs))
            out.append(_line("read%d" % r, names[c], int(rng.integers(1, 100000)), extra=bool(rng.random() < 0.8)))
        if interleave_headers and rng.random() < 0.05:
            out.append("@CO\tcomment inside the body")   # skipped, does not end the open group
    eol = "\r\n" if crlf else "\n"
    return eol.join(out) + (eol if final_newline else "")


# (name, text, expected status): "ok" or "arg" (the Go code panics: LXY_ERR_ARG)
def edge_cases():
    L = _line
    cases = [
        ("two_pairs_last_unflushed", "\n".join([L("r1", "A"), L("r1", "B"), L("r2", "B"), L("r2", "C")]) + "\n", "ok"),
        ("flushed_by_single", "\n".join([L("r1", "A"), L("r1", "B"), L("r2", "B"), L("r2", "C"), L("r3", "A")]) + "\n", "ok"),
        ("no_final_newline", "\n".join([L("r1", "A"), L("r1", "B"), L("r2", "B"), L("r2", "A"), L("r3", "A")]), "ok"),
        ("triple_is_dropped", "\n".join([L("r1", "A"), L("r1", "B"), L("r1", "C"), L("r2", "B"), L("r2", "C"), L("r3", "A")]) + "\n", "ok"),
        ("same_contig_pair", "\n".join([L("r1", "A"), L("r1", "A"), L("r2", "A"), L("r2", "B"), L("r3", "B")]) + "\n", "ok"),
        ("id_order_follows_flushes", "\n".join([L("r0", "Z"), L("r1", "Y"), L("r1", "X"), L("r2", "X"), L("r2", "Z"), L("r3", "Q")]) + "\n", "ok"),
        ("read_id_returns_later", "\n".join([L("r1", "A"), L("r1", "B"), L("r2", "C"), L("r1", "A"), L("r1", "C"), L("r4", "D")]) + "\n", "ok"),
        ("header_inside_group", "\n".join([L("r1", "A"), "@CO\tx", L("r1", "B"), L("r2", "C")]) + "\n", "ok"),
        ("empty_read_id_first", "\n".join([L("", "A"), L("", "B"), L("r2", "C"), L("r2", "A"), L("", "B")]) + "\n", "ok"),
        ("empty_contig_name", "\n".join([L("r1", ""), L("r1", "B"), L("r2", "C")]) + "\n", "ok"),
        ("three_fields_only", "r1\t0\tA\nr1\t0\tB\nr2\t0\tC\n", "ok"),
        ("prefix_read_ids", "\n".join([L("r1", "A"), L("r11", "B"), L("r11", "C"), L("r1", "A"), L("r1", "B"), L("r", "C")]) + "\n", "ok"),
        ("crlf", "\r\n".join([L("r1", "A"), L("r1", "B"), L("r2", "B\r"), L("r2", "C"), L("r3", "A")]) + "\r\n", "ok"),
        ("only_headers", "@HD\tVN:1.0\n@SQ\tSN:A\tLN:5\n", "ok"),
        ("single_pair_never_added", "\n".join([L("r1", "A"), L("r1", "B")]) + "\n", "ok"),
        ("spaces_are_not_separators", "\n".join([L("r 1", "A b"), L("r 1", "A  b"), L("r 2", "A b")]) + "\n", "ok"),
        ("empty_line", "\n".join([L("r1", "A"), "", L("r1", "B")]) + "\n", "arg"),
        ("blank_last_line", L("r1", "A") + "\n\n", "arg"),
        ("two_fields", "\n".join([L("r1", "A"), "r1\t16", L("r2", "B")]) + "\n", "arg"),
        ("one_field", "justaname\n", "arg"),
    ]
    return cases
