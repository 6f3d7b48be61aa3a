"""Host-side mirror of the reference interface for the fitness path (Python above the C ABI; the
reference's own host language, Go, has no toolchain in this image — INTEGRATION.md shows the cgo
binding).  Names, argument meaning and error behaviour follow the reference:

  Links, LoadLinks                     /root/reference/src/sequtil/links.go:14-28, 427-491
  GAOrderedIntGenome                   /root/reference/src/lxy/scaff/genome.go:20-209
  GAFixedBitstringGenome               /root/reference/src/lxy/phase/genome.go:19-131
  ScaffScorer.score / PhaseScorer.score   the `sfunc` handed to the genomes: scaff.go:246-310,
                                          phase.go:171-207 — evaluated on the GPU
  ScoreBatch                           new: fills score/hasscore of every unscored genome of a
                                       generation in one device call; Score() is then a memo hit.

Nothing here computes a fitness on the CPU: the scorers call liblxyb200 and fail loudly without it.
"""
import random

import numpy as np

from . import _native as N


class LinksError(Exception):
    pass


class Links:
    """sequtil.Links (links.go:14-28): name<->id maps and the upper-triangular weight table."""

    def __init__(self):
        self.idKey = {}        # name -> id            (links.go:17)
        self.idKeyRev = {}     # id -> name            (links.go:20)
        self.maxid = 0         #                       (links.go:23)
        self.data = {}         # id -> {id -> float}   (links.go:27), smaller id on top

    # addKey (links.go:51-62)
    def addKey(self, key):
        i = self.maxid
        self.idKey[key] = i
        self.idKeyRev[i] = key
        self.maxid = i + 1

    # ID (links.go:100-110): create on first sight
    def ID(self, key):
        if key not in self.idKey:
            self.addKey(key)
        return self.idKey[key]

    # IntIDs / StringIDs (links.go:66-97); the Go map iteration order is random, ascending here
    def IntIDs(self):
        return sorted(self.idKey.values())

    def StringIDs(self):
        return [self.idKeyRev[i] for i in self.IntIDs()]

    def _check(self, id1, id2):
        if id1 not in self.idKeyRev or id2 not in self.idKeyRev:
            raise LinksError("an id is not present in the Links object")     # links.go:116-127
        return (id2, id1) if id1 > id2 else (id1, id2)                       # links.go:130-132

    def Set(self, id1, id2, val):       # links.go:113-145 (overwrite)
        a, b = self._check(id1, id2)
        self.data.setdefault(a, {})[b] = float(val)

    def Add(self, id1, id2, val):       # links.go:149-177 (accumulate)
        a, b = self._check(id1, id2)
        row = self.data.setdefault(a, {})
        row[b] = row.get(b, 0.0) + float(val)

    def Get(self, id1, id2):            # links.go:190-221: (-1, error) unknown id; 0 absent pair
        if id1 not in self.idKeyRev or id2 not in self.idKeyRev:
            return -1.0, LinksError("an id is not present in the Links object")
        a, b = (id2, id1) if id1 > id2 else (id1, id2)
        return self.data.get(a, {}).get(b, 0.0), None

    def Size(self):                     # links.go:267-269
        return len(self.idKey)

    def Decode(self, ids):              # links.go:276-286
        try:
            return [self.idKeyRev[int(i)] for i in ids]
        except KeyError:
            raise LinksError("Error decoding links, an input id was not recognized.")

    def DecodePhasing(self, bits):      # links.go:288-299
        try:
            return {self.idKeyRev[i]: bool(b) for i, b in enumerate(bits)}
        except KeyError:
            raise LinksError("Error decoding links, an input id was not recognized.")

    def Write(self, out):               # links.go:242-264 (header line, then `name name %f`)
        out.write("#" + "".join(f" {k}:{v}" for k, v in self.idKey.items()) + "\n")
        for a, row in self.data.items():
            for b, v in row.items():
                out.write(f"{self.idKeyRev[a]} {self.idKeyRev[b]} {v:f}\n")

    def records(self):
        """Flat (a, b, w) arrays, ascending (a, b): what lxy_ctx_set_links ingests."""
        keys = sorted((a, b) for a, row in self.data.items() for b in row)
        a = np.fromiter((k[0] for k in keys), np.uint32, len(keys))
        b = np.fromiter((k[1] for k in keys), np.uint32, len(keys))
        w = np.fromiter((self.data[k[0]][k[1]] for k in keys), np.float64, len(keys))
        return a, b, w


def _parse_float_or_zero(tok):
    try:                                # `val, _ := strconv.ParseFloat` (links.go:473): error -> 0
        return float(tok)
    except ValueError:
        return 0.0


def LoadLinks(path):
    """links.go:427-491: skip lines starting '#', split on single spaces, ids by first appearance,
    Set (the last row of a pair wins, (a,b) == (b,a)); error when nothing was loaded.  An empty
    line or a line with fewer than three fields makes the Go code panic (line[0], arr[2]); here it
    raises LinksError."""
    l = Links()
    with open(path) as f:
        for raw in f:
            line = raw.rstrip("\n").rstrip("\r")
            if line == "":
                raise LinksError("empty line in links file (the reference panics here)")
            if line[0] == "#":
                continue
            arr = line.split(" ")
            if len(arr) < 3:
                raise LinksError("links line has fewer than 3 fields (the reference panics here)")
            for name in arr[:2]:
                if name not in l.idKey:
                    l.addKey(name)
            l.Set(l.idKey[arr[0]], l.idKey[arr[1]], _parse_float_or_zero(arr[2]))
    if l.Size() <= 0:
        raise LinksError("Loaded empty links object...")                      # links.go:485-487
    return l


def LoadLinksDevice(path, device=0):
    """LoadLinks with the text parsed on the GPU (lxy_links_load_gpu, lxy_b200/csrc/links_parse.cu):
    the same Links object as LoadLinks(path) - ids by first appearance, the last row of a pair wins -
    for files whose size makes the line-by-line host loop the longest step of a run."""
    try:
        t = N.NativeLinks(path, device=device)
    except N.LxyError as e:
        raise LinksError(str(e))
    l = Links()
    for name in t.Decode(range(t.Size())):
        l.addKey(name)
    a, b, w = t.records()
    for x, y, v in zip(a.tolist(), b.tolist(), w.tolist()):
        l.data.setdefault(x, {})[y] = v
    t.close()
    return l


def ScaffoldLinksFromSam(samPath, outPath, device=0):
    """sequtil.ScaffoldLinksFromSam (alignment.go:624-676) with the reference's signature: the SAM text
    is grouped by read id and counted into contig pairs on the GPU (lxy_scaffold_links_from_sam), the
    table is written in the `.links` format LoadLinks reads (links.go:242-264)."""
    try:
        return N.scaffold_links_from_sam(samPath, outPath, device=device)
    except N.LxyError as e:
        raise LinksError(str(e))


# ---------------------------------------------------------------------------------------------
# genomes
# ---------------------------------------------------------------------------------------------
class GAOrderedIntGenome:
    """scaff/genome.go:20-209: a permutation, a memoised score, the link table pointer."""

    def __init__(self, gene, sfunc):
        self.Gene = list(int(x) for x in gene)
        self.score = 0.0
        self.hasscore = False
        self.sfunc = sfunc
        self.data = None

    def Copy(self):                                   # genome.go:140-149 (memo travels along)
        n = GAOrderedIntGenome(self.Gene, self.sfunc)
        n.score, n.hasscore, n.data = self.score, self.hasscore, self.data
        return n

    def Len(self):
        return len(self.Gene)

    def Score(self):                                  # genome.go:153-159
        if not self.hasscore:
            self.score = self.sfunc(self)
            self.hasscore = True
        return self.score

    def Reset(self):                                  # genome.go:179
        self.hasscore = False

    def Switch(self, x, y):                           # genome.go:125-128
        self.Gene[x], self.Gene[y] = self.Gene[y], self.Gene[x]
        self.Reset()

    def Invert(self, p1, p2):                         # genome.go:161-177 (segment reversal)
        if p1 > p2:
            p1, p2 = p2, p1
        while p1 < p2:
            self.Switch(p1, p2)
            p1 += 1
            p2 -= 1

    def Randomize(self, rng=random):                  # genome.go:130-138
        l = len(self.Gene)
        for _ in range(l):
            x, y = rng.randrange(l), rng.randrange(l)
            self.Gene[x], self.Gene[y] = self.Gene[y], self.Gene[x]
        self.Reset()

    def Splice(self, b, frm, to, length):             # genome.go:93-97
        self.Gene[to:to + length] = b.Gene[frm:frm + length]
        self.Reset()

    def Valid(self):                                  # genome.go:111-123
        return len(set(self.Gene)) == len(self.Gene)

    def Crossover(self, b, p1, p2):
        """Partially mapped crossover, same result as genome.go:47-91 but O(n): the reference's
        pmxmap is a linear scan per element (O(n^2)); an index of the positions
        outside the window replaces it."""
        ca, cb = self.Copy(), b.Copy()
        ca.Gene[p1:p2 + 1] = b.Gene[p1:p2 + 1]
        cb.Gene[p1:p2 + 1] = self.Gene[p1:p2 + 1]
        out_a = {v: i for i, v in enumerate(ca.Gene) if i < p1 or i > p2}
        out_b = {v: i for i, v in enumerate(cb.Gene) if i < p1 or i > p2}
        amap, bmap = [], []

        def swap():
            i1, i2 = amap.pop(), bmap.pop()
            va, vb = ca.Gene[i1], cb.Gene[i2]
            ca.Gene[i1], cb.Gene[i2] = vb, va
            out_a.pop(va, None)
            out_b.pop(vb, None)
            out_a[vb] = i1
            out_b[va] = i2

        for i in range(p1, p2 + 1):
            ma = out_a.get(ca.Gene[i])
            if ma is not None:
                amap.append(ma)
                if bmap:
                    swap()
            mb = out_b.get(cb.Gene[i])
            if mb is not None:
                bmap.append(mb)
                if amap:
                    swap()
        ca.Reset()
        cb.Reset()
        return ca, cb

    def String(self):
        return "[" + " ".join(str(x) for x in self.Gene) + "]"


def NewOrderedIntGenome(i, sfunc):                    # genome.go:28-34
    return GAOrderedIntGenome(i, sfunc)


class GAFixedBitstringGenome:
    """phase/genome.go:19-131."""

    def __init__(self, gene, sfunc):
        self.Gene = [bool(x) for x in gene]
        self.score = 0.0
        self.hasscore = False
        self.sfunc = sfunc
        self.data = None

    def Copy(self):                                   # phase/genome.go:74-83
        n = GAFixedBitstringGenome(self.Gene, self.sfunc)
        n.score, n.hasscore, n.data = self.score, self.hasscore, self.data
        return n

    def Len(self):
        return len(self.Gene)

    def Score(self):                                  # phase/genome.go:87-93
        if not self.hasscore:
            self.score = self.sfunc(self)
            self.hasscore = True
        return self.score

    def Reset(self):
        self.hasscore = False

    def Valid(self):
        return True

    def Switch(self, x, y):                           # phase/genome.go:56-59
        self.Gene[x], self.Gene[y] = self.Gene[y], self.Gene[x]
        self.Reset()

    def Invert(self, p1, p2):
        """Bit flip of [p1, p2) (phase/genome.go:95-109).  The reference forgets Reset() here, so a
        mutated copy keeps its parent's memoised score; that is a host-side bug of
        the reference, not part of the scorer, and is NOT reproduced: the memo is cleared."""
        if p1 > p2:
            p1, p2 = p2, p1
        for i in range(p1, p2):
            self.Gene[i] = not self.Gene[i]
        self.Reset()

    def Randomize(self, rng=random):                  # phase/genome.go:61-72
        self.Gene = [rng.randrange(2) == 1 for _ in self.Gene]
        self.Reset()

    def Splice(self, b, frm, to, length):
        self.Gene[to:to + length] = b.Gene[frm:frm + length]
        self.Reset()

    def Crossover(self, b, p1, p2):                   # phase/genome.go:37-46 (2-point)
        ca, cb = self.Copy(), b.Copy()
        ca.Gene[p1:p2 + 1] = b.Gene[p1:p2 + 1]
        cb.Gene[p1:p2 + 1] = self.Gene[p1:p2 + 1]
        ca.Reset()
        cb.Reset()
        return ca, cb

    def String(self):
        return "[" + " ".join("true" if x else "false" for x in self.Gene) + "]"


def NewFixedBitstringGenome(i, sfunc):                # phase/genome.go:27-34
    return GAFixedBitstringGenome(i, sfunc)


# ---------------------------------------------------------------------------------------------
# GPU-backed score functions
# ---------------------------------------------------------------------------------------------
class ScaffScorer:
    """`score` of scaff.go:246-310 on the GPU.  `score(g)` has the reference's signature (it is
    what NewOrderedIntGenome receives as sfunc); ScoreBatch is the per-generation pre-pass."""

    def __init__(self, links, device=0, acc_mode=N.ACC_AUTO):
        self.links = links
        self.ctx = N.Context(device, acc_mode=acc_mode)
        self.n = links.Size()
        a, b, w = links.records()
        self.ctx.set_links(self.n, a, b, w)
        self.scores = 0                                # `var scores int` (scaff.go:18)

    def score(self, g):
        self.scores += 1
        try:
            return float(self.ctx.scaff_score_batch(np.asarray([g.Gene], np.int32))[0])
        except N.LxyError:
            return float("inf")                        # never elite; the reference has no error path

    def ScoreBatch(self, genomes):
        todo = [g for g in genomes if not g.hasscore]
        if not todo:
            return 0
        out = self.ctx.scaff_score_batch(np.asarray([g.Gene for g in todo], np.int32))
        for g, s in zip(todo, out):
            g.score, g.hasscore = float(s), True
        self.scores += len(todo)
        return len(todo)

    def close(self):
        self.ctx.close()


class PhaseScorer:
    """`score` of phase.go:171-207 on the GPU (gene index == link id, phase.go:185)."""

    def __init__(self, links, device=0):
        self.links = links
        self.ctx = N.Context(device)
        self.n = links.Size()
        a, b, w = links.records()
        self.ctx.set_phase_links(self.n, a, b, w)
        self.scores = 0

    def score(self, g):
        self.scores += 1
        try:
            return float(self.ctx.phase_score_batch(np.asarray([g.Gene], np.uint8))[1][0])
        except N.LxyError:
            return float("inf")

    def ScoreBatch(self, genomes):
        todo = [g for g in genomes if not g.hasscore]
        if not todo:
            return 0
        _, out = self.ctx.phase_score_batch(np.asarray([g.Gene for g in todo], np.uint8))
        for g, s in zip(todo, out):
            g.score, g.hasscore = float(s), True
        self.scores += len(todo)
        return len(todo)

    def close(self):
        self.ctx.close()


def Rank(genomes):
    """Ascending fitness (lower is better, scaff.go:308), ties keep population order."""
    return [genomes[i] for i in N.rank(np.asarray([g.Score() for g in genomes], np.float64))]


# ---------------------------------------------------------------------------------------------
# workflows: the generation loops of scaff.Scaffold / phase.Phase on the device-resident GA
# ---------------------------------------------------------------------------------------------
class Intermediate:
    """scaff.go:59-63."""

    def __init__(self, iteration, score, order):
        self.iteration, self.score, self.order = iteration, score, order


def WriteScaffolding(scaffolding, path):              # scaff.go:24-39: one contig name per line
    with open(path, "w") as f:
        for v in scaffolding:
            f.write(v + "\n")


def ReadScaffolding(path):                            # scaff.go:41-57
    with open(path) as f:
        return [line.rstrip("\n") for line in f]


def Scaffold(links, outPath, iterations, optReportFreq, device=0, seed=None, popsize=40):
    """scaff.Scaffold (scaff.go:68-141) with the same arguments and return values: the generation
    loop `gao.Optimize(1); best := gao.Best()` runs on the GPU (lxy_ga_*), with the reference's
    parameters (population 40, PMutate .6, PBreed .2, tournament(0.7, 5), shift + switch + invert
    mutators; scaff.go:72-94).  The reference seeds from the clock (scaff.go:70); `seed` makes a run
    reproducible.  Returns (scaffolding, intermediateSolutions)."""
    import math
    import time
    n = links.Size()
    a, b, w = links.records()
    if seed is None:
        seed = time.time_ns() & 0xFFFFFFFFFFFF
    with N.Context(device) as ctx:
        ctx.set_links(n, a, b, w)
        with N.GA(ctx, N.GA_SCAFF, popsize=popsize, p_breed=0.2, p_mutate=0.6, tour_p=0.7, tour_k=5,
                  mutators=N.MUT_SHIFT | N.MUT_SWITCH | N.MUT_INVERT, seed=seed) as ga:
            ga.init(np.asarray(links.IntIDs(), np.int32), randomize=True)     # GARandomInitializer
            reportEvery = int(math.floor(iterations * optReportFreq) + 1.0)   # scaff.go:96
            intermediates, ct, intermed = [], 0, 0
            while True:                                                       # scaff.go:105-126
                ct += 1
                ga.optimize(1)
                if ct >= iterations:
                    break
                if intermed > reportEvery:
                    gene, sc = ga.best()
                    intermediates.append(Intermediate(ct, sc, links.Decode(gene)))
                    intermed = 0
                intermed += 1
            gene, _ = ga.best()
            scores = ga.stats()["scored"]                                     # `Calls to score`
    scaffolding = links.Decode(gene)
    WriteScaffolding(scaffolding, outPath)
    Scaffold.scores = scores
    return scaffolding, intermediates


def writePhasing(phasing, path):                      # phase.go:151-167: `name 1` / `name 0`
    with open(path, "w") as f:
        for k, v in phasing.items():
            f.write(f"{k} {1 if v else 0}\n")


def Phase(links, outPath, iterations, optReportFreq=0.01, device=0, seed=None, popsize=10):
    """phase.Phase (phase.go:34-111): all-false start genome of links.Size() bits, population 10,
    PMutate .7, PBreed .7, tournament(0.7, 5), invert + switch mutators (phase.go:42-63), the
    generation loop on the GPU.  Returns the phasing as {name: bool} (DecodePhasing)."""
    import time
    n = links.Size()
    a, b, w = links.records()
    if seed is None:
        seed = time.time_ns() & 0xFFFFFFFFFFFF
    with N.Context(device) as ctx:
        ctx.set_phase_links(n, a, b, w)
        with N.GA(ctx, N.GA_PHASE, popsize=popsize, p_breed=0.7, p_mutate=0.7, tour_p=0.7, tour_k=5,
                  mutators=N.MUT_INVERT | N.MUT_SWITCH, seed=seed) as ga:
            ga.init(np.zeros(n, np.uint8), randomize=True)
            ga.optimize(iterations)
            gene, _ = ga.best()
    phasing = links.DecodePhasing(gene)
    writePhasing(phasing, outPath)
    return phasing


# ---- solution-quality metrics with the reference's signatures (maps / lists of names) ----------------
def _atoi_or_zero(tok):                               # strconv.Atoi with the error ignored (phase.go:279-280)
    body = tok[1:] if tok[:1] in ("+", "-") else tok
    if not body or not all("0" <= ch <= "9" for ch in body):
        return 0
    return max(-2**31, min(2**31 - 1, int(tok)))


def EvalPhasingDev(phasing, key, device=0):
    """EvalPhasingDev(phasing, key map[string]bool) (phase.go:266-325) on the device: (score over all
    ordered pairs, over pairs whose name indices differ by <= 2, by <= 9).  Names missing from `key`
    are skipped as in the reference (:289-299); a name without '_' makes the reference panic (:279) and
    raises here."""
    names = [nm for nm in phasing if nm in key]
    if not names:
        return float("nan"), float("nan"), float("nan")                  # 0/0 in the reference
    index = []
    for nm in names:
        parts = nm.split("_")
        if len(parts) < 2:
            raise IndexError(f"variant block name {nm!r} has no '_' (the reference panics here)")
        index.append(_atoi_or_zero(parts[1]))
    genes = np.array([[1 if phasing[nm] else 0 for nm in names]], np.uint8)
    keyb = np.array([1 if key[nm] else 0 for nm in names], np.uint8)
    with N.Context(device) as ctx:
        out = N.eval_phasing_dev(ctx, genes, keyb, np.array(index, np.int32))
    return float(out[0, 0]), float(out[0, 1]), float(out[0, 2])


def EvalScaffolding(scaff, key, device=0):
    """EvalScaffolding(scaff, key []string) (scaff.go:329-384) on the device: (score, neighbour score).
    The device metric is defined for a scaffolding that is a permutation of the key (what Scaffold
    returns for the key's link set); anything else raises."""
    rank = {nm: i for i, nm in enumerate(key)}
    if len(rank) != len(key) or len(scaff) != len(key) or any(nm not in rank for nm in scaff):
        raise ValueError("EvalScaffolding: the scaffolding must be a permutation of the key")
    ids = np.array([[rank[nm] for nm in scaff]], np.int32)      # contig id := its index in the key
    with N.Context(device) as ctx:
        out = N.eval_scaffolding(ctx, ids, np.arange(len(key), dtype=np.int32))
    return float(out[0, 0]), float(out[0, 1])


def EvalPhasing(phasing, key, device=0):
    """EvalPhasing(phasing, key []bool) (phase.go:211-263): (score, firstScore, secondScore, thirdScore)."""
    with N.Context(device) as ctx:
        out = N.eval_phasing(ctx, np.array([[1 if v else 0 for v in phasing]], np.uint8),
                             np.array([1 if v else 0 for v in key], np.uint8))
    return tuple(float(v) for v in out[0])


def readPhasing(path):                                # phase.go:119-149: `name 0|1` per line
    phasing = {}
    with open(path) as f:
        for line in f.read().splitlines():
            arr = line.split(" ")
            if len(arr) != 2 or arr[1] not in ("0", "1"):
                raise ValueError("Tried to read phasing file with more than two haplotypes, not currently supported.")
            phasing[arr[0]] = arr[1] == "1"
    return phasing


def ScaffoldInfer(linksPath, outputPrefix, keyPath, iterations=1000, optReportFreq=0.01, device=0, seed=None):
    """`lxy scaff infer` (scaff/cmd.go:135-202) without the plots: LoadLinks (text parsed on the device),
    Scaffold -> <prefix>.scaff.txt, EvalScaffolding -> <prefix>.qscore.txt (`score=%f;score_neighbor=%f`),
    the intermediates' scores -> <prefix>.opt.txt (`iteration fitness qscore` per line).  Returns
    (scaffolding, score, neighbour score)."""
    links = LoadLinksDevice(linksPath, device=device)
    key = ReadScaffolding(keyPath)
    scaffolding, intermediates = Scaffold(links, outputPrefix + ".scaff.txt", iterations, optReportFreq,
                                          device=device, seed=seed)
    score, nscore = EvalScaffolding(scaffolding, key, device=device)
    with open(outputPrefix + ".qscore.txt", "w") as f:
        f.write("score=%f;score_neighbor=%f\n" % (score, nscore))
    with open(outputPrefix + ".opt.txt", "w") as f:
        if intermediates:                                           # one device call for all of them
            rank = {nm: i for i, nm in enumerate(key)}
            ids = np.array([[rank[nm] for nm in im.order] for im in intermediates], np.int32)
            with N.Context(device) as ctx:
                q = N.eval_scaffolding(ctx, ids, np.arange(len(key), dtype=np.int32))
            for im, row in zip(intermediates, q):
                f.write("%d %f %f\n" % (im.iteration, im.score, row[0]))
    return scaffolding, score, nscore


def PhaseInfer(linksPath, outpath, keyPath, iterations=1000, subruntag="", optReportFreq=0.01, device=0, seed=None):
    """`lxy phase infer` (phase/cmd.go:114-165): LoadLinks, Phase -> <outpath>.<tag>.phasing.txt,
    EvalPhasingDev against the key -> <outpath>.<tag>.qscore.txt
    (`score=%f;300kbp_score=%f;1mbp_score=%f`).  Returns (phasing, (score, score3k, score1mbp))."""
    links = LoadLinksDevice(linksPath, device=device)
    phasing = Phase(links, outpath + "." + subruntag + ".phasing.txt", iterations, optReportFreq,
                    device=device, seed=seed)
    key = readPhasing(keyPath)
    scores = EvalPhasingDev(phasing, key, device=device)
    with open(outpath + "." + subruntag + ".qscore.txt", "w") as f:
        f.write("score=%f;300kbp_score=%f;1mbp_score=%f\n" % scores)
    return phasing, scores
