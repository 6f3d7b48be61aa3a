"""ctypes binding of oracle/lxy_oracle.cpp.  TEST INFRASTRUCTURE: the checker, never the product.

Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
only.  lxy_b200/ must not import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "_build", "liblxy_oracle.so")


def build():
    subprocess.run(["make", "-C", os.path.join(_ROOT, "oracle")], check=True,
                   stdout=subprocess.DEVNULL)
    return _SO


def _load():
    src = os.path.join(_ROOT, "oracle", "lxy_oracle.cpp")
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        build()
    lib = C.CDLL(_SO)
    vp, i32, i64, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_double
    sigs = {
        "lxyo_links_new": (vp, []),
        "lxyo_links_free": (None, [vp]),
        "lxyo_links_add_key": (None, [vp, C.c_char_p]),
        "lxyo_links_id": (i32, [vp, C.c_char_p]),
        "lxyo_links_set": (i32, [vp, i32, i32, f64]),
        "lxyo_links_add": (i32, [vp, i32, i32, f64]),
        "lxyo_links_get": (i32, [vp, i32, i32, C.POINTER(f64)]),
        "lxyo_links_size": (i32, [vp]),
        "lxyo_links_has_name": (i32, [vp, C.c_char_p]),
        "lxyo_links_load": (i32, [C.c_char_p, C.POINTER(vp)]),
        "lxyo_links_from_sam": (i32, [C.c_char_p, C.POINTER(vp)]),
        "lxyo_links_name": (i32, [vp, i32, C.c_char_p, i32]),
        "lxyo_links_npairs": (i64, [vp]),
        "lxyo_links_export": (i64, [vp, vp, vp, vp, i64]),
        "lxyo_links_from_records": (vp, [i32, i64, vp, vp, vp]),
        "lxyo_links_tabulate_varblocks": (None, [vp, C.c_char_p, i64, vp, vp, vp, vp, vp, vp]),
        "lxyo_scaff_score_ref": (f64, [vp, vp, i32]),
        "lxyo_phase_score_ref": (f64, [vp, vp, i32]),
        "lxyo_scaff_score_ref_batch": (None, [vp, vp, i32, i32, i32, vp]),
        "lxyo_phase_score_ref_batch": (None, [vp, vp, i32, i32, i32, vp]),
        "lxyo_scaff_positions": (None, [vp, i32, vp]),
        "lxyo_scaff_score_links": (f64, [i32, i64, vp, vp, vp, vp, vp, vp]),
        "lxyo_scaff_score_links_batch": (None, [i32, i64, vp, vp, vp, vp, i32, i32, vp, vp]),
        "lxyo_scaffg_layout": (None, [i32, vp, vp, vp, vp]),
        "lxyo_scaffg_score_links": (f64, [i32, i64, vp, vp, vp, vp, vp, vp, C.c_uint32, vp, vp,
                                          vp, vp]),
        "lxyo_scaffg_score_links_batch": (None, [i32, i64, vp, vp, vp, vp, vp, vp, C.c_uint32,
                                                 vp, vp, i32, i32, vp, vp]),
        "lxyo_phase_score_links": (f64, [i64, vp, vp, vp, vp, vp]),
        "lxyo_phase_score_links_batch": (None, [i32, i64, vp, vp, vp, vp, i32, i32, vp, vp]),
        "lxyo_rank": (None, [vp, i32, vp]),
        "lxyo_hardware_threads": (i32, []),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


lib = _load()
INT64_MIN = -(1 << 63)


def _p(arr):
    return None if arr is None else arr.ctypes.data_as(C.c_void_p)


def _c(arr, dtype):
    return None if arr is None else np.ascontiguousarray(arr, dtype=dtype)


class Links:
    """sequtil.Links as restated by the oracle (map of maps)."""

    def __init__(self, handle=None):
        self.h = C.c_void_p(handle) if handle else C.c_void_p(lib.lxyo_links_new())

    def __del__(self):
        if getattr(self, "h", None):
            lib.lxyo_links_free(self.h)
            self.h = None

    @classmethod
    def load(cls, path):
        out = C.c_void_p()
        rc = lib.lxyo_links_load(os.fsencode(path), C.byref(out))
        if rc in (-1, -3):
            raise OSError(f"LoadLinks({path}) failed rc={rc}")
        obj = cls(out.value)
        if rc == -2:
            raise ValueError("empty links object")
        return obj

    @classmethod
    def from_sam(cls, path):
        """ScaffoldLinksFromSam's table before Write (alignment.go:624-672)."""
        out = C.c_void_p()
        rc = lib.lxyo_links_from_sam(os.fsencode(path), C.byref(out))
        if rc != 0:
            raise OSError(f"ScaffoldLinksFromSam({path}) failed rc={rc}")
        return cls(out.value)

    @classmethod
    def from_records(cls, n, a, b, w):
        a, b, w = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64)
        return cls(lib.lxyo_links_from_records(n, len(a), _p(a), _p(b), _p(w)))

    def tabulate_varblocks(self, chrom, off1, blk1, off2, blk2):
        """TabulateVariantBlockLinks (links.go:372-420) over a batch of read pairs -> (ct, balance)."""
        off1, off2 = _c(off1, np.int64), _c(off2, np.int64)
        blk1, blk2 = _c(blk1, np.int32), _c(blk2, np.int32)
        ct, bal = C.c_int64(0), C.c_int64(0)
        lib.lxyo_links_tabulate_varblocks(self.h, chrom.encode(), len(off1) - 1, _p(off1), _p(blk1), _p(off2),
                                          _p(blk2), C.cast(C.byref(ct), C.c_void_p), C.cast(C.byref(bal), C.c_void_p))
        return ct.value, bal.value

    def by_name(self):
        """{(name_lo, name_hi): value} with the two names sorted - tables whose ids were assigned in a
        different order (Go map iteration is random) compare equal through this."""
        a, b, w = self.export()
        return {tuple(sorted((self.name(int(x)), self.name(int(y))))): float(v) for x, y, v in zip(a, b, w)}

    def add_key(self, name):
        lib.lxyo_links_add_key(self.h, name.encode())

    def id(self, name):
        return lib.lxyo_links_id(self.h, name.encode())

    def has_name(self, name):
        return bool(lib.lxyo_links_has_name(self.h, name.encode()))

    def set(self, a, b, v):
        return lib.lxyo_links_set(self.h, a, b, v)

    def add(self, a, b, v):
        return lib.lxyo_links_add(self.h, a, b, v)

    def get(self, a, b):
        out = C.c_double()
        rc = lib.lxyo_links_get(self.h, a, b, C.byref(out))
        return out.value, rc

    def size(self):
        return lib.lxyo_links_size(self.h)

    def name(self, i):
        buf = C.create_string_buffer(4096)
        n = lib.lxyo_links_name(self.h, i, buf, 4096)
        return None if n < 0 else buf.value.decode()

    def decode(self, ids):
        names = [self.name(i) for i in ids]
        if any(n is None for n in names):
            raise KeyError("unknown id")
        return names

    def export(self):
        n = lib.lxyo_links_npairs(self.h)
        a, b, w = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.float64)
        lib.lxyo_links_export(self.h, _p(a), _p(b), _p(w), n)
        return a, b, w

    def scaff_score(self, gene):
        g = _c(gene, np.int64)
        return lib.lxyo_scaff_score_ref(self.h, _p(g), len(g))

    def phase_score(self, gene):
        g = _c(gene, np.uint8)
        return lib.lxyo_phase_score_ref(self.h, _p(g), len(g))

    def scaff_score_batch(self, genes, nthreads=1):
        g = _c(genes, np.int32)
        out = np.zeros(g.shape[0], np.float64)
        lib.lxyo_scaff_score_ref_batch(self.h, _p(g), g.shape[0], g.shape[1], nthreads, _p(out))
        return out

    def phase_score_batch(self, genes, nthreads=1):
        g = _c(genes, np.uint8)
        out = np.zeros(g.shape[0], np.float64)
        lib.lxyo_phase_score_ref_batch(self.h, _p(g), g.shape[0], g.shape[1], nthreads, _p(out))
        return out


def scaff_score_links(n, a, b, w, perm, want_dist=False):
    a, b, w, perm = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64), _c(perm, np.int32)
    dist = np.zeros(len(a), np.int64) if want_dist else None
    x100 = C.c_int64()
    s = lib.lxyo_scaff_score_links(n, len(a), _p(a), _p(b), _p(w), _p(perm), _p(dist),
                                   C.cast(C.byref(x100), C.c_void_p))
    return s, x100.value, dist


def scaff_score_links_batch(n, a, b, w, perms, nthreads=1):
    a, b, w = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64)
    perms = _c(perms, np.int32)
    P = perms.shape[0]
    out, x100 = np.zeros(P, np.float64), np.zeros(P, np.int64)
    lib.lxyo_scaff_score_links_batch(n, len(a), _p(a), _p(b), _p(w), _p(perms), P, nthreads,
                                     _p(out), _p(x100))
    return out, x100


def scaffg_layout(lens, perm):
    lens, perm = _c(lens, np.uint32), _c(perm, np.int32)
    n = len(lens)
    start, pos = np.zeros(n, np.uint64), np.zeros(n, np.int32)
    lib.lxyo_scaffg_layout(n, _p(lens), _p(perm), _p(start), _p(pos))
    return start, pos


def scaffg_score_links(lens, unit, a, xa, b, xb, w, perm, ori=None, want_dist=False):
    lens, perm = _c(lens, np.uint32), _c(perm, np.int32)
    a, xa, b, xb = (_c(v, np.uint32) for v in (a, xa, b, xb))
    w, ori = _c(w, np.float64), _c(ori, np.uint8)
    dist = np.zeros(len(a), np.uint64) if want_dist else None
    x100 = C.c_int64()
    s = lib.lxyo_scaffg_score_links(len(lens), len(a), _p(a), _p(xa), _p(b), _p(xb), _p(w),
                                    _p(lens), unit, _p(perm), _p(ori), _p(dist),
                                    C.cast(C.byref(x100), C.c_void_p))
    return s, x100.value, dist


def scaffg_score_links_batch(lens, unit, a, xa, b, xb, w, perms, oris=None, nthreads=1):
    lens, perms = _c(lens, np.uint32), _c(perms, np.int32)
    a, xa, b, xb = (_c(v, np.uint32) for v in (a, xa, b, xb))
    w, oris = _c(w, np.float64), _c(oris, np.uint8)
    P = perms.shape[0]
    out, x100 = np.zeros(P, np.float64), np.zeros(P, np.int64)
    lib.lxyo_scaffg_score_links_batch(len(lens), len(a), _p(a), _p(xa), _p(b), _p(xb), _p(w),
                                      _p(lens), unit, _p(perms), _p(oris), P, nthreads,
                                      _p(out), _p(x100))
    return out, x100


def phase_score_links(a, b, w, gene):
    a, b, w, gene = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64), _c(gene, np.uint8)
    ex = C.c_int64()
    s = lib.lxyo_phase_score_links(len(a), _p(a), _p(b), _p(w), _p(gene),
                                   C.cast(C.byref(ex), C.c_void_p))
    return s, ex.value


def phase_score_links_batch(n, a, b, w, genes, nthreads=1):
    a, b, w, genes = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64), _c(genes, np.uint8)
    P = genes.shape[0]
    out, ex = np.zeros(P, np.float64), np.zeros(P, np.int64)
    lib.lxyo_phase_score_links_batch(n, len(a), _p(a), _p(b), _p(w), _p(genes), P, nthreads,
                                     _p(out), _p(ex))
    return out, ex


def rank(scores):
    s = _c(scores, np.float64)
    order = np.zeros(len(s), np.int32)
    lib.lxyo_rank(_p(s), len(s), _p(order))
    return order


def hardware_threads():
    return lib.lxyo_hardware_threads()
