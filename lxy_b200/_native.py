"""ctypes binding of liblxyb200.so (the C ABI declared in include/lxy_b200.h).

The library is built in-tree (lxy_b200/_lib/liblxyb200.so) by `make -C lxy_b200/csrc` for sm_100a.
There is no CPU fallback anywhere in this package: if the library is missing, or no CUDA device
is present when a context is created, the call raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LXY_B200_LIB") or os.path.join(_HERE, "_lib", "liblxyb200.so")
CSRC = os.path.join(_HERE, "csrc")
INT64_MIN = -(1 << 63)

# every symbol include/lxy_b200.h declares (tests check the library exports each of them)
SYMBOLS = [
    "lxy_abi_version", "lxy_last_error", "lxy_device_count", "lxy_ctx_create", "lxy_ctx_destroy",
    "lxy_ctx_get_timing", "lxy_ctx_stream", "lxy_ctx_sync", "lxy_ctx_set_links",
    "lxy_ctx_set_links_dev",
    "lxy_scaff_score_batch", "lxy_scaff_score_batch_dev", "lxy_scaff_positions",
    "lxy_scaff_link_distances", "lxy_links_get", "lxy_ctx_set_links_g", "lxy_ctx_set_links_g_dev",
    "lxy_scaffg_score_batch",
    "lxy_scaffg_score_batch_dev",
    "lxy_scaffg_layout", "lxy_scaffg_link_distances", "lxy_ctx_set_phase_links",
    "lxy_phase_score_batch", "lxy_phase_score_batch_dev", "lxy_rank",
    "lxy_ga_create", "lxy_ga_destroy", "lxy_ga_init", "lxy_ga_set_population", "lxy_ga_optimize",
    "lxy_ga_step_begin", "lxy_ga_step_end", "lxy_ga_score_children", "lxy_ga_best",
    "lxy_ga_population", "lxy_ga_children", "lxy_ga_get_stats",
    "lxy_links_new", "lxy_links_load", "lxy_links_free", "lxy_links_write", "lxy_links_size",
    "lxy_links_npairs", "lxy_links_id", "lxy_links_name", "lxy_links_set", "lxy_links_add",
    "lxy_links_value", "lxy_links_records", "lxy_ctx_set_links_from",
    "lxy_links_load_gpu", "lxy_ctx_load_links",
    "lxy_links_from_sam", "lxy_links_from_sam_gpu", "lxy_scaffold_links_from_sam", "lxy_ctx_load_sam",
    "lxy_scaffold_file", "lxy_phase_file", "lxy_eval_scaffolding", "lxy_eval_phasing",
    "lxy_eval_phasing_dev",
    "lxy_ctx_path", "lxy_phase_score_batch_bits", "lxy_phase_score_batch_bits_dev",
    "lxy_comm_unique_id", "lxy_comm_init_rank", "lxy_comm_init_all", "lxy_comm_destroy", "lxy_comm_info",
    "lxy_comm_slice", "lxy_scaff_score_sharded", "lxy_scaff_score_sharded_dev", "lxy_ga_set_sharding",
    "lxy_links_tabulate_varblocks", "lxy_comm_link_owners",
    "lxy_phase_score_batch_bits_noalloc_dev", "lxy_scaff_score_batch_u16",
]

ACC_AUTO, ACC_F64, ACC_INT = 0, 1, 2
PATH_AUTO, PATH_STREAM, PATH_PAIRS = 0, 1, 2
SHARD_POPULATION, SHARD_LINKS = 0, 1
COMM_ID_BYTES = 128
ABI_VERSION = 2
FLAG_NO_VALIDATE = 1
FLAG_NO_STREAM1 = 2
FLAG_NO_GRAPH = 8
FLAG_DEFER_ERRORS = 16


class LxyError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"liblxyb200 status {status}: {message}")
        self.status = status


class Cfg(C.Structure):
    _fields_ = [("device", C.c_int32), ("acc_mode", C.c_int32), ("flags", C.c_uint32),
                ("path", C.c_int32)]


class Timing(C.Structure):
    _fields_ = [("ms_total", C.c_float), ("ms_layout", C.c_float), ("ms_score", C.c_float),
                ("ms_finish", C.c_float), ("ms_h2d", C.c_float), ("ms_d2h", C.c_float),
                ("launches", C.c_int32), ("grid", C.c_int32), ("link_bytes", C.c_int64),
                ("table_bytes", C.c_int64)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class LoadStats(C.Structure):
    """lxy_load_stats (include/lxy_b200.h): what the device loader read and where its time went."""
    _fields_ = [("bytes", C.c_int64), ("lines", C.c_int64), ("records", C.c_int64), ("pairs", C.c_int64),
                ("names", C.c_int32), ("slow_values", C.c_int32), ("attempts", C.c_int32),
                ("reserved", C.c_int32), ("ms_read", C.c_float), ("ms_h2d", C.c_float),
                ("ms_device", C.c_float), ("ms_total", C.c_float)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class GaParams(C.Structure):
    """lxy_ga_params (include/lxy_b200.h); defaults = the reference's Scaffold (scaff.go:72-94)."""
    _fields_ = [("kind", C.c_int32), ("popsize", C.c_int32), ("p_breed", C.c_double),
                ("p_mutate", C.c_double), ("tour_p", C.c_double), ("tour_k", C.c_int32),
                ("mutators", C.c_uint32), ("seed", C.c_uint64)]


class GaChild(C.Structure):
    """lxy_ga_child (include/lxy_ga_plan.h)."""
    _fields_ = [("kind", C.c_int32), ("pa", C.c_int32), ("pb", C.c_int32), ("p1", C.c_int32),
                ("p2", C.c_int32), ("slot", C.c_int32)]


class GaStats(C.Structure):
    _fields_ = [("generations", C.c_int64), ("scored", C.c_int64), ("last_children", C.c_int32),
                ("ms_last", C.c_float), ("ms_operators", C.c_float), ("ms_score", C.c_float),
                ("ms_select", C.c_float)]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


GA_SCAFF, GA_PHASE = 0, 1
MUT_SHIFT, MUT_SWITCH, MUT_INVERT = 1, 2, 4
CHILD_DTYPE = np.dtype([("kind", "<i4"), ("pa", "<i4"), ("pb", "<i4"), ("p1", "<i4"), ("p2", "<i4"),
                        ("slot", "<i4")])


def build(verbose=False):
    """Compile the CUDA library in-tree (cross-compiles for sm_100a without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.run(["make", "-C", CSRC, "-j4"], check=True, stdout=out)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `make -C {CSRC}` (or __graft_entry__.build()). "
            "lxy_b200 has no CPU fallback.")
    if "LXY_NCCL_LIB" not in os.environ:
        # comm.cu binds NCCL with dlopen on first use.  In a Python process that may import torch LATER,
        # it must be torch's bundled libnccl.so.2: the dynamic loader resolves torch's dependency by
        # soname to whatever copy is already mapped, and an older system NCCL lacks symbols torch needs.
        try:
            import importlib.util
            spec = importlib.util.find_spec("nvidia.nccl")
            for d in (spec.submodule_search_locations or []) if spec else []:
                cand = os.path.join(d, "lib", "libnccl.so.2")
                if os.path.exists(cand):
                    os.environ["LXY_NCCL_LIB"] = cand
                    break
        except Exception:
            pass
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u32, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_double
    sig = {
        "lxy_abi_version": (i32, []),
        "lxy_last_error": (C.c_char_p, []),
        "lxy_device_count": (i32, [C.POINTER(i32)]),
        "lxy_ctx_create": (i32, [C.POINTER(Cfg), C.POINTER(vp)]),
        "lxy_ctx_destroy": (None, [vp]),
        "lxy_ctx_get_timing": (i32, [vp, C.POINTER(Timing)]),
        "lxy_ctx_stream": (i32, [vp, C.POINTER(vp)]),
        "lxy_ctx_sync": (i32, [vp]),
        "lxy_ctx_set_links": (i32, [vp, i32, i64, vp, vp, vp]),
        "lxy_ctx_set_links_dev": (i32, [vp, i32, i64, vp, vp, vp]),
        "lxy_scaff_score_batch": (i32, [vp, vp, i32, vp, vp]),
        "lxy_scaff_score_batch_dev": (i32, [vp, vp, i32, vp, vp]),
        "lxy_scaff_positions": (i32, [vp, vp, i32, vp]),
        "lxy_scaff_link_distances": (i32, [vp, i32, i64, vp, vp, vp, vp]),
        "lxy_links_get": (i32, [vp, i32, i32, C.POINTER(f64)]),
        "lxy_ctx_set_links_g": (i32, [vp, i32, i64, vp, vp, vp, vp, vp, vp, u32]),
        "lxy_ctx_set_links_g_dev": (i32, [vp, i32, i64, vp, vp, vp, vp, vp, vp, u32]),
        "lxy_scaffg_score_batch": (i32, [vp, vp, vp, i32, vp, vp]),
        "lxy_scaffg_score_batch_dev": (i32, [vp, vp, vp, i32, vp, vp]),
        "lxy_scaffg_layout": (i32, [vp, vp, i32, vp, vp]),
        "lxy_scaffg_link_distances": (i32, [vp, i64, vp, vp, vp, vp, vp, vp, vp]),
        "lxy_ctx_set_phase_links": (i32, [vp, i32, i64, vp, vp, vp]),
        "lxy_phase_score_batch": (i32, [vp, vp, i32, vp, vp]),
        "lxy_phase_score_batch_dev": (i32, [vp, vp, i32, vp, vp]),
        "lxy_rank": (i32, [vp, i32, vp]),
        "lxy_ga_create": (i32, [vp, C.POINTER(GaParams), C.POINTER(vp)]),
        "lxy_ga_destroy": (None, [vp]),
        "lxy_ga_init": (i32, [vp, vp, i32]),
        "lxy_ga_set_population": (i32, [vp, vp]),
        "lxy_ga_optimize": (i32, [vp, i32]),
        "lxy_ga_step_begin": (i32, [vp, C.POINTER(i32), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
        "lxy_ga_step_end": (i32, [vp]),
        "lxy_ga_score_children": (i32, [vp, i32, i32]),
        "lxy_ga_best": (i32, [vp, vp, C.POINTER(f64)]),
        "lxy_ga_population": (i32, [vp, vp, vp]),
        "lxy_ga_children": (i32, [vp, vp, vp]),
        "lxy_ga_get_stats": (i32, [vp, C.POINTER(GaStats)]),
        "lxy_links_new": (i32, [C.POINTER(vp)]),
        "lxy_links_load": (i32, [C.c_char_p, C.POINTER(vp)]),
        "lxy_links_free": (None, [vp]),
        "lxy_links_write": (i32, [vp, C.c_char_p]),
        "lxy_links_size": (i32, [vp]),
        "lxy_links_npairs": (i64, [vp]),
        "lxy_links_id": (i32, [vp, C.c_char_p, i32]),
        "lxy_links_name": (C.c_char_p, [vp, i32]),
        "lxy_links_set": (i32, [vp, i32, i32, f64]),
        "lxy_links_add": (i32, [vp, i32, i32, f64]),
        "lxy_links_value": (i32, [vp, i32, i32, C.POINTER(f64)]),
        "lxy_links_records": (i32, [vp, vp, vp, vp]),
        "lxy_ctx_set_links_from": (i32, [vp, vp, i32]),
        "lxy_links_load_gpu": (i32, [C.c_char_p, i32, C.POINTER(vp), C.POINTER(LoadStats)]),
        "lxy_ctx_load_links": (i32, [vp, C.c_char_p, i32, C.POINTER(vp), C.POINTER(LoadStats)]),
        "lxy_links_from_sam": (i32, [C.c_char_p, C.POINTER(vp)]),
        "lxy_links_from_sam_gpu": (i32, [C.c_char_p, i32, C.POINTER(vp), C.POINTER(LoadStats)]),
        "lxy_scaffold_links_from_sam": (i32, [C.c_char_p, C.c_char_p, i32, C.POINTER(LoadStats)]),
        "lxy_ctx_load_sam": (i32, [vp, C.c_char_p, C.POINTER(vp), C.POINTER(LoadStats)]),
        "lxy_scaffold_file": (i32, [C.c_char_p, C.c_char_p, i32, C.POINTER(GaParams), i32, C.POINTER(f64)]),
        "lxy_phase_file": (i32, [C.c_char_p, C.c_char_p, i32, C.POINTER(GaParams), i32, C.POINTER(f64)]),
        "lxy_eval_scaffolding": (i32, [vp, vp, vp, i32, i32, vp]),
        "lxy_eval_phasing": (i32, [vp, vp, vp, i32, i32, vp]),
        "lxy_eval_phasing_dev": (i32, [vp, vp, vp, vp, i32, i32, vp]),
        "lxy_ctx_path": (i32, [vp, C.POINTER(i32), C.POINTER(i64)]),
        "lxy_phase_score_batch_bits": (i32, [vp, vp, i32, vp, vp]),
        "lxy_phase_score_batch_bits_dev": (i32, [vp, vp, i32, vp, vp]),
        "lxy_phase_score_batch_bits_noalloc_dev": (i32, [vp, vp, i32, vp, vp]),
        "lxy_scaff_score_batch_u16": (i32, [vp, vp, i32, vp, vp]),
        "lxy_comm_unique_id": (i32, [vp]),
        "lxy_comm_init_rank": (i32, [vp, i32, i32, vp]),
        "lxy_comm_init_all": (i32, [C.POINTER(vp), i32]),
        "lxy_comm_destroy": (i32, [vp]),
        "lxy_comm_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32)]),
        "lxy_comm_slice": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]),
        "lxy_scaff_score_sharded": (i32, [vp, i32, vp, i32, vp, vp]),
        "lxy_scaff_score_sharded_dev": (i32, [vp, i32, vp, i32, vp, vp]),
        "lxy_ga_set_sharding": (i32, [vp, i32]),
        "lxy_comm_link_owners": (i32, [i32, i64, vp, vp, i32, vp]),
        "lxy_links_tabulate_varblocks": (i32, [vp, C.c_char_p, i64, vp, vp, vp, vp, i32, C.POINTER(i64), C.POINTER(i64)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise LxyError(rc, lib().lxy_last_error().decode(errors="replace"))


def _p(arr):
    return None if arr is None else arr.ctypes.data_as(C.c_void_p)


def _c(arr, dtype):
    return None if arr is None else np.ascontiguousarray(arr, dtype=dtype)


def pack_bits(genes):
    """genes[P][n] of 0/1 -> row-packed uint32 words [P][ceil(n / 32)], LSB first (what
    lxy_phase_score_batch_bits takes; the Go shim does the same while it copies the []bool rows)."""
    genes = np.ascontiguousarray(genes, np.uint8)
    P, n = genes.shape
    W = (n + 31) // 32
    padded = np.zeros((P, W * 32), np.uint8)
    padded[:, :n] = genes != 0
    return np.ascontiguousarray(np.packbits(padded, axis=1, bitorder="little")).view(np.uint32).reshape(P, W)


def comm_unique_id():
    buf = C.create_string_buffer(COMM_ID_BYTES)
    _check(lib().lxy_comm_unique_id(buf))
    return buf.raw


def comm_init_all(ctxs):
    """One process, one context per device: rank i = ctxs[i] (lxy_comm_init_all)."""
    arr = (C.c_void_p * len(ctxs))(*[c._h.value for c in ctxs])
    _check(lib().lxy_comm_init_all(arr, len(ctxs)))


def comm_slice(total, world, rank):
    f, n = C.c_int32(0), C.c_int32(0)
    _check(lib().lxy_comm_slice(total, world, rank, C.byref(f), C.byref(n)))
    return f.value, n.value


def comm_link_owners(n_ids, a, b, world):
    """lxy_comm_link_owners: the rank of every record under the tile-range split of LXY_SHARD_LINKS."""
    a, b = _c(a, np.uint32), _c(b, np.uint32)
    owner = np.zeros(len(a), np.int32)
    _check(lib().lxy_comm_link_owners(n_ids, len(a), _p(a), _p(b), world, _p(owner)))
    return owner


def device_count():
    n = C.c_int32(0)
    rc = lib().lxy_device_count(C.byref(n))
    return n.value if rc == 0 else 0


def rank(scores):
    s = _c(scores, np.float64)
    order = np.zeros(len(s), np.int32)
    _check(lib().lxy_rank(_p(s), len(s), _p(order)))
    return order


class Context:
    """One lxy_ctx: a CUDA device, its stream, the resident link stores."""

    def __init__(self, device=0, acc_mode=ACC_AUTO, flags=0, path=PATH_AUTO):
        self._h = C.c_void_p()
        cfg = Cfg(device, acc_mode, flags, path)
        _check(lib().lxy_ctx_create(C.byref(cfg), C.byref(self._h)))
        self.device = device
        self.n = 0
        self.phase_n = 0

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().lxy_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- plumbing --------------------------------------------------------------------------
    def timing(self):
        t = Timing()
        _check(lib().lxy_ctx_get_timing(self._h, C.byref(t)))
        return t.asdict()

    def stream(self):
        s = C.c_void_p()
        _check(lib().lxy_ctx_stream(self._h, C.byref(s)))
        return s.value or 0

    def sync(self):
        _check(lib().lxy_ctx_sync(self._h))

    def path(self):
        """(LXY_PATH_STREAM | LXY_PATH_PAIRS chosen for the current scaffolding store, records the
        streaming kernels read per candidate)."""
        p, r = C.c_int32(0), C.c_int64(0)
        _check(lib().lxy_ctx_path(self._h, C.byref(p), C.byref(r)))
        return p.value, r.value

    # -- multi-GPU (comm.cu) -----------------------------------------------------------------
    def comm_init_rank(self, world, rank, id128):
        buf = C.create_string_buffer(bytes(id128), COMM_ID_BYTES)
        _check(lib().lxy_comm_init_rank(self._h, world, rank, buf))

    def comm_destroy(self):
        _check(lib().lxy_comm_destroy(self._h))

    def comm_info(self):
        w, r = C.c_int32(0), C.c_int32(0)
        _check(lib().lxy_comm_info(self._h, C.byref(w), C.byref(r)))
        return w.value, r.value

    def scaff_score_sharded(self, mode, perms, want_x100=False):
        """The whole population on every rank (host memory) -> the whole score vector on every rank."""
        perms = _c(perms, np.int32)
        if perms.ndim != 2 or perms.shape[1] != self.n:
            raise ValueError(f"perms must be [P][{self.n}]")
        P = perms.shape[0]
        out, x100 = np.zeros(P, np.float64), np.zeros(P, np.int64)
        _check(lib().lxy_scaff_score_sharded(self._h, mode, _p(perms), P, _p(out), _p(x100)))
        return (out, x100) if want_x100 else out

    def scaff_score_sharded_dev(self, mode, perm_ptr, P, out_ptr, x100_ptr=None):
        _check(lib().lxy_scaff_score_sharded_dev(self._h, mode, C.c_void_p(perm_ptr), P, C.c_void_p(out_ptr),
                                                 C.c_void_p(x100_ptr) if x100_ptr else None))

    # -- scaffolding -----------------------------------------------------------------------
    def set_links(self, n_ids, a, b, w=None):
        a, b, w = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64)
        if len(a) != len(b) or (w is not None and len(w) != len(a)):
            raise ValueError("a, b, w must have equal length")
        _check(lib().lxy_ctx_set_links(self._h, n_ids, len(a), _p(a), _p(b), _p(w)))
        self.n = n_ids

    def set_links_dev(self, n_ids, n_links, a_ptr, b_ptr, w_ptr=None):
        """a_ptr / b_ptr / w_ptr: device addresses (uint32, uint32, float64) on this device."""
        _check(lib().lxy_ctx_set_links_dev(self._h, n_ids, n_links, C.c_void_p(a_ptr),
                                           C.c_void_p(b_ptr),
                                           C.c_void_p(w_ptr) if w_ptr else None))
        self.n = n_ids

    def scaff_score_batch(self, perms, want_x100=False):
        perms = _c(perms, np.int32)
        if perms.ndim != 2 or perms.shape[1] != self.n:
            raise ValueError(f"perms must be [P][{self.n}]")
        P = perms.shape[0]
        out = np.zeros(P, np.float64)
        x100 = np.zeros(P, np.int64)
        _check(lib().lxy_scaff_score_batch(self._h, _p(perms), P, _p(out), _p(x100)))
        return (out, x100) if want_x100 else out

    def scaff_score_batch_u16(self, perms, want_x100=False):
        """perms[P][n] as uint16 (n <= 65,536): half the bytes over PCIe, identical scores."""
        perms = _c(perms, np.uint16)
        if perms.ndim != 2 or perms.shape[1] != self.n:
            raise ValueError(f"perms must be [P][{self.n}]")
        P = perms.shape[0]
        out = np.zeros(P, np.float64)
        x100 = np.zeros(P, np.int64)
        _check(lib().lxy_scaff_score_batch_u16(self._h, _p(perms), P, _p(out), _p(x100)))
        return (out, x100) if want_x100 else out

    def scaff_score_batch_dev(self, perm_ptr, P, out_ptr, x100_ptr=None):
        _check(lib().lxy_scaff_score_batch_dev(self._h, C.c_void_p(perm_ptr), P,
                                               C.c_void_p(out_ptr),
                                               C.c_void_p(x100_ptr) if x100_ptr else None))

    def scaff_positions(self, perms):
        perms = _c(perms, np.int32)
        pos = np.zeros_like(perms)
        _check(lib().lxy_scaff_positions(self._h, _p(perms), perms.shape[0], _p(pos)))
        return pos

    def scaff_link_distances(self, n_ids, a, b, perm):
        a, b, perm = _c(a, np.uint32), _c(b, np.uint32), _c(perm, np.int32)
        dist = np.zeros(len(a), np.int64)
        _check(lib().lxy_scaff_link_distances(self._h, n_ids, len(a), _p(a), _p(b), _p(perm),
                                              _p(dist)))
        return dist

    def links_get(self, id1, id2):
        out = C.c_double()
        rc = lib().lxy_links_get(self._h, id1, id2, C.byref(out))
        return out.value, rc

    # -- G-profile -------------------------------------------------------------------------
    def set_links_g(self, lens, unit, a, xa, b, xb, w=None):
        lens = _c(lens, np.uint32)
        a, xa, b, xb = (_c(v, np.uint32) for v in (a, xa, b, xb))
        w = _c(w, np.float64)
        _check(lib().lxy_ctx_set_links_g(self._h, len(lens), len(a), _p(a), _p(xa), _p(b), _p(xb),
                                         _p(w), _p(lens), unit))
        self.n = len(lens)

    def set_links_g_dev(self, lens, unit, n_links, a_ptr, xa_ptr, b_ptr, xb_ptr, w_ptr=None):
        lens = _c(lens, np.uint32)
        _check(lib().lxy_ctx_set_links_g_dev(self._h, len(lens), n_links, C.c_void_p(a_ptr),
                                             C.c_void_p(xa_ptr), C.c_void_p(b_ptr),
                                             C.c_void_p(xb_ptr),
                                             C.c_void_p(w_ptr) if w_ptr else None, _p(lens), unit))
        self.n = len(lens)

    def scaffg_score_batch(self, perms, oris=None, want_x100=False):
        perms, oris = _c(perms, np.int32), _c(oris, np.uint8)
        P = perms.shape[0]
        out, x100 = np.zeros(P, np.float64), np.zeros(P, np.int64)
        _check(lib().lxy_scaffg_score_batch(self._h, _p(perms), _p(oris), P, _p(out), _p(x100)))
        return (out, x100) if want_x100 else out

    def scaffg_score_batch_dev(self, perm_ptr, ori_ptr, P, out_ptr, x100_ptr=None):
        _check(lib().lxy_scaffg_score_batch_dev(self._h, C.c_void_p(perm_ptr),
                                                C.c_void_p(ori_ptr) if ori_ptr else None, P,
                                                C.c_void_p(out_ptr),
                                                C.c_void_p(x100_ptr) if x100_ptr else None))

    def scaffg_layout(self, perms):
        perms = _c(perms, np.int32)
        start = np.zeros(perms.shape, np.uint64)
        pos = np.zeros(perms.shape, np.int32)
        _check(lib().lxy_scaffg_layout(self._h, _p(perms), perms.shape[0], _p(start), _p(pos)))
        return start, pos

    def scaffg_link_distances(self, a, xa, b, xb, perm, ori=None):
        a, xa, b, xb = (_c(v, np.uint32) for v in (a, xa, b, xb))
        perm, ori = _c(perm, np.int32), _c(ori, np.uint8)
        dist = np.zeros(len(a), np.uint64)
        _check(lib().lxy_scaffg_link_distances(self._h, len(a), _p(a), _p(xa), _p(b), _p(xb),
                                               _p(perm), _p(ori), _p(dist)))
        return dist

    # -- phasing ---------------------------------------------------------------------------
    def set_phase_links(self, n_ids, a, b, w):
        a, b, w = _c(a, np.uint32), _c(b, np.uint32), _c(w, np.float64)
        _check(lib().lxy_ctx_set_phase_links(self._h, n_ids, len(a), _p(a), _p(b), _p(w)))
        self.phase_n = n_ids

    def phase_score_batch(self, genes):
        genes = _c(genes, np.uint8)
        if genes.ndim != 2 or genes.shape[1] != self.phase_n:
            raise ValueError(f"genes must be [P][{self.phase_n}]")
        P = genes.shape[0]
        oi, of = np.zeros(P, np.int64), np.zeros(P, np.float64)
        _check(lib().lxy_phase_score_batch(self._h, _p(genes), P, _p(oi), _p(of)))
        return oi, of

    def phase_score_batch_bits(self, words):
        """words[P][ceil(n / 32)] uint32, bit (i & 31) of word [p][i >> 5] = gene i (pack_bits)."""
        words = _c(words, np.uint32)
        W = (self.phase_n + 31) // 32
        if words.ndim != 2 or words.shape[1] != W:
            raise ValueError(f"words must be [P][{W}]")
        P = words.shape[0]
        oi, of = np.zeros(P, np.int64), np.zeros(P, np.float64)
        _check(lib().lxy_phase_score_batch_bits(self._h, _p(words), P, _p(oi), _p(of)))
        return oi, of

    def phase_score_batch_bits_dev(self, words_ptr, P, out_i64_ptr, out_f64_ptr):
        _check(lib().lxy_phase_score_batch_bits_dev(self._h, C.c_void_p(words_ptr), P,
                                                    C.c_void_p(out_i64_ptr) if out_i64_ptr else None,
                                                    C.c_void_p(out_f64_ptr) if out_f64_ptr else None))

    def phase_score_batch_bits_noalloc_dev(self, words_ptr, P, out_i64_ptr, out_f64_ptr):
        """The same scores with the scratch in the library's static device array: no allocation in the call."""
        _check(lib().lxy_phase_score_batch_bits_noalloc_dev(self._h, C.c_void_p(words_ptr), P,
                                                            C.c_void_p(out_i64_ptr) if out_i64_ptr else None,
                                                            C.c_void_p(out_f64_ptr) if out_f64_ptr else None))

    def phase_score_batch_dev(self, genes_ptr, P, out_i64_ptr, out_f64_ptr):
        _check(lib().lxy_phase_score_batch_dev(self._h, C.c_void_p(genes_ptr), P,
                                               C.c_void_p(out_i64_ptr) if out_i64_ptr else None,
                                               C.c_void_p(out_f64_ptr) if out_f64_ptr else None))


class GA:
    """lxy_ga: the device-resident generation loop (go-galib's GAParallel as the reference uses it,
    scaff.go:80-126 / phase.go:48-95).  Defaults are Scaffold's: population 40, PBreed .2,
    PMutate .6, tournament(0.7, 5), shift + switch + invert mutators."""

    def __init__(self, ctx, kind=GA_SCAFF, popsize=40, p_breed=0.2, p_mutate=0.6, tour_p=0.7, tour_k=5,
                 mutators=MUT_SHIFT | MUT_SWITCH | MUT_INVERT, seed=1):
        self.ctx = ctx
        self.kind = kind
        self.params = GaParams(kind, popsize, p_breed, p_mutate, tour_p, tour_k, mutators, seed)
        self.popsize = popsize
        self.n = ctx.phase_n if kind == GA_PHASE else ctx.n
        self.dtype = np.uint8 if kind == GA_PHASE else np.int32
        self._h = C.c_void_p()
        _check(lib().lxy_ga_create(ctx._h, C.byref(self.params), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().lxy_ga_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def init(self, first=None, randomize=True):
        first = _c(first, self.dtype)
        _check(lib().lxy_ga_init(self._h, _p(first), 1 if randomize else 0))

    def set_population(self, genes):
        genes = _c(genes, self.dtype)
        if genes.shape != (self.popsize, self.n):
            raise ValueError(f"genes must be [{self.popsize}][{self.n}]")
        _check(lib().lxy_ga_set_population(self._h, _p(genes)))

    def optimize(self, generations=1):
        _check(lib().lxy_ga_optimize(self._h, generations))

    def set_sharding(self, mode):
        """mode SHARD_POPULATION / SHARD_LINKS over the context's communicator, -1: local scoring."""
        _check(lib().lxy_ga_set_sharding(self._h, mode))

    def step_begin(self):
        """-> (n_children, children device address, scores device address, x100 device address)"""
        n = C.c_int32(0)
        ch, sc, xx = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _check(lib().lxy_ga_step_begin(self._h, C.byref(n), C.byref(ch), C.byref(sc), C.byref(xx)))
        return n.value, ch.value or 0, sc.value or 0, xx.value or 0

    def score_children(self, first, count):
        _check(lib().lxy_ga_score_children(self._h, first, count))

    def step_end(self):
        _check(lib().lxy_ga_step_end(self._h))

    def children(self, n_children):
        genes = np.zeros((n_children, self.n), self.dtype)
        plan = np.zeros(n_children, CHILD_DTYPE)
        _check(lib().lxy_ga_children(self._h, _p(genes), _p(plan)))
        return genes, plan

    def best(self):
        gene = np.zeros(self.n, self.dtype)
        s = C.c_double(0)
        _check(lib().lxy_ga_best(self._h, _p(gene), C.byref(s)))
        return gene, s.value

    def population(self):
        genes = np.zeros((self.popsize, self.n), self.dtype)
        scores = np.zeros(self.popsize, np.float64)
        _check(lib().lxy_ga_population(self._h, _p(genes), _p(scores)))
        return genes, scores

    def stats(self):
        st = GaStats()
        _check(lib().lxy_ga_get_stats(self._h, C.byref(st)))
        return st.asdict()


class NativeLinks:
    """lxy_links: the library's own sequtil.Links (links.go:14-28) for the fitness path."""

    def __init__(self, path=None, device=None, ctx=None, phase=False):
        """path None: an empty table; path: LoadLinks on the host (lxy_links_load); with device=d the
        text is parsed by the GPU loader (lxy_links_load_gpu); with ctx=... the parsed table also
        becomes that context's link store without leaving the device (lxy_ctx_load_links)."""
        self._h = C.c_void_p()
        self.load_stats = None
        if path is None:
            _check(lib().lxy_links_new(C.byref(self._h)))
        elif ctx is not None:
            st = LoadStats()
            _check(lib().lxy_ctx_load_links(ctx._h, os.fsencode(path), 1 if phase else 0, C.byref(self._h),
                                            C.byref(st)))
            self.load_stats = st.asdict()
            if phase:
                ctx.phase_n = self.Size()
            else:
                ctx.n = self.Size()
        elif device is not None:
            st = LoadStats()
            _check(lib().lxy_links_load_gpu(os.fsencode(path), int(device), C.byref(self._h), C.byref(st)))
            self.load_stats = st.asdict()
        else:
            _check(lib().lxy_links_load(os.fsencode(path), C.byref(self._h)))     # LoadLinks

    @classmethod
    def from_sam(cls, sam_path, device=None, ctx=None):
        """ScaffoldLinksFromSam's table (alignment.go:624-672): the host loop (lxy_links_from_sam), the
        device pipeline with device=d (lxy_links_from_sam_gpu), or with ctx=... the device pipeline
        feeding that context's link store directly (lxy_ctx_load_sam)."""
        self = cls.__new__(cls)
        self._h = C.c_void_p()
        self.load_stats = None
        if ctx is not None:
            st = LoadStats()
            _check(lib().lxy_ctx_load_sam(ctx._h, os.fsencode(sam_path), C.byref(self._h), C.byref(st)))
            self.load_stats = st.asdict()
            ctx.n = self.Size()
        elif device is not None:
            st = LoadStats()
            _check(lib().lxy_links_from_sam_gpu(os.fsencode(sam_path), int(device), C.byref(self._h), C.byref(st)))
            self.load_stats = st.asdict()
        else:
            _check(lib().lxy_links_from_sam(os.fsencode(sam_path), C.byref(self._h)))
        return self

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().lxy_links_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def Size(self):
        return lib().lxy_links_size(self._h)

    def npairs(self):
        return lib().lxy_links_npairs(self._h)

    def ID(self, name, create=True):
        return lib().lxy_links_id(self._h, name.encode(), 1 if create else 0)

    def Decode(self, ids):
        out = []
        for i in ids:
            s = lib().lxy_links_name(self._h, int(i))
            if s is None:
                raise LxyError(-1, lib().lxy_last_error().decode(errors="replace"))
            out.append(s.decode())
        return out

    def Set(self, id1, id2, v):
        _check(lib().lxy_links_set(self._h, id1, id2, float(v)))

    def Add(self, id1, id2, v):
        _check(lib().lxy_links_add(self._h, id1, id2, float(v)))

    def Get(self, id1, id2):
        """(value, error-or-None) like Links.Get: (-1, error) for an unknown id."""
        v = C.c_double(0)
        rc = lib().lxy_links_value(self._h, id1, id2, C.byref(v))
        return v.value, (None if rc == 0 else LxyError(rc, lib().lxy_last_error().decode(errors="replace")))

    def Write(self, path):
        _check(lib().lxy_links_write(self._h, os.fsencode(path)))

    def records(self):
        m = self.npairs()
        a, b, w = np.zeros(m, np.uint32), np.zeros(m, np.uint32), np.zeros(m, np.float64)
        _check(lib().lxy_links_records(self._h, _p(a), _p(b), _p(w)))
        return a, b, w

    def TabulateVariantBlockLinks(self, chrom, off1, blk1, off2, blk2, device=None):
        """TabulateVariantBlockLinks (links.go:372-420) over a batch of read pairs of one chromosome ->
        (ct, balance).  off*: int64 CSR offsets [n_pairs + 1]; blk*: int32 [entries][4] = (block, #ref,
        #alt, #other) as GetVariantBlocks builds them.  device=None: the host loop; device=d: on GPU d."""
        off1, off2 = _c(off1, np.int64), _c(off2, np.int64)
        blk1, blk2 = _c(blk1, np.int32).reshape(-1, 4), _c(blk2, np.int32).reshape(-1, 4)
        if len(off1) != len(off2) or len(off1) < 1:
            raise ValueError("off1 and off2 must both have n_pairs + 1 entries")
        ct, bal = C.c_int64(0), C.c_int64(0)
        _check(lib().lxy_links_tabulate_varblocks(self._h, chrom.encode(), len(off1) - 1, _p(off1), _p(blk1),
                                                  _p(off2), _p(blk2), -1 if device is None else int(device),
                                                  C.byref(ct), C.byref(bal)))
        return ct.value, bal.value

    def by_name(self):
        """{(name, name) sorted: value}: how tables with differently assigned ids compare."""
        a, b, w = self.records()
        names = self.Decode(range(self.Size()))
        return {tuple(sorted((names[int(x)], names[int(y)]))): float(v) for x, y, v in zip(a, b, w)}

    def attach(self, ctx, phase=False):
        _check(lib().lxy_ctx_set_links_from(ctx._h, self._h, 1 if phase else 0))
        if phase:
            ctx.phase_n = self.Size()
        else:
            ctx.n = self.Size()


def scaffold_links_from_sam(sam_path, out_path, device=0):
    """lxy_scaffold_links_from_sam: ScaffoldLinksFromSam(samPath, outPath) (alignment.go:624-676) with
    the table built on the device; returns the load statistics."""
    st = LoadStats()
    _check(lib().lxy_scaffold_links_from_sam(os.fsencode(sam_path), os.fsencode(out_path), int(device), C.byref(st)))
    return st.asdict()


def scaffold_file(links_path, out_path, iterations, params=None, device=0):
    """lxy_scaffold_file: `lxy scaff infer` core inside the library; returns Best().Score()."""
    s = C.c_double(0)
    _check(lib().lxy_scaffold_file(os.fsencode(links_path), os.fsencode(out_path), iterations,
                                   C.byref(params) if params is not None else None, device, C.byref(s)))
    return s.value


def phase_file(links_path, out_path, iterations, params=None, device=0):
    s = C.c_double(0)
    _check(lib().lxy_phase_file(os.fsencode(links_path), os.fsencode(out_path), iterations,
                                C.byref(params) if params is not None else None, device, C.byref(s)))
    return s.value


def eval_scaffolding(ctx, perms, key_order):
    """EvalScaffolding (scaff.go:329-384) for every row of perms against key_order (contig ids in the
    true order): -> [P][2] (qscore, neighbour score)."""
    perms = _c(perms, np.int32)
    key_order = _c(key_order, np.int32)
    n = perms.shape[1]
    rank = np.empty(n, np.int32)
    rank[key_order] = np.arange(n, dtype=np.int32)
    out = np.zeros((perms.shape[0], 2))
    _check(lib().lxy_eval_scaffolding(ctx._h, _p(perms), _p(rank), n, perms.shape[0], _p(out)))
    return out


def eval_phasing(ctx, genes, key):
    """EvalPhasing (phase.go:211-263): -> [P][4]."""
    genes, key = _c(genes, np.uint8), _c(key, np.uint8)
    out = np.zeros((genes.shape[0], 4))
    _check(lib().lxy_eval_phasing(ctx._h, _p(genes), _p(key), genes.shape[1], genes.shape[0], _p(out)))
    return out


def eval_phasing_dev(ctx, genes, key, index):
    """EvalPhasingDev (phase.go:266-325): -> [P][3] (all pairs, index distance <= 2, <= 9)."""
    genes, key, index = _c(genes, np.uint8), _c(key, np.uint8), _c(index, np.int32)
    if genes.ndim != 2 or genes.shape[1] != len(key) or len(index) != len(key):
        raise ValueError("genes must be [P][n]; key and index [n]")
    out = np.zeros((genes.shape[0], 3))
    _check(lib().lxy_eval_phasing_dev(ctx._h, _p(genes), _p(key), _p(index), genes.shape[1], genes.shape[0], _p(out)))
    return out
