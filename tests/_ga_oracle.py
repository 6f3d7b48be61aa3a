"""ctypes binding of oracle/ga_oracle.cpp.  TEST INFRASTRUCTURE: the checker, never the product.
lxy_b200/ must not import this module."""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "_build", "liblxy_ga_oracle.so")
CHILD_DTYPE = np.dtype([("kind", "<i4"), ("pa", "<i4"), ("pb", "<i4"), ("p1", "<i4"), ("p2", "<i4"),
                        ("slot", "<i4")])
CHILD_BREED_A, CHILD_BREED_B, CHILD_SHIFT, CHILD_SWITCH, CHILD_INVERT = 1, 2, 3, 4, 5


def _load():
    srcs = [os.path.join(_ROOT, "oracle", "ga_oracle.cpp"), os.path.join(_ROOT, "include", "lxy_ga_plan.h")]
    if not os.path.exists(_SO) or any(os.path.getmtime(_SO) < os.path.getmtime(s) for s in srcs):
        subprocess.run(["make", "-C", os.path.join(_ROOT, "oracle")], check=True, stdout=subprocess.DEVNULL)
    lib = C.CDLL(_SO)
    vp, i32, u32, u64, f64 = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_double
    lib.lxyo_pmx.argtypes = [vp, vp, i32, i32, i32, vp, vp]
    lib.lxyo_mutate_perm.argtypes = [i32, vp, i32, i32, i32, vp]
    lib.lxyo_mutate_bits.argtypes = [i32, vp, i32, i32, i32, vp]
    lib.lxyo_ga_plan.argtypes = [u64, u64, i32, i32, f64, f64, f64, i32, u32, vp]
    lib.lxyo_ga_plan.restype = i32
    lib.lxyo_ga_children.argtypes = [u64, u64, i32, i32, f64, f64, f64, i32, u32, i32, vp, vp]
    lib.lxyo_ga_children.restype = i32
    lib.lxyo_ga_select.argtypes = [i32, i32, vp, vp, vp]
    lib.lxyo_eval_scaffolding.argtypes = [vp, vp, i32, vp]
    lib.lxyo_eval_phasing.argtypes = [vp, vp, i32, vp]
    lib.lxyo_eval_phasing_dev.argtypes = [vp, vp, vp, i32, vp]
    return lib


_lib = _load()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pmx(a, b, p1, p2):
    """GAOrderedIntGenome.Crossover (scaff/genome.go:47-91), the reference's O(n^2) form."""
    a, b = np.ascontiguousarray(a, np.int32), np.ascontiguousarray(b, np.int32)
    oa, ob = np.empty_like(a), np.empty_like(b)
    _lib.lxyo_pmx(_p(a), _p(b), len(a), p1, p2, _p(oa), _p(ob))
    return oa, ob


def mutate(kind, parent, p1, p2, bits=False):
    parent = np.ascontiguousarray(parent, np.uint8 if bits else np.int32)
    out = np.empty_like(parent)
    (_lib.lxyo_mutate_bits if bits else _lib.lxyo_mutate_perm)(kind, _p(parent), len(parent), p1, p2, _p(out))
    return out


def plan(seed, gen, popsize, length, p_breed, p_mutate, tour_p, tour_k, mutators):
    out = np.zeros(3 * popsize, CHILD_DTYPE)
    c = _lib.lxyo_ga_plan(seed, gen, popsize, length, p_breed, p_mutate, tour_p, tour_k, mutators, _p(out))
    return out[:c]


def children(seed, gen, pop, p_breed, p_mutate, tour_p, tour_k, mutators, bits=False):
    pop = np.ascontiguousarray(pop, np.uint8 if bits else np.int32)
    P, n = pop.shape
    out = np.zeros((3 * P, n), pop.dtype)
    c = _lib.lxyo_ga_children(seed, gen, P, n, p_breed, p_mutate, tour_p, tour_k, mutators, 1 if bits else 0,
                              _p(pop), _p(out))
    return out[:c]


def select(pop_scores, child_scores):
    keys = np.ascontiguousarray(np.concatenate([pop_scores, child_scores]), np.float64)
    order = np.zeros(len(pop_scores), np.int32)
    _lib.lxyo_ga_select(len(pop_scores), len(child_scores), None, _p(keys), _p(order))
    return order


def eval_scaffolding(scaff, key):
    """EvalScaffolding (scaff.go:329-384), the reference's O(n^3) loop; ids instead of names."""
    scaff, key = np.ascontiguousarray(scaff, np.int32), np.ascontiguousarray(key, np.int32)
    out = np.zeros(2)
    _lib.lxyo_eval_scaffolding(_p(scaff), _p(key), len(scaff), _p(out))
    return out


def eval_phasing(phasing, key):
    """EvalPhasing (phase.go:211-263)."""
    phasing, key = np.ascontiguousarray(phasing, np.uint8), np.ascontiguousarray(key, np.uint8)
    out = np.zeros(4)
    _lib.lxyo_eval_phasing(_p(phasing), _p(key), len(phasing), _p(out))
    return out


def eval_phasing_dev(phasing, key, index):
    """EvalPhasingDev (phase.go:266-325); index[i] = the integer after '_' in the i-th block name."""
    phasing, key = np.ascontiguousarray(phasing, np.uint8), np.ascontiguousarray(key, np.uint8)
    index = np.ascontiguousarray(index, np.int32)
    out = np.zeros(3)
    _lib.lxyo_eval_phasing_dev(_p(phasing), _p(key), _p(index), len(phasing), _p(out))
    return out
