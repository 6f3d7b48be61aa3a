"""CPU-side checks of the C-ABI boundary: the library builds, loads, exports every symbol that
include/lxy_b200.h declares, and refuses to work without a CUDA device (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

import lxy_b200._native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def native():
    if not os.path.exists(N.LIB_PATH):
        N.build()
    return N.lib()


def _header_functions():
    text = open(os.path.join(ROOT, "include", "lxy_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lxy_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _header_functions() == sorted(N.SYMBOLS)


def test_library_exports_every_declared_symbol(native):
    for name in _header_functions():
        assert hasattr(native, name), name


def test_abi_version(native):
    assert native.lxy_abi_version() == N.ABI_VERSION == 2


def test_rank_is_stable_ascending(native):
    s = np.array([3.0, -1.0, 3.0, -7.5, -1.0, 0.0])
    assert N.rank(s).tolist() == [3, 1, 4, 5, 0, 2]
    assert N.rank(np.zeros(0)).tolist() == []


def test_comm_slice_is_a_balanced_contiguous_split(native):
    """lxy_comm_slice needs no device: every row exactly once, sizes differ by at most one."""
    for total, world in [(1024, 8), (1028, 8), (5, 8), (0, 3), (129, 1), (1000, 7)]:
        at, sizes = 0, []
        for r in range(world):
            f, c = N.comm_slice(total, world, r)
            assert f == at
            at += c
            sizes.append(c)
        assert at == total and max(sizes) - min(sizes) <= 1
    with pytest.raises(N.LxyError):
        N.comm_slice(10, 0, 0)


def test_pack_bits_layout():
    g = np.zeros((2, 70), np.uint8)
    g[0, [0, 31, 32, 69]] = 1
    g[1, 33] = 1
    w = N.pack_bits(g)
    assert w.shape == (2, 3) and w.dtype == np.uint32
    assert w[0].tolist() == [1 | (1 << 31), 1, 1 << 5] and w[1].tolist() == [0, 2, 0]


def test_no_cpu_fallback_without_a_device(native):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(N.LxyError) as e:
        N.Context(0)
    assert e.value.status == -2          # LXY_ERR_CUDA
    assert "no CPU fallback" in str(e.value)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "lxy_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "lxy_oracle" not in src and "_oracle" not in src, os.path.join(dirpath, f)


def test_link_owners_split_whole_tiles_and_match_the_bench_mirror(native):
    """lxy_comm_link_owners: contiguous tile ranges of the sorted store; bench.py's torch mirror of
    the formula (used to cut the cfg4 shards on the device) must agree record for record."""
    import torch
    import bench
    rng = np.random.default_rng(7)
    for n, world in [(5000, 8), (50_000, 8), (300, 2), (257, 3), (1000, 1)]:
        a = rng.integers(0, n, 20_000).astype(np.uint32)
        b = rng.integers(0, n, 20_000).astype(np.uint32)
        own = N.comm_link_owners(n, a, b, world)
        assert own.min() >= 0 and own.max() <= world - 1
        mirror = bench.link_owners(torch.from_numpy(a.astype(np.int32)), torch.from_numpy(b.astype(np.int32)), n, world)
        assert own.tolist() == mirror.tolist()
        nblk = (n + 255) // 256
        lo, hi = np.minimum(a, b) >> 8, np.maximum(a, b) >> 8
        t = lo.astype(np.int64) * nblk - lo.astype(np.int64) * (lo.astype(np.int64) - 1) // 2 + (hi - lo)
        order = np.argsort(t, kind="stable")
        assert (np.diff(own[order]) >= 0).all()                      # owners are monotone in the tile index
        for tile in np.unique(t)[:50]:
            assert len(set(own[t == tile].tolist())) == 1            # a tile is never split
    with pytest.raises(N.LxyError):
        N.comm_link_owners(10, [11], [0], 2)


def test_go_shim_binds_only_what_the_header_declares():
    """go/**/*.go (the cgo files a maintainer copies into the reference; no Go toolchain here to compile
    them): every C.lxy_* / C.LXY_* they use must exist in include/*.h."""
    hdr = open(os.path.join(ROOT, "include", "lxy_b200.h")).read() + open(os.path.join(ROOT, "include", "lxy_ga_plan.h")).read()
    declared = set(re.findall(r"\b(lxy_[a-z0-9_]+|LXY_[A-Z0-9_]+)\b", hdr))
    used = set()
    for dirpath, _, files in os.walk(os.path.join(ROOT, "go")):
        for f in files:
            if f.endswith(".go"):
                used |= set(re.findall(r"\bC\.((?:lxy|LXY)_[A-Za-z0-9_]+)", open(os.path.join(dirpath, f)).read()))
    assert used, "no cgo files found under go/"
    assert used <= declared, sorted(used - declared)
