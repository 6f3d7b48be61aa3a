"""Randomised parity sweep of the scaffolding scorers (tools/stress_scaff.py): random contig counts, record
counts down to a single record, weight modes, population sizes and context flags, each case against the
oracle's exact integers (scaff.go:246-310 restated; a third as many G-profile cases on top) — the record rings of the kernels see partial slices,
warps without packets, one-record tiles and shares that end inside a packet."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("seed", [11, 12])
def test_random_shapes_against_the_oracle(seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "stress_scaff.py"), "60", str(seed)],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "80 cases, 0 failures" in r.stdout
