import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "fixtures.json")) as f:
        return json.load(f)


@pytest.fixture()
def golden_files(golden, tmp_path):
    """The reference's fixture files, re-materialised from fixtures.json."""
    paths = {}
    for key, entry in golden.items():
        if not isinstance(entry, dict):
            continue
        p = tmp_path / (key + ".links")
        p.write_text(entry["text"])
        paths[key] = str(p)
    return paths
