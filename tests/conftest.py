import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_built():
    """Build the C restatement (and oracle/_ref when /root/reference is present)."""
    from oracle import run as R
    R.build()
    assert os.path.exists(R.PORT)
    return R


def golden_names():
    d = os.path.join(ROOT, "tests", "golden")
    return sorted(f[:-5] for f in os.listdir(d) if f.endswith(".json") and not f.startswith("wide_"))


def wide_golden_names():
    """goldens over alphabets of 5..7 letters (3-bit Occ layout, level-synchronous engine only)"""
    d = os.path.join(ROOT, "tests", "golden")
    return sorted(f[:-5] for f in os.listdir(d) if f.endswith(".json") and f.startswith("wide_"))


def load_golden(name):
    import json
    with open(os.path.join(ROOT, "tests", "golden", name + ".json")) as f:
        g = json.load(f)
    g["patterns"] = [p.encode() for p in g["patterns"]]
    g["rows_reference_order"] = [tuple(r) for r in g["rows_reference_order"]]
    return g
