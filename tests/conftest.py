import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_PARENT = os.path.join(ROOT, "6d-pose-estimation-for-unseen-categories_b200")
for p in (ROOT, PKG_PARENT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    path = os.path.join(ROOT, "tests", "golden", "f2p_golden.npz")
    return np.load(path, allow_pickle=False)


def load_case(golden, name):
    """Rebuilds a golden case's inputs; skips when this torch build's RNG no longer reproduces them."""
    from golden.cases import CASES, build_inputs, inputs_digest
    c = CASES[name]
    diam_override = float(golden[name + "/diam"]) if c.get("diam_scale") == "fallback" else None
    inp = build_inputs(name, diam_override)
    if inputs_digest(inp) != str(golden[name + "/digest"]):
        pytest.skip(f"{name}: synthetic inputs differ from the ones the golden outputs were made from (torch RNG changed)")
    return c, inp
