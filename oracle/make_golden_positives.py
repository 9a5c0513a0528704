"""Generates tests/golden/positives_golden.npz by running the reference's own `find_positives` source
(dataset/object.py:281-288, compiled from the unmodified file by oracle/ref_loader.load_reference_method; the module
itself cannot be imported here: open3d / robust_laplacian are absent).  Build container only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from golden.positives_cases import POS_CASES, build_positives_inputs  # noqa: E402
from oracle.ref_loader import load_reference_method  # noqa: E402


def main():
    fp = load_reference_method("dataset/object.py", "base_object_dataset", "find_positives")
    out = {}
    for name in POS_CASES:
        a, b, r = build_positives_inputs(name)
        p = fp(None, a, b, r=r)
        assert p.dtype == np.int64 and p.ndim == 2 and p.shape[1] == 2
        out[name + "/P"] = p
        out[name + "/in_sum"] = np.float64(a.astype(np.float64).sum() + b.astype(np.float64).sum())
        print(f"{name:24s} -> P {p.shape}")
    path = os.path.join(ROOT, "tests", "golden", "positives_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
