"""Generates tests/golden/lstsq_golden.npz by running the reference's own `utils.C_from_sparse_P`
(utils/utils.py:67-80, the unmodified module loaded by oracle/ref_loader.load_reference_utils) on the seeded cases of
tests/golden/lstsq_cases.py, called the way scripts/eval.py:74 calls it (int32 P, column-sliced bases).
Build container only."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from golden.lstsq_cases import LSQ_CASES, build_lstsq_inputs  # noqa: E402
from oracle.ref_loader import load_reference_utils  # noqa: E402


def main():
    utils = load_reference_utils()
    out = {}
    for name in LSQ_CASES:
        P, e1, e2, k = build_lstsq_inputs(name)
        C = utils.C_from_sparse_P(torch.from_numpy(P).type(torch.int), torch.from_numpy(e1)[:, :k], torch.from_numpy(e2)[:, :k])
        assert C.dtype == torch.float32 and tuple(C.shape) == (k, k)
        out[name + "/C"] = C.numpy()
        out[name + "/in_sum"] = np.float64(e1.astype(np.float64).sum() + e2.astype(np.float64).sum() + P.sum())
        print(f"{name:28s} -> C {tuple(C.shape)} |C|max={float(C.abs().max()):.4g}")
    path = os.path.join(ROOT, "tests", "golden", "lstsq_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
