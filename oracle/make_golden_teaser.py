"""Generates tests/golden/teaser_golden.npz from the reference's OWN test fixture for the scale-consistency inliers:
TEASER-plusplus/test/teaser/data/registration_test/{objectIn,sceneIn,fixed_scale_inliers}.csv, the data of
RegistrationTest.SolveRegistrationProblemDecoupled "Registration problem 2 / Fixed scale"
(test/teaser/registration-test.cc:300-362: noise_bound = 0.0067364, cbar2 = 1, estimate_scaling = false — the
configuration scripts/test_teaser.py uses).  The expected mask in that file enumerates all ORDERED pairs i != j
(N(N-1) entries, row-major off-diagonal; the C++ test's comparison is commented out for that reason), so it is stored
as the symmetric [N,N] matrix it describes.  Build container only (reads /root/reference)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/TEASER-plusplus/test/teaser/data/registration_test"


def main():
    src = np.loadtxt(os.path.join(SRC, "objectIn.csv"), delimiter=",")
    dst = np.loadtxt(os.path.join(SRC, "sceneIn.csv"), delimiter=",")
    exp = np.loadtxt(os.path.join(SRC, "fixed_scale_inliers.csv"), delimiter=",").astype(bool).ravel()
    n = src.shape[1]
    assert src.shape == dst.shape == (3, n) and exp.size == n * (n - 1)
    full = np.zeros((n, n), bool)
    full[~np.eye(n, dtype=bool)] = exp
    assert (full == full.T).all()
    path = os.path.join(ROOT, "tests", "golden", "teaser_golden.npz")
    np.savez_compressed(path, src=src, dst=dst, inlier_matrix=full, noise_bound=np.float64(0.0067364), cbar2=np.float64(1.0))
    print("wrote", path, os.path.getsize(path), "bytes; N =", n, "inlier pairs =", int(full.sum()) // 2)


if __name__ == "__main__":
    main()
