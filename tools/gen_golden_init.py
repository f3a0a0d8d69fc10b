#!/usr/bin/env python
"""Generates tests/golden/init_scoring_demo.npz: outputs of the reference's own Initializer::CheckHomography /
CheckFundamental bodies (oracle/_ref, cut from Initialization/Initializer.cpp:268-438 by oracle/build_ref.sh) on the
reference's demo pair, i.e. the scoring step of demo_initialization (demo/demo_initialization.cpp:105-121).

Inputs: the undistorted keys and vnMatches12 of tests/golden/demo_pair_th0.npz (reference outputs themselves); hypothesis 0
of each model is what FindHomographyCV / FindFundamentalCV feed to the scoring (Initializer.cpp:127-167):
cv::findHomography(RANSAC, 2, mask, 10000, 0.999) and cv::findFundamentalMat(FM_RANSAC, 1.95, 0.999) from the cv2 4.13.0
wheel; the other hypotheses are perturbations of them plus degenerate matrices.  Run in the build container only.
"""
import sys
from pathlib import Path

import cv2
import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import ref_orb as R  # noqa: E402

OUT = ROOT / "tests" / "golden"


def main():
    assert R.available() or R.build().exists(), "needs /root/reference to build oracle/_ref"
    g = np.load(OUT / "demo_pair_th0.npz", allow_pickle=True)
    k1, k2, m12 = g["kps_un0"], g["kps_un1"], g["matches_0_w100"].astype(np.int32)
    i1 = np.nonzero(m12 >= 0)[0]
    src = np.stack([k1["x"][i1], k1["y"][i1]], 1).astype(np.float32)
    dst = np.stack([k2["x"][m12[i1]], k2["y"][m12[i1]]], 1).astype(np.float32)
    cv2.setRNGSeed(0)
    H, _ = cv2.findHomography(src, dst, cv2.RANSAC, 2, None, 10000, 0.999)
    F, _ = cv2.findFundamentalMat(src, dst, cv2.FM_RANSAC, 1.95, 0.999)
    nh = 64
    rng = np.random.default_rng(11)
    H21 = np.zeros((nh, 3, 3), np.float32); H12 = np.zeros_like(H21); F21 = np.zeros_like(H21)
    for h in range(nh):
        s = 0.0 if h == 0 else 10.0 ** rng.uniform(-5, -1)
        Hh = H * (1 + rng.normal(0, s, (3, 3)))
        H21[h] = Hh.astype(np.float32)                      # Converter::toMatrix3f: double -> float
        H12[h] = np.linalg.inv(H21[h].astype(np.float64)).astype(np.float32)
        F21[h] = (F * (1 + rng.normal(0, s, (3, 3)))).astype(np.float32)
    H21[3] = 0; F21[3] = 0
    H21[5] *= np.float32(1e30); F21[5] *= np.float32(1e30)
    H21[7] = H21[0]; H12[7] = H12[0]; F21[7] = F21[0]
    d = {"source": "tests/golden/demo_pair_th0.npz (kps_un0, kps_un1, matches_0_w100); cv2 %s" % cv2.__version__,
         "kps_un0": k1, "kps_un1": k2, "matches12": m12, "H21": H21.reshape(nh, 9), "H12": H12.reshape(nh, 9),
         "F21": F21.reshape(nh, 9)}
    with np.errstate(all="ignore"):
        for sigma in (1.0, 0.5):
            sH, iH = R.check_homography(k1, k2, m12, d["H21"], d["H12"], sigma)
            sF, iF = R.check_fundamental(k1, k2, m12, d["F21"], sigma)
            tag = "s%d" % int(sigma * 10)
            d.update({"scoresH_" + tag: sH, "inliersH_" + tag: np.packbits(iH, axis=1), "scoresF_" + tag: sF,
                      "inliersF_" + tag: np.packbits(iF, axis=1)})
            print("sigma %.1f: N = %d; H (cv) score %.3f, %d inliers; F (cv) score %.3f, %d inliers; RH = %.3f" % (
                sigma, iH.shape[1], sH[0], iH[0].sum(), sF[0], iF[0].sum(), sH[0] / (sH[0] + sF[0])))
    np.savez_compressed(OUT / "init_scoring_demo.npz", **d)
    print("init_scoring_demo: %.0f KB" % ((OUT / "init_scoring_demo.npz").stat().st_size / 1024))


if __name__ == "__main__":
    main()
