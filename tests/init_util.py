"""Synthetic inputs for the Initializer scoring tests: undistorted keys of two frames related by a homography plus
noise and outliers, a match vector, and hypothesis sets around the true model (test infrastructure)."""
from pathlib import Path

import numpy as np

from oracle.orb_oracle import KP_DTYPE


def scene(seed, n1=600, n2=650, frac_matched=0.5, outliers=0.2, noise=0.7):
    r = np.random.default_rng(seed)
    k1 = np.zeros(n1, KP_DTYPE)
    k2 = np.zeros(n2, KP_DTYPE)
    k1["x"], k1["y"] = r.uniform(0, 640, n1).astype(np.float32), r.uniform(0, 480, n1).astype(np.float32)
    a = np.deg2rad(r.uniform(-4, 4))
    H = np.array([[np.cos(a), -np.sin(a), r.uniform(-15, 15)], [np.sin(a), np.cos(a), r.uniform(-15, 15)],
                  [r.uniform(-2e-5, 2e-5), r.uniform(-2e-5, 2e-5), 1.0]])
    k2["x"], k2["y"] = r.uniform(0, 640, n2).astype(np.float32), r.uniform(0, 480, n2).astype(np.float32)
    m12 = np.full(n1, -1, np.int32)
    pick = r.permutation(n1)[: int(frac_matched * min(n1, n2))]
    tgt = r.permutation(n2)[: len(pick)]
    for i, j in zip(pick, tgt):
        m12[i] = j
        if r.uniform() > outliers:
            p = H @ np.array([k1["x"][i], k1["y"][i], 1.0])
            k2["x"][j] = np.float32(p[0] / p[2] + r.normal(0, noise))
            k2["y"][j] = np.float32(p[1] / p[2] + r.normal(0, noise))
    return k1, k2, m12, H


def hypotheses(seed, H, n_hyp):
    """n_hyp homographies (and float32 inverses) and fundamental-like matrices around the true model; a few are
    degenerate on purpose (zero matrix, huge entries, NaN) to pin the IEEE corner cases."""
    r = np.random.default_rng(seed + 77)
    H21 = np.zeros((n_hyp, 3, 3), np.float32)
    H12 = np.zeros((n_hyp, 3, 3), np.float32)
    F21 = np.zeros((n_hyp, 3, 3), np.float32)
    for h in range(n_hyp):
        s = 0.0 if h == 0 else 10.0 ** r.uniform(-4, -0.5)   # hypothesis 0 is the true model
        Hh = H * (1 + r.normal(0, s, (3, 3)))
        Hh[:2, 2] += r.normal(0, 20 * s, 2)
        H21[h] = Hh.astype(np.float32)
        H12[h] = np.linalg.inv(H21[h].astype(np.float64)).astype(np.float32)
        t = r.normal(0, 1, 3)
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        Kinv = np.linalg.inv(np.array([[500.0, 0, 320], [0, 500, 240], [0, 0, 1]]))
        F21[h] = (Kinv.T @ tx @ (np.eye(3) + r.normal(0, s, (3, 3))) @ Kinv).astype(np.float32)
    if n_hyp >= 8:
        H21[3] = 0; F21[3] = 0
        H21[5] *= np.float32(1e30); F21[5] *= np.float32(1e30)
        H12[6, 2, 2] = np.nan; F21[6, 0, 0] = np.nan
        H21[7] = H21[0]; H12[7] = H12[0]; F21[7] = F21[0]      # an exact tie: the first one must win
    return H21.reshape(n_hyp, 9), H12.reshape(n_hyp, 9), F21.reshape(n_hyp, 9)


def same_scores(a, b):
    """Bit-equal float32 scores; a NaN score (degenerate hypothesis) must be NaN on both sides -- the payload and sign of a
    generated NaN are an x86 / PTX convention (0xFFC00000 vs 0x7FFFFFFF), not something the reference's logic can see."""
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na].view(np.uint32), b[~nb].view(np.uint32))


def golden_scoring():
    return np.load(Path(__file__).resolve().parent / "golden" / "init_scoring_demo.npz", allow_pickle=True)
