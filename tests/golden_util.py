"""Loader for tests/golden/*.npz (reference outputs written by tools/gen_golden.py)."""
from pathlib import Path

import cv2
import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden"
CASES = sorted(p.stem for p in GOLDEN.glob("*.npz") if not p.stem.startswith("init_scoring"))  # front-end cases; the Initializer fixture has its own tests


class Case:
    def __init__(self, name):
        z = np.load(GOLDEN / (name + ".npz"))
        self.z = z
        self.name = name
        self.nf, self.ini, self.mn = int(z["nfeatures"]), int(z["ini_th"]), int(z["min_th"])
        self.K4, self.D4 = z["K4"], z["D4"]
        self.n_frames = int(z["n_frames"])
        self.windows = [int(w) for w in z["windows"]]
        self.lapping = tuple(int(v) for v in z["lapping"])
        self.bounds = z["bounds"]
        self.imgs = [cv2.imdecode(z["png%d" % i], cv2.IMREAD_GRAYSCALE) for i in range(self.n_frames)]

    def frame(self, i):
        z = self.z
        return z["kps%d" % i], z["desc%d" % i], z["kps_un%d" % i], z["grid_off%d" % i], z["grid_idx%d" % i]

    def pairs(self):
        return range(0, self.n_frames - 1, 2)

    def matches(self, a, w):
        return int(self.z["nmatches_%d_w%d" % (a, w)]), self.z["matches_%d_w%d" % (a, w)]


def same_records(a, b):
    return len(a) == len(b) and np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))
