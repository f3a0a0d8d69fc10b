"""Loader for tests/golden/*.npz (reference outputs written by tools/gen_golden.py)."""
from pathlib import Path

import cv2
import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden"
LIGHT = ("kitti_1241x376_nf2000", "uhd_3840x2160_nf8000")
OLD_POLICY = "old_policy"   # ComputeKeyPointsOld outputs (CRC32), tools/gen_golden.py:case_old_policy  # outputs only: the synthetic inputs are regenerated at test time
CASES = sorted(p.stem for p in GOLDEN.glob("*.npz") if not p.stem.startswith("init_scoring") and p.stem not in LIGHT and p.stem != OLD_POLICY)  # the Initializer fixture has its own tests


class LightCase:
    """BASELINE.json configs[3] / configs[4]: reference outputs (counts, CRC32 of the records, vnMatches12) of a synthetic pair."""

    def __init__(self, name):
        import zlib
        from orb_slam_tracking_b200 import synth
        z = np.load(GOLDEN / (name + ".npz"))
        self.z, self.W, self.H, self.nf = z, int(z["W"]), int(z["H"]), int(z["nfeatures"])
        self.windows = [int(w) for w in z["windows"]]
        self.imgs = synth.frame_pair(int(z["pair"]), self.W, self.H)
        assert [zlib.crc32(im.tobytes()) for im in self.imgs] == [int(v) for v in z["img_crc"]], "synthetic inputs differ from the fixture's"
        self.K4, self.D4 = synth.camera_for(self.W, self.H)

    def check_frame(self, i, kps, desc, kps_un):
        import zlib
        z = self.z
        assert len(kps) == int(z["n%d" % i])
        assert zlib.crc32(np.ascontiguousarray(kps).tobytes()) == int(z["kps_crc%d" % i]), "keypoints of frame %d" % i
        assert zlib.crc32(np.ascontiguousarray(desc).tobytes()) == int(z["desc_crc%d" % i]), "descriptors of frame %d" % i
        assert zlib.crc32(np.ascontiguousarray(kps_un).tobytes()) == int(z["kun_crc%d" % i]), "undistorted keys of frame %d" % i

    def matches(self, w):
        return int(self.z["nmatches_w%d" % w]), self.z["matches_w%d" % w]



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


def old_policy_cases():
    """-> [(name, image, nfeatures, iniTh, minTh, n, kps_crc, desc_crc)] of the ComputeKeyPointsOld fixture."""
    from orb_slam_tracking_b200 import synth
    z = np.load(GOLDEN / (OLD_POLICY + ".npz"))
    zd = np.load(GOLDEN / "dbow_640x480_nf1000.npz")
    imgs = {"dbow%d" % i: cv2.imdecode(zd["png%d" % i], cv2.IMREAD_GRAYSCALE) for i in range(4)}
    pr = synth.frame_pair(3, 640, 480)
    imgs.update({"synth0": pr[0], "synth1": pr[1], "synth752_th0": synth.frame_pair(1, 752, 480)[0], "kitti": synth.frame_pair(1, 1241, 376)[0]})
    out = []
    for name, im in imgs.items():
        w, h, nf, ini, mn = (int(v) for v in z[name + "_cfg"])
        assert im.shape == (h, w)
        out.append((name, im, nf, ini, mn, int(z[name + "_n"]), int(z[name + "_kps_crc"]), int(z[name + "_desc_crc"])))
    return out
