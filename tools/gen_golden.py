#!/usr/bin/env python
"""Generates tests/golden/*.npz: outputs of the REFERENCE ITSELF (oracle/_ref/libref_orb.so = the
reference's own ORBextractor.cpp / Frame.cpp / ORBmatcher.cpp / FORB.cpp compiled in place, see
oracle/build_ref.sh) on the reference's bundled images and on a synthetic pair.

Run in the build container only (needs /root/reference):  python tools/gen_golden.py
Each file holds the input images (PNG-encoded, lossless), the configuration, and per frame the
keypoints (cv::KeyPoint records), descriptors, undistorted keypoints, the 64x48 grid and per pair
vnMatches12 + the returned count.
"""
import glob
import sys
from pathlib import Path

import cv2
import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import ref_orb as R  # noqa: E402
from orb_slam_tracking_b200 import synth  # noqa: E402

OUT = ROOT / "tests" / "golden"
REF = Path("/root/reference")


def case(name, imgs, nf, ini, mn, K4, D4, windows=(100,), lapping=(0, 0), source=""):
    E = R.Extractor(nf, 1.2, 8, ini, mn)
    d = {"nfeatures": nf, "scale": np.float32(1.2), "nlevels": 8, "ini_th": ini, "min_th": mn, "K4": K4, "D4": D4,
         "n_frames": len(imgs), "windows": np.array(windows, np.int32), "lapping": np.array(lapping, np.int32),
         "source": source}
    sc, isc, s2, is2, q = E.tables()
    d.update(tab_scale=sc, tab_inv_scale=isc, tab_sigma2=s2, tab_inv_sigma2=is2, tab_per_level=q)
    F = []
    for i, im in enumerate(imgs):
        ok, png = cv2.imencode(".png", im, [cv2.IMWRITE_PNG_COMPRESSION, 9])
        assert ok
        d["png%d" % i] = png
        f = R.Frame(E, im, K4, D4, first=(i == 0))
        k, ku, ds = f.get()
        off, idx = f.grid()
        d.update({"kps%d" % i: k, "kps_un%d" % i: ku, "desc%d" % i: ds, "grid_off%d" % i: off, "grid_idx%d" % i: idx})
        if i == 0:
            d["bounds"] = f.bounds
            for l in (1, 7):
                d["level%d_crc" % l] = np.uint32(cv2.norm(E.level(l, True), cv2.NORM_L1))  # cheap checksum of the padded level
        if tuple(lapping) != (0, 0):
            k2, d2, mono = E(im, lapping)
            d.update({"lap_kps%d" % i: k2, "lap_desc%d" % i: d2, "lap_mono%d" % i: mono})
        F.append(f)
    for a in range(0, len(imgs) - 1, 2):
        for w in windows:
            n, m = R.search_for_initialization(F[a], F[a + 1], 0.9, True, w)
            d["matches_%d_w%d" % (a, w)] = m
            d["nmatches_%d_w%d" % (a, w)] = n
            print("  %s pair %d window %d: N = %d, %d; nmatches = %d" % (name, a, w, F[a].n, F[a + 1].n, n))
    OUT.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(OUT / (name + ".npz"), **d)
    print("%s: %.0f KB" % (name, (OUT / (name + ".npz")).stat().st_size / 1024))


def case_light(name, W, H, nf, pair, windows):
    """Large synthetic shapes: the inputs are regenerated from orb_slam_tracking_b200.synth at test time (deterministic
    numpy), so the fixture holds only the reference's outputs: keypoint counts, CRC32 of the keypoint records / descriptors /
    undistorted keys, and vnMatches12 + the returned count per window."""
    import zlib
    K4, D4 = synth.camera_for(W, H)
    E = R.Extractor(nf, 1.2, 8, 20, 7)
    imgs = synth.frame_pair(pair, W, H)
    d = {"W": W, "H": H, "nfeatures": nf, "pair": pair, "ini_th": 20, "min_th": 7, "windows": np.array(windows, np.int32),
         "img_crc": np.array([zlib.crc32(im.tobytes()) for im in imgs], np.uint32),
         "source": "orb_slam_tracking_b200.synth.frame_pair(%d, %d, %d) through oracle/_ref (the reference's own sources)" % (pair, W, H)}
    F = []
    for i, im in enumerate(imgs):
        f = R.Frame(E, im, K4, D4, first=(i == 0))
        k, ku, ds = f.get()
        d.update({"n%d" % i: len(k), "kps_crc%d" % i: np.uint32(zlib.crc32(k.tobytes())), "desc_crc%d" % i: np.uint32(zlib.crc32(ds.tobytes())),
                  "kun_crc%d" % i: np.uint32(zlib.crc32(ku.tobytes()))})
        F.append(f)
    for w in windows:
        n, m = R.search_for_initialization(F[0], F[1], 0.9, True, w)
        d["matches_w%d" % w] = m
        d["nmatches_w%d" % w] = n
        print("  %s window %d: N = %d, %d; nmatches = %d" % (name, w, F[0].n, F[1].n, n))
    np.savez_compressed(OUT / (name + ".npz"), **d)
    print("%s: %.0f KB" % (name, (OUT / (name + ".npz")).stat().st_size / 1024))


def case_old_policy(dbow):
    """ComputeKeyPointsOld through the reference's own function body (oracle/build_ref.sh swaps the one call
    in a generated copy): keypoint counts and CRC32 of the keypoint records / descriptors for the four DBoW2 images (their PNGs
    are in dbow_640x480_nf1000.npz) and two synthetic frames (regenerated at test time)."""
    import zlib
    d = {"source": "R.OldPolicyExtractor: Features/ORBextractor.cpp with operator() calling ComputeKeyPointsOld (:1207-1496)"}
    cases = [("dbow%d" % i, im, 1000, 20, 7) for i, im in enumerate(dbow)]
    cases += [("synth%d" % i, im, 1000, 20, 7) for i, im in enumerate(synth.frame_pair(3, 640, 480))]
    cases += [("synth752_th0", synth.frame_pair(1, 752, 480)[0], 2000, 0, 0), ("kitti", synth.frame_pair(1, 1241, 376)[0], 2000, 20, 7)]
    for name, im, nf, ini, mn in cases:
        k, ds, mono = R.OldPolicyExtractor(nf, 1.2, 8, ini, mn)(im)
        d.update({name + "_n": len(k), name + "_kps_crc": np.uint32(zlib.crc32(k.tobytes())), name + "_desc_crc": np.uint32(zlib.crc32(ds.tobytes())),
                  name + "_cfg": np.array([im.shape[1], im.shape[0], nf, ini, mn], np.int32)})
        print("  old policy %s: %d keypoints" % (name, len(k)))
    np.savez_compressed(OUT / "old_policy.npz", **d)


def main():
    assert R.available() or R.build().exists(), "needs /root/reference to build oracle/_ref"
    demo = [cv2.cvtColor(cv2.imread(p), cv2.COLOR_BGR2GRAY) for p in sorted(glob.glob(str(REF / "demo/initImages/*.png")))]
    dbow = [cv2.imread(p, 0) for p in sorted(glob.glob(str(REF / "Thirdparty/DBoW2/demo/images/*.png")))]
    assert len(demo) == 2 and len(dbow) == 4
    # configs[0]: the demo pair as demo_initialization runs it (2 x nFeatures, thresholds parsed as 0/0: SURVEY.md D5, D6)
    case("demo_pair_th0", demo, 2000, 0, 0, synth.DEMO_K4, synth.DEMO_D4, windows=(100,),
         source="demo/initImages/*.png -> cv2 BGR2GRAY; Settings.yaml camera; demo_initialization.cpp:72-77,105-108")
    case("demo_pair_th20_7", demo, 2000, 20, 7, synth.DEMO_K4, synth.DEMO_D4, windows=(100, 30), lapping=(200, 500),
         source="demo/initImages/*.png -> cv2 BGR2GRAY; thresholds 20/7")
    K, D = synth.camera_for(640, 480)
    case("dbow_640x480_nf1000", dbow, 1000, 20, 7, K, D, windows=(100,),
         source="Thirdparty/DBoW2/demo/images/image{0..3}.png (640x480 gray); pairs (0,1), (2,3)")
    case("dbow_640x480_nf1250", dbow[:2], 1250, 20, 7, K, D, windows=(100,),
         source="Thirdparty/DBoW2/demo/images/image{0,1}.png; demo_ORBextractor.cpp:20 parameters")
    case("synth_640x480_nf1000", list(synth.frames(2, 640, 480, first_pair=3)), 1000, 20, 7, K, D, windows=(100, 640),
         source="orb_slam_tracking_b200.synth.frames(2, 640, 480, first_pair=3); window 640 = every level-0 pair")
    case_old_policy(dbow)
    # BASELINE.json configs[3] and configs[4]: outputs only (the inputs are regenerated from synth)
    case_light("kitti_1241x376_nf2000", 1241, 376, 2000, 4, (100, 8192))
    case_light("uhd_3840x2160_nf8000", 3840, 2160, 8000, 2, (100, 8192))


if __name__ == "__main__":
    main()
