"""Ad-hoc fuzz of the Frame glue with RANDOM cameras (mild to far outside what a calibration would give: the radial factor of
cv::undistortPoints turns negative inside the image for many of them): image bounds, undistorted keys, 64x48 grid, radius queries and
SearchForInitialization against the oracle.  The keypoints of each size are extracted once.  python tools/gpu/fuzz_camera.py <N> <seed>"""
import sys, numpy as np
sys.path.insert(0, '.')
from oracle import orb_oracle as orc
from orb_slam_tracking_b200 import frontend as fe, synth
from orb_slam_tracking_b200.capi import OrbxError
N, seed = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(seed)
bad = ran = refused = 0
for W, H, nf in ((640, 480, 1000), (752, 480, 1500), (320, 240, 500)):
    E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H)
    ims = synth.frames(2, W, H, first_pair=1200 + W)
    kd = [E(im)[:2] for im in ims]
    for it in range(N):
        wild = it % 3 == 2
        K4 = np.array([rng.uniform(0.3, 1.5) * W, rng.uniform(0.3, 1.5) * W, rng.uniform(0.2, 0.8) * W, rng.uniform(0.2, 0.8) * H], np.float32)
        D4 = np.array([rng.uniform(-0.8, 0.4), rng.uniform(-0.3, 0.4), rng.uniform(-5e-3, 5e-3), rng.uniform(-5e-3, 5e-3)], np.float32)
        if it % 5 == 0: D4[:] = 0
        if wild: D4 *= 4
        try:
            b = E.set_camera(K4, D4)
        except OrbxError as e:
            refused += 1; continue
        ob = orc.image_bounds(W, H, K4, D4)
        why = []
        if not np.array_equal(b, ob): why.append(("bounds", b.tolist(), ob.tolist()))
        G, O = [], []
        try:
            for k, d in kd:
                G.append(fe.Frame(E, kps=k, desc=d)); O.append(orc.Frame(k, d, K4, D4, ob))
        except (OrbxError, RuntimeError) as e:
            why.append(("frame", repr(e)))
        if len(G) == 2 and len(O) == 2:
            for j, (g, o) in enumerate(zip(G, O)):
                off, gi = o.grid()
                if not (np.array_equal(g.mvKeysUn.view(np.uint8), o.keys_un().view(np.uint8)) and np.array_equal(g.grid_offsets, off) and np.array_equal(g.grid_indices, gi)):
                    why.append(("glue", j))
                for _ in range(3):
                    x, y, r = float(rng.uniform(-20, W + 20)), float(rng.uniform(-20, H + 20)), float(rng.uniform(1, 150))
                    if not np.array_equal(g.GetFeaturesInArea(x, y, r), o.features_in_area(x, y, r)): why.append(("area", x, y, r))
            win = int(rng.choice([10, 50, 100, 400]))
            n, m12, cnt = fe.ORBmatcher(0.9, True).SearchForInitialization(G[0], G[1], win)
            rn, rm12, rcnt = orc.search_for_initialization(O[0], O[1], 0.9, True, win)
            if not (n == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt)): why.append(("match", win, n, rn))
        ran += 1
        if why:
            bad += 1; print("MISMATCH", (W, H), K4.tolist(), D4.tolist(), why[:3])
    E.close()
print("cameras", 3 * N, "ran", ran, "refused", refused, "mismatches", bad)
