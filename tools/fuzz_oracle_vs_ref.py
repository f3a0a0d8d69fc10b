"""CPU only (needs /root/reference: oracle/_ref): the oracle's restatement against the reference's own sources compiled in place on RANDOM
configurations -- extraction with a random lapping area (both keypoint policies), the Frame constructor (undistorted keys, bounds, grid),
GetFeaturesInArea and SearchForInitialization.  The committed tests pin the oracle on a handful of shapes; this widens that."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import test_gpu_fuzz as T
from oracle import orb_oracle as orc, ref_orb
from orb_slam_tracking_b200 import synth
N, seed = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(seed)
bad = ran = refused = 0
for cfg in T._configs(N, seed):
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    if W * H > 600 * 420: W, H = W // 2 + 90, H // 2 + 70
    nf = min(nf, 1200)
    ims = synth.frames(2, W, H, first_pair=1200 + idx)
    K4 = np.array([0.8 * W, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25], np.float32)
    D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if dist else [0, 0, 0, 0], np.float32)
    lap = (0, 0) if idx % 3 == 0 else (int(rng.integers(0, W // 2)), int(rng.integers(W // 2, W + 40)))
    why = []
    O = orc.Extractor(nf, sf, nl, ini, mn)
    try:
        ok = [O(im, lapping=lap) for im in ims]
    except RuntimeError as e:
        refused += 1; continue          # aspect ratio the reference cannot handle (it indexes an empty vector there)
    Rf = ref_orb.Extractor(nf, sf, nl, ini, mn)
    for j, im in enumerate(ims):
        rk, rd, rm = Rf(im, lapping=lap)
        k, d, m = ok[j]
        if not (len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd) and m == rm): why.append(("extract", j, lap))
    # Frame constructor + queries + matcher (the reference's Frame extracts with lapping {0, 0})
    fo, fr = [], []
    for j, im in enumerate(ims):
        F = ref_orb.Frame(Rf, im, K4, D4, first=(j == 0)); fr.append(F)
        k, d, _ = O(im)
        b = orc.image_bounds(W, H, K4, D4)
        if not np.array_equal(b, F.bounds): why.append(("bounds", b, F.bounds))
        G = orc.Frame(k, d, K4, D4, b); fo.append(G)
        rk, rku, rd = F.get()
        off, gi = G.grid(); roff, rgi = F.grid()
        if not (np.array_equal(G.keys_un().view(np.uint8), rku.view(np.uint8)) and np.array_equal(off, roff) and np.array_equal(gi, rgi)): why.append(("frame", j))
    for _ in range(12):
        x, y, r = float(rng.uniform(-50, W + 50)), float(rng.uniform(-50, H + 50)), float(rng.uniform(1, 250))
        l0, l1 = int(rng.integers(-1, nl)), int(rng.integers(-1, nl + 1))
        if not np.array_equal(fo[1].features_in_area(x, y, r, l0, l1), fr[1].features_in_area(x, y, r, l0, l1)): why.append(("area", x, y, r, l0, l1))
    for win in (100, int(rng.integers(1, 80)), 4096):
        for ratio, ori in ((0.9, True), (0.6, False)):
            n, m12, _ = orc.search_for_initialization(fo[0], fo[1], ratio, ori, win)
            rn, rm12 = ref_orb.search_for_initialization(fr[0], fr[1], ratio, ori, win)
            if not (n == rn and np.array_equal(m12, rm12)): why.append(("match", win, ratio, ori, n, rn))
    # ComputeKeyPointsOld
    O1 = orc.Extractor(nf, sf, nl, ini, mn, policy=1)
    try:
        o1 = [O1(im, lapping=lap) for im in ims]
        R1 = ref_orb.OldPolicyExtractor(nf, sf, nl, ini, mn)
        for j, im in enumerate(ims):
            rk, rd, rm = R1(im, lapping=lap)
            k, d, m = o1[j]
            if not (len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd) and m == rm): why.append(("old policy", j, lap))
    except RuntimeError:
        pass   # outside the function's working range (-7, -8): the reference divides by zero or reads outside its images there
    ran += 1
    if why:
        bad += 1; print("MISMATCH", (W, H, nf, sf, nl, ini, mn, dist, idx), why[:4])
print("configs", N, "ran", ran, "refused", refused, "mismatches", bad)
