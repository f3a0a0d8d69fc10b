"""Ad-hoc fuzz of the single-frame paths: Frame::Frame in one call (orbx_frame_construct, CUDA graph) with random lapping areas and
a camera change in between (graph rebuilds), against orbx_extract + orbx_frame_glue and against the oracle; strided input rows."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import test_gpu_fuzz as T
from oracle import orb_oracle as orc
from orb_slam_tracking_b200 import frontend as fe, synth
from orb_slam_tracking_b200.capi import OrbxError
N, seed = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(seed)
bad = ran = 0
for cfg in T._configs(N, seed):
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    try:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H)
    except OrbxError:
        continue
    R = orc.Extractor(nf, sf, nl, ini, mn)
    ims = synth.frames(2, W, H, first_pair=900 + idx)
    why = []
    for rep in range(3):
        K4 = np.array([0.8 * W + rep, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25 * rep], np.float32)
        D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if (dist ^ (rep == 1)) else [0, 0, 0, 0], np.float32)
        b = E.set_camera(K4, D4)
        lap = (0, 0) if rep == 0 else (int(rng.integers(0, W // 2)), int(rng.integers(W // 2, W + 50)))
        for j, im in enumerate(ims):
            if rep == 2 and j == 1:   # strided rows: a view into a wider buffer
                big = np.zeros((H, W + 13), np.uint8); big[:, :W] = im; im_in = big[:, :W]
            else:
                im_in = im
            rk, rd, rm = R(im, lapping=lap)
            Fo = orc.Frame(rk, rd, K4, D4, b)
            for fused in (True, False):
                F = fe.Frame(E, im_in, fused=fused, vLappingArea=lap) if fused else None
                if F is None:
                    k, d, m = E(im_in, vLappingArea=lap)
                    g = len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd) and m == rm
                else:
                    off, gi = Fo.grid()
                    g = (F.N == Fo.N and np.array_equal(F.mvKeys.view(np.uint8), rk.view(np.uint8)) and np.array_equal(F.mDescriptors, rd)
                         and np.array_equal(F.mvKeysUn.view(np.uint8), Fo.keys_un().view(np.uint8)) and np.array_equal(F.grid_offsets, off)
                         and np.array_equal(F.grid_indices, gi))
                if not g: why.append((rep, j, fused, lap))
    ran += 1
    if why:
        bad += 1; print("MISMATCH", cfg, why[:4])
    E.close()
print("configs", N, "ran", ran, "mismatches", bad)
