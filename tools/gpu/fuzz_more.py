"""Ad-hoc wider fuzz: N random configurations (tests/test_gpu_fuzz._configs with another seed), extraction + matching against the
oracle, each also with the quadtree's tight node table (squeeze path) and through a batch of 4 (k_resize path instead of the chain)."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import test_gpu_fuzz as T
from oracle import orb_oracle as orc
from orb_slam_tracking_b200 import frontend as fe, synth
from orb_slam_tracking_b200.capi import OrbxError
N, seed = int(sys.argv[1]), int(sys.argv[2])
BIG = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0   # scales the shapes (and triples the feature counts): large levels, quotas >= 1000
bad = 0
for cfg in T._configs(N, seed):
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    if BIG != 1.0: W, H, nf = int(W * BIG), int(H * BIG), nf * 3
    R = orc.Extractor(nf, sf, nl, ini, mn)
    try:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H, max_batch=4)
    except OrbxError as e:
        print("skip", cfg[:7], e); continue
    K4 = np.array([0.8 * W, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25], np.float32)
    D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if dist else [0, 0, 0, 0], np.float32)
    b = E.set_camera(K4, D4)
    ims = synth.frame_pair(300 + idx, W, H)
    ok = True
    ref = [R(im) for im in ims]
    for tight in (0, 1):
        E.set_option("qt_tight", tight)
        for im, (rk, rd, rm) in zip(ims, ref):
            k, d, m = E(im)
            ok &= len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd)
    out = E.extract_batch([ims[0], ims[1], ims[0], ims[1]])
    for j in range(4):
        rk, rd, _ = ref[j & 1]
        kb, db, _ = out[j]
        ok &= len(kb) == len(rk) and np.array_equal(kb.view(np.uint8), rk.view(np.uint8)) and np.array_equal(db, rd)
    G = [fe.Frame(E, im) for im in ims]
    O = [orc.Frame(rk, rd, K4, D4, b) for rk, rd, _ in ref]
    for nnr, ori, win in ((0.9, True, 100), (0.8, True, 30), (0.9, False, 4096)):
        n, m12, cnt = fe.ORBmatcher(nnr, ori).SearchForInitialization(G[0], G[1], win)
        rn, rm12, rcnt = orc.search_for_initialization(O[0], O[1], nnr, ori, win)
        ok &= n == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt)
    if not ok:
        bad += 1; print("MISMATCH", cfg)
    E.close()
print("configs", N, "mismatches", bad)
