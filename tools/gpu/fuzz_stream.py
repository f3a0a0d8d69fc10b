"""Ad-hoc fuzz of the batched paths: N random configurations; a batch of 2 * P distinct frames (P random, 33..70 pairs: both the
whole-batch schedule with its two pieces and odd chunkings) through orbx_stream_upload / _run / _download and through
orbx_stream_process, against the oracle frame by frame and pair by pair; then the same frames under ComputeKeyPointsOld
(policy 1) and with a random lapping area through the single-frame path."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import test_gpu_fuzz as T
from oracle import orb_oracle as orc
from orb_slam_tracking_b200 import frontend as fe, synth
from orb_slam_tracking_b200.capi import OrbxError
N, seed = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(seed)
bad = 0
for cfg in T._configs(N, seed):
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    if W * H > 500 * 400: W, H = W // 2 + 80, H // 2 + 60     # keep the oracle side of a whole batch short
    nf = min(nf, 900)
    P = int(rng.integers(33, 71))
    try:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H, max_batch=2 * P)
    except OrbxError as e:
        continue
    R = orc.Extractor(nf, sf, nl, ini, mn)
    K4 = np.array([0.8 * W, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25], np.float32)
    D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if dist else [0, 0, 0, 0], np.float32)
    b = E.set_camera(K4, D4)
    base = synth.frames(8, W, H, first_pair=500 + 4 * idx)
    imgs = np.ascontiguousarray(np.stack([base[(3 * i + (i // 8)) % 8] for i in range(2 * P)]))
    ref = {}
    def oracle_frame(i):
        key = (3 * i + (i // 8)) % 8
        if key not in ref:
            k, d, _ = R(base[key]); ref[key] = (k, d, orc.Frame(k, d, K4, D4, b))
        return ref[key]
    win = int(rng.choice([100, 40, 4096]))
    S = fe.Stream(E, 0.9, True, win)
    ok = True
    why = []
    for mode in ("run", "process"):
        if mode == "run":
            S.upload(imgs); S.run(2 * P); S.sync(); out = S.download(2 * P)
        else:
            out = S.process(imgs, chunk=int(rng.choice([0, 6, 10, 32])))
        for i in range(2 * P):
            rk, rd, _ = oracle_frame(i)
            n = int(out["n"][i])
            g = n == len(rk) and np.array_equal(out["kps"][i][:n].view(np.uint8), rk.view(np.uint8)) and np.array_equal(out["desc"][i][:n], rd)
            if not g: why.append((mode, 'frame', i, n, len(rk)))
            ok &= g
        for p in range(0, P, 7):
            f1, f2 = oracle_frame(2 * p)[2], oracle_frame(2 * p + 1)[2]
            rn, rm12, rcnt = orc.search_for_initialization(f1, f2, 0.9, True, win)
            g = int(out["nmatches"][p]) == rn and np.array_equal(out["matches12"][p][:f1.N], rm12) and np.array_equal(out["counters"][p], rcnt)
            if not g: why.append((mode, 'pair', p, int(out['nmatches'][p]), rn))
            ok &= g
    # lapping area + the other keypoint policy on single frames
    lap = (int(rng.integers(0, W // 2)), int(rng.integers(W // 2, W)))
    for pol in (0, 1):
        try:
            E.set_keypoint_policy(pol)
        except OrbxError:
            continue
        Rp = orc.Extractor(nf, sf, nl, ini, mn, policy=pol)
        for im in base[:2]:
            try:
                rk, rd, rm = Rp(im, lapping=lap)
            except RuntimeError:
                continue
            k, d, m = E(im, vLappingArea=lap)
            g = m == rm and len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd)
            if not g: why.append(('policy', pol, 'lap', lap, len(k), len(rk), m, rm))
            ok &= g
    if not ok:
        bad += 1; print("MISMATCH", cfg, "W,H used", W, H, "P", P, "win", win, why[:4])
    E.close()
print("configs", N, "mismatches", bad)
