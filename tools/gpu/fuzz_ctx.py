"""Several contexts alive in one process with interleaved calls (per-function attributes, graphs, streams are per context or per
device), and 150 create / destroy cycles with the free device memory watched."""
import sys, ctypes as C, numpy as np
sys.path.insert(0, '.')
from oracle import orb_oracle as orc
from orb_slam_tracking_b200 import frontend as fe, synth
import torch
cfgs = [(640, 480, 1000, 1.2, 8, 20, 7, 0), (1241, 376, 2000, 1.2, 8, 20, 7, 1), (333, 251, 300, 1.3, 5, 7, 3, 0), (3840, 2160, 8000, 1.2, 8, 20, 7, 0),
        (752, 480, 1200, 1.2, 8, 0, 0, 1)]
ctx = []
for (W, H, nf, sf, nl, ini, mn, pol) in cfgs:
    E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H, max_batch=4)
    E.set_camera(*synth.camera_for(W, H))
    if pol: E.set_keypoint_policy(1)
    R = orc.Extractor(nf, sf, nl, ini, mn, policy=pol)
    ims = synth.frames(4, W, H, first_pair=40)
    ref = [R(im) for im in ims]
    ctx.append((E, ims, ref))
bad = 0
rng = np.random.default_rng(3)
for it in range(60):
    E, ims, ref = ctx[int(rng.integers(0, len(ctx)))]
    mode = int(rng.integers(0, 3))
    if mode == 0:
        j = int(rng.integers(0, 4)); k, d, m = E(ims[j]); rk, rd, rm = ref[j]
        ok = len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd)
    elif mode == 1:
        out = E.extract_batch(list(ims))
        ok = all(len(o[0]) == len(r[0]) and np.array_equal(o[0].view(np.uint8), r[0].view(np.uint8)) and np.array_equal(o[1], r[1]) for o, r in zip(out, ref))
    else:
        j = int(rng.integers(0, 4)); F = fe.Frame(E, ims[j], fused=True); rk, rd, rm = ref[j]
        ok = F.N == len(rk) and np.array_equal(F.mvKeys.view(np.uint8), rk.view(np.uint8)) and np.array_equal(F.mDescriptors, rd)
    if not ok:
        bad += 1; print("MISMATCH iteration", it, mode)
for E, _, _ in ctx: E.close()
free0 = torch.cuda.mem_get_info()[0]
for it in range(150):
    W, H = int(rng.integers(200, 900)), int(rng.integers(150, 600))
    try:
        E = fe.ORBextractor(int(rng.integers(50, 2000)), 1.2, int(rng.integers(2, 9)), 20, 7, W, H, max_batch=int(rng.integers(1, 16)))
    except Exception:
        continue
    E.set_camera(*synth.camera_for(W, H))
    if it % 10 == 0: E(synth.frames(1, W, H)[0])
    E.close()
free1 = torch.cuda.mem_get_info()[0]
print("interleaved mismatches", bad, "| device memory not returned after 150 create / destroy cycles: %.1f MB" % ((free0 - free1) / 1e6))
