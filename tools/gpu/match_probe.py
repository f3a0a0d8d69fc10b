"""SearchForInitialization on host frames, one pair per call (the reference's call shape): wall clock per call."""
import sys, time, numpy as np
sys.path.insert(0, '.')
from orb_slam_tracking_b200 import frontend as fe, synth
W, H, nf = 640, 480, 1000
E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H, max_batch=2)
K4, D4 = synth.camera_for(W, H); E.set_camera(K4, D4)
fr = synth.frames(4, W, H)
F = [fe.Frame(E, im, fused=True) for im in fr]
M = fe.ORBmatcher(0.9, True)
for i in range(6):
    M.SearchForInitialization(F[0], F[1], 100)
t = []
for i in range(40):
    t0 = time.perf_counter(); M.SearchForInitialization(F[2 * (i & 1)], F[2 * (i & 1) + 1], 100); t.append((time.perf_counter() - t0) * 1e6)
print('pair match us p50', np.percentile(t, 50))
