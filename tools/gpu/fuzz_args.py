"""Error behaviour: bad arguments must come back as ORBX_ERR_* codes (or defined results), never crash the process or poison the context."""
import sys, ctypes as C, numpy as np
sys.path.insert(0, '.')
from orb_slam_tracking_b200 import frontend as fe, synth, capi
from orb_slam_tracking_b200.capi import OrbxError, ptr
L = capi.lib()
W, H = 640, 480
codes = []
def expect_err(name, fn):
    try:
        r = fn()
        codes.append((name, "ok", r if isinstance(r, int) else None))
    except OrbxError as e:
        codes.append((name, "err", e.code))
st = C.c_int(0)
for args in [(0, -5, 480, 1000, 1.2, 8, 20, 7, 1), (0, 640, 480, 0, 1.2, 8, 20, 7, 1), (0, 640, 480, 1000, 1.0, 8, 20, 7, 1), (0, 640, 480, 1000, 1.2, 0, 20, 7, 1),
             (0, 640, 480, 1000, 1.2, 40, 20, 7, 1), (99, 640, 480, 1000, 1.2, 8, 20, 7, 1), (0, 40, 40, 100, 1.2, 8, 20, 7, 1), (0, 640, 480, 1000, 0.5, 8, 20, 7, 1),
             (0, 640, 480, 1000, float('nan'), 8, 20, 7, 1), (0, 640, 480, 1000, 1.2, 8, -3, 999, 1), (0, 100, 1000, 500, 1.2, 4, 20, 7, 1)]:
    h = L.orbx_create(*args, C.addressof(st))
    codes.append(("create%r" % (args,), "handle" if h else "null", st.value))
    if h: L.orbx_destroy(h)
E = fe.ORBextractor(1000, 1.2, 8, 20, 7, W, H, max_batch=4)
im = synth.frames(2, W, H)
k0, d0, _ = E(im[0])
kp = np.zeros(10, capi.KP_DTYPE); de = np.zeros((10, 32), np.uint8); n = C.c_int(0); mono = C.c_int(0)
codes.append(("extract cap too small", L.orbx_extract(E._h, ptr(im[0]), W, None, ptr(kp), ptr(de), 10, C.addressof(n), C.addressof(mono)), n.value))
codes.append(("extract null image", L.orbx_extract(E._h, None, W, None, ptr(kp), ptr(de), 10, C.addressof(n), C.addressof(mono)), None))
codes.append(("extract step < width", L.orbx_extract(E._h, ptr(im[0]), W - 1, None, ptr(kp), ptr(de), 10, C.addressof(n), C.addressof(mono)), None))
expect_err("glue before camera", lambda: fe.Frame(E, im[0]))
expect_err("stream run before camera", lambda: fe.Stream(E).run(2))
E.set_camera(*synth.camera_for(W, H))
expect_err("stream run n > max_batch", lambda: L.orbx_stream_run(E._h, 6, 0.9, 1, 100))
expect_err("stream run n = 0", lambda: L.orbx_stream_run(E._h, 0, 0.9, 1, 100))
expect_err("stream wait bad ticket", lambda: capi.check(L.orbx_stream_wait(E._h, 12345), E._h))
expect_err("policy 7", lambda: E.set_keypoint_policy(7))
expect_err("trig mode 9", lambda: E.set_trig_mode(9))
expect_err("option unknown", lambda: E.set_option("nonsense", 1))
F = fe.Frame(E, im[0]); G = fe.Frame(E, im[1])
for x, y, r in [(float('nan'), 10.0, 50.0), (1e30, -1e30, 1e30), (100.0, 100.0, -5.0), (100.0, 100.0, 0.0), (float('inf'), 0.0, 10.0)]:
    try:
        res = F.GetFeaturesInArea(x, y, r)
        codes.append(("area %r" % ((x, y, r),), "ok", len(res)))
    except OrbxError as e:
        codes.append(("area %r" % ((x, y, r),), "err", e.code))
for win in (-5, 0, 2 ** 30):
    try:
        nm, m12, cnt = fe.ORBmatcher(0.9, True).SearchForInitialization(F, G, win)
        codes.append(("match window %d" % win, "ok", nm))
    except OrbxError as e:
        codes.append(("match window %d" % win, "err", e.code))
for ratio in (float('nan'), -1.0, 1e9):
    nm, m12, cnt = fe.ORBmatcher(ratio, True).SearchForInitialization(F, G, 100)
    codes.append(("match ratio %r" % ratio, "ok", nm))
# the defined edge cases against the oracle
from oracle import orb_oracle as orc
K4o, D4o = synth.camera_for(W, H)
bo = orc.image_bounds(W, H, K4o, D4o)
Ro = orc.Extractor(1000, 1.2, 8, 20, 7)
Fo = []
for j in range(2):
    kk, dd, _ = Ro(im[j]); Fo.append(orc.Frame(kk, dd, K4o, D4o, bo))
edge_bad = 0
for win in (-5, 0, 1, 2 ** 30):
    for ratio in (float('nan'), -1.0, 0.0, 1e9, 0.9):
        nm, m12, cnt = fe.ORBmatcher(ratio, True).SearchForInitialization(F, G, win)
        rn, rm12, rcnt = orc.search_for_initialization(Fo[0], Fo[1], ratio, True, win)
        if not (nm == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt)): edge_bad += 1; print("EDGE MISMATCH match", win, ratio, nm, rn)
for x, y, r in [(float('nan'), 10.0, 50.0), (1e30, -1e30, 1e30), (100.0, 100.0, -5.0), (100.0, 100.0, 0.0), (float('inf'), 0.0, 10.0), (-1e9, 3e9, 5e9), (320.0, 240.0, 1e9)]:
    a = F.GetFeaturesInArea(x, y, r); bq = Fo[0].features_in_area(x, y, r)
    if not np.array_equal(a, bq): edge_bad += 1; print("EDGE MISMATCH area", x, y, r, len(a), len(bq))
print("edge cases against the oracle: mismatches", edge_bad)
# NaN / degenerate camera
for K4, D4 in [((0, 0, 0, 0), (0, 0, 0, 0)), ((float('nan'), 500, 320, 240), (0.1, 0, 0, 0)), ((500, 500, 320, 240), (1e20, 1e20, 1e20, 1e20))]:
    try:
        b = E.set_camera(np.array(K4, np.float32), np.array(D4, np.float32)); F2 = fe.Frame(E, im[0]); codes.append(("camera %r" % (K4,), "ok", F2.N))
    except (OrbxError, AssertionError) as e:
        codes.append(("camera %r" % (K4,), "err", getattr(e, "code", "assert")))
# the context still works
E.set_camera(*synth.camera_for(W, H))
k1, d1, _ = E(im[0])
print("\n".join("%-70s %s %s" % c for c in codes))
print("context intact after all of it:", len(k1) == len(k0) and np.array_equal(k1.view(np.uint8), k0.view(np.uint8)) and np.array_equal(d1, d0))
