"""Determinism soak (no compute-sanitizer on this pool): the same 512-frame batch through orbx_stream_run N times and through the
streaming host path, every result block compared by CRC with the first one.  A race in the tile ring, the per-warp lists, the
quadtree's atomics or the matcher's ordered replay would show up as a differing block sooner or later."""
import sys, zlib, numpy as np
sys.path.insert(0, '.')
from orb_slam_tracking_b200 import frontend as fe, synth
N = int(sys.argv[1]) if len(sys.argv) > 1 else 200
W, H, B = 640, 480, 512
E = fe.ORBextractor(1000, 1.2, 8, 20, 7, W, H, max_batch=B)
E.set_camera(*synth.camera_for(W, H))
imgs = synth.frames_parallel(B, W, H)
S = fe.Stream(E, 0.9, True, 100)
def crc(out): return tuple(zlib.crc32(np.ascontiguousarray(out[k]).tobytes()) for k in ("n", "kps", "kps_un", "desc", "matches12", "nmatches", "counters"))
S.upload(imgs)
first = None; bad = 0
for it in range(N):
    S.run(B); S.sync()
    c = crc(S.download(B))
    if first is None: first = c
    elif c != first: bad += 1; print("resident iteration", it, "differs", [a == b for a, b in zip(c, first)])
for it in range(N // 4):
    c = crc(S.process(imgs, chunk=[0, 32, 50, 128][it % 4]))
    if c != first: bad += 1; print("process iteration", it, "differs", [a == b for a, b in zip(c, first)])
t = [S.process_async(imgs), S.process_async(imgs)]
for it in range(N // 4):
    c = crc(S.wait(t.pop(0))); t.append(S.process_async(imgs))
    if c != first: bad += 1; print("async iteration", it, "differs")
for x in t: S.wait(x)
print("iterations", N, "+", N // 4, "+", N // 4, "differing result blocks:", bad)
