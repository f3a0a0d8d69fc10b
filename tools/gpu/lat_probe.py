import sys, time, numpy as np
sys.path.insert(0, '.')
from orb_slam_tracking_b200 import frontend as fe, synth
W,H,nf=640,480,1000
E=fe.ORBextractor(nf,1.2,8,20,7,W,H,max_batch=2)
K4,D4=synth.camera_for(W,H); E.set_camera(K4,D4)
fr=synth.frames(4,W,H)
E.set_option("frame_graph", int(sys.argv[1]) if len(sys.argv)>1 else 1)
for i in range(6):
    F=fe.Frame(E, fr[i%4], fused=True)
t=[]
for i in range(40):
    t0=time.perf_counter(); F=fe.Frame(E, fr[i%4], fused=True); t.append((time.perf_counter()-t0)*1e6)
print('frame us p50', np.percentile(t,50))
