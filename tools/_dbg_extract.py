import sys, numpy as np
sys.path.insert(0, '.')
from orb_slam_tracking_b200 import frontend as fe, synth
E = fe.ORBextractor(1000, 1.2, 8, 20, 7, 640, 480)
im = synth.frames(2, 640, 480)[0]
k, d, m = E(im)
print(len(k), m)
