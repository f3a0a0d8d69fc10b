#!/bin/bash
# Single-frame latency through the C++ drop-in classes (tests/csrc/dropin_demo.cpp: Frame::Frame per image, SearchForInitialization
# per pair, no Python in the loop): builds the demo against liborbx.so and runs it on a synthetic 640x480 and 752x480 pair.
set -e
mkdir -p gpurun_out
g++ -std=c++17 -O2 -Iinclude tests/csrc/dropin_demo.cpp -o gpurun_out/dropin_demo -Lorb_slam_tracking_b200 -lorbx -Wl,-rpath,$PWD/orb_slam_tracking_b200
for shape in "640 480 1000" "752 480 1200"; do
  set -- $shape
  python -c "
from orb_slam_tracking_b200 import synth
synth.frame_pair(5, $1, $2).tofile('gpurun_out/pair_$1.raw')"
  echo "== $1 x $2, $3 features"
  ./gpurun_out/dropin_demo gpurun_out/pair_$1.raw $1 $2 $3 | grep -E "RESULT|LATENCY"
done
rm -f gpurun_out/pair_*.raw gpurun_out/dropin_demo
