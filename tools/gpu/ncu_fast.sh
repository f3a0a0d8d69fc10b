#!/bin/bash
# one ncu --set full capture of k_fast (whole-batch launch, single stream) with source correlation
mkdir -p gpurun_out
TAG=${1:-r02}
ORBX_SINGLE_STREAM=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_fast --launch-skip 4 --launch-count 1 \
  -o gpurun_out/${TAG}_fast -f python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu_fast.log 2>&1
echo "ncu rc $?"; tail -3 gpurun_out/${TAG}_ncu_fast.log
