#!/bin/bash
# usage: build_variants.sh name1 "flags1" name2 "flags2" ...   -> orb_slam_tracking_b200/variants/liborbx_<name>.so (git-ignored, travels to the GPU box)
# On the box: cp orb_slam_tracking_b200/variants/liborbx_<name>.so orb_slam_tracking_b200/liborbx.so before the run.
set -e
cd "$(dirname "$0")/.."
mkdir -p orb_slam_tracking_b200/variants
while [ $# -ge 2 ]; do
  ORBX_NVCC_EXTRA="$2" python -m orb_slam_tracking_b200.build --force > /dev/null
  cp orb_slam_tracking_b200/liborbx.so orb_slam_tracking_b200/variants/liborbx_$1.so
  echo "built variant $1 ($2)"
  shift 2
done
python -m orb_slam_tracking_b200.build --force > /dev/null
