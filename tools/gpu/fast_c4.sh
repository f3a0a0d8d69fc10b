#!/bin/bash
mkdir -p gpurun_out
run() { # name env...
  n=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-variants --steps 10 > gpurun_out/fc4_$n.json 2> gpurun_out/fc4_$n.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/fc4_$n.json").read().strip().splitlines()[-1])
    print("$n resident", round(d["value"]), d["stage_ms_per_step"])
except Exception as e: print("$n fail", e)
PY
}
run base A=1
cp orb_slam_tracking_b200/liborbx.so /tmp/base.so
cp orb_slam_tracking_b200/variants/liborbx_c4.so orb_slam_tracking_b200/liborbx.so
run c4_15k A=1
run c4_12k ORBX_FAST_TILE=12288
run c4_11k ORBX_FAST_TILE=11264
run c4_10k ORBX_FAST_TILE=10240
cp /tmp/base.so orb_slam_tracking_b200/liborbx.so
timeout 600 python -m pytest tests -m gpu -x -q -k "parity or golden" 2>&1 | tail -2
