#!/bin/bash
# usage: variants.sh TAG name1 name2 ...  -- bench line (stage times) per prebuilt library variant (tools/build_variants.sh); "base" = the tree's own
TAG=$1; shift
mkdir -p gpurun_out
cp orb_slam_tracking_b200/liborbx.so /tmp/liborbx_base.so
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/liborbx_base.so orb_slam_tracking_b200/liborbx.so; else cp orb_slam_tracking_b200/variants/liborbx_$v.so orb_slam_tracking_b200/liborbx.so; fi
  timeout 300 python bench.py --no-cpu --no-variants --steps 10 $ORBX_VARIANT_ARGS > gpurun_out/${TAG}_${v}.json 2> gpurun_out/${TAG}_${v}.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_${v}.json").read().strip().splitlines()[-1])
    print("$v resident", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["stage_ms_per_step"])
except Exception as e: print("$v fail", e)
PY
done
cp /tmp/liborbx_base.so orb_slam_tracking_b200/liborbx.so
