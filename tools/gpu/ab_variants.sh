#!/bin/bash
# A/B of prebuilt library variants on one box: tools/gpu/ab_variants.sh <rounds> build_variants/a.so build_variants/b.so ...
# Each variant is copied over orb_slam_tracking_b200/liborbx.so and bench.py prints its value and stage times (rounds interleaved).
rounds=$1; shift
cp orb_slam_tracking_b200/liborbx.so /tmp/liborbx.keep
for r in $(seq $rounds); do
  for v in "$@"; do
    cp "$v" orb_slam_tracking_b200/liborbx.so
    python bench.py --steps 30 --warmup 5 --no-cpu --no-variants 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', round(d['value']), round(d['e2e']['value']), d['stage_ms_per_step'])"
  done
done
cp /tmp/liborbx.keep orb_slam_tracking_b200/liborbx.so
