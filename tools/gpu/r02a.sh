#!/bin/bash
# first GPU run of the strip-based persistent k_fast: parity suite, then bench at several tile budgets
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version --format=csv,noheader > gpurun_out/r02a_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r02a_pytest.log
tail -5 gpurun_out/r02a_pytest.log
for t in 15360 11000 22000 30000; do
  ORBX_FAST_TILE=$t timeout 300 python bench.py --no-cpu --steps 10 > gpurun_out/r02a_bench_$t.json 2> gpurun_out/r02a_bench_$t.err
  echo "tile $t rc $?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r02a_bench_$t.json").read().strip().splitlines()[-1])
    print($t, round(d["value"]), round(d["e2e"]["value"]), d["stage_ms_per_step"])
except Exception as e: print("fail", e)
PY
done
