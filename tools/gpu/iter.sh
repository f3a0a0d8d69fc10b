#!/bin/bash
# one development iteration on the GPU box: parity suite, bench line, ncu capture of one kernel
# usage: iter.sh TAG [kernel-regex] [extra bench args]
TAG=${1:-r02x}; KRE=${2:-k_fast}; shift 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --no-cpu --steps 10 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("resident", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["stage_ms_per_step"])
except Exception as e: print("bench parse fail", e)
PY
if [ "$KRE" != "none" ]; then
ORBX_SINGLE_STREAM=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:$KRE --launch-skip 4 --launch-count 1 \
  -o gpurun_out/${TAG}_ncu -f python bench.py --no-cpu --steps 2 --warmup 3 "$@" > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc $?"
fi
