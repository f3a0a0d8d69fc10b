#!/bin/bash
# usage: sweep.sh TAG ENVVAR v1 v2 ...   -- bench line per value of one ORBX_* option
TAG=$1; VAR=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v timeout 300 python bench.py --no-cpu --steps 10 > gpurun_out/${TAG}_${v}.json 2> gpurun_out/${TAG}_${v}.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_${v}.json").read().strip().splitlines()[-1])
    print("$VAR=$v resident", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["stage_ms_per_step"])
except Exception as e: print("$VAR=$v fail", e)
PY
done
