#!/bin/bash
# brute-force matcher: UHD bench line (auto path and forced k_match_cta), ncu of the matrix kernels
TAG=${1:-r02bf}
mkdir -p gpurun_out
for path in 0 1; do
ORBX_MATCH_PATH=$path timeout 600 python bench.py --workload uhd --no-cpu --steps 5 > gpurun_out/${TAG}_uhd_path$path.json 2> gpurun_out/${TAG}_uhd_path$path.err; echo "uhd path $path rc $?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_uhd_path$path.json").read().strip().splitlines()[-1])
    print("resident", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["stage_ms_per_step"], d["next_rows"]["brute_force_hamming"])
except Exception as e: print("parse fail", e)
PY
done
ORBX_SINGLE_STREAM=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:'k_bf_topk|k_hamming_topk|k_bf_resolve' --launch-skip 3 --launch-count 4 \
  -o gpurun_out/${TAG}_ncu -f python bench.py --workload uhd --no-cpu --steps 1 --warmup 3 > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc $?"
