#!/bin/bash
# bench lines of the named workloads at N GPUs (one process per GPU under torchrun for N > 1)
# usage: workloads.sh TAG N workload...
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
for w in "$@"; do
  out=gpurun_out/${TAG}_${w}_n${N}.json
  if [ "$N" = "1" ]; then
    timeout 900 python bench.py --gpus 1 --steps 10 --workload $w --no-cpu --no-variants > $out 2> ${out%.json}.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --steps 10 --workload $w --no-cpu --no-variants > $out 2> ${out%.json}.err
  fi
  echo "$w n=$N rc $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$out").read().splitlines() if l.startswith("{")][-1])
    print("  resident", round(d["value"]), "e2e", round(d["e2e"]["value"]), "link", d["e2e"].get("host_link", {}).get("frac_of_e2e_step"), d["stage_ms_per_step"])
except Exception as e: print("  parse fail", e)
PY
done
