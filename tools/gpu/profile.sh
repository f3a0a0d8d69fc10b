#!/bin/bash
# The round's evidence for profiles/: parity suite, smoke, bench line (no profiler), reference arm, then the ncu launch list and
# ONE ncu --set full capture of a whole step (single stream, every kernel once) of the same command.
# usage: profile.sh TAG
TAG=${1:-r02x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc $?"; tail -1 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --steps 20 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref.json 2> gpurun_out/${TAG}_ref.err; echo "ref rc $?"
ORBX_SINGLE_STREAM=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu --no-variants > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu launches rc $?"
ORBX_SINGLE_STREAM=1 timeout 900 ncu --set full --import-source on --clock-control none --launch-skip 45 --launch-count 16 -o gpurun_out/${TAG}_full -f \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-variants > gpurun_out/${TAG}_ncu2.log 2>&1; echo "ncu full rc $?"
