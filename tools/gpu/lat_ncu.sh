#!/bin/bash
# per-kernel durations of the single-frame path (Frame::Frame through orbx_frame_construct): ncu launch list of tools/gpu/lat_probe.py
TAG=${1:-lat}
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 150 -c 60 --csv --log-file gpurun_out/${TAG}_lat_launches.csv \
  python tools/gpu/lat_probe.py 0 > gpurun_out/${TAG}_lat_ncu.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_lat_launches.csv")) if len(r)>10 and r[0].isdigit()]
agg={}
order=[]
for r in rows:
    k=r[4].split("(")[0]
    if k not in agg: order.append(k)
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=float(r[-1])/1e3
for k in order: print("%-28s n=%3d mean %.1f us"%(k,agg[k][0],agg[k][1]/agg[k][0]))
PY
