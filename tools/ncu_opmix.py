#!/usr/bin/env python
"""Executed warp instructions of one kernel by opcode (from the source page of an .ncu-rep).  usage: ncu_opmix.py rep kernel [units]
units: divide the counts by this number (e.g. warp-rows) to print instructions per unit."""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1:3]
units = float(sys.argv[3]) if len(sys.argv) > 3 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[h]
end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] in ("Kernel Name", "Address")), len(rows))
ie, so = hdr.index("Instructions Executed"), hdr.index("Source")
mix = collections.Counter()
for r in rows[h + 1:end]:
    t = r[so].split()
    if not t: continue
    op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    mix[op.split(".")[0] if not op.startswith("IMAD.MOV") else "IMAD.MOV"] += float(r[ie] or 0)
tot = sum(mix.values())
print("total %.4g" % tot)
for op, n in mix.most_common(25):
    print("%-12s %6.2f%%%s" % (op, 100 * n / tot, ("  %7.2f per unit" % (n / units)) if units else ""))
