#!/usr/bin/env python
"""Per-source-line instruction/stall shares of one kernel from an .ncu-rep (SASS page joined
with nvdisasm line info of the object file).  usage: ncu_lines.py rep kernel obj.o [min_pct]"""
import csv, io, re, subprocess, sys, tempfile, os, glob
rep, kern, obj = sys.argv[1:4]
minp = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# first kernel instance only
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
h = starts[0]
end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] == "Kernel Name"), len(rows))
hdr = rows[h]
ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
sass = [(r[1], float(r[ie] or 0), float(r[ss] or 0)) for r in rows[h + 1:end] if len(r) > ie]
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
dis = subprocess.run(["nvdisasm", "-g", "-c", glob.glob(d + "/*.cubin")[0]], capture_output=True, text=True).stdout
lines, cur, infn = [], None, False
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        infn = kern in m.group(1); continue
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
n = min(len(lines), len(sass))
if len(lines) != len(sass): print("warning: %d disasm vs %d ncu instructions" % (len(lines), len(sass)))
agg = {}
for k in range(n):
    a = agg.setdefault(lines[k], [0, 0]); a[0] += sass[k][1]; a[1] += sass[k][2]
ti, ts = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
print("total warp instr %.3g, samples %d" % (ti, ts))
for key in sorted(agg, key=lambda k: (k is None, k)):
    a = agg[key]
    if 100 * a[0] / ti >= minp or 100 * a[1] / max(ts, 1) >= minp:
        print("%-22s instr %5.1f%%  samples %5.1f%%" % ("%s:%d" % key if key else "?", 100 * a[0] / ti, 100 * a[1] / max(ts, 1)))
