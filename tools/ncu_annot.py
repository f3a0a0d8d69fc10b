#!/usr/bin/env python
"""Annotated SASS of one kernel from an .ncu-rep: execution share, stall samples, active lanes and source line per
instruction.  usage: ncu_annot.py rep kernel obj.o out.txt"""
import csv, io, re, subprocess, sys, tempfile, os, glob
rep, kern, obj, dst = sys.argv[1:5]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[h]
ie, ss, te = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] in ("Kernel Name", "Address")), len(rows))  # first instance only
sass = [(r[1], float(r[ie] or 0), float(r[ss] or 0), float(r[te] or 0)) for r in rows[h + 1:end] if len(r) > ie]
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
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l): lines.append(cur)
tot = sum(s[1] for s in sass)
with open(dst, "w") as f:
    for k, (s, l) in enumerate(zip(sass, lines)):
        f.write("%5d %6.3f%% %5d lanes %4.1f  %-22s %s\n" % (k, 100 * s[1] / tot, s[2], (s[3] / s[1] if s[1] else 0), ("%s:%d" % l if l else "?"), s[0]))
print(len(sass), "instructions, total warp instr %.4g" % tot)
