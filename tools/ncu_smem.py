#!/usr/bin/env python
"""Shared-memory wavefronts per SASS instruction of one kernel (ideal / actual / excessive), with the source line.
usage: ncu_smem.py rep kernel obj.o [top]"""
import csv, io, re, subprocess, sys, tempfile, os, glob
rep, kern, obj = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[h]
end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] in ("Kernel Name", "Address")), len(rows))
ix = {k: hdr.index(k) for k in ("Source", "Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "L1 Wavefronts Shared Excessive", "# Samples")}
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
body = rows[h + 1:end]
assert len(lines) == len(body), (len(lines), len(body))
f = lambda r, k: float(r[ix[k]] or 0)
tot = sum(f(r, "L1 Wavefronts Shared") for r in body); exc = sum(f(r, "L1 Wavefronts Shared Excessive") for r in body)
print("shared wavefronts %.4g, excessive %.4g (%.1f%%)" % (tot, exc, 100 * exc / max(tot, 1)))
rk = sorted(range(len(body)), key=lambda i: -f(body[i], "L1 Wavefronts Shared"))[:top]
for i in rk:
    r = body[i]
    print("%-16s %6.2f%% wf  %9.4g instr  wf/instr %5.2f (ideal %5.2f)  samples %5d  %s" % (
        "%s:%d" % lines[i] if lines[i] else "?", 100 * f(r, "L1 Wavefronts Shared") / tot, f(r, "Instructions Executed"),
        f(r, "L1 Wavefronts Shared") / max(f(r, "Instructions Executed"), 1), f(r, "L1 Wavefronts Shared Ideal") / max(f(r, "Instructions Executed"), 1),
        int(f(r, "# Samples")), r[ix["Source"]].strip()[:60]))
