#!/usr/bin/env python
"""Per-phase instruction / stall-sample shares of one kernel from an .ncu-rep.  Phases are the source lines that hold
a `// ---- phase NAME` marker in the kernel's .cu file; SASS instructions inlined from helpers inherit the phase of the
surrounding code (program order).  Also prints the headline metrics.
usage: ncu_phase.py rep kernel obj.o src.cu [pixels]"""
import csv, io, re, subprocess, sys, tempfile, os, glob
rep, kern, obj, src = sys.argv[1:5]
pixels = float(sys.argv[5]) if len(sys.argv) > 5 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "sass"],
                     capture_output=True, text=True).stdout
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
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
assert len(lines) == len(sass), (len(lines), len(sass))
text = open(src).read().splitlines()
marks = [(i + 1, re.search(r"// ---- phase (\S+)", l).group(1)) for i, l in enumerate(text) if "// ---- phase " in l]
base = os.path.basename(src)
first = marks[0][0] if marks else 1 << 30
def phase_of(ln):
    p = None
    for a, n in marks:
        if ln >= a: p = n
    return p
agg, cur = {}, "pre"
for (txt, n, s, t), ln in zip(sass, lines):
    if ln and ln[0] == base and ln[1] >= first:
        cur = phase_of(ln[1])
    a = agg.setdefault(cur, [0, 0, 0]); a[0] += n; a[1] += s; a[2] += t
ti, ts = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
print("total warp instr %.4g%s, samples %d" % (ti, (" = %.3f per pixel" % (ti / pixels)) if pixels else "", ts))
for k, a in agg.items():
    print("%-10s instr %5.1f%%%s  samples %5.1f%%  lanes/instr %4.1f" % (k, 100 * a[0] / ti, (" (%.3f/px)" % (a[0] / pixels)) if pixels else "",
                                                                      100 * a[1] / max(ts, 1), a[2] / max(a[0], 1)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
want = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"]
want += [x for x in rr[0] if x.startswith("smsp__average_warps_issue_stalled") and x.endswith("per_issue_active.ratio")]
for w in want:
    if w in rr[0]:
        i = rr[0].index(w)
        v = rr[2][i]
        try:
            if float(v) < 0.05 and "stalled" in w: continue
        except ValueError: pass
        print("%-90s %s %s" % (w.replace("smsp__average_warps_issue_stalled_", "stall ").replace("_per_issue_active.ratio", ""), v, rr[1][i]))
