#!/usr/bin/env python
"""Writes profiles/<tag>_summary.md from gpurun_out/<tag>_launches.csv (launch list),
gpurun_out/<tag>_full.ncu-rep (one --set full capture) and gpurun_out/<tag>_bench.json."""
import csv, io, json, subprocess, sys, shutil
from collections import OrderedDict
from pathlib import Path
tag = sys.argv[1]
note = sys.argv[2] if len(sys.argv) > 2 else ""
G, P = Path("gpurun_out"), Path("profiles")
rows = [r for r in csv.reader(open(G / f"{tag}_launches.csv")) if len(r) > 10 and r[0].isdigit()]
agg = OrderedDict()
for r in rows:
    k = r[4].split("(")[0]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(r[-1]) / 1e3
tot = sum(a[1] for a in agg.values())
out = [f"# {tag}: {note}", "", "## launch list (ncu --metrics gpu__time_duration.sum --clock-control none; `bench.py --steps 2 --warmup 1 --no-cpu`)", "",
       "| kernel | launches | total us | mean us | share |", "|---|---|---|---|---|"]
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    out.append(f"| {k} | {a[0]} | {a[1]:.1f} | {a[1]/a[0]:.1f} | {100*a[1]/tot:.1f}% |")
b = G / f"{tag}_bench.json"
if b.exists():
    line = json.loads([l for l in b.read_text().splitlines() if l.startswith("{")][-1])
    st = line["stage_ms_per_step"]; s = sum(st.values())
    out += ["", "## bench.py line of the same build (no profiler): value %.0f frames/s resident, e2e %.0f frames/s, cpu_baseline %s" %
            (line["value"], line["e2e"]["value"], json.dumps(line.get("cpu_baseline"))), "",
            "CUDA-event stage times (ms per step of %d frames): " % line["config"]["frames_per_step_per_gpu"] +
            ", ".join(f"{k} {v:.3f} ({100*v/s:.0f}%)" for k, v in st.items()), "",
            "roofline: " + json.dumps(line["roofline"]), "", "clocks: " + json.dumps(line["clocks"])]
    shutil.copy(b, P / f"{tag}_bench.json")
rep = G / f"{tag}_full.ncu-rep"
if rep.exists():
    raw = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw))); hdr = rr[0]
    cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "ms"), ("dram__bytes_read.sum", "dram rd MB"), ("dram__bytes_write.sum", "dram wr MB"),
            ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
            ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"), ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1 %"),
            ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %"), ("launch__registers_per_thread", "regs"),
            ("smsp__inst_executed.sum", "warp instr"), ("launch__grid_size", "grid"), ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem conflicts")]
    idx = [(hdr.index(c), n) for c, n in cols if c in hdr]
    out += ["", "## ncu --set full (one launch each)", "", "| " + " | ".join(n for _, n in idx) + " |", "|" + "---|" * len(idx)]
    # per-launch DRAM traffic (read + write) of each kernel at the bench batch, for bench.py's roofline.traffic
    ir, iw, ig = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("launch__grid_size")
    unit = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per = {}
    for r in rr[2:]:
        k = r[idx[0][0]].split("(")[0]
        b = float(r[ir]) * unit[rr[1][ir]] + float(r[iw]) * unit[rr[1][iw]]
        g = per.setdefault(k, {})
        g[r[ig]] = max(g.get(r[ig], 0.0), b)
    traffic = {k: (sum(g.values()) if "k_resize" in k else max(g.values())) for k, g in per.items()}
    # instruction-issue side of the same launches: warp instructions, issue-slot and pipe utilisation (what bounds the integer kernels)
    ii, iis = hdr.index("smsp__inst_executed.sum"), hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")
    pipes = {n: hdr.index(c) for n, c in (("alu", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"), ("fma", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                                          ("xu", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"), ("lsu", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active")) if c in hdr}
    issue = {}
    for r in rr[2:]:
        k = r[idx[0][0]].split("(")[0]
        if k in issue and "k_resize" not in k: continue
        e = issue.setdefault(k, {"warp_instr": 0.0, "issue_active_pct": 0.0, "pipes_pct": {}, "launches": 0})
        e["warp_instr"] += float(r[ii]); e["launches"] += 1
        e["issue_active_pct"] = max(e["issue_active_pct"], float(r[iis]))
        for n, i in pipes.items(): e["pipes_pct"][n] = max(e["pipes_pct"].get(n, 0.0), float(r[i]))
    (P / f"{tag}_traffic.json").write_text(json.dumps({"source": f"ncu --set full, {tag}_full.ncu-rep, dram__bytes_read.sum + dram__bytes_write.sum per launch "
                                                       "(k_resize: sum over its 7 launches of a step)", "bytes_per_launch": traffic, "issue": issue}, indent=1))
    seen = {}
    for r in rr[2:]:
        k = r[idx[0][0]].split("(")[0]
        seen[k] = seen.get(k, 0) + 1
        if seen[k] > (7 if "k_resize" in k else 1): continue
        scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "msecond": 1.0, "ms": 1.0, "s": 1e3, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
        vals = [k]
        for i, _ in idx[1:]:
            try:
                vals.append("%.4g" % (float(r[i]) * scale.get(rr[1][i], 1.0)))
            except ValueError:
                vals.append(r[i])
        out.append("| " + " | ".join(vals) + " |")
shutil.copy(G / f"{tag}_launches.csv", P / f"{tag}_launches.csv")
(P / f"{tag}_summary.md").write_text("\n".join(out) + "\n")
print("\n".join(out))
