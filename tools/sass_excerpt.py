#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that identify the hand-written sm_100a paths (TMA tile loads and their mbarrier
waits, DPX 3-input min/max, byte-SIMD absolute differences, integer dot products, population counts, warp reductions),
with one sample line each, from the built library.  usage: sass_excerpt.py liborbx.so out.md"""
import re, subprocess, sys
lib, dst = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = ["UTMALDG", "SYNCS", "VIMNMX3", "VIMNMX", "VABSDIFF4", "IDP", "POPC", "CREDUX", "ATOMS", "ATOMG", "SHFL", "VOTE", "LDS", "STS", "LDG", "STG", "BAR"]
out = ["# SASS excerpt of %s (cuobjdump -sass, sm_100a)" % lib, "", "| kernel | instructions | " + " | ".join(pats) + " |", "|---|---|" + "---|" * len(pats)]
samples = {}
cur, body = None, {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); body[cur] = []; continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        body[cur].append(line)
for k, ls in body.items():
    name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip().split("(")[0]
    row = []
    for p in pats:
        hits = [l for l in ls if re.search(r"\b%s(\.|\b)" % p, l.split("*/")[1] if "*/" in l else l)]
        row.append(str(len(hits)))
        if hits and p in ("UTMALDG", "SYNCS", "VIMNMX3", "VABSDIFF4", "IDP", "POPC", "CREDUX"):
            samples.setdefault(p, (name, re.sub(r"\s+/\*[0-9a-f]+\*/\s*$", "", hits[0].split("*/", 1)[1]).strip()))
    out.append("| %s | %d | %s |" % (name, len(ls), " | ".join(row)))
out += ["", "## one sample line per mnemonic", ""] + ["* `%s` in %s: `%s`" % (p, n, l) for p, (n, l) in samples.items()]
open(dst, "w").write("\n".join(out) + "\n")
print("\n".join(out[:30]))
