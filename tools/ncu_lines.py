#!/usr/bin/env python
"""Per-source-line stall samples of one kernel from an .ncu-rep (read on the CPU box).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep <kernel-regex> <mangled-substring> <file.cu> [top]

ncu's CSV source page is per SASS instruction; the line mapping comes from
`nvdisasm -g` of the cubin extracted from libbmfm.so (compiled with -lineinfo).
"""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

rep, kregex, mangled, cufile = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "llm-driven-multi-factor-model_b200", "bmfm", "libbmfm.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
base = os.path.basename(cufile)[:-3]
cubin = [c for c in glob.glob(os.path.join(tmp, "*.cubin")) if os.path.basename(c).startswith(base)][0]
sass = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(sass) if l.strip().startswith(".section") and ".text." in l and mangled in l][0]
end = next((i for i in range(start + 1, len(sass)) if sass[i].strip().startswith(".section")), len(sass))
addr_line, cur = {}, None
for l in sass[start:end]:
    m = re.search(r'//## File "(.*?)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
    if m and cur:
        addr_line[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kregex],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
h = rows[1]
ci = {x: i for i, x in enumerate(h)}


def gi(r, k):
    try:
        return int(r[ci[k]] or 0)
    except Exception:
        return 0


data = [r for r in rows[2:] if len(r) > 10]
first = None
seen = set()
bylines, tot, tote = {}, 0, 0
stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
agg = {s: 0 for s in stalls}
for r in data:
    try:
        ad = int(r[ci["Address"]], 16)
    except Exception:
        continue
    if first is None:
        first = ad
    if ad in seen:      # later launches repeat the listing
        break
    seen.add(ad)
    ln = addr_line.get(ad - first, ("?", 0))
    s, e = gi(r, "# Samples"), gi(r, "Instructions Executed")
    tot += s
    tote += e
    d = bylines.setdefault(ln, [0, 0])
    d[0] += s
    d[1] += e
    for k in stalls:
        agg[k] += gi(r, k)
print(f"samples {tot}  warp-instructions {tote}")
print("stalls:", ", ".join(f"{k[6:]} {v / max(tot,1) * 100:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
src = {}
for ln, (s, e) in sorted(bylines.items(), key=lambda kv: -kv[1][0])[:top]:
    f = ln[0]
    if f not in src:
        cand = glob.glob(os.path.join(root, "llm-driven-multi-factor-model_b200", "csrc", f))
        src[f] = open(cand[0]).read().split("\n") if cand else None
    txt = src[f][ln[1] - 1].strip()[:72] if src[f] and ln[1] > 0 else ""
    print(f"{f[:20]:20s} {ln[1]:4d} {s:6d} {s / max(tot,1) * 100:5.1f}% {e:9d} | {txt}")
