#!/usr/bin/env python
"""Print the key figures of a bench.py JSON line (file argument or stdin)."""
import json
import sys

txt = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
d = json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
print(f"value {d['value']:.4g} {d['unit']}  ms/step {d['ms_per_step']:.4f}  n_gpus {d['n_gpus']}  launches {d.get('gpu_launches')}")
print("phase_ms", d.get("config", {}).get("phase_ms"))
if "roofline" in d:
    r = d["roofline"]
    print(f"roofline frac {r['frac']:.4f}  achieved {r['achieved']:.0f} {r['unit']}  traffic {r.get('traffic')}")
if "e2e" in d:
    print("e2e", d["e2e"].get("value"), d["e2e"].get("matches_device_path"))
for k in ("eigen_mc", "c3", "c5"):
    b = d.get(k)
    if b:
        print(k, b.get("value"), b.get("unit"), b.get("ms_per_step"), b.get("phase_ms", ""),
              (b.get("roofline") or {}).get("frac", ""))
if d.get("df_to_df"):
    print("df_to_df", {k: v for k, v in d["df_to_df"].items() if k != "config"})
print("cpu_baseline", d.get("cpu_baseline", {}).get("value"), d.get("cpu_baseline", {}).get("kind"), "clocks", d.get("clocks"))
