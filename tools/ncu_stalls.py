#!/usr/bin/env python
"""Per-instruction stall summary of one kernel from an ncu report captured with --import-source on.
    python tools/ncu_stalls.py gpurun_out/prof.ncu-rep k_ngp_mlp_bwd_tc [top_n]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
start = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hdr = rows[start[0]]
end = start[1] - 1 if len(start) > 1 else len(rows)
data = [r for r in rows[start[0] + 1:end] if len(r) == len(hdr)]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix["# Samples"]]) for r in data)
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(int(r[ix[h]]) for r in data) for h in reasons}
print(f"{kern}: {tot} samples, {len(data)} instructions")
for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
    print(f"  {h:28s} {100 * v / tot:5.1f}%")
print("  top instructions:")
for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:top_n]:
    n = int(r[ix["# Samples"]])
    why = ",".join(f"{h[6:]}={r[ix[h]]}" for h in reasons if int(r[ix[h]]) > 0.3 * n)
    print(f"  {data.index(r):5d} {100 * n / tot:5.1f}%  {r[ix['Source']].strip()[:70]:70s} {why}")
