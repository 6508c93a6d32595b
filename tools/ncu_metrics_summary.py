#!/usr/bin/env python
"""Summarise an `ncu --metrics ... --csv` log (one row per kernel launch and metric) into a per-kernel table:

    python tools/ncu_metrics_summary.py gpurun_out/p_train_kernels.csv profiles/r2_train_kernels.json [--min-us 2]

mean per launch of every metric over the launches of each kernel (launches below --min-us are dropped: warm-up fills)."""
import collections
import csv
import json
import re
import sys


def short(name):
    n = re.sub(r"^void\s+", "", name)
    n = n.split("(")[0].replace("acn::", "").strip()
    return re.sub(r"\s+", "", n)[:80]


def main():
    src, dst = sys.argv[1], sys.argv[2]
    min_us = float(sys.argv[sys.argv.index("--min-us") + 1]) if "--min-us" in sys.argv else 0.0
    rows = list(csv.DictReader(l for l in open(src) if l.startswith('"')))
    per = collections.OrderedDict()
    for r in rows:
        key = (r["ID"], short(r["Kernel Name"]))
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        u = r["Metric Unit"]
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
        d = per.setdefault(key, {"grid": r["Grid Size"], "block": r["Block Size"]})
        d[r["Metric Name"]] = v
    agg = collections.OrderedDict()
    for (_, name), d in per.items():
        if d.get("gpu__time_duration.sum", 0.0) < min_us:
            continue
        a = agg.setdefault(name, {"launches": 0, "grid": d["grid"], "block": d["block"], "_sum": collections.Counter()})
        a["launches"] += 1
        for k, v in d.items():
            if k not in ("grid", "block"):
                a["_sum"][k] += v
    out = collections.OrderedDict()
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["_sum"]["gpu__time_duration.sum"]):
        n = a["launches"]
        m = {k: v / n for k, v in a["_sum"].items()}
        out[name] = {"launches": n, "grid": a["grid"], "block": a["block"], "us": round(m.get("gpu__time_duration.sum", 0.0), 2),
                     "dram_read_bytes": round(m.get("dram__bytes_read.sum", 0.0)), "dram_write_bytes": round(m.get("dram__bytes_write.sum", 0.0)),
                     **{k.split(".")[0]: round(v, 2) for k, v in m.items() if "pct" in k},
                     "warp_inst": round(m.get("smsp__inst_executed.sum", 0.0)), "regs": m.get("launch__registers_per_thread")}
    json.dump(out, open(dst, "w"), indent=1)
    for name, d in list(out.items())[:40]:
        print(f"{name[:46]:46s} x{d['launches']:<3d} {d['us']:9.1f} us  dram {d['dram_read_bytes'] / 1e6:9.1f} + {d['dram_write_bytes'] / 1e6:9.1f} MB  "
              f"issue {d.get('smsp__issue_active', 0):5.1f}%  l1 {d.get('l1tex__throughput', 0):5.1f}%  lts {d.get('lts__throughput', 0):5.1f}%  "
              f"dram {d.get('gpu__dram_throughput', 0):5.1f}%  tensor {d.get('sm__pipe_tensor_cycles_active', 0):5.1f}%")


if __name__ == "__main__":
    main()
