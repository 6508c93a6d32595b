#!/usr/bin/env python
"""Turn the two ncu captures of the profiling recipe into the committed summaries under profiles/.

    python tools/ncu_summary.py --launches gpurun_out/launches.csv --rep gpurun_out/prof.ncu-rep --tag r1_v1 [--steps 2]

* <tag>_launches.md   per-kernel share of one step from the `--metrics gpu__time_duration.sum` launch list
* <tag>_kernels.json  per-launch metrics of the `--set full` capture (duration, DRAM bytes, pipe utilisation, ...)
* traffic.json        kernel-group -> measured DRAM bytes per launch (bench.py's roofline.traffic)
"""
import argparse
import collections
import csv
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GROUP = {"k_ngp_hash_fwd": "hash_fwd", "k_ngp_hash_bwd": "hash_bwd", "k_ngp_mlp_fwd": "mlp_fwd", "k_ngp_mlp_bwd": "mlp_bwd",
         "k_wgrad_partial": "mlp_bwd", "k_wgrad_reduce": "mlp_bwd", "k_ngp_mlp_fwd_tc": "mlp_fwd", "k_ngp_mlp_bwd_tc": "mlp_bwd",
         "k_composite_fwd": "composite_fwd", "k_composite_bwd": "composite_bwd", "k_adam": "adam", "k_adam_multi": "adam", "k_sumsq_multi": "sumsq"}
METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_tensor.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
           "smsp__inst_executed.sum", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
           "smsp__average_warp_latency_issue_stalled_long_scoreboard.pct", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]


def short(name):
    n = name.split("(")[0]
    return n.replace("acn::", "").replace("void ", "").strip()


def launches(path, steps, marker="k_strat_t"):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    idx = [i for i, r in enumerate(rows) if marker in r["Kernel Name"]]
    start = idx[-steps]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for r in rows[start:]:
        v = float(r["Metric Value"])
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r["Metric Unit"], 1.0)
        tot[short(r["Kernel Name"])] += v
        cnt[short(r["Kernel Name"])] += 1
    T = sum(tot.values())
    out = ["| kernel | launches/step | us/step | share |", "|---|---:|---:|---:|"]
    for n, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        if v / T < 0.0005:
            continue
        out.append(f"| `{n[:70]}` | {cnt[n] / steps:.0f} | {v / steps:.1f} | {100 * v / T:.1f}% |")
    out.append(f"| **total (serialised, cold-cache: compare shares, not absolutes)** | | {T / steps:.1f} | 100% |")
    return "\n".join(out), len(rows)


def kernels(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    out = []
    for d in data:
        k = {"kernel": short(d[ix["Kernel Name"]])}
        for m in METRICS:
            if m in ix and d[ix[m]] != "":
                try:
                    k[m] = float(d[ix[m]])
                except ValueError:
                    k[m] = d[ix[m]]
                k[m + "__unit"] = units[ix[m]]
        out.append(k)
    return out


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    ap.add_argument("--tag", required=True)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--cmd", default="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render")
    a = ap.parse_args()
    prof = os.path.join(ROOT, "profiles")
    os.makedirs(prof, exist_ok=True)
    if a.launches:
        table, n = launches(a.launches, a.steps)
        with open(os.path.join(prof, f"{a.tag}_launches.md"), "w") as f:
            f.write(f"# {a.tag}: kernel launch list of one training step\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` over `{a.cmd}` "
                    f"({n} launches captured; the last {a.steps} steps averaged).\n\n{table}\n")
        print(table)
    if a.rep:
        ks = kernels(a.rep)
        json.dump(ks, open(os.path.join(prof, f"{a.tag}_kernels.json"), "w"), indent=1)
        traffic = collections.defaultdict(list)
        for k in ks:
            g = GROUP.get(k["kernel"].split("<")[0])
            if g == "mlp_bwd" and "wgrad" in k["kernel"]:
                continue
            if g and "dram__bytes_read.sum" in k:
                traffic[g].append(to_bytes(k["dram__bytes_read.sum"], k["dram__bytes_read.sum__unit"])
                                  + to_bytes(k["dram__bytes_write.sum"], k["dram__bytes_write.sum__unit"]))
        tj = {g: sum(v) / len(v) for g, v in traffic.items()}
        json.dump(tj, open(os.path.join(prof, "traffic.json"), "w"), indent=1)
        for k in ks:
            print(k["kernel"][:40], k.get("gpu__time_duration.sum"), k.get("gpu__time_duration.sum__unit"), "dram MB",
                  round((to_bytes(k.get("dram__bytes_read.sum", 0), k.get("dram__bytes_read.sum__unit", "byte"))
                         + to_bytes(k.get("dram__bytes_write.sum", 0), k.get("dram__bytes_write.sum__unit", "byte"))) / 1e6, 1))


if __name__ == "__main__":
    main()
