#!/usr/bin/env python
"""Summarise a gpurun visit into profiles/: per-kernel totals from the ncu launch list
(`--metrics gpu__time_duration.sum`) and a metric subset of the `ncu --set full` captures.

    python scripts/ncu_summarise.py gpurun_out/r01a profiles/r01a
"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

KEEP = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]


def launch_list(path):
    with open(path) as f:
        lines = f.readlines()
    start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
    agg, total = collections.OrderedDict(), 0.0
    for row in csv.DictReader(lines[start:]):
        if row["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}[row["Metric Unit"]]
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        total += v
    return agg, total


def traffic_list(path, skip_frac=0.0):
    """Per-kernel DRAM bytes per launch (read + write, averaged over the launches of the capture) from
    `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv`."""
    with open(path) as f:
        lines = f.readlines()
    start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
    per = collections.OrderedDict()
    for row in csv.DictReader(lines[start:]):
        m = row["Metric Name"]
        if not m.startswith("dram__bytes"):
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[row["Metric Unit"]]
        name = re.sub(r"[<(].*", "", row["Kernel Name"].replace("<unnamed>::", "")).replace("void ", "").split("::")[-1]
        d = per.setdefault(name, {})
        d[row["ID"]] = d.get(row["ID"], 0.0) + v
    return {k: {"launches": len(d), "avg_bytes_per_launch": sum(d.values()) / len(d), "max": max(d.values())}
            for k, d in per.items()}


def full_capture(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [i for i, h in enumerate(hdr) if h in KEEP]
    res = []
    for r in rows[2:]:
        res.append({hdr[i] + (f" [{units[i]}]" if units[i] else ""): r[i] for i in idx})
    return res


def main(src, dst):
    os.makedirs(os.path.dirname(dst) or ".", exist_ok=True)
    md = [f"# ncu summary of `{src}`\n"]
    ll = os.path.join(src, "launches.csv")
    if os.path.exists(ll):
        agg, total = launch_list(ll)
        md.append("## launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`, "
                  "`python bench.py --steps 2 --warmup 1`): per-kernel totals\n")
        md.append("Per-launch times under ncu are cold-cache and serialised: read the SHARE column, not the absolutes.\n")
        md.append("| kernel | launches | total us | share |\n|---|---:|---:|---:|")
        for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            md.append(f"| `{k[:90]}` | {n} | {v:.0f} | {100 * v / total:.1f}% |")
        md.append(f"| **total** | | {total:.0f} | |\n")
    tl = os.path.join(src, "traffic.csv")
    if os.path.exists(tl):
        t = traffic_list(tl)
        md.append("## DRAM traffic per launch (`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`, same command)\n")
        md.append("| kernel | launches | avg MB / launch | max MB |\n|---|---:|---:|---:|")
        for k, d in sorted(t.items(), key=lambda kv: -kv[1]["avg_bytes_per_launch"] * kv[1]["launches"]):
            md.append(f"| `{k[:70]}` | {d['launches']} | {d['avg_bytes_per_launch'] / 1e6:.1f} | {d['max'] / 1e6:.1f} |")
        md.append("")
        flat = {k: {"avg_bytes_per_launch": d["avg_bytes_per_launch"], "launches": d["launches"]} for k, d in t.items()}
        for path in (dst + "_traffic_avg.json", os.path.join(os.path.dirname(dst) or ".", "traffic.json")):
            with open(path, "w") as f:
                json.dump(flat, f, indent=1)
    traffic = {}
    for f in sorted(os.listdir(src)):
        if f.endswith(".ncu-rep"):
            md.append(f"## `{f}` (`ncu --set full --clock-control none --import-source on`)\n")
            for i, d in enumerate(full_capture(os.path.join(src, f))):
                md.append(f"launch {i}: " + "; ".join(f"{k} = {v}" for k, v in d.items()) + "\n")
                name = re.sub(r"[<(].*", "", d.get("Kernel Name", "").replace("<unnamed>::", "")).replace("void ", "")
                try:
                    rd = [float(v) for k, v in d.items() if k.startswith("dram__bytes_read.sum")][0]
                    wr = [float(v) for k, v in d.items() if k.startswith("dram__bytes_write.sum")][0]
                    unit = [k for k in d if k.startswith("dram__bytes_read.sum")][0]
                    mult = 1e6 if "Mbyte" in unit else 1e9 if "Gbyte" in unit else 1e3 if "Kbyte" in unit else 1.0
                    traffic.setdefault(name, []).append((rd + wr) * mult)
                except (IndexError, ValueError):
                    pass
    with open(dst + "_ncu.md", "w") as f:
        f.write("\n".join(md) + "\n")
    if traffic:
        with open(dst + "_traffic.json", "w") as f:
            json.dump({k: {"per_launch_bytes": v, "max": max(v)} for k, v in traffic.items()}, f, indent=1)
    print("wrote", dst + "_ncu.md")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
