#!/usr/bin/env python
"""Turn the raw captures of tools/profile_round.sh (gpurun_out/) into the tracked summaries under
profiles/: the launch list summed per kernel and the headline metrics of each --set full capture.

  python tools/profile_summary.py r1
"""
import csv
import io
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

FULL_METRICS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum",
    "sm__instruction_throughput.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__cycles_active.avg",
    "sm__cycles_elapsed.avg",
    "lts__t_sector_hit_rate.pct",
    "launch__registers_per_thread",
    "launch__grid_size",
    "launch__block_size",
    "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem",
]


SCALE = {"Gbyte": (1e9, "byte"), "Mbyte": (1e6, "byte"), "Kbyte": (1e3, "byte"), "byte": (1, "byte"),
         "s": (1e6, "us"), "ms": (1e3, "us"), "us": (1, "us"), "ns": (1e-3, "us")}


def normalise(val, unit):
    """ncu picks a unit per value (Kbyte / Mbyte, us / ms): bring them to bytes and microseconds"""
    val = val.replace(",", "")
    if unit in SCALE:
        k, u = SCALE[unit]
        return (f"{float(val) * k:.6g}", u)
    return (val, unit)


def launch_list(tag):
    src = os.path.join(OUT, "launches.csv")
    lines = open(src).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
    shutil.copy(src, os.path.join(PROF, f"{tag}_launches_raw.csv"))
    agg = {}
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        name = r["Kernel Name"].split("(")[0]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(PROF, f"{tag}_launches_summary.csv"), "w") as f:
        f.write("kernel,launches,total_us,avg_us,share_of_device_time\n")
        for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{name},{n},{us:.1f},{us / n:.2f},{us / tot:.4f}\n")
    return agg, tot


def full_captures(tag):
    reps = sorted(p for p in os.listdir(OUT) if p.startswith("full_") and p.endswith(".ncu-rep"))
    table = {}
    for rep in reps:
        kernel = rep[len("full_"):-len(".ncu-rep")]
        txt = subprocess.run(["ncu", "-i", os.path.join(OUT, rep), "--page", "raw", "--csv"], capture_output=True,
                             text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        head, units, vals = rows[0], rows[1], rows[2]
        col = {h: i for i, h in enumerate(head)}
        table[kernel] = {m: normalise(vals[col[m]], units[col[m]]) for m in FULL_METRICS if m in col}
    with open(os.path.join(PROF, f"{tag}_ncu_full_summary.csv"), "w") as f:
        f.write("metric,unit," + ",".join(table) + "\n")
        for m in FULL_METRICS:
            unit = next((t[m][1] for t in table.values() if m in t), "")
            f.write(f"{m},{unit}," + ",".join(t.get(m, ("", ""))[0].replace(",", "") for t in table.values()) + "\n")
    return table


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    agg, tot = launch_list(tag)
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name:32s} {n:4d} launches {us:10.1f} us  {us / tot:6.1%}")
    for k, t in full_captures(tag).items():
        print(k, {m.split(".")[0].split("__")[-1]: v[0] for m, v in t.items()})
