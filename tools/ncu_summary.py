"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches_c2.csv profiles/r01_launches_c2.txt
    python tools/ncu_summary.py full gpurun_out/prof_decode_c2.ncu-rep profiles/r01_decode_c2_full.txt

`launches`: per-kernel count / total / share / average of gpu__time_duration.sum (cold-cache,
serialised: shares are meaningful, absolutes are not). `full`: the metrics the roofline line of
bench.py quotes (dram bytes per launch, duration, throughput %, occupancy limits, registers).
"""
from __future__ import annotations

import collections
import csv
import subprocess
import sys

FULL_METRICS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu.sum",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "launch__grid_size",
    "launch__block_size",
    "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem",
    "launch__waves_per_multiprocessor",
    "smsp__cycles_active.avg",
    "sm__cycles_elapsed.max",
]


def launches(src: str, dst: str) -> None:
    with open(src) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v * 1e6 if u == "s" else v
        a = agg[row["Kernel Name"]]
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values()) or 1.0
    out = [f"# source: {src}  (ncu --metrics gpu__time_duration.sum --clock-control none)",
           f"# {sum(v[0] for v in agg.values())} launches, {tot:.1f} us total; cold-cache and "
           "serialised: compare SHARES", f"{'count':>7} {'total_us':>12} {'share':>7} {'avg_us':>10}  kernel"]
    for name, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        out.append(f"{c:7d} {t:12.1f} {t / tot * 100:6.2f}% {t / c:10.2f}  {name[:150]}")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out[:12]))


def full(src: str, dst: str) -> None:
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True,
                         text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    head, units, data = rows[0], rows[1], rows[2:]
    out = [f"# source: {src}  (ncu --set full --clock-control none --import-source on)"]
    kcol = head.index("Kernel Name")
    for i, r in enumerate(data):
        out.append(f"launch {i}: {r[kcol][:140]}")
    for m in FULL_METRICS:
        if m in head:
            j = head.index(m)
            out.append(f"{m} [{units[j]}]: " + ", ".join(r[j] for r in data))
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
