"""Turn ncu artefacts brought back in gpurun_out/ into small text summaries under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches.csv profiles/launches_<tag>.md
    python tools/ncu_summary.py report   gpurun_out/prof.ncu-rep  profiles/ncu_<tag>.md

`launches` aggregates a `--metrics gpu__time_duration.sum --csv` log into per-kernel time
shares; `report` extracts the roofline-relevant counters of every captured launch from a
`--set full` report (read with `ncu -i ... --page raw --csv`, no GPU needed).
"""
from __future__ import annotations

import collections
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__cycles_active.avg", "sm__cycles_active.avg",
    "smsp__inst_executed.sum",
]


def _to_us(v: float, unit: str) -> float:
    return {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(unit, v)


def launches(src: str, dst: str, top: int = 45):
    with open(src) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = _to_us(float(row["Metric Value"].replace(",", "")), row["Metric Unit"])
        a = agg[row["Kernel Name"]]
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    n = sum(v[0] for v in agg.values())
    out = [f"# kernel launches: {n} launches, {tot / 1e3:.2f} ms of device time (serialised, cold cache: compare shares)",
           "", "| share | total us | launches | avg us | kernel |", "|---|---|---|---|---|"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        out.append(f"| {v[1] / tot * 100:.2f}% | {v[1]:.1f} | {v[0]} | {v[1] / v[0]:.1f} | `{k[:120]}` |")
    own = sum(v[1] for k, v in agg.items() if "b200::" in k)
    out += ["", f"own kernels (b200::*): {own / tot * 100:.1f}% of device time"]
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out[:25]))


def report(src: str, dst: str):
    if src.endswith(".csv"):     # already exported on the GPU box: ncu -i x.ncu-rep --page raw --csv
        text = open(src).read()
    else:
        text = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(text.splitlines()) if len(r) > 10]
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    out = [f"# ncu --set full summary of {src}", ""]
    for row in data:
        out.append(f"## {row[idx['Kernel Name']][:140]}")
        out.append("")
        out.append("| metric | value | unit |")
        out.append("|---|---|---|")
        for k in KEYS:
            if k in idx:
                out.append(f"| {k} | {row[idx[k]]} | {units[idx[k]]} |")
        out.append("")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out[:60]))


if __name__ == "__main__":
    {"launches": launches, "report": report}[sys.argv[1]](sys.argv[2], sys.argv[3])
