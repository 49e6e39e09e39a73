#!/usr/bin/env python
"""Summarise .ncu-rep captures (read on the CPU box with `ncu -i`) into profiles/<name>.md."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio" ]


def summarize(rep: str, out_md: str, title: str, note: str = "") -> None:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, data = rows[0], rows[1], rows[2:]
    idx = {n: i for i, n in enumerate(head)}
    with open(out_md, "w") as f:
        f.write(f"# {title}\n\nSource: `{os.path.basename(rep)}` (ncu --set full --clock-control none, B200, read with `ncu -i ... --page raw --csv`).\n")
        f.write(f"Roofline denominators (MEASURED_PEAKS.json): HBM {PEAKS.get('hbm_gbs')} GB/s, bf16 {PEAKS.get('bf16_tflops')} TFLOP/s.\n\n{note}\n\n")
        for d in data:
            name = d[idx["Kernel Name"]].split("(")[0]
            f.write(f"## `{name}`\n\n| metric | value |\n|---|---|\n")
            for k in KEYS:
                if k in idx:
                    f.write(f"| {k} | {d[idx[k]]} {units[idx[k]]} |\n")
            try:
                t_us = float(d[idx["gpu__time_duration.sum"]].replace(",", ""))
                unit = units[idx["gpu__time_duration.sum"]]
                t_s = t_us * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "usecond": 1e-6, "nsecond": 1e-9}.get(unit, 1e-6)
                def b(k):
                    v, u = float(d[idx[k]].replace(",", "")), units[idx[k]]
                    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                tot = b("dram__bytes_read.sum") + b("dram__bytes_write.sum")
                f.write(f"| DRAM traffic / duration | {tot / t_s / 1e9:.0f} GB/s = {tot / t_s / 1e9 / PEAKS['hbm_gbs']:.2f} of measured HBM copy peak |\n")
            except Exception:
                pass
            f.write("\n")


if __name__ == "__main__":
    rep, out_md, title = sys.argv[1:4]
    summarize(rep, out_md, title, sys.argv[4] if len(sys.argv) > 4 else "")
    print(open(out_md).read()[:1500])
