#!/usr/bin/env python
"""One table row per captured launch from an `ncu --set full` raw page exported as CSV on the GPU box
(`ncu -i x.ncu-rep --page raw --csv > x_raw.csv`; the .ncu-rep itself is often larger than what gpurun copies back).

    python bench_tools/ncu_raw_table.py gpurun_out/prof_r2_raw.csv profiles/ncu_top_kernels_r2_final.md "title" ["note"]

Roofline columns: tensor-pipe activity is the fraction of the tcgen05 issue rate (what the cuBLAS bf16 number in
MEASURED_PEAKS.json also runs on); DRAM traffic / duration is set against the measured HBM copy bandwidth.
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PEAKS = {"hbm_gbs": 6553.9, "bf16_tflops": 1669.7, "bf16_tflops_sustained": 1403.6}
if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
    PEAKS.update(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))))

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "nsecond": 1e-9, "us": 1e-6,
         "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "second": 1.0}


def short(name):
    name = re.sub(r"\(.*$", "", name).replace("void ", "").replace("dps::", "")
    return name.strip()


def main():
    src, out, title = sys.argv[1:4]
    note = sys.argv[4] if len(sys.argv) > 4 else ""
    rows = [r for r in csv.reader(open(src, newline="")) if r and not r[0].startswith("==")]
    head, units, data = rows[0], rows[1], rows[2:]
    idx = {n: i for i, n in enumerate(head)}

    def val(d, key, default=float("nan")):
        if key not in idx:
            return default
        try:
            v = float(d[idx[key]].replace(",", ""))
        except ValueError:
            return default
        return v * SCALE.get(units[idx[key]].split("/")[0], 1.0)

    stall_keys = [k for k in head if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")]
    with open(out, "w") as f:
        f.write(f"# {title}\n\n{note}\n\n")
        f.write(f"Source: `{os.path.basename(src)}` = raw page of an `ncu --set full --clock-control none` capture, exported on the "
                f"GPU box.  Denominators (MEASURED_PEAKS.json): HBM copy {PEAKS['hbm_gbs']:.0f} GB/s, cuBLAS bf16 "
                f"{PEAKS['bf16_tflops_sustained']:.0f} TFLOP/s sustained ({PEAKS['bf16_tflops']:.0f} burst).  ncu serialises "
                f"launches and replays each ~40 times with cold-ish caches: durations are longer than inside the CUDA-graph step.\n\n")
        f.write("| # | kernel | grid x block | regs | smem KB | us | tensor pipe active (of elapsed) | SM busy | L2 | DRAM | DRAM GB/s "
                "(of measured peak) | warps active | top issue stalls (warps per issue) |\n")
        f.write("|---:|---|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---|\n")
        agg = {}
        for i, d in enumerate(data):
            name = short(d[idx["Kernel Name"]])
            t = val(d, "gpu__time_duration.sum")
            tp_a = val(d, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
            tp_e = val(d, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed")
            smb = val(d, "sm__throughput.avg.pct_of_peak_sustained_elapsed")
            l2 = val(d, "lts__throughput.avg.pct_of_peak_sustained_elapsed")
            dr = val(d, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                     val(d, "dram__throughput.avg.pct_of_peak_sustained_elapsed"))
            byts = val(d, "dram__bytes_read.sum", 0.0) + val(d, "dram__bytes_write.sum", 0.0)
            gbs = byts / t / 1e9 if t == t and t > 0 else float("nan")
            wa = val(d, "sm__warps_active.avg.pct_of_peak_sustained_active")
            regs = val(d, "launch__registers_per_thread")
            smem = (val(d, "launch__shared_mem_per_block_dynamic", 0.0) + val(d, "launch__shared_mem_per_block_static", 0.0)) / 1e3
            stalls = sorted(((val(d, k, 0.0), k) for k in stall_keys), reverse=True)[:3]
            st = ", ".join(f"{k.split('issue_stalled_')[1].replace('_per_issue_active.ratio', '')} {v:.1f}" for v, k in stalls if v > 0)
            f.write(f"| {i} | `{name}` | {d[idx['Grid Size']]} x {d[idx['Block Size']]} | {regs:.0f} | {smem:.0f} | {t * 1e6:.2f} | "
                    f"{tp_a:.1f}% ({tp_e:.1f}%) | {smb:.1f}% | {l2:.1f}% | {dr:.1f}% | {gbs:.0f} ({gbs / PEAKS['hbm_gbs']:.2f}) | "
                    f"{wa:.1f}% | {st} |\n")
            a = agg.setdefault(name, [0, 0.0, 0.0, 0.0, 0.0])
            a[0] += 1
            a[1] += t * 1e6
            a[2] += tp_a
            a[3] += gbs
            a[4] += tp_e
        nominal = 2250.0       # dense bf16 TFLOP/s of the tensor pipe at 100 % activity on every SM
        f.write("\n## per kernel\n\nAchieved rate = tensor-pipe activity over the ELAPSED cycles of all 148 SMs x the nominal dense bf16 peak "
                f"({nominal:.0f} TFLOP/s); cuBLAS itself sustains {PEAKS['bf16_tflops_sustained']:.0f} on a large GEMM.\n\n"
                "| kernel | launches | mean us | tensor pipe active (while the SM is busy) | over elapsed | achieved TFLOP/s | of the sustained "
                "cuBLAS rate | mean DRAM GB/s |\n|---|---:|---:|---:|---:|---:|---:|---:|\n")
        for name, (n, t, tp, gb, te) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            ach = te / n / 100 * nominal
            f.write(f"| `{name}` | {n} | {t / n:.2f} | {tp / n:.1f}% | {te / n:.1f}% | {ach:.0f} | "
                    f"{ach / PEAKS['bf16_tflops_sustained']:.2f} | {gb / n:.0f} |\n")
    print(open(out).read()[:2500])


if __name__ == "__main__":
    main()
