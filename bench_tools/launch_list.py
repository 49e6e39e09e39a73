#!/usr/bin/env python
"""Turn an `ncu --csv --log-file` launch list (long format: one row per launch and metric) into a markdown table per
kernel: launches, total / mean duration, mean tensor-pipe activity, L2 and DRAM throughput.

    python bench_tools/launch_list.py gpurun_out/launches_v6.csv profiles/launches_v6_batch128.md "title" ["note"]
"""
import collections
import csv
import re
import sys


def parse(path):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.reader(lines)
    head = None
    for r in rd:
        if head is None:
            if "Kernel Name" in r and "Metric Name" in r:
                head = {n: i for i, n in enumerate(r)}
            continue
        if len(r) < len(head):
            continue
        rows.append(r)
    launches = collections.OrderedDict()
    for r in rows:
        key = r[head["ID"]]
        d = launches.setdefault(key, {"name": r[head["Kernel Name"]], "grid": r[head.get("Grid Size", head["ID"])],
                                      "block": r[head.get("Block Size", head["ID"])]})
        try:
            v = float(r[head["Metric Value"]].replace(",", ""))
        except ValueError:
            continue
        unit = r[head["Metric Unit"]]
        name = r[head["Metric Name"]]
        if name == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
        d[name] = v
    return list(launches.values())


def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("dps::", "").replace("void ", "")
    return name.strip()


def main():
    src, out, title = sys.argv[1:4]
    note = sys.argv[4] if len(sys.argv) > 4 else ""
    L = parse(src)
    T, TP, L2, DR = ("gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                     "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed")
    agg = collections.OrderedDict()
    for d in L:
        a = agg.setdefault(short(d["name"]), {"n": 0, "t": 0.0, "tp": 0.0, "l2": 0.0, "dr": 0.0, "min_tp": 1e9})
        a["n"] += 1
        a["t"] += d.get(T, 0.0)
        a["tp"] += d.get(TP, 0.0)
        a["l2"] += d.get(L2, 0.0)
        a["dr"] += d.get(DR, 0.0)
        a["min_tp"] = min(a["min_tp"], d.get(TP, 0.0))
    tot = sum(a["t"] for a in agg.values())
    with open(out, "w") as f:
        f.write(f"# {title}\n\n{note}\n\nkernels in the window: {len(L)}, sum of durations: {tot:.0f} us\n\n")
        f.write("| kernel | launches | total us | share | mean us | mean tensor-pipe active | min tensor-pipe | mean L2 throughput | mean DRAM throughput |\n")
        f.write("|---|---:|---:|---:|---:|---:|---:|---:|---:|\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
            n = a["n"]
            f.write(f"| `{k}` | {n} | {a['t']:.1f} | {100 * a['t'] / tot:.1f}% | {a['t'] / n:.2f} | {a['tp'] / n:.1f}% | "
                    f"{a['min_tp']:.1f}% | {a['l2'] / n:.1f}% | {a['dr'] / n:.1f}% |\n")
        f.write("\n## every launch, in order\n\n| # | kernel | grid | us | tensor-pipe active |\n|---:|---|---|---:|---:|\n")
        for i, d in enumerate(L):
            f.write(f"| {i} | `{short(d['name'])}` | {d['grid']} | {d.get(T, 0.0):.2f} | {d.get(TP, 0.0):.1f}% |\n")
    print(open(out).read()[:3000])


if __name__ == "__main__":
    main()
