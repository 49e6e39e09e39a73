"""Result aggregation CLI + headless charts (PNG files only).

Covers the role of the reference's scripts/parse_cloudwatch_logs.py (collect METRICS_JSON records into a
result file) and scripts/visualize_results.py (sync-vs-async comparison per worker count; scaling figure
with time / epoch time / accuracy / speed-up vs workers on log2 axes with the ideal-linear line; console
summary table) -- reference: scripts/visualize_results.py:77-296.  No CloudWatch, no matplotlib, no window.

    python -m dps_b200.utils.charts collect --metrics-dir out/sync8 --name sync_8workers --output results/sync_8workers.json
    python -m dps_b200.utils.charts plot --results-dir results --output-dir charts
"""
from __future__ import annotations

import argparse
import glob
import json
import math
import os
from typing import Dict, List, Optional, Sequence, Tuple

from . import metrics as M
from .png import Canvas

PALETTE = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189)]


def load_results(results_dir: Optional[str] = None, files: Optional[Sequence[str]] = None) -> List[dict]:
    paths = list(files or []) + (sorted(glob.glob(os.path.join(results_dir, "*.json"))) if results_dir else [])
    out = []
    for p in paths:
        try:
            with open(p) as f:
                r = json.load(f)
        except (OSError, json.JSONDecodeError):
            continue
        if "worker_metrics_aggregated" in r:
            out.append(r)
    return out


def result_key(res: dict) -> Tuple[str, int]:
    """(mode, num_workers).  The reference needs server_metrics for this and silently plots nothing when
    it is null (every shipped result); fall back to the worker aggregate and the experiment name."""
    sm = res.get("server_metrics") or {}
    agg = res.get("worker_metrics_aggregated") or {}
    name = res.get("experiment_name", "")
    mode = sm.get("mode") or ("async" if "async" in name else "sync")
    return mode, int(sm.get("total_workers") or agg.get("num_workers") or 0)


def summary_rows(results: List[dict]) -> List[dict]:
    rows = []
    for r in results:
        mode, w = result_key(r)
        agg = r.get("worker_metrics_aggregated") or {}
        rows.append({"experiment": r.get("experiment_name", "?"), "mode": mode, "workers": w,
                     "total_time_s": agg.get("total_training_time_seconds", 0.0),
                     "avg_epoch_s": agg.get("average_epoch_time_seconds", 0.0),
                     "final_acc": agg.get("final_test_accuracy_percent", 0.0),
                     "images_per_s": agg.get("images_per_second_sum", 0.0)})
    return sorted(rows, key=lambda x: (x["mode"], x["workers"]))


def print_summary(results: List[dict]) -> str:
    lines = ["%-28s %-6s %7s %12s %12s %9s %12s" % ("experiment", "mode", "workers", "total_s", "epoch_s", "acc_%",
                                                      "img/s")]
    for r in summary_rows(results):
        lines.append("%-28s %-6s %7d %12.2f %12.2f %9.2f %12.1f" % (r["experiment"][:28], r["mode"], r["workers"],
                                                                   r["total_time_s"], r["avg_epoch_s"], r["final_acc"],
                                                                   r["images_per_s"]))
    text = "\n".join(lines)
    print(text)
    return text


def _axes(cv: Canvas, x0, y0, x1, y1, title: str, xlabel: str, ylabel: str):
    cv.rect(x0, y0, x1, y1, (250, 250, 250))
    cv.line(x0, y1, x1, y1, (0, 0, 0))
    cv.line(x0, y0, x0, y1, (0, 0, 0))
    cv.text(x0 + 4, y0 - 14, title, scale=1)
    cv.text((x0 + x1) // 2 - 3 * len(xlabel), y1 + 16, xlabel)
    cv.text(x0 - 50, y0 - 14, ylabel)


def _plot_series(cv: Canvas, box, series: Dict[str, List[Tuple[float, float]]], title, xlabel, ylabel,
                 logx: bool = False, ideal: Optional[List[Tuple[float, float]]] = None):
    x0, y0, x1, y1 = box
    _axes(cv, x0, y0, x1, y1, title, xlabel, ylabel)
    pts = [p for s in series.values() for p in s] + (ideal or [])
    if not pts:
        return
    fx = (lambda v: math.log2(max(v, 1e-9))) if logx else (lambda v: v)
    xs, ys = [fx(p[0]) for p in pts], [p[1] for p in pts]
    xmin, xmax, ymin, ymax = min(xs), max(xs), min(0.0, min(ys)), max(ys) * 1.1 + 1e-9
    if xmax == xmin:
        xmax = xmin + 1
    px = lambda v: x0 + 10 + (fx(v) - xmin) / (xmax - xmin) * (x1 - x0 - 20)
    py = lambda v: y1 - (v - ymin) / (ymax - ymin) * (y1 - y0 - 10)
    for t in range(5):
        yv = ymin + (ymax - ymin) * t / 4
        cv.line(x0, py(yv), x1, py(yv), (225, 225, 225))
        cv.text(x0 - 48, int(py(yv)) - 3, ("%.3g" % yv))
    for xv in sorted({p[0] for p in pts}):
        cv.text(int(px(xv)) - 6, y1 + 4, ("%g" % xv))
    if ideal:
        for a, b in zip(ideal, ideal[1:]):
            cv.line(px(a[0]), py(a[1]), px(b[0]), py(b[1]), (160, 160, 160))
    for i, (name, s) in enumerate(series.items()):
        col = PALETTE[i % len(PALETTE)]
        s = sorted(s)
        for a, b in zip(s, s[1:]):
            cv.line(px(a[0]), py(a[1]), px(b[0]), py(b[1]), col, 2)
        for p in s:
            cv.marker(px(p[0]), py(p[1]), col)
        cv.rect(x1 - 120, y0 + 6 + 12 * i, x1 - 110, y0 + 14 + 12 * i, col)
        cv.text(x1 - 106, y0 + 6 + 12 * i, name)


def plot_scaling(results: List[dict], path: str) -> str:
    """2x2: total time, epoch time, accuracy, speed-up vs workers (log2 x axis, ideal-linear line)."""
    rows = summary_rows(results)
    cv = Canvas(1000, 760)
    modes = sorted({r["mode"] for r in rows})
    def ser(key):
        return {m: [(r["workers"], r[key]) for r in rows if r["mode"] == m and r["workers"] > 0] for m in modes}
    boxes = [(70, 40, 480, 330), (570, 40, 980, 330), (70, 420, 480, 710), (570, 420, 980, 710)]
    _plot_series(cv, boxes[0], ser("total_time_s"), "TOTAL TRAINING TIME (S)", "WORKERS", "S", logx=True)
    _plot_series(cv, boxes[1], ser("avg_epoch_s"), "AVG EPOCH TIME (S)", "WORKERS", "S", logx=True)
    _plot_series(cv, boxes[2], ser("final_acc"), "FINAL TEST ACCURACY (%)", "WORKERS", "%", logx=True)
    speed = {}
    for m in modes:
        pts = sorted((r["workers"], r["images_per_s"] or (1.0 / max(r["avg_epoch_s"], 1e-9))) for r in rows
                     if r["mode"] == m and r["workers"] > 0)
        if pts:
            base_w, base_v = pts[0]
            speed[m] = [(w, v / base_v * base_w) for w, v in pts]
    ws = sorted({w for s in speed.values() for w, _ in s})
    _plot_series(cv, boxes[3], speed, "SPEED-UP VS WORKERS", "WORKERS", "X", logx=True,
                 ideal=[(w, float(w)) for w in ws])
    cv.save(path)
    return path


def plot_sync_vs_async(results: List[dict], workers: int, path: str) -> Optional[str]:
    """2x2 comparison for one worker count: epoch times, accuracy by epoch, total time and throughput bars."""
    sel = {result_key(r)[0]: r for r in results if result_key(r)[1] == workers}
    if not sel:
        return None
    cv = Canvas(1000, 760)
    ep, acc = {}, {}
    for mode, r in sel.items():
        agg = r["worker_metrics_aggregated"]
        ep[mode] = [(e["epoch"], e["max_time"]) for e in agg.get("epoch_times_by_epoch", [])]
        acc[mode] = [(e["epoch"], e["avg_accuracy"]) for e in agg.get("accuracy_by_epoch", [])]
    boxes = [(70, 40, 480, 330), (570, 40, 980, 330), (70, 420, 480, 710), (570, 420, 980, 710)]
    _plot_series(cv, boxes[0], ep, "EPOCH TIME, MAX OVER WORKERS (S) - %d WORKERS" % workers, "EPOCH", "S")
    _plot_series(cv, boxes[1], acc, "TEST ACCURACY BY EPOCH (%)", "EPOCH", "%")
    for bi, key, title in ((2, "total_training_time_seconds", "TOTAL TIME (S)"), (3, "images_per_second_sum", "IMAGES/S")):
        x0, y0, x1, y1 = boxes[bi]
        _axes(cv, x0, y0, x1, y1, title, "MODE", "")
        vals = {m: float(r["worker_metrics_aggregated"].get(key, 0.0) or 0.0) for m, r in sel.items()}
        top = max(vals.values()) * 1.15 + 1e-9
        for i, (m, v) in enumerate(sorted(vals.items())):
            bx = x0 + 60 + i * 140
            cv.rect(bx, int(y1 - v / top * (y1 - y0 - 10)), bx + 80, y1, PALETTE[i])
            cv.text(bx + 10, y1 + 4, m)
            cv.text(bx, int(y1 - v / top * (y1 - y0 - 10)) - 10, "%.4g" % v)
    cv.save(path)
    return path


def main(argv=None) -> int:
    p = argparse.ArgumentParser(description="collect METRICS_JSON records / draw charts")
    sub = p.add_subparsers(dest="cmd", required=True)
    c = sub.add_parser("collect")
    c.add_argument("--metrics-dir", required=True)
    c.add_argument("--name", "--experiment-name", dest="name", required=True)
    c.add_argument("--output", required=True)
    q = sub.add_parser("plot")
    q.add_argument("--results-dir", default=None)
    q.add_argument("--experiments", nargs="*", default=None)
    q.add_argument("--output-dir", default="charts")
    q.add_argument("--compare-workers", type=int, nargs="*", default=None)
    a = p.parse_args(argv)
    if a.cmd == "collect":
        res = M.collect_dir(a.metrics_dir, a.name, a.output)
        print(f"wrote {a.output}: {(res['worker_metrics_aggregated'] or {}).get('num_workers', 0)} workers")
        return 0
    results = load_results(a.results_dir, a.experiments)
    if not results:
        print("no result files found")
        return 1
    os.makedirs(a.output_dir, exist_ok=True)
    print_summary(results)
    out = [plot_scaling(results, os.path.join(a.output_dir, "scaling.png"))]
    counts = a.compare_workers or sorted({result_key(r)[1] for r in results})
    for w in counts:
        pth = plot_sync_vs_async(results, w, os.path.join(a.output_dir, f"sync_vs_async_{w}workers.png"))
        if pth:
            out.append(pth)
    print("charts:", ", ".join(out))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
