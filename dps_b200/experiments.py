"""Experiment pipeline on one box: baseline -> {sync, async} x worker counts -> result files -> charts.

The reference's pipeline is human-driven and multi-host (reference: EXPERIMENT_GUIDE.md; terraform/deploy.sh to
bring up 1 server + W Fargate workers, scripts/parse_cloudwatch_logs.py to scrape METRICS_JSON lines into
experiment_results/<mode>_<W>workers.json, scripts/visualize_results.py for the sync-vs-async and scaling figures,
baseline/baseline_training.py for the single-machine number).  Here the same sequence is one command on one
NVSwitch box; every run is a `dps_b200.launch` job, its per-rank records are aggregated into the reference's
result-file schema, and the figures are headless PNGs.

    python -m dps_b200.experiments --workers 1,2,4,8 --modes sync,async --epochs 3 --out experiment_results
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys
import time
from typing import List, Optional

from .utils import charts as C


def run_one(mode: str, workers: int, out: str, extra: List[str], port: int, timeout_s: float) -> Optional[dict]:
    name = f"{mode}_{workers}workers"
    mdir = os.path.join(out, "raw", name)
    shutil.rmtree(mdir, ignore_errors=True)
    os.makedirs(mdir, exist_ok=True)
    cmd = [sys.executable, "-m", "dps_b200.launch", "--gpus", str(workers), "--master-port", str(port), "--",
           "--mode", mode, "--workers", str(workers), "--metrics-dir", mdir, "--experiment-name", name] + extra
    t0 = time.time()
    with open(os.path.join(out, "raw", name + ".stdout.log"), "w") as log:
        try:
            rc = subprocess.call(cmd, stdout=log, stderr=subprocess.STDOUT, timeout=timeout_s)
        except subprocess.TimeoutExpired:
            rc = -1
    res_path = os.path.join(mdir, name + ".result.json")
    if rc != 0 or not os.path.exists(res_path):
        print(f"[experiments] {name}: FAILED rc={rc} (see {out}/raw/{name}.stdout.log)", flush=True)
        return None
    dst = os.path.join(out, name + ".json")
    shutil.copyfile(res_path, dst)
    res = json.load(open(dst))
    agg = res.get("worker_metrics_aggregated") or {}
    print(f"[experiments] {name}: {agg.get('images_per_second_sum', 0):.0f} img/s, "
          f"final acc {agg.get('final_test_accuracy_percent', float('nan')):.2f}% "
          f"({time.time() - t0:.1f}s wall incl. start-up)", flush=True)
    return res


def main(argv=None) -> int:
    p = argparse.ArgumentParser(description="run the sync/async x workers experiment grid on this box")
    p.add_argument("--workers", default="1,2,4,8")
    p.add_argument("--modes", default="sync,async")
    p.add_argument("--epochs", type=int, default=3)
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--lr", type=float, default=0.1)
    p.add_argument("--train-samples", type=int, default=50000)
    p.add_argument("--test-samples", type=int, default=10000)
    p.add_argument("--dataset", default="synthetic")
    p.add_argument("--out", default="experiment_results")
    p.add_argument("--no-baseline", action="store_true")
    p.add_argument("--timeout-s", type=float, default=900.0)
    p.add_argument("--master-port", type=int, default=29611)
    a, rest = p.parse_known_args(argv)
    os.makedirs(a.out, exist_ok=True)
    extra = ["--epochs", str(a.epochs), "--batch-size", str(a.batch_size), "--lr", str(a.lr),
             "--train-samples", str(a.train_samples), "--test-samples", str(a.test_samples),
             "--dataset", a.dataset] + [r for r in rest if r != "--"]
    if a.dataset == "synthetic":                     # generate once here; every rank of every run then loads the cache
        from .utils import data as D
        D.load_dataset("synthetic", True, a.train_samples, 100, 32, 0)
        D.load_dataset("synthetic", False, a.test_samples, 100, 32, 0)
    if not a.no_baseline:
        from . import baseline_trainer as BT
        BT.main(["--epochs", str(a.epochs), "--batch-size", str(a.batch_size), "--lr", str(a.lr),
                 "--train-samples", str(a.train_samples), "--test-samples", str(a.test_samples),
                 "--dataset", a.dataset, "--output-dir", os.path.join(a.out, "baseline")])
    port = a.master_port
    done = []
    for mode in [m for m in a.modes.split(",") if m]:
        for w in [int(t) for t in a.workers.split(",") if t]:
            port += 1
            if run_one(mode, w, a.out, extra, port, a.timeout_s) is not None:
                done.append((mode, w))
    results = C.load_results(a.out)
    if results:
        cdir = os.path.join(a.out, "charts")
        os.makedirs(cdir, exist_ok=True)
        C.plot_scaling(results, os.path.join(cdir, "scaling_analysis.png"))
        for w in sorted({w for _, w in done}):
            C.plot_sync_vs_async(results, w, os.path.join(cdir, f"comparison_{w}workers.png"))
        with open(os.path.join(a.out, "summary.txt"), "w") as f:
            f.write(C.print_summary(results))
    return 0 if done else 1


if __name__ == "__main__":
    raise SystemExit(main())
