"""JSON metrics export and aggregation.

Record formats are kept field-compatible with the reference so existing tooling
keeps working (reference: src/parameter_server/server.py:346-367 SERVER_FINAL_METRICS,
src/workers/worker.py:419-434 WORKER_FINAL_METRICS, and the aggregate result schema
of scripts/parse_cloudwatch_logs.py:125-203).  Records are printed as one
``METRICS_JSON: {...}`` line (the reference's scraping contract,
parse_cloudwatch_logs.py:100) and, unlike the reference, also written to files so
no CloudWatch round-trip is needed.  Extra B200 fields: images_per_second,
push/pull GB/s, roofline fractions, staleness histogram.
"""
from __future__ import annotations

import json
import os
import re
import time
from datetime import datetime, timezone
from typing import Dict, Iterable, List, Optional

METRICS_PREFIX = "METRICS_JSON:"
_METRICS_RE = re.compile(r"METRICS_JSON:\s*(\{.*\})")


def server_final_metrics(*, mode, total_workers, total_time, global_step, total_updates,
                         gradients_processed, avg_update_time, learning_rate,
                         extra: Optional[dict] = None) -> dict:
    rec = {
        "type": "SERVER_FINAL_METRICS",
        "mode": mode,
        "total_workers": total_workers,
        "total_training_time_seconds": round(total_time, 2),
        "global_steps_completed": global_step,
        "total_parameter_updates": total_updates,
        "gradients_processed": gradients_processed,
        "average_update_time_seconds": round(avg_update_time, 4),
        "updates_per_second": round(total_updates / total_time, 2) if total_time > 0 else 0.0,
        "learning_rate": learning_rate,
    }
    if extra:
        rec.update(extra)
    return rec


def worker_final_metrics(*, worker_id, total_workers, total_time, epoch_times, accuracies,
                         local_steps, batch_size, learning_rate, num_epochs,
                         extra: Optional[dict] = None) -> dict:
    avg_epoch = sum(epoch_times) / len(epoch_times) if epoch_times else 0.0
    rec = {
        "type": "WORKER_FINAL_METRICS",
        "worker_id": worker_id,
        "total_workers": total_workers,
        "total_training_time_seconds": round(total_time, 2),
        "average_epoch_time_seconds": round(avg_epoch, 2),
        "epoch_times_seconds": [round(t, 2) for t in epoch_times],
        "final_test_accuracy_percent": round(accuracies[-1], 2) if accuracies else 0,
        "all_accuracies_percent": [round(a, 2) for a in accuracies],
        "local_steps_completed": local_steps,
        "batch_size": batch_size,
        "learning_rate": learning_rate,
        "num_epochs": num_epochs,
    }
    if extra:
        rec.update(extra)
    return rec


def emit(record: dict, metrics_dir: Optional[str] = None, tag: str = "") -> str:
    """Print the METRICS_JSON line and optionally persist it. Returns the line."""
    line = f"{METRICS_PREFIX} {json.dumps(record)}"
    print(line, flush=True)
    if metrics_dir:
        os.makedirs(metrics_dir, exist_ok=True)
        name = f"{tag or record.get('type', 'metrics').lower()}"
        if "worker_id" in record:
            name += f"_{record['worker_id']}"
        with open(os.path.join(metrics_dir, name + ".json"), "w") as f:
            json.dump(record, f, indent=2)
    return line


def parse_metrics_lines(lines: Iterable[str]) -> Dict[str, List[dict]]:
    """Split METRICS_JSON lines into server / worker records (parse_cloudwatch_logs.py:92-123)."""
    out = {"server": [], "worker": []}
    for line in lines:
        m = _METRICS_RE.search(line)
        if not m:
            continue
        try:
            rec = json.loads(m.group(1))
        except json.JSONDecodeError:
            continue
        t = rec.get("type")
        if t == "SERVER_FINAL_METRICS":
            out["server"].append(rec)
        elif t == "WORKER_FINAL_METRICS":
            out["worker"].append(rec)
    return out


def aggregate_workers(worker_records: List[dict]) -> Optional[dict]:
    """Aggregate schema of the reference (parse_cloudwatch_logs.py:125-177), minus its
    phantom-worker pollution: duplicate ids keep the last record and zero-step workers
    (ids >= W that died instantly) are excluded from the means."""
    if not worker_records:
        return None
    by_id: Dict[int, dict] = {}
    for r in worker_records:
        by_id[int(r.get("worker_id", len(by_id)))] = r
    recs = [r for _, r in sorted(by_id.items()) if r.get("local_steps_completed", 0) > 0] \
        or [r for _, r in sorted(by_id.items())]
    n = len(recs)
    max_epochs = max(len(r.get("epoch_times_seconds", [])) for r in recs)
    epoch_rows, acc_rows = [], []
    for e in range(max_epochs):
        ts = [r["epoch_times_seconds"][e] for r in recs if len(r.get("epoch_times_seconds", [])) > e]
        ac = [r["all_accuracies_percent"][e] for r in recs if len(r.get("all_accuracies_percent", [])) > e]
        if ts:
            epoch_rows.append({"epoch": e + 1, "max_time": max(ts), "avg_time": sum(ts) / len(ts),
                               "min_time": min(ts)})
        if ac:
            acc_rows.append({"epoch": e + 1, "avg_accuracy": sum(ac) / len(ac), "max_accuracy": max(ac),
                             "min_accuracy": min(ac)})
    agg = {
        "num_workers": n,
        "total_training_time_seconds": max(r["total_training_time_seconds"] for r in recs),
        "average_epoch_time_seconds": sum(r["average_epoch_time_seconds"] for r in recs) / n,
        "final_test_accuracy_percent": sum(r["final_test_accuracy_percent"] for r in recs) / n,
        "total_local_steps": sum(r["local_steps_completed"] for r in recs),
        "per_worker_metrics": recs,
        "epoch_times_by_epoch": epoch_rows,
        "accuracy_by_epoch": acc_rows,
    }
    ips = [r["images_per_second"] for r in recs if "images_per_second" in r]
    if ips:
        agg["images_per_second_sum"] = sum(ips)
    return agg


def build_result(experiment_name: str, server_records: List[dict], worker_records: List[dict]) -> dict:
    """Result-file schema (parse_cloudwatch_logs.py:197-203)."""
    return {
        "experiment_name": experiment_name,
        "timestamp": datetime.now(timezone.utc).isoformat(),
        "server_metrics": server_records[-1] if server_records else None,
        "worker_metrics_aggregated": aggregate_workers(worker_records),
        "raw_worker_metrics": worker_records,
    }


def collect_dir(metrics_dir: str, experiment_name: str, out_path: Optional[str] = None) -> dict:
    """Gather the per-process JSON files (or *.log files with METRICS_JSON lines) of a run."""
    servers, workers = [], []
    for fn in sorted(os.listdir(metrics_dir)):
        p = os.path.join(metrics_dir, fn)
        if fn.endswith(".json"):
            with open(p) as f:
                try:
                    rec = json.load(f)
                except json.JSONDecodeError:
                    continue
            if rec.get("type") == "SERVER_FINAL_METRICS":
                servers.append(rec)
            elif rec.get("type") == "WORKER_FINAL_METRICS":
                workers.append(rec)
        elif fn.endswith(".log") or fn.endswith(".txt"):
            with open(p, errors="replace") as f:
                parsed = parse_metrics_lines(f)
            servers += parsed["server"]
            workers += parsed["worker"]
    result = build_result(experiment_name, servers, workers)
    if out_path:
        os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
        with open(out_path, "w") as f:
            json.dump(result, f, indent=2)
    return result


class PhaseTimer:
    """Wall-clock phase accumulator for the CPU path (device path uses CUDA events)."""

    def __init__(self):
        self.totals: Dict[str, float] = {}
        self.counts: Dict[str, int] = {}

    def add(self, name: str, dt: float) -> None:
        self.totals[name] = self.totals.get(name, 0.0) + dt
        self.counts[name] = self.counts.get(name, 0) + 1

    class _Ctx:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __enter__(self):
            self.t0 = time.perf_counter()
            return self

        def __exit__(self, *exc):
            self.owner.add(self.name, time.perf_counter() - self.t0)

    def phase(self, name: str) -> "PhaseTimer._Ctx":
        return PhaseTimer._Ctx(self, name)

    def summary(self) -> Dict[str, dict]:
        return {k: {"total_s": round(v, 6), "count": self.counts[k],
                    "mean_s": round(v / self.counts[k], 6)} for k, v in self.totals.items()}
