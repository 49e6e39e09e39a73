"""`bench.py --impl reference`: the reference's own gRPC parameter server + workers, unmodified, run
offline from baseline/_ref on localhost (CPU: the reference has no GPU path in its distributed code).

1 server process + N worker processes (N = --gpus), batch 128, synthetic CIFAR-100-shaped data sized so
that one epoch is exactly K steps per worker; with warm-up an extra first epoch is run and discarded.
images/sec = N * K * batch / max over workers of the timed epoch's `epoch_times_seconds` entry from the
workers' own METRICS_JSON lines (reference: src/workers/worker.py:386-387,419-434).
"""
from __future__ import annotations

import json
import os
import re
import socket
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
PUBLISHED = {"sync": 82.7, "async": 168.6}


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_reference(args) -> dict:
    server_py = os.path.join(REF, "src", "parameter_server", "server.py")
    worker_py = os.path.join(REF, "src", "workers", "worker.py")
    if not (os.path.exists(server_py) and os.path.exists(worker_py)):
        return {"impl": "reference", "unavailable": "baseline/_ref is missing (reference is not pip-installable: "
                "no setup.py/pyproject.toml; vendored copy absent)"}
    try:
        import grpc  # noqa: F401
        import torchvision  # noqa: F401
        from PIL import Image  # noqa: F401
    except Exception as e:  # pragma: no cover
        return {"impl": "reference", "unavailable": f"reference dependency missing: {e}"}
    N, K, W, B = args.gpus, args.steps, args.warmup, args.batch
    epochs = 2 if W > 0 else 1
    # the reference needs seconds per step (CPU + pickle over gRPC): cap the timed epoch so the arm finishes
    steps_per_epoch = min(max(K, W), 10 if N >= 4 else 20)
    train_samples = N * B * steps_per_epoch
    port = _free_port()
    launch = os.path.join(ROOT, "bench_tools", "ref_launch.py")
    env = dict(os.environ)
    ncpu = os.cpu_count() or 8
    env["OMP_NUM_THREADS"] = env["MKL_NUM_THREADS"] = str(max(1, ncpu // max(1, N)))
    env["PYTHONPATH"] = os.path.join(REF, "src") + os.pathsep + env.get("PYTHONPATH", "")
    env["PYTHONUNBUFFERED"] = "1"
    logdir = os.path.join(ROOT, "gpurun_out", "reference_arm")
    os.makedirs(logdir, exist_ok=True)
    common = [sys.executable, launch, str(train_samples), str(B)]
    srv_log = open(os.path.join(logdir, "server.log"), "w")
    srv = subprocess.Popen(common + [server_py, "--mode", args.mode, "--workers", str(N), "--port", str(port),
                                     "--lr", "0.1"], env=env, stdout=srv_log, stderr=subprocess.STDOUT, cwd=logdir)
    time.sleep(8.0)      # model construction + port bind (the reference has no readiness signal)
    workers, logs = [], []
    t0 = time.time()
    for i in range(N):
        lp = os.path.join(logdir, f"worker{i}.log")
        logs.append(lp)
        workers.append(subprocess.Popen(
            common + [worker_py, "--server", f"localhost:{port}", "--worker-name", f"bench-{i}", "--epochs",
                      str(epochs), "--batch-size", str(B), "--lr", "0.1", "--sync-steps", "1"],
            env=env, stdout=open(lp, "w"), stderr=subprocess.STDOUT, cwd=logdir))
    rc = []
    deadline = time.time() + 3600
    for p in workers:
        try:
            rc.append(p.wait(timeout=max(1, deadline - time.time())))
        except subprocess.TimeoutExpired:
            p.kill()
            rc.append(-9)
    wall = time.time() - t0
    srv.terminate()
    try:
        srv.wait(10)
    except subprocess.TimeoutExpired:
        srv.kill()
    recs = []
    pat = re.compile(r"METRICS_JSON:\s*(\{.*\})")
    for lp in logs:
        for line in open(lp, errors="replace"):
            m = pat.search(line)
            if m:
                try:
                    r = json.loads(m.group(1))
                    if r.get("type") == "WORKER_FINAL_METRICS" and r.get("local_steps_completed", 0) > 0:
                        recs.append(r)
                except json.JSONDecodeError:
                    pass
    if len(recs) != N:
        tail = ""
        try:
            tail = open(logs[0], errors="replace").read()[-300:].replace("\n", " | ")
        except Exception:
            pass
        return {"impl": "reference", "unavailable": f"reference run produced {len(recs)}/{N} worker records "
                f"(exit codes {rc}); worker0 log tail: {tail}"}
    timed = max(r["epoch_times_seconds"][-1] for r in recs)
    value = N * steps_per_epoch * B / timed
    return {
        "metric": "ResNet-18 images/sec (%s mode, parameter server)" % args.mode, "impl": "reference",
        "value": round(value, 2), "unit": "images/sec", "n_gpus": N, "steps": steps_per_epoch, "warmup": W,
        "ms_per_step": round(timed / steps_per_epoch * 1e3, 2), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": round(value / PUBLISHED[args.mode], 3), "dtype": "fp32", "data": "synthetic",
        "config": {"model": "ResNet-18 (CIFAR-100 shape, 32x32, random init)", "global_batch": N * B,
                   "batch_per_worker": B, "parallelism": f"dp{N} via 1 central gRPC parameter server (CPU)",
                   "mode": args.mode, "transport": "gRPC localhost + pickle, fp16 gradient compression",
                   "device": "CPU (the reference's distributed path has no GPU support)",
                   "timing": "reference's own per-epoch wall clock, max over workers, warm-up epoch discarded",
                   "omp_threads_per_worker": int(env["OMP_NUM_THREADS"])},
        "e2e": {"value": round(N * steps_per_epoch * epochs * B / wall, 2), "unit": "images/sec",
                "note": "whole job wall clock incl. evaluation, registration and both epochs"},
        "gpu_launches": 0,
    }
