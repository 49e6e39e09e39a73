"""BASELINE.json config 1: sync mode, 1 server + N workers over gRPC localhost on CPU (plumbing, no GPU).

Runs THIS framework's gRPC path (`dps_b200.rpc.servicer` + N x `dps_b200.rpc.client`) with the same protocol as the
reference arm (`bench_tools/reference_arm.py`): batch 128, synthetic CIFAR-100-shaped data sized so one epoch is K
steps per worker, a warm-up epoch that is discarded, images/sec = N * K * batch / max over workers of the timed
epoch.  `--with-reference` also runs the reference arm back to back and prints the ratio.

    python bench_tools/grpc_cpu_bench.py --workers 2 --steps 10 [--mode async] [--compat-pickle] [--with-reference]
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _wait_port(port: int, timeout: float = 60.0) -> bool:
    t0 = time.time()
    while time.time() - t0 < timeout:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.5).close()
            return True
        except OSError:
            time.sleep(0.2)
    return False


def run_ours(workers: int, steps: int, batch: int, mode: str, compat_pickle: bool, outdir: str,
             amp: str = "none", pull_compress: str = "none") -> dict:
    N, K, B = workers, steps, batch
    train_samples = N * B * K
    port = _free_port()
    env = dict(os.environ)
    ncpu = os.cpu_count() or 8
    env["OMP_NUM_THREADS"] = env["MKL_NUM_THREADS"] = str(max(1, ncpu // max(1, N)))
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    env["PYTHONUNBUFFERED"] = "1"
    env["DPS_DATA_CACHE"] = "0"
    os.makedirs(outdir, exist_ok=True)
    for f in os.listdir(outdir):
        if f.endswith(".json"):
            os.unlink(os.path.join(outdir, f))
    common = ["--train-samples", str(train_samples), "--test-samples", str(B), "--device", "cpu",
              "--metrics-dir", outdir, "--pull-compress", pull_compress] + (["--compat-pickle"] if compat_pickle else [])
    srv = subprocess.Popen([sys.executable, "-m", "dps_b200.rpc.servicer", "--mode", mode, "--workers", str(N),
                            "--port", str(port), "--lr", "0.1"] + common, env=env, cwd=ROOT,
                           stdout=open(os.path.join(outdir, "server.log"), "w"), stderr=subprocess.STDOUT)
    try:
        if not _wait_port(port):
            return {"impl": "ours-grpc-cpu", "unavailable": "server did not come up"}
        t0 = time.time()
        procs = [subprocess.Popen([sys.executable, "-m", "dps_b200.rpc.client", "--server", f"localhost:{port}",
                                   "--worker-name", f"bench-{i}", "--epochs", "2", "--batch-size", str(B), "--lr", "0.1",
                                   "--sync-steps", "1", "--amp", amp] + common, env=env, cwd=ROOT,
                                  stdout=open(os.path.join(outdir, f"worker{i}.log"), "w"), stderr=subprocess.STDOUT)
                 for i in range(N)]
        rc = [p.wait(timeout=3600) for p in procs]
        wall = time.time() - t0
    finally:
        time.sleep(0.5)
        srv.terminate()
        try:
            srv.wait(10)
        except subprocess.TimeoutExpired:
            srv.kill()
    recs = []
    for f in sorted(os.listdir(outdir)):
        if f.startswith("worker") and f.endswith(".json"):
            r = json.load(open(os.path.join(outdir, f)))
            if r.get("type") == "WORKER_FINAL_METRICS" and r.get("local_steps_completed", 0) > 0:
                recs.append(r)
    if len(recs) != N:
        return {"impl": "ours-grpc-cpu", "unavailable": f"{len(recs)}/{N} worker records (exit codes {rc})"}
    timed = max(r["epoch_times_seconds"][-1] for r in recs)
    value = N * K * B / timed
    phases = {}
    for r in recs:
        for name, ph in (r.get("phases") or {}).items():
            if isinstance(ph, dict) and "mean_s" in ph:
                phases.setdefault(name, []).append(ph["mean_s"])
    return {
        "metric": f"ResNet-18 images/sec ({mode} mode, parameter server)", "impl": "ours-grpc-cpu",
        "value": round(value, 2), "unit": "images/sec", "workers": N, "steps": K,
        "ms_per_step": round(timed / K * 1e3, 2), "dtype": "fp32" if amp == "none" else "bf16 autocast (fp32 master)",
        "data": "synthetic",
        "config": {"model": "ResNet-18 (CIFAR-100 shape, 32x32, random init)", "global_batch": N * B,
                   "transport": "gRPC localhost, " + ("reference pickle payloads" if compat_pickle else "DPS1 frames")
                                + ", fp16 gradient compression, pull " + pull_compress, "device": "CPU", "mode": mode,
                   "timing": "worker epoch wall clock, max over workers, warm-up epoch discarded",
                   "omp_threads_per_worker": int(env["OMP_NUM_THREADS"])},
        "phase_mean_seconds": {k: round(sum(v) / len(v), 4) for k, v in phases.items()},
        "job_wall_seconds": round(wall, 2),
    }


def main(argv=None) -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--workers", type=int, default=2)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--batch", type=int, default=128)
    p.add_argument("--mode", choices=["sync", "async"], default="sync")
    p.add_argument("--compat-pickle", action="store_true")
    p.add_argument("--with-reference", action="store_true")
    p.add_argument("--amp", choices=["none", "bf16"], default="none")
    p.add_argument("--pull-compress", choices=["none", "fp16"], default="none")
    p.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "grpc_cpu_bench"))
    a = p.parse_args(argv)
    ours = run_ours(a.workers, a.steps, a.batch, a.mode, a.compat_pickle, a.out, a.amp, a.pull_compress)
    print(json.dumps(ours))
    if a.with_reference:
        from bench_tools.reference_arm import run_reference
        ns = argparse.Namespace(gpus=a.workers, steps=a.steps, warmup=1, batch=a.batch, mode=a.mode)
        ref = run_reference(ns)
        print(json.dumps(ref))
        if "value" in ours and "value" in ref:
            print(json.dumps({"ours_over_reference": round(ours["value"] / ref["value"], 3)}))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
