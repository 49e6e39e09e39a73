"""One-box launcher: N processes (one per GPU), each a Worker plus a ParameterServer shard.

    python -m dps_b200.launch --gpus 8 -- --mode sync --epochs 3 --batch-size 128 --metrics-dir out/sync8

Plays the role of the reference's deployment tier (reference: terraform/main.tf ECS services,
Dockerfile.server / Dockerfile.worker, terraform/deploy.sh): "start 1 server + W workers and collect
their METRICS_JSON".  Here the server is sharded over the same W processes, rank == worker id, and
the per-process records are aggregated into the reference's result-file schema
(scripts/parse_cloudwatch_logs.py:197-203) without CloudWatch.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from typing import List, Optional


def _worker_main(argv: List[str]) -> int:
    import torch
    from . import config as CFG
    from .api import Job
    from .utils import data as D
    from .utils import metrics as M

    p = argparse.ArgumentParser(description="dps_b200 on-box rank")
    CFG.add_server_args(p)
    p.add_argument("--epochs", type=int, default=3)
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--sync-steps", type=int, default=1)
    p.add_argument("--max-steps", type=int, default=None)
    p.add_argument("--eval-batches", type=int, default=None)
    p.add_argument("--experiment-name", default=None)
    CFG.add_common_args(p)
    ns = p.parse_args(argv)
    cfg = CFG.from_args(ns)
    job = Job(cfg)
    train, test, part = D.make_loaders(job.cfg, job.rank, job.world, pin=True, device=job.device)
    if job.rank == 0:
        print(f"[dps_b200] {job.world} GPU(s), mode={cfg.mode}, batch={cfg.batch_size}/worker, "
              f"partition of rank 0 = {part}", flush=True)
    fault = None
    if cfg.inject_fault:
        r, s = cfg.inject_fault.split(":")
        fault = (int(r), int(s))
    if fault and fault[0] == job.rank:
        # test hook: this rank's PROCESS dies at `step` (os._exit, no JobFinished, no clean-up); the
        # supervising launcher (--supervise) restarts the job from the last periodic checkpoint
        job.worker.fault_step = fault[1]
    if job.rank == 0 and job.worker.resumed_from_step:
        print(f"[dps_b200] resumed from {cfg.resume} at step {job.worker.resumed_from_step}", flush=True)
    from .api import JobFailure
    try:
        job.worker.run_training(train, test, epochs=cfg.epochs, max_steps=ns.max_steps, eval_batches=ns.eval_batches)
    except JobFailure as e:
        print(f"[dps_b200] rank {job.rank} gives up: {e}", flush=True)
        os._exit(18)                       # peers may be gone: no collective clean-up; the supervisor takes over
    if job.rank == 0:
        print(f"[dps_b200] FIRST_STEP_DONE_AT {getattr(job.worker, 'first_step_wall', 0.0):.3f}", flush=True)
    if cfg.checkpoint:
        job.server.save_checkpoint(cfg.checkpoint)
    job.server.finish()
    job.barrier()
    if job.rank == 0 and cfg.metrics_dir:
        name = ns.experiment_name or f"{cfg.mode}_{job.world}workers"
        res = M.collect_dir(cfg.metrics_dir, name, os.path.join(cfg.metrics_dir, name + ".result.json"))
        agg = res["worker_metrics_aggregated"] or {}
        print(f"[dps_b200] {name}: {agg.get('num_workers')} workers, "
              f"{agg.get('images_per_second_sum', 0):.0f} img/s (sum of per-worker rates)", flush=True)
    job.close()
    if job.world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


def main(argv: Optional[List[str]] = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if "WORLD_SIZE" in os.environ and "--gpus" not in argv:
        return _worker_main(argv)
    p = argparse.ArgumentParser(description="dps_b200 one-box launcher")
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--master-port", type=int, default=29517)
    p.add_argument("--supervise", action="store_true",
                   help="restart the job from its last checkpoint (--checkpoint, --checkpoint-every) when a rank dies "
                        "or the device path reports a missing peer; same rank <-> partition mapping")
    p.add_argument("--max-restarts", type=int, default=3)
    ns, rest = p.parse_known_args(argv)
    rest = [a for a in rest if a != "--"]
    if ns.gpus == 1 and not ns.supervise:
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
        os.environ.setdefault("LOCAL_RANK", "0")
        return _worker_main(rest)

    def run_once(args: List[str], attempt: int) -> int:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={ns.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(ns.master_port + attempt), "-m",
               "dps_b200.launch"] + args
        return subprocess.call(cmd)

    if not ns.supervise:
        return run_once(rest, 0)
    return supervise(run_once, rest, ns.max_restarts)


def _strip_flag(args: List[str], flag: str) -> List[str]:
    out, skip = [], False
    for a in args:
        if skip:
            skip = False
            continue
        if a == flag:
            skip = True
            continue
        if a.startswith(flag + "="):
            continue
        out.append(a)
    return out


def _flag_value(args: List[str], flag: str) -> Optional[str]:
    for i, a in enumerate(args):
        if a == flag and i + 1 < len(args):
            return args[i + 1]
        if a.startswith(flag + "="):
            return a.split("=", 1)[1]
    return None


def supervise(run_once, args: List[str], max_restarts: int = 3) -> int:
    """Elastic recovery for the on-box job (the reference claims "automatic failover" and a < 30 s recovery target,
    README.md:51, baseline/results/baseline_summary.json:39, and relies on ECS restarting tasks that then get a fresh,
    out-of-range worker id): run the job; when it fails -- a rank's process died, or a rank raised JobFailure after a
    missed barrier -- restart ALL ranks with --resume <last checkpoint>.  Rank i keeps GPU i, shard i and data
    partition i.  Prints RECOVERY lines with the wall time from the failure to the restarted job being launched."""
    import time
    ckpt = _flag_value(args, "--checkpoint")
    attempt = 0
    while True:
        t0 = time.time()
        rc = run_once(args, attempt)
        if rc == 0:
            return 0
        t_fail = time.time()
        attempt += 1
        if attempt > max_restarts:
            print(f"[dps_b200] job failed (exit {rc}) and the restart budget is spent", flush=True)
            return rc
        args = _strip_flag(args, "--inject-fault")               # the injected fault happens once
        if ckpt and os.path.exists(ckpt):
            args = _strip_flag(args, "--resume") + ["--resume", ckpt]
            src = f"checkpoint {ckpt}"
        else:
            src = "scratch (no checkpoint on disk)"
        print(f"[dps_b200] RECOVERY attempt {attempt}: job exited with {rc} after {t_fail - t0:.1f}s; restarting from "
              f"{src}; failure_detected_at={t_fail:.3f}", flush=True)


if __name__ == "__main__":
    raise SystemExit(main())
