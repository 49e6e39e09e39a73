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
        # test hook: this rank stops stepping at `step`; peers hit the bounded barrier wait
        ns.max_steps = fault[1]
    job.worker.run_training(train, test, epochs=cfg.epochs, max_steps=ns.max_steps, eval_batches=ns.eval_batches)
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
    ns, rest = p.parse_known_args(argv)
    rest = [a for a in rest if a != "--"]
    if ns.gpus == 1:
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
        os.environ.setdefault("LOCAL_RANK", "0")
        return _worker_main(rest)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={ns.gpus}",
           "--master-addr", "127.0.0.1", "--master-port", str(ns.master_port), "-m", "dps_b200.launch"] + rest
    return subprocess.call(cmd)


if __name__ == "__main__":
    raise SystemExit(main())
