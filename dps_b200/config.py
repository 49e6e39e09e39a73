"""Single configuration object for every role (server, worker, on-box job).

The reference spreads settings over three uncoordinated tiers: Terraform vars ->
container env vars -> argparse defaults (reference: src/parameter_server/server.py:405-433,
src/workers/worker.py:455-482, terraform/main.tf:259-272).  Here there is one dataclass;
the CLI flag names and env-var names of the reference are kept so existing launch
scripts keep working, and the whole config is echoed into the metrics JSON.

Fixes relative to the reference:
  * ``--staleness-bound`` is actually forwarded (reference drops it, server.py:428-433).
  * server and worker share one ``lr`` (reference configures them independently).
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass
from typing import Any, Dict, Optional, Sequence

# Env-var names used by the reference's container definitions.
ENV_MODE = "SERVER_MODE"
ENV_WORKERS = "TOTAL_WORKERS_EXPECTED"
ENV_PORT = "SERVER_PORT"
ENV_SERVER_ADDR = "PARAMETER_SERVER_ADDRESS"

MAX_WORKERS = 32  # reference validates 1 <= workers <= 32 (server.py:424)


@dataclass
class Config:
    # ---- roles / topology -------------------------------------------------
    mode: str = "sync"                    # 'sync' | 'async'
    workers: int = 4                      # expected number of workers (W)
    port: int = 8000
    server: str = "localhost:8000"        # address the worker dials (gRPC path)
    worker_name: str = "worker-node"
    # ---- optimisation ------------------------------------------------------
    lr: float = 0.1
    momentum: float = 0.0                 # 0 == reference-exact plain SGD (server.py:133)
    weight_decay: float = 0.0
    staleness_bound: int = 5              # async: reject pushes with staleness > bound
    staleness_alpha: float = 0.1          # weight = max(floor, 1/(1+alpha*s))
    staleness_floor: float = 0.1
    # ---- training loop -----------------------------------------------------
    epochs: int = 3
    batch_size: int = 128
    sync_steps: int = 1                   # fetch/push every k batches
    local_step_policy: str = "accumulate"  # 'reference' (discard k-1 grads) | 'accumulate'
    num_classes: int = 100
    image_size: int = 32
    seed: int = 0
    # ---- behaviour switches ---------------------------
    nonblocking_sync: bool = False        # True = reference Q1 behaviour (RPC never blocks)
    barrier_timeout_s: float = 60.0       # Q3 fix: sync barrier timeout (drop missing ranks)
    worker_timeout_s: float = 0.0         # gRPC server: evict workers silent for this long (0 = off; clients beat every 30 s)
    bn_buffer_policy: str = "mean"        # 'reference' | 'local' | 'mean'   (Q5)
    compat_pickle: bool = False           # accept/emit reference pickle payloads (Q15)
    compress: str = "fp16"                # gradient wire format: 'fp16' | 'none'
    pull_compress: str = "none"           # gRPC path, FetchParameters payload: 'none' (fp32, reference) | 'fp16'
    amp: str = "none"                     # gRPC/CPU worker compute: 'none' (fp32, reference-exact) | 'bf16' (autocast)
    # ---- data --------------------------------------------------------------
    dataset: str = "synthetic"            # 'synthetic' | 'cifar100'
    data_root: str = "./data"
    train_samples: int = 50000
    test_samples: int = 10000
    data_on_device: bool = True           # GPU path: dataset resident in HBM, augmentation kernel per batch
    # ---- device path -------------------------------------------------------
    device: str = "auto"                  # 'auto' | 'cpu' | 'cuda'
    engine: str = "native"                # the sm_100a kernels; the ATen/cuDNN baseline lives in bench_tools/torch_engine.py
    use_cuda_graph: bool = True
    overlap_pull: bool = True             # sync mode: pull at step start, overlapped with forward per layer group
    overlap_push: bool = True             # sync mode: weight gradients + per-group push on a side stream during backward
    async_ring_slots: int = 4             # ring depth per (owner, worker) pair
    async_server: bool = True             # async mode: the owner role runs on its own high-priority stream (apply on arrival)
    async_server_period_us: int = 100     # pause between two scans of the arrival ring by the server thread
    staleness_policy: str = "reference"   # 'reference' | 'world' (bound + W-1) | 'batch' (one version per drained batch)
    nvls: str = "off"                     # pull broadcast through an NVLS multicast window: 'off' | 'auto' | 'on'
    nvls_reduce: bool = False             # sync push: sources stage locally, owners sum the replicas with multimem.ld_reduce
    serve_port: Optional[int] = None      # device path: rank 0 also serves the gRPC contract on this port (0 = any free)
    remote_slots: int = 0                 # source slots for remote workers attached through that door (sync mode)
    # ---- outputs -----------------------------------------------------------
    metrics_dir: Optional[str] = None     # write METRICS_JSON records to files here
    checkpoint: Optional[str] = None
    checkpoint_every: int = 0             # gRPC server: also save every N global steps (0 = only at job end)
    resume: Optional[str] = None
    inject_fault: Optional[str] = None    # "rank:step" test hook
    watchdog: bool = True                 # device path: host thread watching error bits + peer heartbeats
    heartbeat_timeout_s: float = 30.0     # device path: a peer whose heartbeat word stands still this long is stale

    def validate(self) -> "Config":
        if self.mode not in ("sync", "async"):
            raise ValueError(f"mode must be 'sync' or 'async', got {self.mode!r}")
        if not (1 <= self.workers <= MAX_WORKERS):
            raise ValueError(f"Number of workers must be between 1 and {MAX_WORKERS}")
        if self.bn_buffer_policy not in ("reference", "local", "mean"):
            raise ValueError("bn_buffer_policy must be reference|local|mean")
        if self.local_step_policy not in ("reference", "accumulate"):
            raise ValueError("local_step_policy must be reference|accumulate")
        if self.compress not in ("fp16", "none"):
            raise ValueError("compress must be fp16|none")
        if self.pull_compress not in ("none", "fp16"):
            raise ValueError("pull_compress must be none|fp16")
        if self.amp not in ("none", "bf16"):
            raise ValueError("amp must be none|bf16")
        if self.staleness_policy not in ("reference", "world", "batch"):
            raise ValueError("staleness_policy must be reference|world|batch")
        if self.nvls not in ("off", "auto", "on"):
            raise ValueError("nvls must be off|auto|on")
        if self.sync_steps < 1:
            raise ValueError("sync_steps must be >= 1")
        return self

    def to_dict(self) -> Dict[str, Any]:
        return dataclasses.asdict(self)

    def replace(self, **kw: Any) -> "Config":
        return dataclasses.replace(self, **kw)


def _env_default(name: str, fallback: str) -> str:
    return os.environ.get(name, fallback)


def add_server_args(p: argparse.ArgumentParser) -> None:
    """Reference server flags (server.py:406-419), same names and env defaults."""
    p.add_argument("--mode", choices=["sync", "async"], default=_env_default(ENV_MODE, "sync"))
    p.add_argument("--workers", type=int, default=int(_env_default(ENV_WORKERS, "4")))
    p.add_argument("--lr", type=float, default=0.1)
    p.add_argument("--port", type=int, default=int(_env_default(ENV_PORT, "8000")))
    p.add_argument("--staleness-bound", type=int, default=5)


def add_worker_args(p: argparse.ArgumentParser) -> None:
    """Reference worker flags (worker.py:456-469), same names and env defaults."""
    p.add_argument("--server", default=_env_default(ENV_SERVER_ADDR, "localhost:8000"))
    p.add_argument("--worker-name", default="worker-node")
    p.add_argument("--epochs", type=int, default=3)
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--lr", type=float, default=0.1)
    p.add_argument("--sync-steps", type=int, default=1)


def add_common_args(p: argparse.ArgumentParser) -> None:
    p.add_argument("--momentum", type=float, default=0.0)
    p.add_argument("--weight-decay", type=float, default=0.0)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--dataset", choices=["synthetic", "cifar100"], default="synthetic")
    p.add_argument("--data-root", default="./data")
    p.add_argument("--train-samples", type=int, default=50000)
    p.add_argument("--test-samples", type=int, default=10000)
    p.add_argument("--image-size", type=int, default=32)
    p.add_argument("--data-on-host", dest="data_on_device", action="store_false",
                   help="GPU path: produce batches on the host (loader thread + pinned H2D) instead of on the device")
    p.add_argument("--device", choices=["auto", "cpu", "cuda"], default="auto")
    p.add_argument("--bn-buffer-policy", choices=["reference", "local", "mean"], default="mean")
    p.add_argument("--local-step-policy", choices=["reference", "accumulate"], default="accumulate")
    p.add_argument("--nonblocking-sync", action="store_true")
    p.add_argument("--barrier-timeout-s", type=float, default=60.0)
    p.add_argument("--heartbeat-timeout-s", type=float, default=30.0)
    p.add_argument("--no-watchdog", dest="watchdog", action="store_false")
    p.add_argument("--worker-timeout-s", type=float, default=0.0)
    p.add_argument("--compat-pickle", action="store_true")
    p.add_argument("--compress", choices=["fp16", "none"], default="fp16")
    p.add_argument("--pull-compress", choices=["none", "fp16"], default="none",
                   help="gRPC server: send parameters as fp16 (master stays fp32); halves the pull payload")
    p.add_argument("--amp", choices=["none", "bf16"], default="none",
                   help="gRPC/CPU worker: bf16 autocast + channels_last for the local step (master weights stay fp32)")
    p.add_argument("--metrics-dir", default=None)
    p.add_argument("--checkpoint", default=None)
    p.add_argument("--checkpoint-every", type=int, default=0)
    p.add_argument("--resume", default=None)
    p.add_argument("--inject-fault", default=None)
    p.add_argument("--staleness-policy", choices=["reference", "world", "batch"], default="reference",
                   help="async mode: reference = reject staleness > bound; world = bound widened by W-1 (W workers put "
                        "W-1 updates between a pull and the next push by construction); batch = arrivals drained "
                        "together count as one version")
    p.add_argument("--nvls", choices=["off", "auto", "on"], default="off",
                   help="on-box job: pull broadcast through an NVLS multicast window (one multimem.st per vector "
                        "instead of W peer stores); auto = when the box supports it")
    p.add_argument("--nvls-reduce", action="store_true", default=None,
                   help="on-box sync job (with --nvls auto|on): in-switch reduction of the push (multimem.ld_reduce)")
    p.add_argument("--serve-port", type=int, default=None,
                   help="on-box job: rank 0 also serves the gRPC contract on this port (remote / CPU workers)")
    p.add_argument("--remote-slots", type=int, default=0,
                   help="on-box sync job: source slots for remote workers attached through --serve-port")


def from_args(ns: argparse.Namespace) -> Config:
    """Build a Config from an argparse namespace; unknown attributes are ignored."""
    fields = {f.name for f in dataclasses.fields(Config)}
    kw = {k: v for k, v in vars(ns).items() if k in fields and v is not None}
    return Config(**kw).validate()


def parse_server_cli(argv: Optional[Sequence[str]] = None) -> Config:
    p = argparse.ArgumentParser(description="dps_b200 parameter server (gRPC / CPU path)")
    add_server_args(p)
    add_common_args(p)
    return from_args(p.parse_args(argv))


def parse_worker_cli(argv: Optional[Sequence[str]] = None) -> Config:
    p = argparse.ArgumentParser(description="dps_b200 worker (gRPC / CPU path)")
    add_worker_args(p)
    add_common_args(p)
    return from_args(p.parse_args(argv))
