"""Public on-box API: ``Job`` -> ``ParameterServer`` shard + ``Worker`` on every GPU.

One process per GPU (``torchrun`` / ``dps_b200.launch``); rank == worker id == shard id.
The roles keep the reference's names and responsibilities (reference:
src/parameter_server/server.py:81 ParameterServerServicer, src/workers/worker.py:78 DistributedWorker)
but the transport between them is the sm_100a Push / Pull / Async kernels over NVLink peer
memory instead of pickle-over-gRPC; ``torch.distributed`` is used for bootstrap and for
checkpoint gathers only.

    job = Job(Config(mode="sync", workers=W, batch_size=128))
    for x, y in loader:                       # pinned host tensors
        loss = job.worker.train_step(x, y)    # H2D -> fwd/bwd -> push -> reduce+SGD -> pull, one graph replay
    job.server.save_checkpoint("ckpt.pt")     # loads into the reference ResNet18 (strict=True)
"""
from __future__ import annotations

import os
import time
from typing import Dict, Iterable, List, Optional, Tuple

import torch

from .config import Config
from .models.resnet import build_model
from .store.shardmap import Layout, build_layout
from .utils import checkpoint as CK
from .utils import metrics as M


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class JobFailure(RuntimeError):
    """The on-box job cannot continue (a peer missed a barrier / stopped sending heartbeats).  The process exits
    non-zero and the supervising launcher restarts the job from the last checkpoint (dps_b200.launch --supervise)."""


class Job:
    """Bootstraps one rank: process group, symmetric arenas, parameter shard, worker engine."""

    def __init__(self, cfg: Config, rank: Optional[int] = None, world: Optional[int] = None,
                 device_index: Optional[int] = None, init_state: Optional[Dict[str, torch.Tensor]] = None,
                 engine_factory=None):
        """engine_factory: baseline harnesses only (bench_tools/torch_engine.py runs the same arenas through
        ATen/cuDNN); the product has ONE engine, the sm_100a kernels of models/engine.py."""
        from .parallel.ps_device import DevicePS
        self.rank = _env_int("RANK", 0) if rank is None else rank
        self.world = _env_int("WORLD_SIZE", 1) if world is None else world
        self.device_index = _env_int("LOCAL_RANK", self.rank) if device_index is None else device_index
        self.cfg = cfg.replace(workers=self.world).validate()
        if self.cfg.serve_port is not None and self.cfg.remote_slots == 0 and self.cfg.mode == "sync":
            self.cfg = self.cfg.replace(remote_slots=1)        # same decision on every rank: it sizes the staging arena
        if not torch.cuda.is_available():
            raise RuntimeError(
                "dps_b200.api.Job is the on-box device path (one process per B200, NVLink peer memory) and needs CUDA. "
                "On a machine without a GPU use the gRPC roles: `python -m dps_b200.rpc.servicer` + "
                "`python -m dps_b200.rpc.client` (or dps_b200.rpc.servicer.build_server / "
                "dps_b200.rpc.client.DistributedWorker), see INTEGRATION.md.")
        torch.cuda.set_device(self.device_index)
        self.device = torch.device("cuda", self.device_index)
        self.engine_factory = engine_factory
        self.group = None
        if self.world > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29500")
                dist.init_process_group("nccl", rank=self.rank, world_size=self.world,
                                        device_id=self.device)
        model = build_model(self.cfg.num_classes, seed=self.cfg.seed)     # same seed on every rank (Q13)
        if init_state is not None:
            model.load_state_dict(init_state)
        elif self.cfg.resume:
            sd, side = CK.load(self.cfg.resume)
            model.load_state_dict(sd)
        self.model_cpu = model
        self.layout: Layout = build_layout(model, self.world)
        self.ps = DevicePS(self.layout, self.cfg, self.rank, self.world, self.device_index, group=self.group)
        flat = torch.zeros(self.layout.total)
        self.layout.flatten_into(flat, dict(model.named_parameters()))
        self.ps.load_master(flat)
        self.server = ParameterServer(self)
        self.worker = Worker(self)
        if self.cfg.resume:
            _, side = CK.load(self.cfg.resume)
            self.server.restore_sidecar(side)
        self.version_offset = self.ps.version() - self.ps.step      # shard version - round number (resume step)
        # failure detection: a host thread watches this rank's flag block -- sticky error bits raised by
        # the bounded waits of the kernels, and peers' heartbeat words -- and marks the job failed; the training loop
        # then raises JobFailure instead of training on with a rank missing
        self.failed: Optional[str] = None
        self.watchdog = None
        if self.cfg.watchdog and self.world > 1:
            from .parallel.watchdog import decode_error_bits, for_device_ps, missing_ranks
            from .ops import _lib as L

            wd_stream = {}

            def on_poll(stale):
                if "s" not in wd_stream:
                    wd_stream["s"] = torch.cuda.Stream(device=self.device)
                with torch.cuda.stream(wd_stream["s"]):
                    return _on_poll(stale)

            def _on_poll(stale):
                bits = self.ps.error_bits()
                if bits and self.failed is None:
                    miss = missing_ranks(int(self.ps.flag_words()[L.F_MISSING]) & 0xFFFFFFFF, self.world)
                    self.failed = f"device error bits {decode_error_bits(bits)}, missing ranks {miss}"
                elif stale and self.failed is None and self.worker.local_step_counter > 0:
                    self.failed = f"no heartbeat from ranks {stale} for {self.cfg.heartbeat_timeout_s:.0f}s"

            self.watchdog = for_device_ps(self.ps, timeout_s=self.cfg.heartbeat_timeout_s, on_stale=on_poll)
            self.watchdog.always_notify = True
            self.watchdog.start(period_s=min(1.0, max(0.05, self.cfg.heartbeat_timeout_s / 10.0)))
        # gRPC front door onto the device store (reference: the four RPCs of server.py:190-318): rank 0 hosts it
        self.door = self.door_server = None
        self.door_port: Optional[int] = None
        if self.cfg.serve_port is not None and self.rank == 0:
            from .rpc.device_door import start_door
            self.door_server, self.door, self.door_port = start_door(self, int(self.cfg.serve_port))
        self.barrier()

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(device_ids=[self.device_index])

    def check_alive(self) -> None:
        if self.failed is not None:
            raise JobFailure(f"rank {self.rank}: {self.failed}")

    def close(self) -> None:
        if self.watchdog is not None:
            self.watchdog.stop()
            self.watchdog = None
        if self.door_server is not None:
            self.door_server.stop(0)
            self.door_server = None
        self.barrier()
        self.ps.close()


class ParameterServer:
    """This rank's shard of the store: fp32 master + momentum + version counter."""

    def __init__(self, job: Job):
        self.job, self.ps, self.cfg = job, job.ps, job.cfg
        self.start_time = time.time()
        self._final: Optional[dict] = None

    @property
    def global_step(self) -> int:
        return self.ps.version()

    def parameters(self) -> Dict[str, torch.Tensor]:
        """Full fp32 parameters, state_dict-shaped (collective when world > 1)."""
        flat = self.ps.gather_master(self.job.group)
        return {k: v.cpu() for k, v in self.job.layout.unflatten(flat).items()}

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """All 122 reference keys: trainable tensors from the shards + BN buffers per policy (Q5)."""
        params = self.parameters()
        bufs = self.job.worker.bn_buffers_for_checkpoint()
        sd = {}
        for k, v in self.job.model_cpu.state_dict().items():
            sd[k] = params[k] if k in params else bufs.get(k, v.detach().clone())
        return sd

    def save_checkpoint(self, path: str) -> None:
        sd = self.state_dict()
        mom_flat = self.ps.gather_momentum(self.job.group)
        mom = {k: v.cpu() for k, v in self.job.layout.unflatten(mom_flat).items()}
        versions = self._all_versions()
        if self.job.rank == 0:
            CK.save(path, sd, global_step=max(versions), momentum=mom, shard_versions=versions,
                    config=self.cfg.to_dict())
        self.job.barrier()

    def restore_sidecar(self, side: dict) -> None:
        mom = side.get("momentum") or {}
        if mom:
            flat = torch.zeros(self.job.layout.total)
            self.job.layout.flatten_into(flat, mom)
            shards = self.job.layout.to_shards(flat.to(self.job.device))
            self.ps.mom[self.job.rank].copy_(shards[self.job.rank])
        # every shard has its own version counter (async mode advances them independently): restore this shard's,
        # falling back to the job-wide step for checkpoints written with a different shard count
        from .ops import _lib as L
        step = int(side.get("global_step", 0))
        versions = side.get("shard_versions") or []
        mine = int(versions[self.job.rank]) if len(versions) == self.job.world else step
        self.ps.flags[self.job.rank][L.F_VERSION] = mine
        torch.cuda.synchronize(self.job.device)
        # a resumed job continues the step count, so `max_steps` keeps meaning "steps of the whole job"
        self.job.worker.local_step_counter = step
        self.job.worker.resumed_from_step = step

    def _all_versions(self) -> List[int]:
        v = torch.tensor([self.ps.version()], device=self.job.device, dtype=torch.int64)
        if self.job.world == 1:
            return [int(v)]
        import torch.distributed as dist
        out = [torch.zeros_like(v) for _ in range(self.job.world)]
        dist.all_gather(out, v)
        return [int(x) for x in out]

    def metrics(self) -> dict:
        total = time.time() - self.start_time
        steps = self.global_step
        extra = {"shard": self.job.rank, "shards": self.job.world, "transport": "nvlink-p2p",
                 "error_bits": self.ps.error_bits(), "config": self.cfg.to_dict()}
        if self.cfg.mode == "async":
            st = self.ps.async_stats()
            last = [s for s in st["last_staleness_per_worker"]]
            extra.update({"async_updates": st["async_updates"], "rejected_updates": st["rejected_updates"],
                          "average_gradient_staleness": round(sum(last) / max(1, len(last)), 2),
                          "max_staleness_observed": max(last) if last else 0,
                          "running_mean_staleness": round(st["running_mean_staleness"], 4),
                          "running_max_staleness": st["running_max_staleness"],
                          "staleness_histogram": st["staleness_histogram"], "torn_pulls": st["torn_pulls"],
                          "dropped_pushes": st["dropped_pushes"],
                          "staleness_policy": self.cfg.staleness_policy, "staleness_bound": self.cfg.staleness_bound,
                          "rejection_rate": round(st["rejected_updates"] / max(1, st["async_updates"] + st["rejected_updates"]), 4),
                          "owner_stream": "dedicated high-priority stream" if self.job.worker.async_server else "worker stream",
                          "server_passes": self.ps.server_passes})
        # update time measured ON THE DEVICE by the reduce / apply kernels themselves (DPS_F_UPDATE_NS)
        from .ops import _lib as L
        f = self.ps.flag_words()
        upd_ns = (int(f[L.F_UPDATE_NS]) & 0xFFFFFFFF) | ((int(f[L.F_UPDATE_NS + 1]) & 0xFFFFFFFF) << 32)
        done = max(1, steps - self.job.version_offset)
        extra["average_update_time_us"] = round(upd_ns / done / 1e3, 2)
        # bytes this shard's owner moved per update and the rate it sustained
        wk = self.job.worker
        if wk.local_step_counter > wk.resumed_from_step and sum(wk.epoch_times) > 0:
            per_step = sum(wk.epoch_times) / (wk.local_step_counter - wk.resumed_from_step)
            extra["step_time_ms"] = round(per_step * 1e3, 4)
            for name, nbytes in (("push", self.ps.bytes_per_push()), ("pull", self.ps.bytes_per_pull())):
                extra[f"{name}_bytes_per_step"] = nbytes
                # both transfers are overlapped with compute: the rate over the whole step is what the links see
                extra[f"{name}_GBps_over_step"] = round(nbytes / per_step / 1e9, 2)
                extra[f"{name}_link_roofline_us"] = round(nbytes / 770e9 * 1e6, 2)      # measured peer-copy rate
        return M.server_final_metrics(mode=self.cfg.mode, total_workers=self.job.world, total_time=total,
                                      global_step=steps, total_updates=steps,
                                      gradients_processed=steps * (self.job.world if self.cfg.mode == "sync" else 1),
                                      avg_update_time=upd_ns / done / 1e9, learning_rate=self.cfg.lr, extra=extra)

    def finish(self) -> dict:
        """Emit SERVER_FINAL_METRICS exactly once."""
        if self._final is None:
            self._final = self.metrics()
            M.emit(self._final, self.cfg.metrics_dir, f"server_shard{self.job.rank}")
        return self._final


class LossHandle:
    """Result of a step: the loss lands in pinned host memory; ``item()`` waits for that copy only."""

    def __init__(self, host: torch.Tensor, event: torch.cuda.Event):
        self.host, self.event = host, event

    def item(self) -> float:
        self.event.synchronize()
        return float(self.host[0])

    def correct(self) -> int:
        self.event.synchronize()
        return int(self.host[1])


class Worker:
    """This rank's worker: forward/backward on the local replica + push/pull against all shards."""

    def __init__(self, job: Job):
        self.job, self.ps, self.cfg = job, job.ps, job.cfg
        self.rank, self.world = job.rank, job.world
        self.worker_id, self.total_workers = job.rank, job.world      # rank == id: never out of range (Q10)
        self.local_step_counter = 0
        self.resumed_from_step = 0
        self.fault_step: Optional[int] = None             # test hook (--inject-fault rank:step): die at this step
        self.images_seen = 0
        self.epoch_times: List[float] = []
        self.accuracies: List[float] = []
        self.training_start_time: Optional[float] = None
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._ring = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(4)]
        self._events = [torch.cuda.Event() for _ in range(4)]
        self.engine_kind = "native" if job.engine_factory is None else "baseline"
        if self.engine_kind == "native":
            from .models.engine import ResNetEngine
            self.engine = ResNetEngine(job.layout, self.ps.wbf16[self.rank], self.ps.small[self.rank],
                                       self.ps.grad[self.rank], batch=self.cfg.batch_size,
                                       image_size=self.cfg.image_size, num_classes=self.cfg.num_classes,
                                       device=job.device)
            self.engine.load_bn_buffers(job.model_cpu.state_dict())
        else:
            self.engine = job.engine_factory(job.layout, self.ps.wbf16[self.rank], self.ps.small[self.rank],
                                             self.ps.grad[self.rank], job.model_cpu, batch=self.cfg.batch_size,
                                             image_size=self.cfg.image_size, device=job.device)
        self._accum = 0
        if self.overlap_pull:
            self._side = torch.cuda.Stream(device=job.device)
            self.ps.set_segments(self.engine.segment_boundaries)
            self.engine.seg_wait = lambda k: self.ps.wait_segment(k)
            self._weights_stale = False
        if self.cfg.mode == "async" and self.engine_kind == "native" and self.cfg.overlap_push:
            # async mode: weight gradients and projection shortcuts still run beside the dgrad / BatchNorm chain
            self._wstream = torch.cuda.Stream(device=job.device)
            self._xstream = torch.cuda.Stream(device=job.device)
            self.engine.side_stream, self.engine.branch_stream = self._wstream, self._xstream
        if self.async_server:
            self.ps.start_async_server()
        if self.overlap_push:
            # weight gradients and the per-group SEND half of the push run on a second stream beside the
            # dgrad / BatchNorm chain of the backward pass; only the first layer group + the reduce stay exposed
            self._wstream = torch.cuda.Stream(device=job.device)
            self.engine.side_stream = self._wstream
            self._xstream = torch.cuda.Stream(device=job.device)     # projection shortcuts beside their block
            self.engine.branch_stream = self._xstream
            self.engine.on_group_done = self._send_group
            nseg = self.ps.nseg
            self._send_ctas = [self.ps.send_ctas(k) for k in range(nseg)]
            self._arrivals = sum(self._send_ctas)

    @property
    def arrivals_per_round(self) -> int:
        """Arrival signals a source sends to every owner per sync round (what the owners' REDUCE counts)."""
        return self._arrivals if self.overlap_push else self.ps.ctas

    # ------------------------------------------------------------------ the step
    @property
    def overlap_pull(self) -> bool:
        """Sync mode, native engine: the pull of update t runs at the START of step t+1 on a side stream, layer
        group by layer group, and the forward pass waits per group (BASELINE north star: pull overlapped with the
        next forward pass)."""
        return (self.cfg.overlap_pull and self.cfg.mode == "sync" and self.engine_kind == "native"
                and self.cfg.sync_steps == 1)

    @property
    def async_server(self) -> bool:
        return self.cfg.mode == "async" and self.cfg.async_server and self.engine_kind == "native"

    @property
    def overlap_push(self) -> bool:
        """Sync mode with the overlapped pull: the push is also split per layer group and sent while the backward
        pass is still running (the reference pushes once, after the whole backward: worker.py:376)."""
        return self.overlap_pull and self.cfg.overlap_push

    def _send_group(self, k: int) -> None:
        """Called by the engine on the weight-gradient stream when every gradient of layer group k is complete."""
        if k == 0:
            return                                       # first group: sent on the main stream with the reduce
        self.ps.push_segment(k, self.ps.SEND, ctas=self._send_ctas[k])

    def _pre(self) -> None:
        """Start of a step: fork the segmented pull onto the side stream."""
        if not self.overlap_pull:
            return
        cur = torch.cuda.current_stream()
        self._side.wait_stream(cur)
        with torch.cuda.stream(self._side):
            self.ps.pull_segments()

    def _exchange(self) -> None:
        """Everything after backward: push -> (barrier, reduce, SGD | ring, staleness apply) -> pull."""
        if self.overlap_push:
            # groups 1.. were sent during backward (engine.on_group_done); BN affine gradients, the stem and layer1
            # finish last: send them now, then reduce the whole shard once every source has signalled all its groups
            self.ps.push_segment(0, self.ps.SEND, ctas=self._send_ctas[0])
            self.ps.push_segment(-1, self.ps.REDUCE, arrivals_per_step=self._arrivals)
            self.ps.bump_step()
            torch.cuda.current_stream().wait_stream(self._side)      # join the pull branch
        elif self.overlap_pull:
            self.ps.push_sync()
            self.ps.bump_step()
            torch.cuda.current_stream().wait_stream(self._side)      # join the pull branch
        elif self.cfg.mode == "sync":
            self.ps.push_sync()
            self.ps.pull()
        elif self.async_server:
            # the owner role runs on its own stream (DevicePS.start_async_server): the worker only pushes and pulls
            self.ps.push_async()
            self.ps.pull_async(bump_step=True)
        else:
            self.ps.push_async()
            self.ps.serve_async()
            self.ps.pull_async()

    def _build_graph(self) -> None:
        self.ps.use_device_counters(True)
        saved = {k: v.clone() for k, v in self.engine.bn_buffers().items()}   # warm-up must not count as a step
        self._ensure_staging()
        fused = self.cfg.sync_steps == 1
        # one graph per staging slot: graph k reads the batch straight from slot k and ends with the D2H copy of
        # (loss, #correct) into pinned slot k -- no eager copies between two replays
        before = self.ps.launches
        self.engine.capture(extra=self._exchange if fused else None, pre=self._pre,
                            inputs=list(zip(self._stage_x, self._stage_y)),
                            post=lambda k: self._ring[k].copy_(self.engine.out, non_blocking=True))
        self.ps_launches_per_step = (self.ps.launches - before) // max(1, len(self.engine.graphs))
        for k, v in self.engine.bn_buffers().items():
            v.copy_(saved[k])
        # capture ran one eager warm-up step on the engine only: clear what it accumulated
        self.ps.grad[self.rank].zero_()
        self.graph = self.engine.graph
        self.graphs = self.engine.graphs
        torch.cuda.synchronize(self.job.device)

    def train_step(self, images: torch.Tensor, labels: torch.Tensor) -> LossHandle:
        """One local batch.  `images` fp32 [B,3,S,S] and `labels` int64 [B]: pinned host tensors (copied on a side
        stream, overlapping the previous step), device tensors, or pageable host tensors.  The returned handle is
        valid until two further steps have been issued (its pinned slot is then reused)."""
        use_graph = self.cfg.use_cuda_graph and self.engine_kind == "native"
        if self.async_server and self.ps._server is None:
            self.ps.start_async_server()                 # drain() stopped it: training goes on, so does the owner role
        if use_graph and self.graph is None:
            self._build_graph()
        fused_exchange = use_graph and self.cfg.sync_steps == 1
        if use_graph:
            k = self._stage_batch(images, labels, handover=False)
            self.graphs[k].replay()
            self._consumed[k].record()
        else:
            k = None
            self._stage_batch(images, labels)
            self._pre()
            self.engine.forward_backward()
        if self.cfg.sync_steps > 1 and self.cfg.local_step_policy == "reference":
            # reference behaviour (worker.py:367-377): fetch + push only on the first batch of every k; the other
            # k-1 batches still run forward/backward but their gradients are thrown away
            if self._accum == 0:
                self._exchange()
                if self.cfg.mode == "sync":
                    self.ps.step += 1
            else:
                self.ps.grad[self.rank].zero_()
            self._accum = (self._accum + 1) % self.cfg.sync_steps
        else:
            self._accum += 1
            if self._accum >= self.cfg.sync_steps:
                if not fused_exchange:
                    if self._accum > 1:
                        self.ps.grad[self.rank].div_(self._accum)       # Q12: accumulate, not discard
                    self._exchange()
                if self.cfg.mode == "sync":
                    self.ps.step += 1
                self._accum = 0
        if k is None:                                    # eager path: read the result back with an eager copy
            slot = 2 + self.local_step_counter % 2       # slots 0, 1 belong to the graphs
            self._ring[slot].copy_(self.engine.out, non_blocking=True)
        else:
            slot = k
        self._events[slot].record()
        if self.local_step_counter == self.resumed_from_step:
            self._events[slot].synchronize()
            self.first_step_wall = time.time()                # wall clock of this incarnation's first finished step
        self.local_step_counter += 1
        self.images_seen += int(labels.numel())
        return LossHandle(self._ring[slot], self._events[slot])

    def _ensure_staging(self) -> None:
        if hasattr(self, "_copy_stream"):
            return
        self._copy_stream = torch.cuda.Stream(device=self.job.device)
        self._stage_x = [torch.zeros_like(self.engine.images) for _ in range(2)]
        self._stage_y = [torch.zeros_like(self.engine.labels) for _ in range(2)]
        self._h2d_done = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        self._stage_uses = 0

    def _stage_batch(self, images: torch.Tensor, labels: torch.Tensor, handover: bool = True) -> int:
        """Bring the batch into staging slot k = (#batches staged) % 2 and return k.  Pinned host batches are copied
        on a side stream (overlaps the previous step's kernels; the slot is reused only after the step that read it
        two batches ago has finished); anything else is copied on the compute stream.  handover=True also copies the
        slot into the engine's own input buffers (eager launches read those; graph k reads slot k directly)."""
        if not hasattr(self.engine, "images"):
            self.engine.load_batch(images, labels)
            return 0
        self._ensure_staging()
        k = self._stage_uses % 2
        main = torch.cuda.current_stream()
        if images.is_pinned() and labels.is_pinned():
            with torch.cuda.stream(self._copy_stream):
                if self._stage_uses >= 2:
                    self._copy_stream.wait_event(self._consumed[k])
                self._stage_x[k].copy_(images, non_blocking=True)
                self._stage_y[k].copy_(labels, non_blocking=True)
                self._h2d_done[k].record(self._copy_stream)
            main.wait_event(self._h2d_done[k])
        else:
            self._stage_x[k].copy_(images, non_blocking=True)
            self._stage_y[k].copy_(labels, non_blocking=True)
        if handover:
            self.engine.images.copy_(self._stage_x[k], non_blocking=True)
            self.engine.labels.copy_(self._stage_y[k], non_blocking=True)
            self._consumed[k].record(main)
        self._stage_uses += 1
        return k

    # ------------------------------------------------------------------ tracing
    def profile_phases(self, images: torch.Tensor, labels: torch.Tensor, reps: int = 10) -> Dict[str, float]:
        """Device time per phase of a step in microseconds (CUDA events on the launching stream, eager launches,
        mean over `reps`, MAX over ranks): input staging, forward (includes the overlapped pull in sync mode),
        backward, exchange (push + reduce/SGD [+ pull]).  The steps are REAL steps (weights move).  The reference
        only has wall-clock timers around the optimizer call and the epoch (server.py:121-143, worker.py:355-387).
        With DPS_NVTX=1 the same phases are NVTX ranges for an external profiler."""
        nvtx = os.environ.get("DPS_NVTX", "0") == "1"
        names = ["stage_batch", "forward", "backward", "exchange"]
        acc = {n: 0.0 for n in names}

        class _R:
            def __init__(self, name): self.name = name
            def __enter__(self):
                if nvtx: torch.cuda.nvtx.range_push("dps/" + self.name)
            def __exit__(self, *a):
                if nvtx: torch.cuda.nvtx.range_pop()

        native = self.engine_kind == "native"
        for rep in range(reps + 1):                       # rep 0 is a warm-up
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            ev[0].record()
            with _R("stage_batch"):
                self._stage_batch(images, labels)
            ev[1].record()
            with _R("forward"):
                self._pre()
                if native:
                    self.engine._run(self.engine.fwd)
            ev[2].record()
            with _R("backward"):
                if native:
                    self.engine.backward()
                else:
                    self.engine.forward_backward()       # autograd: forward and backward are one call
            ev[3].record()
            with _R("exchange"):
                self._exchange()
            ev[4].record()
            if self.cfg.mode == "sync":
                self.ps.step += 1
            self.local_step_counter += 1
            self.images_seen += int(labels.numel())
            torch.cuda.synchronize(self.job.device)
            if rep > 0:
                for i, n in enumerate(names):
                    acc[n] += ev[i].elapsed_time(ev[i + 1]) * 1e3 / reps
        t = torch.tensor([acc[n] for n in names], device=self.job.device, dtype=torch.float64)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out = {n: round(float(v), 2) for n, v in zip(names, t.tolist())}
        out["total"] = round(sum(out.values()), 2)
        return out

    # ------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def evaluate(self, loader: Iterable[Tuple[torch.Tensor, torch.Tensor]], max_batches: Optional[int] = None,
                 shard: bool = False) -> float:
        """Top-1 accuracy in percent with running BN statistics (reference: worker.py:313-331).
        shard=True evaluates batches rank::world and all-reduces the counts."""
        if self.overlap_pull and self.local_step_counter > 0:
            self.ps.refresh_weights(self.job.group)     # the last update is pulled at the start of the NEXT step
        correct = total = 0
        B = self.cfg.batch_size
        for i, (x, y) in enumerate(loader):
            if max_batches is not None and i >= max_batches:
                break
            if shard and i % self.world != self.rank:
                continue
            n = int(y.numel())
            if n != B:                                   # static shapes: pad the tail batch
                xp = torch.zeros(B, *x.shape[1:], dtype=x.dtype, device=x.device)
                yp = torch.full((B,), -1, dtype=y.dtype, device=y.device)
                xp[:n], yp[:n] = x, y
                x, y = xp, yp
            correct += self.engine.eval_batch(x, y)
            total += n
        if shard and self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([correct, total], device=self.job.device, dtype=torch.int64)
            dist.all_reduce(t)
            correct, total = int(t[0]), int(t[1])
        acc = 100.0 * correct / max(1, total)
        self.accuracies.append(acc)
        return acc

    # ------------------------------------------------------------------ loop (reference: worker.py:350-403)
    def run_training(self, train_loader, test_loader=None, epochs: Optional[int] = None,
                     max_steps: Optional[int] = None, eval_batches: Optional[int] = None) -> dict:
        if self.cfg.use_cuda_graph and self.engine_kind == "native" and self.graph is None:
            self._build_graph()                             # one-time capture is set-up, not epoch time
        self.training_start_time = time.time()
        epochs = self.cfg.epochs if epochs is None else epochs
        # Sync mode is a true barrier: every rank must take the same number of steps per epoch.  The last worker's
        # partition also holds the remainder (reference: worker.py:166-179) and can contain one more full batch.
        per_epoch = None
        if self.world > 1 and self.cfg.mode == "sync" and all(hasattr(train_loader, a) for a in ("start", "end")):
            import torch.distributed as dist
            t = torch.tensor([(train_loader.end - train_loader.start) // self.cfg.batch_size],
                             device=self.job.device, dtype=torch.int64)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            per_epoch = int(t.item())
        done = False
        last = None
        for _ in range(epochs):
            t0 = time.time()
            taken = 0
            for x, y in train_loader:
                if int(y.numel()) != self.cfg.batch_size:
                    continue                                # static shapes: drop the ragged tail batch
                if per_epoch is not None and taken >= per_epoch:
                    break
                last = self.train_step(x, y)
                taken += 1
                self.job.check_alive()
                if self.fault_step is not None and self.local_step_counter >= self.fault_step:
                    print(f"[dps_b200] FAULT_INJECTED_AT {time.time():.3f} (rank {self.rank}, step "
                          f"{self.local_step_counter})", flush=True)
                    os._exit(17)                            # injected fault: the process dies without any clean-up
                every = self.cfg.checkpoint_every
                if self.cfg.checkpoint and every > 0 and self.local_step_counter % every == 0:
                    self.job.server.save_checkpoint(self.cfg.checkpoint)       # collective; same step on every rank
                if max_steps is not None and self.local_step_counter >= max_steps:
                    done = True
                    break
            if last is not None:
                last.item()
            torch.cuda.synchronize(self.job.device)
            self.epoch_times.append(time.time() - t0)       # excludes evaluation (Q17)
            if test_loader is not None:
                self.evaluate(test_loader, eval_batches)
            if done:
                break
        return self.finish()

    def drain(self, rounds: int = 3) -> None:
        """Async mode: apply arrivals that landed after this rank's last step (end of job)."""
        if self.cfg.mode != "async":
            return
        self.ps.stop_async_server()                  # from here on this thread plays the owner role itself
        for _ in range(rounds):
            self.job.barrier()
            self.ps.serve_async(bump_step=False)
        self.job.barrier()

    def finish(self) -> dict:
        self.drain()
        total = time.time() - (self.training_start_time or time.time())
        train_time = sum(self.epoch_times)
        rec = M.worker_final_metrics(
            worker_id=self.worker_id, total_workers=self.total_workers, total_time=total,
            epoch_times=self.epoch_times, accuracies=self.accuracies, local_steps=self.local_step_counter,
            batch_size=self.cfg.batch_size, learning_rate=self.cfg.lr, num_epochs=self.cfg.epochs,
            extra={"images_per_second": round(self.images_seen / train_time, 2) if train_time > 0 else 0.0,
                   "transport": "nvlink-p2p", "engine": self.engine_kind, "mode": self.cfg.mode,
                   "push_bytes_per_step": self.ps.bytes_per_push(), "pull_bytes_per_step": self.ps.bytes_per_pull()})
        M.emit(rec, self.cfg.metrics_dir, "worker")
        return rec

    # ------------------------------------------------------------------ BN buffers (Q5)
    def bn_buffers_for_checkpoint(self) -> Dict[str, torch.Tensor]:
        bufs = {k: v.detach().clone() for k, v in self.engine.bn_buffers().items()}
        policy = self.cfg.bn_buffer_policy
        if policy == "reference":                      # never updated on the server: initial values
            init = self.job.model_cpu.state_dict()
            return {k: init[k].clone() for k in bufs}
        if policy == "mean" and self.world > 1:
            import torch.distributed as dist
            for k, v in bufs.items():
                if v.dtype.is_floating_point:
                    dist.all_reduce(v)
                    v /= self.world
                else:
                    dist.all_reduce(v, op=dist.ReduceOp.MAX)
        return {k: v.cpu() for k, v in bufs.items()}
