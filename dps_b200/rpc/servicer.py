"""CPU / remote parameter-server role served over gRPC.

This is the control-plane and slow-path twin of the on-box sharded store
(``dps_b200.store`` + the sm_100a push/pull kernels).  It reproduces the
behaviour of the reference servicer (reference: src/parameter_server/server.py:81-368)
with the defects listed in SURVEY §2.6 fixed behind flags:

  Q1  sync push is a *true* barrier by default (``nonblocking_sync=True`` restores the
      reference's return-immediately behaviour, server.py:264-288)
  Q2  the barrier counts distinct worker ids per round, not arrivals
  Q3  the barrier times out (``barrier_timeout_s``) and proceeds with the ranks present
  Q4  SGD has optional momentum / weight decay (0/0 == reference, server.py:133)
  Q5  BN running statistics can be maintained (``bn_buffer_policy``)
  Q8  the staleness bound is honoured as configured
  Q9  staleness statistics are running mean / max / histogram, not last-value-only
  Q10 worker ids are re-slotted instead of growing without bound
  Q11 SERVER_FINAL_METRICS is emitted exactly once (stdout + optional file)
  Q15 payloads are DPS1 frames; pickles only via the restricted unpickler
  Q16 every shared counter is mutated under a lock
"""
from __future__ import annotations

import threading
import time
from collections import deque
from concurrent import futures
from typing import Dict, List, Optional

import grpc
import numpy as np
import torch

from ..config import Config
from ..utils import metrics as M
from . import codec, proto


def staleness_weight(staleness: int, alpha: float = 0.1, floor: float = 0.1) -> float:
    """max(floor, 1/(1+alpha*s))  (reference: server.py:178)."""
    return max(floor, 1.0 / (1.0 + staleness * alpha))


class HostOptimizer:
    """fp32 SGD(+momentum, +weight decay) over a dict of numpy arrays, in place.

    momentum=0, weight_decay=0 is bit-for-bit ``p = p - lr*g`` (reference: server.py:126-143).
    """

    def __init__(self, params: Dict[str, np.ndarray], lr: float, momentum: float = 0.0,
                 weight_decay: float = 0.0):
        self.params = params
        self.lr, self.momentum, self.weight_decay = float(lr), float(momentum), float(weight_decay)
        self.velocity: Dict[str, np.ndarray] = {}

    def step(self, grads: Dict[str, np.ndarray], scale: float = 1.0) -> None:
        """In place, through torch's vectorised CPU kernels on the numpy buffers (a numpy ``p - lr*g`` allocates two
        temporaries per tensor: 7.7 ms per 2 M elements against 1.5 ms).  Every product and sum is a separate
        fp32 operation, exactly as in the numpy expression -- no fused multiply-add -- so momentum = weight_decay = 0
        stays bit-for-bit the reference's ``p = p - lr*g``."""
        for name, g in grads.items():
            p = self.params.get(name)
            if p is None or not np.issubdtype(p.dtype, np.floating):
                continue
            if not (p.flags.writeable and p.flags.c_contiguous):
                p = self.params[name] = np.array(p, dtype=np.float32, copy=True)
            pt = torch.from_numpy(p)
            g = np.asarray(g, dtype=np.float32)
            gt = torch.from_numpy(g if (g.flags.writeable and g.flags.c_contiguous) else np.array(g, copy=True))
            if scale != 1.0:
                gt = gt * float(np.float32(scale))
            if self.weight_decay:
                gt = gt + pt * float(np.float32(self.weight_decay))
            if self.momentum:
                v = self.velocity.get(name)
                if v is None:
                    v = self.velocity[name] = gt.clone().numpy()
                    gt = torch.from_numpy(v)
                else:
                    vt = torch.from_numpy(v)
                    vt.mul_(float(np.float32(self.momentum)))
                    vt.add_(gt)
                    gt = vt
            pt.sub_(gt * float(np.float32(self.lr)))          # p - lr*g: product rounded, then the difference


def mean_gradients(grad_sets: List[Dict[str, np.ndarray]]) -> Dict[str, np.ndarray]:
    """Element-wise mean over worker gradient dicts (reference: server.py:145-169): sum in worker order, then one
    division by the number of contributors."""
    if not grad_sets:
        return {}
    out: Dict[str, np.ndarray] = {}
    for name in grad_sets[0].keys():
        acc, n = None, 0
        for gs in grad_sets:
            g = gs.get(name)
            if g is None:
                continue
            g = np.ascontiguousarray(g)
            if acc is None:
                acc = np.array(g, dtype=np.float32, copy=True)
                at = torch.from_numpy(acc)
            else:
                at.add_(torch.from_numpy(g if g.flags.writeable else np.array(g, copy=True)).to(torch.float32))
            n += 1
        if n:
            if n > 1:
                at.div_(float(n))
            out[name] = acc
    return out


class ParameterServerServicer(proto.ParameterServerServicer):
    """gRPC servicer. Constructor keeps the reference's positional signature (server.py:84)."""

    def __init__(self, mode: str = "sync", total_workers: int = 4, learning_rate: float = 0.1,
                 staleness_bound: int = 5, config: Optional[Config] = None,
                 initial_state: Optional[Dict[str, np.ndarray]] = None, verbose: bool = False):
        cfg = (config or Config()).replace(mode=mode, workers=total_workers, lr=learning_rate,
                                           staleness_bound=staleness_bound).validate()
        self.cfg = cfg
        self.mode, self.total_workers = cfg.mode, cfg.workers
        self.learning_rate, self.staleness_bound = cfg.lr, cfg.staleness_bound
        self.verbose = verbose

        if initial_state is None:
            from ..models.resnet import build_model
            model = build_model(cfg.num_classes, seed=cfg.seed)
            initial_state = {k: v.detach().cpu().numpy() for k, v in model.state_dict().items()}
        resume_side: dict = {}
        if cfg.resume:                     # PS recovery: the reference loses the model when the server restarts
            from ..utils import checkpoint as CK
            sd, resume_side = CK.load(cfg.resume)
            initial_state = CK.to_numpy_state(sd)
        self.parameters: Dict[str, np.ndarray] = {k: np.array(v, copy=True) for k, v in initial_state.items()}
        self.param_lock = threading.Lock()
        self.opt = HostOptimizer(self.parameters, cfg.lr, cfg.momentum, cfg.weight_decay)
        for k, v in (resume_side.get("momentum") or {}).items():
            self.opt.velocity[k] = np.array(v.numpy() if hasattr(v, "numpy") else v, dtype=np.float32, copy=True)
        self._resume_step = int(resume_side.get("global_step", 0))
        self._last_ckpt_step = self._resume_step
        self._ckpt_lock = threading.Lock()
        self._ckpt_written_step = -1
        self._ckpt_threads: List[threading.Thread] = []
        self.checkpoint_errors = 0

        # membership
        self.registration_lock = threading.Lock()
        self.registered_workers: Dict[int, dict] = {}
        self.registered_workers_count = 0
        self.active_workers: set = set()
        self._free_ids: List[int] = []
        self._departed: set = set()        # ids that finished or were evicted: the sync barrier does not wait for them
        self._evicted: set = set()         # evicted as silent; revived by their next RPC if the id is still free
        self.membership_epoch = 0
        self._monitor: Optional[threading.Thread] = None
        self._monitor_stop = threading.Event()

        # training state
        self.global_step = self._resume_step
        self.total_updates = 0
        self.gradients_processed = 0
        self.stats_lock = threading.Lock()

        # sync barrier
        self.sync_cv = threading.Condition()
        self.round_id = 0
        self.pending_gradients: Dict[int, Dict[str, np.ndarray]] = {}
        self.pending_buffers: Dict[int, Dict[str, np.ndarray]] = {}
        self.round_started_at: Optional[float] = None
        self.barrier_timeouts = 0

        # async stats (Q9)
        self.async_updates = 0
        self.rejected_updates = 0
        self.gradient_staleness: Dict[int, int] = {}       # last value per worker (reference field)
        self.staleness_sum = 0
        self.staleness_max = 0
        self.staleness_hist: Dict[int, int] = {}

        self.update_times = deque(maxlen=100)
        self.start_time = time.time()
        self._fetch_cache: Dict[bool, tuple] = {}
        self._final_emitted = False
        self.final_metrics: Optional[dict] = None

    # ------------------------------------------------------------------ numerics
    def apply_gradients(self, grads: Dict[str, np.ndarray], scale: float = 1.0) -> None:
        t0 = time.time()
        with self.param_lock:
            self.opt.step(grads, scale)
            self.global_step += 1
            self.total_updates += 1
            self._fetch_cache.clear()
            due = (self.cfg.checkpoint and self.cfg.checkpoint_every > 0
                   and self.global_step - self._last_ckpt_step >= self.cfg.checkpoint_every)
            snap = None
            if due:
                # claim this checkpoint while still holding the lock (two concurrent async pushes must not both see
                # `due`) and take the snapshot here; the file is written by a background thread so that neither the
                # push RPC nor -- in sync mode -- the workers parked on the barrier wait for 45 MB of disk I/O
                self._last_ckpt_step = self.global_step
                snap = self._snapshot_locked()
        self.update_times.append(time.time() - t0)
        if snap is not None:
            t = threading.Thread(target=self._write_checkpoint, args=(self.cfg.checkpoint, snap), daemon=True,
                                 name="dps-ckpt")
            self._ckpt_threads = [x for x in self._ckpt_threads if x.is_alive()] + [t]
            t.start()

    def wait_for_checkpoints(self, timeout: Optional[float] = None) -> None:
        """Block until every periodic checkpoint that has been claimed so far is on disk."""
        for t in list(self._ckpt_threads):
            t.join(timeout)

    def _snapshot_locked(self) -> dict:
        return {"state": {k: np.array(v, copy=True) for k, v in self.parameters.items()},
                "momentum": {k: np.array(v, copy=True) for k, v in self.opt.velocity.items()},
                "step": self.global_step, "total_updates": self.total_updates}

    def _write_checkpoint(self, path: str, snap: dict) -> Optional[str]:
        """One writer at a time; an I/O failure is logged and never fails the push that triggered the save."""
        from ..utils import checkpoint as CK
        try:
            with self._ckpt_lock:
                if snap["step"] < self._ckpt_written_step:          # a newer snapshot is already on disk
                    return None
                CK.save(path, snap["state"], global_step=snap["step"], momentum=snap["momentum"],
                        config=self.cfg.to_dict(), extra={"mode": self.mode, "total_updates": snap["total_updates"]})
                self._ckpt_written_step = snap["step"]
            self._log(f"checkpoint written to {path} at global step {snap['step']}")
            return path
        except Exception as e:                                       # noqa: BLE001 - disk full, permissions, ...
            self.checkpoint_errors += 1
            self._log(f"checkpoint to {path} failed: {e!r}")
            return None

    def save_checkpoint(self, path: Optional[str] = None) -> Optional[str]:
        """state_dict-compatible checkpoint (loads into the reference ResNet18 with strict=True) + resume sidecar
        (global step, momentum, config).  The snapshot is taken under the parameter lock, the file is written
        outside it and renamed into place."""
        path = path or self.cfg.checkpoint
        if not path:
            return None
        with self.param_lock:
            snap = self._snapshot_locked()
            self._last_ckpt_step = snap["step"]
        self.wait_for_checkpoints()
        return self._write_checkpoint(path, snap)

    def aggregate_gradients_sync(self, worker_gradients: List[Dict[str, np.ndarray]]):
        return mean_gradients(worker_gradients)

    def apply_gradients_async(self, grads: Dict[str, np.ndarray], worker_id: int, staleness: int) -> bool:
        if staleness > self.staleness_bound:
            with self.stats_lock:
                self.rejected_updates += 1
            return False
        w = staleness_weight(staleness, self.cfg.staleness_alpha, self.cfg.staleness_floor)
        self.apply_gradients(grads, scale=w)
        with self.stats_lock:
            self.async_updates += 1
        return True

    def decompress_gradients(self, grads):
        return codec.decompress_gradients(grads)

    # ------------------------------------------------------------------ RPCs
    def RegisterWorker(self, request, context):
        with self.registration_lock:
            if self._free_ids:
                wid = self._free_ids.pop(0)
            elif self.registered_workers_count < self.total_workers:
                wid = self.registered_workers_count
                self.registered_workers_count += 1
            else:
                if context is not None:
                    context.abort(grpc.StatusCode.RESOURCE_EXHAUSTED,
                                  f"all {self.total_workers} worker slots are taken")
                raise RuntimeError("all worker slots are taken")
            now = time.time()
            self.registered_workers[wid] = {"worker_id": wid, "worker_ip": request.worker_ip,
                                            "registered_at": now, "last_seen": now}
            self.active_workers.add(wid)
            self._departed.discard(wid)
            self._evicted.discard(wid)
            self.membership_epoch += 1
        self._log(f"[RegisterWorker] Worker {wid} from {request.worker_ip} registered "
                  f"({len(self.active_workers)}/{self.total_workers})")
        return proto.RegisterReply(worker_id=wid, total_workers=self.total_workers)

    def _touch(self, wid: int) -> None:
        with self.registration_lock:
            info = self.registered_workers.get(wid)
            if info is not None:
                info["last_seen"] = time.time()
                if wid in self._evicted and wid in self._free_ids:
                    # it was only slow, not dead, and nobody took its slot meanwhile: back in
                    self._evicted.discard(wid)
                    self._free_ids.remove(wid)
                    self._departed.discard(wid)
                    self.active_workers.add(wid)
                    self.membership_epoch += 1

    def _fetch_payload(self) -> tuple:
        """(global_step, payload bytes, serialized FetchReply) of the current parameters, encoded once per update."""
        as_pickle = bool(self.cfg.compat_pickle)
        with self.param_lock:
            cached = self._fetch_cache.get(as_pickle)
            if cached is None or cached[0] != self.global_step:
                params = self.parameters
                if self.cfg.pull_compress == "fp16" and not as_pickle:
                    # floating tensors travel as fp16 (the frame is self-describing; the worker casts back);
                    # integer buffers (num_batches_tracked) and the fp32 master here are untouched
                    params = codec.compress_gradients({k: v for k, v in params.items() if v.dtype == np.float32},
                                                      "fp16") | {k: v for k, v in params.items()
                                                                 if v.dtype != np.float32}
                    params = {k: params[k] for k in self.parameters}          # keep the state_dict order
                payload = codec.encode_any(params, as_pickle, {"global_step": self.global_step})
                cached = (self.global_step, payload, proto.encode_fetch_reply_wire(payload, self.global_step))
                self._fetch_cache[as_pickle] = cached
            return cached

    @staticmethod
    def _is_ping(context) -> bool:
        """A FetchParameters sent with metadata ('dps-ping', '1') is a liveness ping: no payload wanted."""
        try:
            return any(k == "dps-ping" for k, _ in (context.invocation_metadata() or ()))
        except Exception:
            return False

    def FetchParameters(self, request, context):
        self._touch(request.worker_id)
        if context is not None and self._is_ping(context):
            return proto.FetchReply(parameters=b"", global_step=self.global_step)
        step, payload, _ = self._fetch_payload()
        return proto.FetchReply(parameters=payload, global_step=step)

    def FetchParametersWire(self, request, context):
        """Same RPC for the gRPC runtime: the reply is the cached, already-serialized message (every worker of a round
        gets the same bytes; no per-request copy of the 45 MB payload on the server)."""
        self._touch(request.worker_id)
        if context is not None and self._is_ping(context):
            return proto.encode_fetch_reply_wire(b"", self.global_step)
        return self._fetch_payload()[2]

    def PushGradrients(self, request, context):
        return self._push(request.worker_id, request.gradients, request.local_step)

    def PushGradrientsWire(self, raw, context):
        """Same RPC for the gRPC runtime: the request arrives as raw bytes and the 22 MB payload is used in place
        (no protobuf parse, no copy out of the message)."""
        wid, payload, local_step = proto.split_push_request_wire(raw)
        return self._push(wid, payload, local_step)

    def _push(self, wid: int, payload, local_step: int):
        try:
            tensors, meta = codec.decode_any(payload, allow_pickle=self.cfg.compat_pickle)
        except Exception as e:  # malformed payload == reference's received=False path (server.py:260-262)
            self._log(f"Error processing gradients from worker {wid}: {e}")
            return proto.PushReply(received=False)
        buffers = {k[len("buffer/"):]: np.asarray(v) for k, v in tensors.items() if k.startswith("buffer/")}
        grads = self.decompress_gradients({k: v for k, v in tensors.items() if not k.startswith("buffer/")})
        with self.stats_lock:
            self.gradients_processed += 1
        self._touch(wid)
        if self.mode == "sync":
            ok = self._handle_sync_gradients(wid, grads, local_step, buffers)
        else:
            ok = self._handle_async_gradients(wid, grads, local_step, buffers)
        return proto.PushReply(received=ok)

    # sync -----------------------------------------------------------------
    def _expected(self) -> int:
        """Contributions a sync round waits for: W minus the workers that have left (finished their partition, or
        were evicted as dead) and not come back.  The reference waits for W arrivals forever (Q3)."""
        return max(1, self.total_workers - len(self._departed))

    def _complete_round_locked(self) -> None:
        sets = list(self.pending_gradients.values())
        agg = self.aggregate_gradients_sync(sets)
        if agg:
            self.apply_gradients(agg)
        self._merge_buffers(list(self.pending_buffers.values()))
        self.pending_gradients.clear()
        self.pending_buffers.clear()
        self.round_started_at = None
        self.round_id += 1
        self.sync_cv.notify_all()

    def _handle_sync_gradients(self, wid, grads, local_step, buffers=None) -> bool:
        with self.sync_cv:
            my_round = self.round_id
            if not self.pending_gradients:
                self.round_started_at = time.time()
            self.pending_gradients[wid] = grads            # Q2: keyed by id; completion counts ids
            if buffers:
                self.pending_buffers[wid] = buffers
            if len(self.pending_gradients) >= self._expected():
                self._complete_round_locked()
                return True
            if self.cfg.nonblocking_sync:                   # Q1 compat: reference never blocks
                return True
            deadline = (self.round_started_at or time.time()) + self.cfg.barrier_timeout_s
            while self.round_id == my_round:
                if len(self.pending_gradients) >= self._expected():     # somebody left while we were waiting
                    self._complete_round_locked()
                    break
                left = deadline - time.time()
                if left <= 0:                               # Q3: proceed with the ranks present
                    if self.round_id == my_round:
                        self.barrier_timeouts += 1
                        self.membership_epoch += 1
                        self._log(f"sync barrier timed out with {len(self.pending_gradients)}/"
                                  f"{self._expected()} workers; proceeding")
                        self._complete_round_locked()
                    break
                self.sync_cv.wait(timeout=left)
            return True

    # async ----------------------------------------------------------------
    def _handle_async_gradients(self, wid, grads, local_step, buffers=None) -> bool:
        with self.param_lock:
            staleness = self.global_step - local_step       # Q6: read under the lock
        ok = self.apply_gradients_async(grads, wid, staleness)
        if ok and buffers:
            self._merge_buffers([buffers], ema=1.0 / max(1, self.total_workers))
        with self.stats_lock:
            self.gradient_staleness[wid] = staleness
            self.staleness_sum += max(0, staleness)
            self.staleness_max = max(self.staleness_max, staleness)
            self.staleness_hist[staleness] = self.staleness_hist.get(staleness, 0) + 1
        return ok

    # BN buffers (Q5) --------------------------------------------------------
    def _merge_buffers(self, buffer_sets: List[Dict[str, np.ndarray]], ema: Optional[float] = None) -> None:
        if not buffer_sets or self.cfg.bn_buffer_policy == "reference":
            return
        with self.param_lock:
            for name in buffer_sets[0].keys():
                cur = self.parameters.get(name)
                if cur is None:
                    continue
                vals = [bs[name] for bs in buffer_sets if name in bs]
                if np.issubdtype(cur.dtype, np.integer):
                    self.parameters[name] = np.maximum(cur, np.max(np.stack(vals), axis=0)).astype(cur.dtype)
                    continue
                mean = np.mean(np.stack([v.astype(np.float32) for v in vals]), axis=0)
                if ema is None:
                    self.parameters[name] = mean.astype(cur.dtype)
                else:
                    self.parameters[name] = ((1.0 - ema) * cur + ema * mean).astype(cur.dtype)
            self._fetch_cache.clear()

    # finish ---------------------------------------------------------------
    def JobFinished(self, request, context):
        wid = request.worker_id
        emit = False
        with self.registration_lock:
            if wid in self.active_workers:
                self.active_workers.discard(wid)
                self._free_ids.append(wid)
                self._free_ids.sort()
                self._departed.add(wid)
                self.membership_epoch += 1
            if not self.active_workers and not self._final_emitted:
                self._final_emitted = emit = True
        with self.sync_cv:                       # a round may now be complete without this worker
            self.sync_cv.notify_all()
        if emit:
            self._print_final_statistics()
            if self.cfg.checkpoint:
                self.save_checkpoint()
        return proto.JobFinishedReply(message="Acknowledged")

    # liveness -------------------------------------------------
    def stale_workers(self, older_than_s: float) -> List[int]:
        now = time.time()
        with self.registration_lock:
            return [w for w in self.active_workers
                    if now - self.registered_workers[w]["last_seen"] > older_than_s]

    def evict(self, wid: int) -> None:
        """Drop a dead worker: frees its id/partition for re-admission and unblocks the barrier."""
        with self.registration_lock:
            if wid in self.active_workers:
                self.active_workers.discard(wid)
                self._free_ids.append(wid)
                self._free_ids.sort()
                self._departed.add(wid)
                self._evicted.add(wid)
                self.membership_epoch += 1
        with self.sync_cv:
            self.sync_cv.notify_all()

    def start_monitor(self, period_s: Optional[float] = None) -> None:
        """Liveness watchdog (the reference writes `last_seen` on every RPC and never reads it, server.py:200): evict
        workers that have been silent for `worker_timeout_s`; their id and partition become available again."""
        timeout = self.cfg.worker_timeout_s
        if timeout <= 0 or self._monitor is not None:
            return
        period = period_s if period_s is not None else max(0.05, min(5.0, timeout / 3.0))

        def loop():
            while not self._monitor_stop.wait(period):
                for w in self.stale_workers(timeout):
                    self._log(f"worker {w} silent for more than {timeout:.1f}s: evicted")
                    self.evict(w)

        self._monitor = threading.Thread(target=loop, daemon=True, name="dps-liveness")
        self._monitor.start()

    def stop_monitor(self) -> None:
        self._monitor_stop.set()

    # metrics --------------------------------------------------------------
    def server_metrics(self) -> dict:
        total_time = time.time() - self.start_time
        avg_update = float(np.mean(self.update_times)) if self.update_times else 0.0
        extra = {}
        if self.mode == "async" and self.gradient_staleness:
            n = max(1, sum(self.staleness_hist.values()))
            extra = {
                "async_updates": self.async_updates,
                "average_gradient_staleness": round(float(np.mean(list(self.gradient_staleness.values()))), 2),
                "max_staleness_observed": max(self.gradient_staleness.values()),
                "rejected_updates": self.rejected_updates,
                "running_mean_staleness": round(self.staleness_sum / n, 4),
                "running_max_staleness": self.staleness_max,
                "staleness_histogram": {str(k): v for k, v in sorted(self.staleness_hist.items())},
            }
        return M.server_final_metrics(
            mode=self.mode, total_workers=self.total_workers, total_time=total_time,
            global_step=self.global_step, total_updates=self.total_updates,
            gradients_processed=self.gradients_processed, avg_update_time=avg_update,
            learning_rate=self.learning_rate,
            extra={**extra, "barrier_timeouts": self.barrier_timeouts,
                   "membership_epoch": self.membership_epoch, "config": self.cfg.to_dict()})

    def _print_final_statistics(self) -> None:
        self.final_metrics = self.server_metrics()
        M.emit(self.final_metrics, self.cfg.metrics_dir, "server")

    def state_dict_numpy(self) -> Dict[str, np.ndarray]:
        with self.param_lock:
            return {k: np.array(v, copy=True) for k, v in self.parameters.items()}

    def _log(self, msg: str) -> None:
        if self.verbose:
            print(msg, flush=True)


def build_server(cfg: Config, servicer: Optional[ParameterServerServicer] = None,
                 bind: str = "[::]", verbose: bool = False):
    """Create (server, servicer, bound_port) without blocking.  Port 0 picks a free port."""
    servicer = servicer or ParameterServerServicer(cfg.mode, cfg.workers, cfg.lr, cfg.staleness_bound,
                                                   config=cfg, verbose=verbose)
    # a true barrier parks up to W handler threads; leave room for fetches/registrations
    pool = futures.ThreadPoolExecutor(max_workers=max(20, 2 * cfg.workers + 8))
    server = grpc.server(pool, options=proto.GRPC_OPTIONS)
    proto.add_ParameterServerServicer_to_server(servicer, server)
    port = server.add_insecure_port(f"{bind}:{cfg.port}")
    servicer.start_monitor()
    return server, servicer, port


def serve(mode: str = "sync", total_workers: int = 4, learning_rate: float = 0.1, port: int = 8000,
          staleness_bound: int = 5, config: Optional[Config] = None) -> None:
    """Blocking entry point with the reference's signature (server.py:370) + staleness_bound."""
    cfg = (config or Config()).replace(mode=mode, workers=total_workers, lr=learning_rate, port=port,
                                       staleness_bound=staleness_bound).validate()
    server, servicer, port = build_server(cfg, verbose=True)
    server.start()
    print(f"--- Parameter Server listening on port {port} ---")
    print(f"Mode: {cfg.mode.upper()}, Workers: {cfg.workers}, LR: {cfg.lr}, staleness bound: {cfg.staleness_bound}")
    try:
        server.wait_for_termination()
    except KeyboardInterrupt:
        print("Server shutting down...")
        server.stop(0)
        if not servicer._final_emitted:
            servicer._print_final_statistics()


def main(argv=None) -> int:
    from ..config import parse_server_cli
    try:
        cfg = parse_server_cli(argv)
    except ValueError as e:
        print(f"Error: {e}")
        return 1
    serve(cfg.mode, cfg.workers, cfg.lr, cfg.port, cfg.staleness_bound, config=cfg)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
