"""Remote / CPU worker role over gRPC.

Same public surface as the reference worker (reference: src/workers/worker.py:78-453):
``DistributedWorker(server_address, worker_name)`` with ``connect_to_server``,
``setup_model``, ``setup_data``, ``fetch_parameters``, ``push_gradients``,
``train_local_batch``, ``evaluate_model``, ``run_training``, ``cleanup``.

Differences: gradients are serialised once (the reference pickles the
fp32 dict just to log a ratio, worker.py:281-292); ``sync_steps=k`` accumulates the k
local gradients instead of discarding k-1 of them (Q12; ``local_step_policy='reference'``
restores that); a liveness thread actually runs (the reference's is dead code,
worker.py:112-126) and never re-registers under a new id; BN running statistics
travel with the push so evaluation does not use untrained statistics (Q5).
"""
from __future__ import annotations

import threading
import time
import warnings
from typing import Dict, Optional

import grpc
import numpy as np
import torch
from torch import nn

from ..config import Config
from ..models.resnet import build_model
from ..utils import data as D
from ..utils import metrics as M
from . import codec, proto


_TORCH_DTYPES = {"float32": torch.float32, "float16": torch.float16, "float64": torch.float64,
                 "int64": torch.int64, "int32": torch.int32, "uint8": torch.uint8, "int8": torch.int8,
                 "bool": torch.bool, "int16": torch.int16}


class SimulatedCrash(RuntimeError):
    """Raised by the --inject-fault test hook."""


class DistributedWorker:
    def __init__(self, server_address: str = "localhost:8000", worker_name: str = "worker-node",
                 config: Optional[Config] = None, verbose: bool = False):
        self.cfg = (config or Config()).replace(server=server_address, worker_name=worker_name)
        self.server_address, self.worker_name = server_address, worker_name
        self.worker_id: Optional[int] = None
        self.total_workers: Optional[int] = None
        # CLI-compatible knobs (reference CLI defaults 128 / 0.1 / 3 / 1, worker.py:464-469)
        self.batch_size = self.cfg.batch_size
        self.learning_rate = self.cfg.lr
        self.num_epochs = self.cfg.epochs
        self.local_steps_per_sync = self.cfg.sync_steps
        self.model: Optional[nn.Module] = None
        self.criterion = None
        self.train_loader = self.test_loader = None
        self.channel = None
        self.stub: Optional[proto.ParameterServerStub] = None
        self._chan_lock = threading.Lock()
        self.local_step_counter = 0
        self.global_step_cache = 0
        self.training_start_time: Optional[float] = None
        self.epoch_times, self.accuracies = [], []
        self.training_active = True
        self.rejected_pushes = 0
        self.timer = M.PhaseTimer()
        self.verbose = verbose
        self.device = torch.device("cuda" if (self.cfg.device == "cuda" or
                                              (self.cfg.device == "auto" and torch.cuda.is_available()))
                                   else "cpu")
        self._hb_thread: Optional[threading.Thread] = None
        self.max_steps: Optional[int] = None     # test/bench hook: stop after this many local steps
        self.images_seen = 0

    # ------------------------------------------------------------------ connection
    def connect_to_server(self, max_retries: int = 5, backoff_base: float = 2.0) -> None:
        """Insecure channel + RegisterWorker with exponential back-off (worker.py:199-231)."""
        self.channel = proto.make_channel(self.server_address)
        self.stub = proto.ParameterServerStub(self.channel)
        for attempt in range(max_retries):
            try:
                rep = self.stub.RegisterWorker(proto.RegisterRequest(worker_ip=self.worker_name))
                self.worker_id, self.total_workers = rep.worker_id, rep.total_workers
                self._log(f"Registered as Worker {self.worker_id} (Total workers: {self.total_workers})")
                return
            except grpc.RpcError as e:
                if e.code() == grpc.StatusCode.RESOURCE_EXHAUSTED or attempt == max_retries - 1:
                    raise
                time.sleep(backoff_base ** attempt)

    def ping(self, timeout: float = 10.0) -> int:
        """Cheap liveness RPC: a FetchParameters whose request metadata asks for an EMPTY payload (this framework's
        server honours it and still refreshes `last_seen`; the reference server would answer with the full 45 MB).
        Returns the server's global step."""
        rep = self.stub.FetchParameters(proto.FetchRequest(worker_id=self.worker_id), timeout=timeout,
                                        metadata=(("dps-ping", "1"),))
        return int(rep.global_step)

    def start_heartbeat(self, period_s: float = 30.0) -> None:
        """Liveness ping (reference: the health-check loop that is never started, worker.py:112-126); reconnects the
        channel (same worker id) on failure.  The stub is swapped under a lock and the old channel is closed."""
        def loop():
            while self.training_active:
                time.sleep(period_s)
                if not self.training_active:
                    break
                try:
                    self.ping()
                except grpc.RpcError:
                    try:
                        channel = proto.make_channel(self.server_address)
                        stub = proto.ParameterServerStub(channel)
                        with self._chan_lock:
                            old, self.channel, self.stub = self.channel, channel, stub
                        if old is not None:
                            old.close()
                    except Exception:
                        pass
        self._hb_thread = threading.Thread(target=loop, daemon=True)
        self._hb_thread.start()

    # ------------------------------------------------------------------ setup
    def setup_model(self) -> None:
        self.model = build_model(self.cfg.num_classes, seed=self.cfg.seed).to(self.device)
        self.amp = self.cfg.amp == "bf16"
        if self.amp:                                  # oneDNN / cuDNN bf16 kernels want NHWC
            self.model = self.model.to(memory_format=torch.channels_last)
        self.criterion = nn.CrossEntropyLoss()

    def setup_data(self) -> None:
        cfg = self.cfg.replace(batch_size=self.batch_size)
        self.train_loader, self.test_loader, self.partition = D.make_loaders(
            cfg, self.worker_id, self.total_workers)
        self._log(f"Worker {self.worker_id} assigned samples {self.partition[0]}-{self.partition[1] - 1}")

    # ------------------------------------------------------------------ pull / push
    def fetch_parameters(self) -> int:
        with self.timer.phase("pull"):
            req = proto.FetchRequest(worker_id=self.worker_id)
            raw_call = getattr(self.stub, "FetchParametersRaw", None)
            if raw_call is not None:              # zero-copy: no protobuf parse, no copy of the payload out of it
                payload, step = proto.split_fetch_reply_wire(raw_call(req))
            else:
                rep = self.stub.FetchParameters(req)
                payload, step = rep.parameters, rep.global_step
            arrays, _ = codec.decode_any(payload, allow_pickle=self.cfg.compat_pickle)
            self.global_step_cache = step
            sd = self.model.state_dict()
            keep_local_bn = self.cfg.bn_buffer_policy == "local" and self.local_step_counter > 0
            with torch.no_grad(), warnings.catch_warnings():
                warnings.simplefilter("ignore")       # frombuffer on a read-only frame: we only read from it
                for name, arr in arrays.items():
                    dst = sd.get(name)
                    if dst is None:
                        continue
                    if keep_local_bn and ("running_" in name or "num_batches" in name):
                        continue
                    if arr.size == 0:
                        continue
                    # zero-copy view of the received frame (read-only is fine: it is only a copy source)
                    src = torch.frombuffer(arr, dtype=_TORCH_DTYPES[arr.dtype.name]).view(dst.shape) \
                        if arr.dtype.name in _TORCH_DTYPES and arr.flags.c_contiguous \
                        else torch.from_numpy(np.array(arr, copy=True))
                    dst.copy_(src)
        return step

    def compress_gradients(self, grads: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
        return codec.compress_gradients(grads, self.cfg.compress)

    def push_gradients(self) -> bool:
        with self.timer.phase("push"):
            grads = {n: p.grad.detach().cpu().numpy() for n, p in self.model.named_parameters()
                     if p.grad is not None}
            tensors = self.compress_gradients(grads)
            as_pickle = bool(self.cfg.compat_pickle)
            if not as_pickle and self.cfg.bn_buffer_policy != "reference":
                for n, b in self.model.named_buffers():
                    tensors["buffer/" + n] = b.detach().cpu().numpy()
            payload = codec.encode_any(tensors, as_pickle, {"worker": self.worker_id})
            self.last_push_bytes = len(payload)
            raw_call = getattr(self.stub, "PushGradrientsRaw", None)
            if raw_call is not None:
                rep = raw_call(proto.encode_push_request_wire(self.worker_id, payload, self.global_step_cache))
            else:
                rep = self.stub.PushGradrients(proto.PushRequest(
                    worker_id=self.worker_id, gradients=payload, local_step=self.global_step_cache))
            if not rep.received:
                self.rejected_pushes += 1
            return rep.received

    # ------------------------------------------------------------------ compute
    def train_local_batch(self, data: torch.Tensor, targets: torch.Tensor, zero: bool = True) -> float:
        with self.timer.phase("compute"):
            self.model.train()
            if zero:
                self.model.zero_grad(set_to_none=True)
            data, targets = data.to(self.device), targets.to(self.device)
            if self.amp:
                data = data.contiguous(memory_format=torch.channels_last)
            with torch.autocast(self.device.type, dtype=torch.bfloat16, enabled=self.amp):
                logits = self.model(data)
            loss = self.criterion(logits.float(), targets)
            loss.backward()
            self.local_step_counter += 1
            self.images_seen += int(targets.numel())
            return float(loss.item())

    def evaluate_model(self, max_batches: Optional[int] = None) -> float:
        with self.timer.phase("eval"):
            acc = D.evaluate(self.model, self.test_loader, self.device, max_batches)
        self.accuracies.append(acc)
        return acc

    # ------------------------------------------------------------------ loop
    def run_training(self, evaluate: bool = True, eval_batches: Optional[int] = None) -> None:
        self.training_start_time = time.time()
        self._crashed, self._fault_step = False, None
        if self.cfg.inject_fault:
            fid, fstep = self.cfg.inject_fault.split(":")
            if int(fid) == self.worker_id:
                self._fault_step = int(fstep)
        k = max(1, self.local_steps_per_sync)
        accumulate = self.cfg.local_step_policy == "accumulate"
        try:
            done = False
            for epoch in range(self.num_epochs):
                t0 = time.time()
                for batch_idx, (x, y) in enumerate(self.train_loader):
                    first = batch_idx % k == 0
                    if first:
                        self.fetch_parameters()
                    # reference: zero_grad every batch, push only on the first of each k (Q12)
                    self.train_local_batch(x, y, zero=first or not accumulate)
                    is_push = ((batch_idx + 1) % k == 0 or batch_idx + 1 == len(self.train_loader)) \
                        if accumulate else first
                    if is_push:
                        if accumulate and k > 1:
                            n_acc = (batch_idx % k) + 1
                            with torch.no_grad():
                                for p in self.model.parameters():
                                    if p.grad is not None:
                                        p.grad.div_(n_acc)
                        self.push_gradients()
                    if self.max_steps is not None and self.local_step_counter >= self.max_steps:
                        done = True
                        break
                    if self._fault_step is not None and self.local_step_counter >= self._fault_step:
                        # test hook (--inject-fault id:step): die like a killed process -- no JobFinished,
                        # no more heartbeats; the server's barrier timeout / liveness monitor must cope
                        self._crashed = True
                        self.training_active = False
                        raise SimulatedCrash(f"worker {self.worker_id} injected fault at step {self._fault_step}")
                self.epoch_times.append(time.time() - t0)   # excludes evaluation (Q17)
                if evaluate:
                    self.evaluate_model(eval_batches)
                if done:
                    break
        finally:
            if not self._crashed:
                self.cleanup()
            elif self.channel is not None:
                self.channel.close()
            self.print_worker_statistics()

    def print_worker_statistics(self) -> dict:
        total = time.time() - (self.training_start_time or time.time())
        train_time = sum(self.epoch_times)
        rec = M.worker_final_metrics(
            worker_id=self.worker_id, total_workers=self.total_workers, total_time=total,
            epoch_times=self.epoch_times, accuracies=self.accuracies,
            local_steps=self.local_step_counter, batch_size=self.batch_size,
            learning_rate=self.learning_rate, num_epochs=self.num_epochs,
            extra={"images_per_second": round(self.images_seen / train_time, 2) if train_time > 0 else 0.0,
                   "rejected_pushes": self.rejected_pushes, "phases": self.timer.summary(),
                   "transport": "grpc", "device": str(self.device)})
        self.final_metrics = rec
        M.emit(rec, self.cfg.metrics_dir, "worker")
        return rec

    def cleanup(self) -> None:
        self.training_active = False
        try:
            if self.stub is not None and self.worker_id is not None:
                self.stub.JobFinished(proto.JobFinishedRequest(worker_id=self.worker_id), timeout=10)
        except Exception:
            pass
        if self.channel is not None:
            self.channel.close()

    def _log(self, msg: str) -> None:
        if self.verbose:
            print(msg, flush=True)


def main(argv=None) -> int:
    from ..config import parse_worker_cli
    cfg = parse_worker_cli(argv)
    w = DistributedWorker(cfg.server, cfg.worker_name, config=cfg, verbose=True)
    try:
        w.connect_to_server()
        w.setup_model()
        w.setup_data()
        w.start_heartbeat()
        w.run_training()
    except Exception as e:
        print(f"Worker failed: {e}")
        return 1
    print("Worker completed successfully!")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
