"""gRPC front door onto the ON-DEVICE parameter store.

In the reference the four RPCs *are* the store's interface (reference: src/parameter_server/server.py:190-318:
RegisterWorker :190-211, FetchParameters :213-230, PushGradrients :239-262, JobFinished :306-318).  On the box the
workers talk to the sharded store through the push / pull kernels instead, but a remote or CPU worker -- or the
reference's own ``ParameterServerStub`` -- still needs that door.  ``DeviceStoreServicer`` is it: hosted by rank 0 of
a running job (``Config.serve_port``), same proto, same payload codecs, and every RPC acts on the GPU shards:

  RegisterWorker   reserves one of ``Config.remote_slots`` extra SOURCE slots (worker id W+x) and sets its membership
                   bit in every owner's flag block (peer stores).
  FetchParameters  reads every owner's fp32 master shard over the peer mappings (no collective), un-shards it into
                   the reference's 122-key state_dict (BN buffers from rank 0's replica) and encodes it.
  PushGradrients   decodes the gradient dict, lays it out in arena order, copies it to the device and runs the SEND
                   half of the push kernel for source W+x: fp16 tiles into slot W+x of every owner's staging area +
                   arrival signals -- the same path an on-box worker's gradients take.  The owners' REDUCE then
                   averages over W + (remote gradients that arrived since the last round) sources: the on-box
                   workers never wait for a remote one (a CPU worker needs seconds per step), its gradient is
                   folded into whichever round is current.  For the remote worker the push is a true barrier: the
                   reply is sent when every shard has applied the round that consumed its gradient.
  JobFinished      detaches the source (its membership bit is cleared on every owner).
"""
from __future__ import annotations

import threading
import time
from concurrent import futures
from typing import Dict, List

import grpc
import numpy as np
import torch

from ..ops import _lib as L
from . import codec, proto


class DeviceStoreServicer(proto.ParameterServerServicer):
    def __init__(self, job, verbose: bool = False):
        self.job, self.ps, self.cfg = job, job.ps, job.cfg
        if self.cfg.mode != "sync" and self.ps.n_extra:
            raise ValueError("the device front door attaches remote workers to SYNC rounds; async jobs serve "
                             "FetchParameters only (remote_slots must be 0)")
        self.verbose = verbose
        self.lock = threading.Lock()
        self.slots: Dict[int, dict] = {}                       # x -> {ip, pushes, registered_at, last_seen}
        self.stream = torch.cuda.Stream(device=job.device)     # door traffic never sits on the training stream
        self.gradients_processed = 0
        self.rejected = 0
        self.last_reject_reason = ""
        self.version_offset = job.version_offset              # shard version - round number (resume step)
        self._grad_dev = [torch.zeros(self.ps.total, dtype=torch.float32, device=job.device)
                          for _ in range(self.ps.n_extra)]
        self._grad_host = [torch.zeros(self.ps.total, dtype=torch.float32).pin_memory()
                           for _ in range(self.ps.n_extra)]

    # ------------------------------------------------------------------ helpers
    def _log(self, msg: str) -> None:
        if self.verbose:
            print(msg, flush=True)

    def _versions(self) -> List[int]:
        return [int(self.ps.peer_flags(r)[L.F_VERSION]) for r in range(self.ps.world)]

    @property
    def arrivals(self) -> int:
        """Signals one source sends per round and owner (sum of the SEND launches' CTAs)."""
        return int(self.job.worker.arrivals_per_round)

    # ------------------------------------------------------------------ RPCs
    def RegisterWorker(self, request, context):
        torch.cuda.set_device(self.job.device)
        with self.lock:
            free = [x for x in range(self.ps.n_extra) if x not in self.slots]
            if not free:
                if context is not None:
                    context.abort(grpc.StatusCode.RESOURCE_EXHAUSTED,
                                  f"all {self.ps.n_extra} remote slots of this job are taken")
                raise RuntimeError("all remote slots are taken")
            x = free[0]
            with torch.cuda.stream(self.stream):
                self.ps.attach_source(x)
            now = time.time()
            self.slots[x] = {"worker_ip": request.worker_ip, "pushes": 0, "registered_at": now, "last_seen": now}
        wid = self.ps.world + x
        self._log(f"[door] remote worker {wid} from {request.worker_ip} attached")
        return proto.RegisterReply(worker_id=wid, total_workers=self.ps.world + self.ps.n_extra)

    def _state_dict_numpy(self) -> Dict[str, np.ndarray]:
        """122 reference keys: trainable tensors from the shards, BN buffers from rank 0's worker replica."""
        if self.ps.n_extra:
            flat = self.ps.snapshot_master()
        elif self.ps.world == 1:
            flat = self.ps.gather_master()
        else:                      # no peer-visible master without remote slots: the bf16 replica + fp32 small prefix
            flat = self.ps.wbf16[self.ps.rank].float()
            flat[: self.ps.n_small] = self.ps.small[self.ps.rank][: self.ps.n_small]
        params = {k: v.cpu() for k, v in self.job.layout.unflatten(flat).items()}
        bufs = {k: v.detach().cpu() for k, v in self.job.worker.engine.bn_buffers().items()}
        if self.cfg.bn_buffer_policy == "reference":
            bufs = {}
        out = {}
        for k, v in self.job.model_cpu.state_dict().items():
            t = params[k] if k in params else bufs.get(k, v.detach())
            out[k] = t.contiguous().numpy()
        return out

    def FetchParameters(self, request, context):
        torch.cuda.set_device(self.job.device)
        x = request.worker_id - self.ps.world
        with self.lock:
            if x in self.slots:
                self.slots[x]["last_seen"] = time.time()
        with torch.cuda.stream(self.stream):
            step = min(self._versions())
            try:
                ping = context is not None and any(k == "dps-ping" for k, _ in (context.invocation_metadata() or ()))
            except Exception:
                ping = False
            if ping:
                return proto.FetchReply(parameters=b"", global_step=step)
            sd = self._state_dict_numpy()
        as_pickle = bool(self.cfg.compat_pickle)
        if self.cfg.pull_compress == "fp16" and not as_pickle:
            sd = {k: (v.astype(np.float16) if v.dtype == np.float32 else v) for k, v in sd.items()}
        payload = codec.encode_any(sd, as_pickle, {"global_step": step})
        return proto.FetchReply(parameters=payload, global_step=step)

    def PushGradrients(self, request, context):
        torch.cuda.set_device(self.job.device)
        x = request.worker_id - self.ps.world
        with self.lock:
            slot = self.slots.get(x)
        if slot is None:
            self.last_reject_reason = f"worker id {request.worker_id} is not attached"
            return proto.PushReply(received=False)
        try:
            tensors, _ = codec.decode_any(request.gradients, allow_pickle=self.cfg.compat_pickle)
        except Exception as e:                                  # malformed payload: reference's received=False path
            self.last_reject_reason = f"undecodable gradients: {e!r}"
            self._log(f"[door] bad gradients from worker {request.worker_id}: {e}")
            return proto.PushReply(received=False)
        grads = {k: torch.from_numpy(np.array(v, dtype=np.float32, copy=True)) for k, v in tensors.items()
                 if not k.startswith("buffer/") and k in self.job.layout.entries}
        host = self._grad_host[x]
        host.zero_()
        self.job.layout.flatten_into(host, grads)
        k = slot["pushes"] + 1
        with torch.cuda.stream(self.stream):
            alive = int(self.ps.peer_flags(self.ps.rank)[L.F_ALIVE])
            if not (alive >> (self.ps.world + x)) & 1:          # dropped by the owners after a missed barrier
                self.rejected += 1
                self.last_reject_reason = (f"source dropped by the owners (alive mask {alive:#x}, error bits "
                                           f"{self.ps.error_bits():#x})")
                return proto.PushReply(received=False)
            self._grad_dev[x].copy_(host, non_blocking=True)
            self.ps.push_remote(x, self._grad_dev[x], self.arrivals, stream=self.stream)
            self.stream.synchronize()
            slot["pushes"] = k
            slot["last_seen"] = time.time()
            self.gradients_processed += 1
            # true barrier: answer when every shard has folded this gradient into a round AND finished it
            deadline = time.time() + self.cfg.barrier_timeout_s
            while True:
                prog = self.ps.remote_progress(x)
                if all(c >= k and v >= rnd + 1 + self.version_offset for c, rnd, v in prog):
                    break
                if time.time() > deadline:
                    self.last_reject_reason = (f"push {k} not applied in time: (consumed, round, version) per owner = "
                                               f"{prog}, error bits {self.ps.error_bits():#x}")
                    self._log("[door] " + self.last_reject_reason)
                    return proto.PushReply(received=False)
                time.sleep(0.0005)
        return proto.PushReply(received=True)

    def JobFinished(self, request, context):
        torch.cuda.set_device(self.job.device)
        x = request.worker_id - self.ps.world
        with self.lock:
            slot = self.slots.pop(x, None)
        if slot is not None:
            with torch.cuda.stream(self.stream):
                self.ps.detach_source(x)
            self._log(f"[door] remote worker {request.worker_id} finished after {slot['pushes']} pushes")
        return proto.JobFinishedReply(message="Acknowledged")

    def metrics(self) -> dict:
        with self.lock:
            return {"remote_slots": self.ps.n_extra, "remote_attached": len(self.slots),
                    "remote_gradients_processed": self.gradients_processed, "remote_rejected": self.rejected}


def start_door(job, port: int, bind: str = "127.0.0.1", verbose: bool = False):
    """Start the gRPC server of a running job (rank 0).  Returns (server, servicer, bound_port)."""
    servicer = DeviceStoreServicer(job, verbose=verbose)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max(8, 2 * job.ps.n_extra + 4)),
                         options=proto.GRPC_OPTIONS)
    proto.add_ParameterServerServicer_to_server(servicer, server)
    bound = server.add_insecure_port(f"{bind}:{port}")
    server.start()
    return server, servicer, bound
