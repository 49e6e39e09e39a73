"""On-box sharded parameter server: the device-side ParameterServer + Worker transport.

Each rank owns one shard of the store (fp32 master + momentum) and one worker replica
(bf16 weight arena + fp32 gradient arena).  The three data-path operations are the
hand-written sm_100a kernels of csrc/ps_kernels.cu:

    push  (sync)  : dps_push_sync   -- compress + peer store + arrival + reduce + SGD
    pull          : dps_pull_bcast  -- cast + broadcast into every worker's bf16 arena
    async         : dps_push_async / dps_async_apply / dps_pull_async -- arrival ring, per-shard
                    version counter, staleness weight/reject

Reference semantics preserved (reference: src/parameter_server/server.py:126-186, 264-304):
mean over workers then ``p -= lr*g`` in sync mode; ``staleness = version - version_seen``,
reject above the bound, weight ``max(0.1, 1/(1+0.1 s))`` in async mode.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from ..config import Config
from ..ops import _lib as L
from ..store.shardmap import Layout
from .fabric import ArenaSpec, Fabric, MulticastArena


def _default_ctas(shard_elems: int) -> int:
    # every CTA must be co-resident (owner CTAs spin on arrival counters).  256-thread CTAs with no shared
    # memory fit 8 per SM, so 4 per SM (592) is always resident and gives the copy loops enough loads in flight.
    want = max(8, shard_elems // (256 * 8 * 2))
    return int(min(592, want))


class DevicePS:
    def __init__(self, layout: Layout, cfg: Config, rank: int, world: int, device_index: int,
                 emulate: bool = False, group=None, timeout_s: Optional[float] = None):
        self.layout, self.cfg, self.rank, self.world = layout, cfg, rank, world
        self.emulate = emulate
        self.device = torch.device("cuda", device_index)
        self.E, self.total, self.n_small = layout.shard_elems, layout.total, layout.n_small
        self.R = max(1, min(L.MAX_SLOTS, cfg.async_ring_slots * world))
        self.compress = cfg.compress == "fp16"
        self.is_async = cfg.mode == "async"
        self.ctas = _default_ctas(self.E)
        self.async_ctas = max(1, min(self.ctas, 128 // world))
        # The gather may have to wait for an owner's apply to finish (seq-lock): it must never fill the SMs, or the
        # apply it waits for could not be scheduled beside it (measured: 592 CTAs at W = 1 spun out their bound and
        # counted torn pulls).  256 CTAs in all = fewer than two per SM.
        self.pull_async_ctas = max(1, min(self.ctas, 256 // world))
        self.timeout_ns = int((timeout_s if timeout_s is not None else cfg.barrier_timeout_s) * 1e9)
        # remote workers attached through the gRPC front door are extra SOURCES of the sync push: they get a staging
        # slot on every owner after the W on-box workers, and the master shards move into the peer-visible arena so
        # that the rank hosting the door can read all of them without a collective
        self.n_extra = int(cfg.remote_slots) if cfg.mode == "sync" else 0
        if world + self.n_extra > L.MAX_RANKS:
            raise ValueError(f"workers + remote_slots must be <= {L.MAX_RANKS}")

        # NVLS reduction of the sync push (Config.nvls_reduce, needs nvls != 'off'): the staging area moves into the
        # multicast arena, [owner][E] fp16 on every rank; sources store into their own replica, owners multimem.ld_reduce
        want_mc_stage = bool(cfg.nvls_reduce) and cfg.mode == "sync"
        if want_mc_stage and (self.n_extra or not self.compress):
            raise ValueError("nvls_reduce sums every replica in the switch: fp16 wire format and no remote_slots")
        spec = ArenaSpec()
        spec.add("flags", L.FLAG_WORDS * 4)
        stage_bytes = (world + self.n_extra) * self.E * (2 if self.compress else 4)
        if self.n_extra:
            spec.add("master", self.E * 4)
        # NVLS (Config.nvls): the two pull destinations live in a multicast-bound arena instead, so the owner's
        # broadcast is one multimem.st per vector.  'auto' = use it when the box supports multicast, 'on' = require it.
        want_mc = cfg.nvls != "off" and world > 1 and not emulate
        self.mc: Optional[MulticastArena] = None
        if want_mc:
            try:
                if not MulticastArena.supported(device_index):
                    raise RuntimeError("this device reports no multicast support")
                mspec = ArenaSpec()
                mspec.add("wbf16", self.total * 2)
                mspec.add("small", max(self.n_small, 8) * 4)
                if want_mc_stage:
                    mspec.add("stage", stage_bytes)
                self.mc = MulticastArena(rank, world, device_index, mspec, group=group)
                if not self.mc.mc_base:
                    raise RuntimeError("no multicast window could be mapped")
            except Exception as e:
                if cfg.nvls == "on":
                    raise RuntimeError(f"Config.nvls='on': {e}") from e
                import sys
                print(f"[dps_b200] rank {rank}: NVLS pull not available ({e!r}); per-peer stores", file=sys.stderr, flush=True)
                self.mc = None
            want_mc = self.mc is not None
        if not want_mc:
            spec.add("wbf16", self.total * 2)
            spec.add("small", max(self.n_small, 8) * 4)
        if want_mc_stage and not want_mc:
            if cfg.nvls == "off":
                raise ValueError("nvls_reduce needs nvls = 'auto' or 'on'")
            want_mc_stage = False                            # 'auto' on a box without multicast: per-peer push
        if not want_mc_stage:
            spec.add("stage", stage_bytes)
        if self.is_async:
            spec.add("ring", self.R * self.E * 2)
            spec.add("pub", self.E * 2)
            spec.add("pubsmall", max(self.n_small, 8) * 4)
        self.fabric = Fabric(rank, world, device_index, spec, emulate=emulate, group=group)
        self.lib = self.fabric.lib
        self.nvls = bool(self.mc is not None and self.mc.mc_base)
        self.nvls_reduce = bool(self.nvls and want_mc_stage)
        warena = self.mc if self.mc is not None else self.fabric      # where wbf16 / small live
        sarena = self.mc if self.nvls_reduce else self.fabric         # where the sync staging area lives

        # local-only buffers
        self.local = list(self.fabric.local_ranks())
        dev = self.device
        self.grad: Dict[int, torch.Tensor] = {r: torch.zeros(self.total, dtype=torch.float32, device=dev)
                                              for r in self.local}
        if self.n_extra:
            self.master: Dict[int, torch.Tensor] = {r: self.fabric.tensor(r, "master", torch.float32) for r in self.local}
        else:
            self.master = {r: torch.zeros(self.E, dtype=torch.float32, device=dev) for r in self.local}
        self.mom: Dict[int, torch.Tensor] = {r: torch.zeros(self.E, dtype=torch.float32, device=dev)
                                             for r in self.local}
        self.wbf16 = {r: warena.tensor(r, "wbf16", torch.bfloat16) for r in self.local}
        self.small = {r: warena.tensor(r, "small", torch.float32) for r in self.local}
        self.flags = {r: self.fabric.tensor(r, "flags", torch.int32) for r in self.local}
        for r in self.local:                 # membership: every source is waited for until it misses a barrier
            self.flags[r][L.F_ALIVE] = (1 << world) - 1 if world < 32 else -1

        t = L.DpsTables()
        t.world, t.block_elems, t.ring_slots, t.n_extra = world, layout.block_elems, self.R, self.n_extra
        t.shard_elems, t.n_small = self.E, self.n_small
        t.rank_lo, t.rank_hi = (0, world) if emulate else (rank, rank + 1)
        for r in range(world):
            t.stage[r] = sarena.ptr(r, "stage")
            t.wbf16[r] = warena.ptr(r, "wbf16")
            t.small[r] = warena.ptr(r, "small")
            t.flags[r] = self.fabric.ptr(r, "flags")
            if self.is_async:
                t.ring[r] = self.fabric.ptr(r, "ring")
                t.pub[r] = self.fabric.ptr(r, "pub")
                t.pubsmall[r] = self.fabric.ptr(r, "pubsmall")
        for r in self.local:
            t.grad[r] = self.grad[r].data_ptr()
            t.master[r] = self.master[r].data_ptr()
            t.mom[r] = self.mom[r].data_ptr()
        if self.nvls:
            t.mc_wbf16, t.mc_small = self.mc.mc_ptr("wbf16"), self.mc.mc_ptr("small")
        if self.nvls_reduce:
            t.mc_stage = self.mc.mc_ptr("stage")
        # device-resident round counters + learning rate (graph replay needs no baked-in step / lr)
        self.step_ctr = torch.zeros(4, dtype=torch.int32, device=dev)
        self.lr_dev = torch.full((1,), float(cfg.lr), dtype=torch.float32, device=dev)
        self.device_counters = False
        self.tables = t
        self._tp = C.byref(self.tables)
        import threading
        self._serve_lock = threading.Lock()
        self._server = None
        self.server_passes = 0
        self.step = 0                       # sync round / pull sequence number
        self.async_steps = {r: 0 for r in self.local}
        self.launches = 0

    def use_device_counters(self, on: bool = True) -> None:
        """Read round number and lr from device memory (required before CUDA-graph capture)."""
        assert not self.emulate or not on, "device counters are per-process; not used in emulation"
        self.step_ctr[0] = self.step
        self.step_ctr[1] = self.async_steps.get(self.rank, 0)
        self.tables.step_ctr = self.step_ctr.data_ptr() if on else None
        self.tables.lr_ptr = self.lr_dev.data_ptr() if on else None
        self.device_counters = on

    def set_lr(self, lr: float) -> None:
        self.cfg = self.cfg.replace(lr=lr)
        self.lr_dev.fill_(lr)

    # ------------------------------------------------------------------ state
    def load_master(self, flat_global: torch.Tensor) -> None:
        """Scatter a full fp32 arena into the local shard(s) and publish initial weights."""
        shards = self.layout.to_shards(flat_global.to(self.device, torch.float32))
        for r in self.local:
            self.master[r].copy_(shards[r])
            self.mom[r].zero_()
        # initial weights for the local worker replicas (the "pull before first step")
        for r in self.local:
            self.wbf16[r].copy_(flat_global.to(self.device).to(torch.bfloat16))
            self.small[r].copy_(flat_global[: self.n_small].to(self.device, torch.float32))
        if self.is_async:
            for r in self.local:
                L.check(self.lib.dps_publish(self._tp, r, L.stream_ptr()), "publish")
        torch.cuda.synchronize(self.device)

    def gather_master(self, group=None) -> torch.Tensor:
        """Full fp32 arena assembled from all shards (checkpoint path; not on the hot path)."""
        if self.emulate or self.world == 1:
            shards = torch.stack([self.master[r] for r in range(self.world)])
        else:
            import torch.distributed as dist
            parts = [torch.empty_like(self.master[self.rank]) for _ in range(self.world)]
            dist.all_gather(parts, self.master[self.rank], group=group)
            shards = torch.stack(parts)
        return self.layout.from_shards(shards)

    def gather_momentum(self, group=None) -> torch.Tensor:
        if self.emulate or self.world == 1:
            shards = torch.stack([self.mom[r] for r in range(self.world)])
        else:
            import torch.distributed as dist
            parts = [torch.empty_like(self.mom[self.rank]) for _ in range(self.world)]
            dist.all_gather(parts, self.mom[self.rank], group=group)
            shards = torch.stack(parts)
        return self.layout.from_shards(shards)

    # ------------------------------------------------------------------ sync
    def push_sync(self, lr: Optional[float] = None, stream=None) -> None:
        cfg = self.cfg
        L.check(self.lib.dps_push_sync(self._tp, self.step, lr if lr is not None else cfg.lr, cfg.momentum,
                                       cfg.weight_decay, int(self.compress), self.ctas, self.timeout_ns,
                                       L.stream_ptr(stream)), "push_sync")
        self.launches += 1

    # ---- segmented push (overlaps the backward pass) --------------------------------------
    SEND, REDUCE = 1, 2

    def push_segment(self, seg: int, phases: int, arrivals_per_step: Optional[int] = None, stream=None,
                     ctas: Optional[int] = None) -> None:
        """Push kernel on layer group `seg` of every shard (set_segments() defines the groups; seg < 0 = the whole
        shard).  phases: SEND (never waits; launched as soon as the group's weight gradients are complete), REDUCE
        (waits for `arrivals_per_step` signals per live source, then mean + SGD), or both."""
        cfg = self.cfg
        lo, hi = (0, self.E) if seg < 0 else (int(self._seg[seg]), int(self._seg[seg + 1]))
        n = self.ctas if ctas is None else ctas
        if lo == hi and not (phases & self.REDUCE):
            n = 1                                     # empty group on this world size: still signal the arrival
        L.check(self.lib.dps_push_segment(self._tp, self.step, cfg.lr, cfg.momentum, cfg.weight_decay,
                                          int(self.compress), n, self.timeout_ns, lo, hi, phases,
                                          arrivals_per_step if arrivals_per_step is not None else n,
                                          L.stream_ptr(stream)), "push_segment")
        self.launches += 1

    def send_ctas(self, seg: int) -> int:
        """CTAs of the SEND launch of a layer group: sized by the group, small enough to share the SMs with the
        backward kernels it runs beside."""
        n = int(self._seg[seg + 1]) - int(self._seg[seg])
        return int(max(1, min(148, n // (256 * 8 * 4))))

    # ---- remote sources (gRPC front door) ---------------------------------------------------
    def peer_flags(self, r: int) -> torch.Tensor:
        """Flag block of ANY rank as seen from this device (peer mapping): host-side membership commands."""
        return self.fabric.tensor(r, "flags", torch.int32)

    def peer_master(self, r: int) -> torch.Tensor:
        assert self.n_extra, "master shards are peer-visible only when remote_slots > 0"
        return self.fabric.tensor(r, "master", torch.float32)

    def snapshot_master(self) -> torch.Tensor:
        """Full fp32 arena read from every owner's shard over the peer mappings (no collective: callable from the
        thread that serves an RPC).  Consistent when the caller is in lock-step with the round (a remote sync worker
        between its push and its next push); otherwise shards may differ by one round."""
        shards = torch.stack([self.peer_master(r).clone() for r in range(self.world)])
        return self.layout.from_shards(shards)

    def attach_source(self, x: int) -> None:
        """Extra source x (rank W+x) may contribute from now on: every owner's per-round gate (remote_gate_kernel)
        folds its gradient into the round that is current once the gradient has arrived completely.  Runs on the
        caller's stream and waits only for that stream (the training stream may be busy)."""
        bit = 1 << (self.world + x)
        cur = torch.cuda.current_stream()
        for r in range(self.world):
            f = self.peer_flags(r)
            f[L.F_ARRIVE + self.world + x] = 0
            f[L.F_CONSUMED + x] = 0
        cur.synchronize()
        for r in range(self.world):
            self._flag_update(r, L.F_ALIVE, 0xFFFFFFFF, bit)
        cur.synchronize()

    def remote_progress(self, x: int):
        """Per owner: (pushes of extra source x consumed so far, round that consumed the latest one, shard version)."""
        out = []
        for r in range(self.world):
            f = self.peer_flags(r)
            out.append((int(f[L.F_CONSUMED + x]), int(f[L.F_CONSUMED_ROUND + x]), int(f[L.F_VERSION])))
        return out

    def detach_source(self, x: int) -> None:
        """Stop waiting for extra source x.  An owner already polling for it notices the cleared bit."""
        bit = 1 << (self.world + x)
        for r in range(self.world):
            self._flag_update(r, L.F_ALIVE, 0xFFFFFFFF & ~bit, 0)
        torch.cuda.current_stream().synchronize()

    def _flag_update(self, r: int, word: int, and_mask: int, or_mask: int) -> None:
        L.check(self.lib.dps_flag_update(C.c_void_p(self.fabric.ptr(r, "flags") + 4 * word), and_mask, or_mask,
                                         L.stream_ptr()), "flag_update")

    def push_remote(self, x: int, grad_flat: torch.Tensor, arrivals: int, stream=None) -> None:
        """SEND half of the push for extra source x from this device: `grad_flat` (fp32 arena order, device tensor of
        layout.total elements; zeroed by the kernel) goes to slot W+x of every owner, `arrivals` signals per owner."""
        t = L.DpsTables.from_buffer_copy(self.tables)
        t.rank_lo, t.rank_hi = self.world + x, self.world + x + 1
        t.grad[self.world + x] = grad_flat.data_ptr()
        t.step_ctr = None                               # the heartbeat value is informational for a remote source
        cfg = self.cfg
        L.check(self.lib.dps_push_segment(C.byref(t), 0, cfg.lr, cfg.momentum, cfg.weight_decay, int(self.compress),
                                          arrivals, self.timeout_ns, 0, self.E, self.SEND, arrivals,
                                          L.stream_ptr(stream)), "push_remote")

    def pull(self, stream=None, wait: bool = True) -> None:
        L.check(self.lib.dps_pull_bcast(self._tp, self.step, self.ctas, L.stream_ptr(stream)), "pull_bcast")
        self.launches += 1
        if wait:
            for r in self.local:
                L.check(self.lib.dps_wait_pull(self._tp, r, self.step, self.ctas, self.timeout_ns,
                                               int(self.device_counters), L.stream_ptr(stream)), "wait_pull")
                self.launches += 1

    # ---- segmented pull (overlaps the next forward pass) ----------------------------------
    def set_segments(self, global_boundaries) -> None:
        """global_boundaries: increasing arena element offsets [0, ..., total]; each is rounded UP to a multiple of
        W*block (so a layer group is fully inside its segment) and converted to local shard offsets."""
        q = self.world * self.layout.block_elems
        loc = []
        for g in global_boundaries:
            g = min(self.total, (int(g) + q - 1) // q * q)
            loc.append(g // self.world)
        loc[0], loc[-1] = 0, self.E
        for i in range(1, len(loc)):
            loc[i] = max(loc[i], loc[i - 1])
        self.nseg = len(loc) - 1
        self._seg = (C.c_longlong * len(loc))(*loc)

    def pull_segments(self, stream=None) -> None:
        L.check(self.lib.dps_pull_segments(self._tp, self._seg, self.nseg, self.ctas, L.stream_ptr(stream)),
                "pull_segments")
        self.launches += 1

    def wait_segment(self, seg: int, stream=None, rank: Optional[int] = None) -> None:
        r = self.rank if rank is None else rank
        L.check(self.lib.dps_wait_segment(self._tp, r, seg, self.step, self.ctas, self.timeout_ns,
                                          L.stream_ptr(stream)), "wait_segment")
        self.launches += 1

    def bump_step(self, stream=None) -> None:
        L.check(self.lib.dps_bump_step(self._tp, L.stream_ptr(stream)), "bump_step")
        self.launches += 1

    def refresh_weights(self, group=None) -> None:
        """Bring the local bf16 replica up to date with the master outside the step (evaluation / shutdown)."""
        flat = self.gather_master(group)
        for r in self.local:
            self.wbf16[r].copy_(flat.to(torch.bfloat16))
            self.small[r].copy_(flat[: self.n_small])

    def sync_round(self, lr: Optional[float] = None, stream=None) -> None:
        """push -> (barrier, reduce, SGD) -> pull -> wait; advances the round counter."""
        self.push_sync(lr, stream)
        self.pull(stream)
        self.step += 1

    # ------------------------------------------------------------------ async
    def pull_async(self, rank: Optional[int] = None, stream=None, bump_step: bool = False) -> None:
        r = self.rank if rank is None else rank
        L.check(self.lib.dps_pull_async(self._tp, r, self.pull_async_ctas, int(bump_step), L.stream_ptr(stream)),
                "pull_async")
        self.launches += 1

    def push_async(self, rank: Optional[int] = None, stream=None) -> None:
        r = self.rank if rank is None else rank
        L.check(self.lib.dps_push_async(self._tp, r, self.async_steps[r], self.async_ctas, self.timeout_ns,
                                        L.stream_ptr(stream)), "push_async")
        self.async_steps[r] += 1
        self.launches += 1

    POLICY = {"reference": 0, "world": 1, "batch": 2}

    def serve_async(self, owner: Optional[int] = None, lr: Optional[float] = None, stream=None,
                    max_batch: int = L.MAX_PLAN, bump_step: bool = True, ctas: Optional[int] = None) -> None:
        """Owner role: drain the arrival ring (non-blocking) and apply in ticket order."""
        o = self.rank if owner is None else owner
        cfg = self.cfg
        with self._serve_lock:
            for i in range(max(1, (self.world + L.MAX_PLAN - 1) // L.MAX_PLAN)):     # W arrivals per step, 16 per pass
                L.check(self.lib.dps_async_apply(self._tp, o, self.async_ctas, cfg.staleness_bound, cfg.staleness_alpha,
                                                 cfg.staleness_floor, max_batch, lr if lr is not None else cfg.lr,
                                                 cfg.momentum, cfg.weight_decay, self.ctas if ctas is None else ctas,
                                                 self.POLICY[cfg.staleness_policy], int(bump_step and i == 0),
                                                 L.stream_ptr(stream)), "async_apply")
                self.launches += 2

    # ---- dedicated owner stream (apply on arrival) -------------------------------------------
    def start_async_server(self) -> None:
        """The reference's server thread applies a gradient the moment it arrives, whatever the other workers are doing
        (server.py:290-304).  Here a host thread keeps the owner's scan + apply pair queued on a HIGH-PRIORITY stream
        of its own, so arrivals are applied while this rank's own forward / backward is still running instead of when
        its training stream gets there; a slow or stalled worker no longer freezes its shard for everybody else.  The
        apply uses a bounded number of CTAs so it shares the SMs with the compute kernels."""
        import threading
        if self._server is not None:
            return
        lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
        self._server_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self._server_stop = threading.Event()
        period = max(20, int(self.cfg.async_server_period_us)) * 1e-6
        ctas = max(8, min(self.ctas, 64))

        def loop():
            torch.cuda.set_device(self.device)
            ev = torch.cuda.Event()
            while not self._server_stop.is_set():
                self.serve_async(stream=self._server_stream, bump_step=False, ctas=ctas)
                ev.record(self._server_stream)
                ev.synchronize()                       # at most one scan + apply pair in flight
                self.server_passes += 1
                self._server_stop.wait(period)

        self._server = threading.Thread(target=loop, daemon=True, name="dps-async-server")
        self._server.start()

    def stop_async_server(self) -> None:
        if self._server is not None:
            self._server_stop.set()
            self._server.join(5.0)
            self._server = None

    # ------------------------------------------------------------------ introspection
    def flag_words(self, rank: Optional[int] = None) -> torch.Tensor:
        return self.flags[self.rank if rank is None else rank].cpu()

    def error_bits(self, rank: Optional[int] = None) -> int:
        return int(self.flag_words(rank)[L.F_ERROR]) & 0xFFFFFFFF

    def version(self, rank: Optional[int] = None) -> int:
        return int(self.flag_words(rank)[L.F_VERSION])

    def async_stats(self, rank: Optional[int] = None) -> dict:
        f = self.flag_words(rank).tolist()
        acc, rej = f[L.F_STAT_ACCEPTED], f[L.F_STAT_REJECTED]
        n = max(1, acc + rej)
        return {
            "async_updates": acc, "rejected_updates": rej,
            "running_mean_staleness": f[L.F_STAT_SUM] / n, "running_max_staleness": f[L.F_STAT_MAX],
            "staleness_histogram": {str(i): f[L.F_STAT_HIST + i] for i in range(L.HIST_BINS) if f[L.F_STAT_HIST + i]},
            "last_staleness_per_worker": f[L.F_STAT_LAST:L.F_STAT_LAST + self.world],
            "version": f[L.F_VERSION], "torn_pulls": f[L.F_TORN_PULLS], "dropped_pushes": f[L.F_DROPPED],
            "heartbeat": f[L.F_HEARTBEAT:L.F_HEARTBEAT + self.world],
        }

    def bytes_per_push(self) -> int:
        """Bytes this rank sends over NVLink per push (its slices of the other W-1 shards)."""
        return (self.world - 1) * self.E * (2 if self.compress else 4)

    def bytes_per_pull(self) -> int:
        return (self.world - 1) * self.E * 2

    def close(self) -> None:
        self.stop_async_server()
        if self.mc is not None:
            self.mc.close()
        self.fabric.close()
