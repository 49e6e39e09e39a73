"""Failure detection for the on-box job.

The reference claims automatic failover but implements none: `last_seen` is written and never read, the
worker health-check thread is dead code, and a dead worker freezes sync training for ever (reference:
src/parameter_server/server.py:200,219,251,272; src/workers/worker.py:112-126).  Here:

  * every push stores the sender's step into a heartbeat word of each owner's flag block (device side);
  * every cross-GPU wait in the kernels is time-bounded: on expiry the owner averages over the ranks that
    did arrive, records the missing ones (DPS_F_MISSING) and raises a sticky error bit -- the stream never hangs;
  * this host-side watchdog turns those words into decisions: which ranks are stale, whether to abort and
    restart from the last checkpoint (`Config.resume`).
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

ERR_NAMES = {1: "barrier_timeout", 2: "pull_timeout", 4: "ring_full"}


def decode_error_bits(bits: int) -> List[str]:
    return [name for bit, name in ERR_NAMES.items() if bits & bit]


def missing_ranks(mask: int, world: int) -> List[int]:
    return [r for r in range(world) if (mask >> r) & 1]


@dataclass
class HeartbeatWatchdog:
    """Tracks per-rank heartbeat counters; a rank is stale when its counter has not moved for `timeout_s`."""
    read_heartbeats: Callable[[], Sequence[int]]
    world: int
    timeout_s: float = 30.0
    clock: Callable[[], float] = time.monotonic
    _last: Dict[int, int] = field(default_factory=dict)
    _moved_at: Dict[int, float] = field(default_factory=dict)
    _thread: Optional[threading.Thread] = None
    _stop: threading.Event = field(default_factory=threading.Event)
    on_stale: Optional[Callable[[List[int]], None]] = None
    always_notify: bool = False           # call on_stale on every poll (the callback also inspects error bits)

    def poll(self) -> List[int]:
        now = self.clock()
        hb = list(self.read_heartbeats())[: self.world]
        stale = []
        for r, v in enumerate(hb):
            if self._last.get(r) != v:
                self._last[r], self._moved_at[r] = v, now
            elif now - self._moved_at.get(r, now) > self.timeout_s:
                stale.append(r)
        return stale

    def start(self, period_s: float = 1.0) -> None:
        def loop():
            while not self._stop.wait(period_s):
                try:
                    stale = self.poll()
                    if (stale or self.always_notify) and self.on_stale is not None:
                        self.on_stale(stale)
                except Exception:                        # the device went away under us: the training loop will notice
                    pass
        self._thread = threading.Thread(target=loop, daemon=True)
        self._thread.start()

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(2.0)


def for_device_ps(ps, timeout_s: float = 30.0, on_stale=None) -> HeartbeatWatchdog:
    """Watchdog over a DevicePS: reads the local flag block (a tiny D2H copy) on every poll."""
    import torch
    from ..ops import _lib as L
    state = {}

    def read():
        # private stream: the copy must not queue behind a training stream that is itself stuck in a bounded wait
        if "s" not in state:
            torch.cuda.set_device(ps.device)
            state["s"] = torch.cuda.Stream(device=ps.device)
        with torch.cuda.stream(state["s"]):
            f = ps.flag_words()
        return [int(x) for x in f[L.F_HEARTBEAT:L.F_HEARTBEAT + ps.world]]
    return HeartbeatWatchdog(read, ps.world, timeout_s, on_stale=on_stale)
