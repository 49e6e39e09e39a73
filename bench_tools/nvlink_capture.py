#!/usr/bin/env python
"""Push SEND and segmented Pull of ONE rank of a 2-GPU job driven from a single process, so that ncu can profile the
data-path kernels WITH NVLink traffic (ncu must not wrap a multi-rank command, and kernels that wait on a peer must
not be replayed: the SEND half of the push and the pull broadcast never wait).

Rank 0 lives on GPU 0, rank 1 on GPU 1 (peer access enabled, no IPC needed inside one process).  Rank 0
  * pushes: fp32 gradients -> fp16 tiles into its staging slot on BOTH owners (its own shard locally, rank 1's shard
    over NVLink) + arrival signals                                   (dps_push_segment, phases = SEND)
  * pulls : its fp32 master shard -> bf16 into BOTH workers' weight arenas (GPU 1's over NVLink) + delivery counters
                                                                     (dps_pull_segments)
Payload = the real model (11.2 M parameters: 22.4 MB fp16 per push, half of it crosses the link at W = 2).

    python bench_tools/nvlink_capture.py                # timing only (CUDA events)
    ncu --set full --metrics nvltx__bytes.sum,nvlrx__bytes.sum -k regex:push_sync_kernel\\|pull_seg_kernel ... python bench_tools/nvlink_capture.py
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dps_b200.models.resnet import build_model
from dps_b200.ops import _lib as L
from dps_b200.store.shardmap import build_layout

assert torch.cuda.device_count() >= 2, "needs two GPUs"
W = 2
lib = L.lib()
layout = build_layout(build_model(100, seed=0), W)
E, total, n_small = layout.shard_elems, layout.total, layout.n_small
for a, b in ((0, 1), (1, 0)):
    L.check(lib.dps_enable_peer(a, b), "enable_peer")

buf = {}
for d in range(W):
    dev = torch.device("cuda", d)
    buf[d] = {"flags": torch.zeros(L.FLAG_WORDS, dtype=torch.int32, device=dev),
              "stage": torch.zeros(W * E, dtype=torch.float16, device=dev),
              "wbf16": torch.zeros(total, dtype=torch.bfloat16, device=dev),
              "small": torch.zeros(max(n_small, 8), dtype=torch.float32, device=dev)}
    buf[d]["flags"][L.F_ALIVE] = (1 << W) - 1
torch.cuda.set_device(0)
grad = torch.randn(total, device="cuda:0") * 0.01
master = torch.randn(E, device="cuda:0") * 0.05
mom = torch.zeros(E, device="cuda:0")

t = L.DpsTables()
t.world, t.block_elems, t.ring_slots, t.n_extra = W, layout.block_elems, 1, 0
t.shard_elems, t.n_small = E, n_small
t.rank_lo, t.rank_hi = 0, 1                      # this launch drives rank 0 only
for r in range(W):
    t.stage[r] = buf[r]["stage"].data_ptr()
    t.wbf16[r] = buf[r]["wbf16"].data_ptr()
    t.small[r] = buf[r]["small"].data_ptr()
    t.flags[r] = buf[r]["flags"].data_ptr()
t.grad[0], t.master[0], t.mom[0] = grad.data_ptr(), master.data_ptr(), mom.data_ptr()
tp = C.byref(t)
seg = (C.c_longlong * 2)(0, E)
CTAS = 592
st = L.stream_ptr()


def push():
    L.check(lib.dps_push_segment(tp, 0, 0.1, 0.0, 0.0, 1, CTAS, 0, 0, E, 1, CTAS, st), "push SEND")


def pull():
    L.check(lib.dps_pull_segments(tp, seg, 1, CTAS, st), "pull")


res = {}
for name, fn, wire in (("push_send", push, E * 2), ("pull_seg", pull, E * 2)):
    for _ in range(3):
        if name == "push_send":
            grad.normal_()
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    times = []
    for _ in range(reps):
        if name == "push_send":
            grad.normal_()                       # the kernel zeroes the gradients it has sent
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3)
    us = sorted(times)[len(times) // 2]
    res[name] = {"us_median": round(us, 2), "bytes_over_nvlink": wire, "link_GBps": round(wire / us / 1e3, 1),
                 "frac_of_770GBps_measured_peer_copy": round(wire / us / 1e3 / 770, 3)}
# the peer really received the data
sent = buf[1]["stage"][:E].float().abs().sum().item()
got = buf[1]["wbf16"].float().abs().sum().item()
res["peer_stage_nonzero"], res["peer_weights_nonzero"] = sent > 0, got > 0
res["arrivals_on_peer"] = int(buf[1]["flags"][L.F_ARRIVE])
print(json.dumps(res))
