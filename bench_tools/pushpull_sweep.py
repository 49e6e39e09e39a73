#!/usr/bin/env python
"""Push / pull bandwidth sweep, 1 KB - 64 MB per worker, at the current world size (BASELINE config 5).

For each payload size S (bytes of fp16 gradient per worker = bytes of bf16 parameters per worker):
  ours  : dps_push_sync (compress + peer store + arrival + reduce + SGD) and dps_pull_bcast (+wait), device-timed
  nccl  : reduce_scatter(fp16) + all_gather(bf16) of the same payload (what a ZeRO-1 style DDP step moves)
Reported per op: time (us, max over ranks) and algorithm GB/s = S / t; plus the roofline fraction
min-time = bytes that must cross NVLink per GPU ((W-1)/W * S) / link bandwidth (770 GB/s measured, 900 nominal).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench_tools/pushpull_sweep.py --out sweep8.json
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from dps_b200.config import Config
from dps_b200.parallel.ps_device import DevicePS
from dps_b200.store.shardmap import Layout

p = argparse.ArgumentParser()
p.add_argument("--out", default=None)
p.add_argument("--iters", type=int, default=20)
p.add_argument("--no-nccl", action="store_true")
p.add_argument("--nvls", choices=["off", "auto", "on"], default="off", help="pull through the NVLS multicast window")
p.add_argument("--nvls-reduce", action="store_true", help="push summed in the switch (multimem.ld_reduce)")
p.add_argument("--kb", default="1,4,16,64,256,1024,4096,16384,65536", help="payload sizes in KB")
a = p.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
LINK = {"measured": 770e9, "nominal": 900e9}
rows = []


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev, dtype=torch.float64)   # us
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


for kb in [int(v) for v in a.kb.split(",")]:
    nbytes = kb * 1024
    elems = nbytes // 2
    BS = 512 if elems < world * 2048 else 2048
    total = max(world * BS, (elems + world * BS - 1) // (world * BS) * (world * BS))
    layout = Layout(entries={}, order=[], n_small=0, total=total, world=world, block_elems=BS)
    cfg = Config(mode="sync", workers=world, lr=0.01, barrier_timeout_s=20.0, nvls=a.nvls, nvls_reduce=a.nvls_reduce)
    ps = DevicePS(layout, cfg, rank, world, local)
    ps.grad[rank].normal_()

    # push and pull alternate inside a round (arrival / delivery counters advance together)
    def round_():
        ps.push_sync()
        ps.pull()
        ps.step += 1
    t_round = timed(round_, a.iters)
    # split the round: time push alone by running pull untimed between pushes is not possible without
    # breaking the counters, so measure pull = round - push with events inside one loop
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    tp = tq = 0.0
    for _ in range(a.iters):
        ev[0].record(); ps.push_sync(); ev[1].record(); ps.pull(); ev[2].record()
        ps.step += 1
        torch.cuda.synchronize()
        tp += ev[0].elapsed_time(ev[1]) * 1e3
        tq += ev[1].elapsed_time(ev[2]) * 1e3
    t2 = torch.tensor([tp / a.iters, tq / a.iters], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    t_push, t_pull = float(t2[0]), float(t2[1])
    row = {"payload_bytes": total * 2, "world": world, "push_us": round(t_push, 2), "pull_us": round(t_pull, 2),
           "round_us": round(t_round, 2), "push_GBps": round(total * 2 / t_push / 1e3, 2),
           "pull_GBps": round(total * 2 / t_pull / 1e3, 2), "error_bits": ps.error_bits(), "nvls_pull": bool(ps.nvls), "nvls_reduce": bool(ps.nvls_reduce)}
    wire = (world - 1) / world * total * 2
    for k, bw in LINK.items():
        tmin = wire / bw * 1e6
        row[f"push_roofline_frac_{k}"] = round(tmin / t_push, 4) if wire else None
        row[f"pull_roofline_frac_{k}"] = round(tmin / t_pull, 4) if wire else None
    if world > 1 and not a.no_nccl:
        g16 = torch.randn(total, device=dev).half()
        out16 = torch.empty(total // world, device=dev, dtype=torch.float16)
        w16 = torch.empty(total, device=dev, dtype=torch.bfloat16)
        sh16 = torch.empty(total // world, device=dev, dtype=torch.bfloat16)
        row["nccl_reduce_scatter_us"] = round(timed(lambda: dist.reduce_scatter_tensor(out16, g16), a.iters), 2)
        row["nccl_all_gather_us"] = round(timed(lambda: dist.all_gather_into_tensor(w16, sh16), a.iters), 2)
    rows.append(row)
    ps.close()
    if rank == 0:
        print(json.dumps(row), flush=True)
if rank == 0 and a.out:
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump({"world": world, "nvls": a.nvls, "nvls_reduce": a.nvls_reduce, "rows": rows}, f, indent=1)
if world > 1:
    dist.destroy_process_group()
