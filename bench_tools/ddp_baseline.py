#!/usr/bin/env python
"""Baseline (b) of BASELINE.json: the same ResNet-18 / batch / synthetic data trained with torch DDP +
NCCL all-reduce + cuDNN/cuBLAS (bf16 autocast, channels_last, SGD), device-timed, max over ranks.

    python bench_tools/ddp_baseline.py --steps 30 --warmup 5 [--fp32] [--graph]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench_tools/ddp_baseline.py --graph

--graph captures the WHOLE step (forward, backward, gradient all-reduce, SGD) in one CUDA graph at any world size, so
the baseline is not bound by host launch overhead (an eager 8-rank run on one host is: 7.8 ms/step against 4.6 at one
rank).  With world > 1 the captured step is cuDNN forward/backward + ONE flat NCCL all-reduce of all gradients + SGD
("flat-allreduce": the data-parallel math of DDP without its Python-side bucket hooks; `--ddp` tries to capture the real
DistributedDataParallel wrapper instead, which did not return from capture at 2 ranks on this stack).  Without --graph
the real DDP wrapper runs eagerly.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch import nn

from dps_b200.models.resnet import build_model

p = argparse.ArgumentParser()
p.add_argument("--steps", type=int, default=30)
p.add_argument("--warmup", type=int, default=5)
p.add_argument("--batch", type=int, default=128)
p.add_argument("--fp32", action="store_true")
p.add_argument("--graph", action="store_true", help="capture fwd+bwd+all-reduce+step in one CUDA graph")
p.add_argument("--ddp", action="store_true", help="with --graph: try to capture real DistributedDataParallel (its whole-step "
               "capture hung on this stack at 2 ranks; default = the same math with one flat NCCL all-reduce in the graph)")
a = p.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
g = torch.Generator(device=dev).manual_seed(rank)
x = torch.randn(a.batch, 3, 32, 32, device=dev, generator=g).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 100, (a.batch,), device=dev, generator=g)


def build(use_ddp: bool):
    model = build_model(100, seed=0).to(dev).to(memory_format=torch.channels_last)
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.0)
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if use_ddp else model
    flat = None
    if world > 1 and not use_ddp:
        # gradients as views of one flat buffer: a single NCCL all-reduce moves all of them
        n = sum(p.numel() for p in model.parameters())
        flat = torch.zeros(n, device=dev)
        off = 0
        for p in model.parameters():
            p.grad = flat[off: off + p.numel()].view_as(p)
            off += p.numel()

    def step():
        opt.zero_grad(set_to_none=False)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=not a.fp32):
            loss = nn.functional.cross_entropy(net(x).float(), y)
        loss.backward()
        if flat is not None:
            dist.all_reduce(flat)
            flat.div_(world)
        opt.step()
        return loss
    return step


def capture(step):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(11):                      # DDP needs >= 11 side-stream iterations before a whole-step capture
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        step()
    return gr


graph, how = None, "eager"
use_ddp = world > 1 and a.ddp
if a.graph:
    side = torch.cuda.Stream()
    try:
        with torch.cuda.stream(side):            # DDP itself must be constructed on a side stream for graph capture
            step = build(use_ddp)
        torch.cuda.current_stream().wait_stream(side)
        graph = capture(step)
        how = ("DDP" if use_ddp else ("flat-allreduce" if world > 1 else "single")) + " in one CUDA graph"
    except Exception as e:                       # noqa: BLE001
        if rank == 0:
            print(f"# DDP capture refused ({type(e).__name__}: {str(e)[:120]}); falling back to flat all-reduce", flush=True)
        torch.cuda.synchronize()
        step = build(False)
        graph = capture(step)
        how = "flat-allreduce in one CUDA graph (DDP capture refused)"
else:
    use_ddp = world > 1
    step = build(use_ddp)
    how = ("DDP" if use_ddp else "single") + " eager"
    for _ in range(max(3, a.warmup)):
        step()
for _ in range(3):
    graph.replay() if graph is not None else step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    graph.replay() if graph is not None else step()
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"impl": "torch DDP + NCCL + cuDNN", "dtype": "fp32" if a.fp32 else "bf16 autocast",
                      "cuda_graph": graph is not None, "how": how, "n_gpus": world, "steps": a.steps,
                      "ms_per_step": round(float(ms) / a.steps, 4),
                      "value": round(a.steps * a.batch * world / (float(ms) / 1e3), 1), "unit": "images/sec"}))
sys.stdout.flush()
if world > 1:
    # tearing the process group down after NCCL kernels were captured into a graph did not return on this stack
    # (the launcher's timeout had to kill it): the numbers are out, leave without the collective clean-up
    dist.barrier()
    os._exit(0)
