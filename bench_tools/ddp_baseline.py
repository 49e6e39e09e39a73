#!/usr/bin/env python
"""Baseline (b) of BASELINE.json: the same ResNet-18 / batch / synthetic data trained with torch DDP +
NCCL all-reduce + cuDNN/cuBLAS (bf16 autocast, channels_last, SGD), device-timed, max over ranks.

    python bench_tools/ddp_baseline.py --steps 30 --warmup 5 [--fp32] [--graph]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench_tools/ddp_baseline.py
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
p.add_argument("--graph", action="store_true", help="capture fwd+bwd+step in a CUDA graph (world 1 only)")
a = p.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
model = build_model(100, seed=0).to(dev).to(memory_format=torch.channels_last)
opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.0)
net = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
g = torch.Generator(device=dev).manual_seed(rank)
x = torch.randn(a.batch, 3, 32, 32, device=dev, generator=g).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 100, (a.batch,), device=dev, generator=g)


def step():
    opt.zero_grad(set_to_none=False)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=not a.fp32):
        loss = nn.functional.cross_entropy(net(x).float(), y)
    loss.backward()
    opt.step()
    return loss


graph = None
for _ in range(max(3, a.warmup)):
    step()
if a.graph and world == 1:
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        step()
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
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
                      "cuda_graph": graph is not None, "n_gpus": world, "steps": a.steps,
                      "ms_per_step": round(float(ms) / a.steps, 4),
                      "value": round(a.steps * a.batch * world / (float(ms) / 1e3), 1), "unit": "images/sec"}))
if world > 1:
    dist.destroy_process_group()
