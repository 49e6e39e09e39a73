"""Multi-GPU self-check run under torchrun by tests/test_multi_gpu.py: a few steps through the public API,
then cross-rank consistency checks (replicas identical, versions, sync == mean-gradient step)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from dps_b200 import Config
from dps_b200.api import Job

mode, out_path = sys.argv[1], sys.argv[2]
B, steps = 16, (int(sys.argv[3]) if len(sys.argv) > 3 else 3)
cfg = Config(mode=mode, batch_size=B, lr=0.05, barrier_timeout_s=20.0, use_cuda_graph=True,
             nvls=os.environ.get("DPS_NVLS", "off"),        # 'auto' / 'on': pull broadcast through the NVLS window
             nvls_reduce=os.environ.get("DPS_NVLS_REDUCE", "0") == "1")     # push summed in the switch
job = Job(cfg)
rank, world, dev = job.rank, job.world, job.device
g = torch.Generator().manual_seed(100 + rank)
nb = min(steps, 8)
xs = [torch.randn(B, 3, 32, 32, generator=g).pin_memory() for _ in range(nb)]
ys = [torch.randint(0, 100, (B,), generator=g).pin_memory() for _ in range(nb)]

extra = {}
if mode == "sync":
    # oracle for the FIRST step: mean over ranks of the fp16-rounded local gradients, applied to the initial master
    w0 = job.ps.gather_master().clone()
    job.worker.engine.load_batch(xs[0], ys[0])
    eng = job.worker.engine
    hook, eng.seg_wait = eng.seg_wait, None                 # oracle pass: compute only, no pull waits ...
    push_hook, eng.on_group_done = eng.on_group_done, None  # ... and no per-group SEND of the overlapped push
    eng.forward_backward()
    eng.seg_wait, eng.on_group_done = hook, push_hook
    torch.cuda.synchronize()
    g_local = job.ps.grad[rank].clone().half().float()
    job.ps.grad[rank].zero_()
    saved = {k: v.clone() for k, v in job.worker.engine.bn_buffers().items()}
    dist.all_reduce(g_local)
    want = w0 - cfg.lr * (g_local / world)
    for k, v in job.worker.engine.bn_buffers().items():   # undo the oracle pass's running-stat update
        v.copy_(job.model_cpu.state_dict()[k].to(v.device))
losses = []
vers = []
for i in range(steps):
    losses.append(job.worker.train_step(xs[i % nb], ys[i % nb]).item())
    if mode == "async" and i % 5 == 0:
        vers.append(job.ps.version())          # shard version counters only move forward
    if mode == "sync" and i == 0:
        got = job.ps.gather_master()
        extra["max_abs_diff_vs_single_process_mean_step"] = float((got - want).abs().max())
job.worker.drain()
job.barrier()
if job.worker.overlap_pull:
    # the last update is normally pulled at the start of the next step: run that segmented pull now
    job.ps.pull_segments()
    for k in range(job.ps.nseg):
        job.ps.wait_segment(k)
    torch.cuda.synchronize()
    job.barrier()
master = job.ps.gather_master()
mine = job.ps.wbf16[rank].clone()
if mode == "async":
    job.ps.pull_async()
    torch.cuda.synchronize()
    mine = job.ps.wbf16[rank].clone()
ref0 = mine.clone()
dist.broadcast(ref0, 0)
ns = job.ps.n_small                                        # the fp32 prefix (BN affine, fc bias) travels with the weights
small_ok = mode != "sync" or torch.equal(job.ps.small[rank][:ns], master[:ns])
same = torch.tensor([int(torch.equal(ref0, mine) and torch.equal(mine, master.to(torch.bfloat16)) and small_ok)],
                    device=dev)
dist.all_reduce(same, op=dist.ReduceOp.MIN)
st = job.ps.async_stats() if mode == "async" else {"async_updates": 0, "rejected_updates": 0}
hist = st.get("staleness_histogram", {})
mono = int(all(b >= a for a, b in zip(vers, vers[1:])))
info = torch.tensor([job.ps.error_bits(), job.ps.version(), st["async_updates"], st["rejected_updates"],
                     st.get("dropped_pushes", 0), st.get("torn_pulls", 0), sum(hist.values()),
                     st.get("running_max_staleness", 0), mono], device=dev, dtype=torch.int64)
allinfo = [torch.zeros_like(info) for _ in range(world)]
dist.all_gather(allinfo, info)
if rank == 0:
    res = {"mode": mode, "steps": steps, "losses": losses, "replicas_equal": bool(same.item()),
           "finite": bool(torch.isfinite(master).all()), "error_bits": [int(t[0]) for t in allinfo],
           "versions": [int(t[1]) for t in allinfo], "accepted": [int(t[2]) for t in allinfo],
           "rejected": [int(t[3]) for t in allinfo], "dropped": [int(t[4]) for t in allinfo],
           "torn_pulls": [int(t[5]) for t in allinfo], "hist_total": [int(t[6]) for t in allinfo],
           "max_staleness": [int(t[7]) for t in allinfo], "versions_monotone": [int(t[8]) for t in allinfo],
           "world": world, "staleness_bound": cfg.staleness_bound, "nvls_pull": bool(job.ps.nvls),
           "nvls_reduce": bool(job.ps.nvls_reduce), **extra}
    with open(out_path, "w") as f:
        json.dump(res, f)
    print(json.dumps(res))
job.close()
dist.destroy_process_group()
