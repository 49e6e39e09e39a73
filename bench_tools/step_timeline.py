"""In-situ timeline of one training step: CUDA events around every launch of the engine schedule (eager launches,
warm caches, real clocks).  Complements the ncu launch lists, which are serialised and cold-cache."""
import collections, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dps_b200 import Config
from dps_b200.api import Job
from dps_b200.ops.conv import ConvPlan

cfg = Config(mode="sync", workers=1, batch_size=128, use_cuda_graph=False, overlap_pull=False)
job = Job(cfg, rank=0, world=1, device_index=0)
wk, eng = job.worker, job.worker.engine
x = torch.randn(128, 3, 32, 32).pin_memory(); y = torch.randint(0, 100, (128,)).pin_memory()
for _ in range(5):
    wk.train_step(x, y)
torch.cuda.synchronize()


def label(fn):
    s = getattr(fn, "__self__", None)
    if isinstance(s, ConvPlan):
        kind = {0: "fprop", 1: "dgrad", 2: "wgrad"}[s.kind]
        return f"conv_{kind} grid{s.grid} N{s.block_n}"
    names = set(fn.__code__.co_names) if hasattr(fn, "__code__") else set()
    for key, lab in (("_bn_fwd", "bn_fwd"), ("_bn_bwd_apply", "bn_bwd_apply"), ("_bn_bwd", "bn_bwd_full"),
                     ("seg_wait", "wait_seg"), ("zero_", "memset"), ("dps_stem_im2col", "im2col")):
        if key in names:
            return lab
    return "head"


sched = [(label(f), f) for f in eng.fwd + eng.bwd] + [("push_sync", job.ps.push_sync), ("pull", job.ps.pull)]
R = 20
acc = collections.OrderedDict()
per = [0.0] * len(sched)
for rep in range(R):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(sched) + 1)]
    ev[0].record()
    for i, (lab, f) in enumerate(sched):
        f()
        ev[i + 1].record()
    job.ps.step += 1
    torch.cuda.synchronize()
    for i in range(len(sched)):
        per[i] += ev[i].elapsed_time(ev[i + 1]) * 1e3 / R
tot = sum(per)
print(f"sum of per-launch times (eager, events between launches): {tot:.0f} us over {len(sched)} launches")
cat = collections.OrderedDict()
for (lab, _), t in zip(sched, per):
    k = lab.split(" ")[0]
    c = cat.setdefault(k, [0, 0.0]); c[0] += 1; c[1] += t
for k, (n, t) in sorted(cat.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:14s} n={n:3d} {t:8.1f} us {100 * t / tot:5.1f}%  mean {t / n:6.2f}")
print()
for (lab, _), t in zip(sched, per):
    if lab.startswith("conv"):
        print(f"{t:7.2f} us  {lab}")
job.close()
