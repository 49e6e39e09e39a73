"""In-situ timeline of one training step: every launch of the engine schedule timed as a CUDA graph of REPS copies of
that launch (no host launch overhead in the number, warm L2, real clocks, PDL between the copies).  Complements the
ncu launch lists, which are serialised and cold-cache.  Kernels are idempotent up to accumulation (wgrad / statistics
add into buffers that the next real step zeroes), so replaying one launch many times is harmless."""
import collections, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dps_b200 import Config
from dps_b200.api import Job
from dps_b200.ops.conv import ConvPlan

B = int(os.environ.get("DPS_TL_BATCH", 128))
REPS = 20
cfg = Config(mode="sync", workers=1, batch_size=B, use_cuda_graph=False, overlap_pull=False)
job = Job(cfg, rank=0, world=1, device_index=0)
wk, eng = job.worker, job.worker.engine
x = torch.randn(B, 3, 32, 32).pin_memory(); y = torch.randint(0, 100, (B,)).pin_memory()
for _ in range(3):
    wk.train_step(x, y)
torch.cuda.synchronize()


def label(fn):
    s = getattr(fn, "__self__", None)
    if isinstance(s, ConvPlan):
        kind = {0: "fprop", 1: "dgrad", 2: "wgrad"}[s.kind]
        return f"conv_{kind} {s.desc} grid{s.grid} N{s.block_n}"
    names = set(fn.__code__.co_names) if hasattr(fn, "__code__") else set()
    for key, lab in (("_bn_fwd", "bn_fwd"), ("_bn_bwd_apply", "bn_bwd_apply"), ("_bn_bwd", "bn_bwd_full"),
                     ("seg_wait", "wait_seg"), ("zero_", "memset"), ("dps_stem_im2col", "im2col")):
        if key in names:
            return lab
    return "head"


from dps_b200.models.engine import _Marker
sched = [(label(f), f) for f in eng.fwd + eng.bwd if not isinstance(f, _Marker)] + [("push_sync", job.ps.push_sync), ("pull", job.ps.pull)]
per = []
side = torch.cuda.Stream()
for lab, f in sched:
    with torch.cuda.stream(side):
        f(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(REPS):
                f()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side); g.replay(); e1.record(side); torch.cuda.synchronize()
        per.append(e0.elapsed_time(e1) * 1e3 / REPS)
tot = sum(per)
print(f"sum of per-launch times (graph of {REPS} copies each): {tot:.0f} us over {len(sched)} launches, batch {B}")
cat = collections.OrderedDict()
for (lab, _), t in zip(sched, per):
    k = lab.split(" ")[0]
    c = cat.setdefault(k, [0, 0.0]); c[0] += 1; c[1] += t
for k, (n, t) in sorted(cat.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:14s} n={n:3d} {t:8.1f} us {100 * t / tot:5.1f}%  mean {t / n:6.2f}")
print()
for (lab, _), t in zip(sched, per):
    print(f"{t:7.2f} us  {lab}")
job.close()
