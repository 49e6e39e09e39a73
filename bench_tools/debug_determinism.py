"""Run-to-run variation of the native engine (fp32 atomics in BN statistics / wgrad reorder sums)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from test_engine_gpu import _make

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16
eng, ref, layout, grad = _make(N, seed=1)
g = torch.Generator().manual_seed(5)
x = torch.randn(N, 3, 32, 32, generator=g); y = torch.randint(0, 100, (N,), generator=g)
def run():
    grad.zero_(); eng.load_batch(x, y); eng.forward_backward(); torch.cuda.synchronize()
    return grad.clone(), eng.feat.float().clone(), eng.units[0]["c"].float().clone(), eng.units[0]["a"].float().clone(), eng.d_feat.float().clone(), float(eng.out[0])
def rel(a, b): return float((a - b).norm() / (b.norm() + 1e-20))
r = [run() for _ in range(3)]
for i in (1, 2):
    print("eager run %d vs 0: grad %.3e feat %.3e stem_c %.3e stem_a %.3e d_feat %.3e loss %.6f/%.6f" % (
        i, rel(r[i][0], r[0][0]), rel(r[i][1], r[0][1]), rel(r[i][2], r[0][2]), rel(r[i][3], r[0][3]), rel(r[i][4], r[0][4]), r[i][5], r[0][5]))
got0, got1 = layout.unflatten(r[0][0]), layout.unflatten(r[1][0])
for n in list(got0)[:6] + list(got0)[-6:]:
    print("  %-28s %.3e" % (n, rel(got1[n], got0[n])))
