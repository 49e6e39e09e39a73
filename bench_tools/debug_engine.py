"""Per-tensor gradient error of the native engine vs torch fp32, with torch bf16-autocast as calibration."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_engine_gpu import _make

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16
eng, ref, layout, grad = _make(N)
g = torch.Generator().manual_seed(3)
x = torch.randn(N, 3, 32, 32, generator=g)
y = torch.randint(0, 100, (N,), generator=g)
eng.load_batch(x, y)
eng.forward_backward()
torch.cuda.synchronize()
ref.train()
xb = x.cuda().to(torch.bfloat16).float()
loss = torch.nn.functional.cross_entropy(ref(xb), y.cuda())
loss.backward()
g32 = {n: p.grad.clone() for n, p in ref.named_parameters()}
ref.zero_grad()
with torch.autocast("cuda", dtype=torch.bfloat16):
    l2 = torch.nn.functional.cross_entropy(ref(xb).float(), y.cuda())
l2.backward()
g16 = {n: p.grad.clone() for n, p in ref.named_parameters()}
print("loss engine %.4f torch32 %.4f torch-bf16 %.4f" % (float(eng.out[0]), float(loss), float(l2)))
got = layout.unflatten(grad)
for n in g32:
    a, b, c = got[n].float().flatten(), g32[n].flatten(), g16[n].flatten()
    cos = float(torch.nn.functional.cosine_similarity(a, b, dim=0))
    rel = float((a - b).norm() / (b.norm() + 1e-12))
    cos2 = float(torch.nn.functional.cosine_similarity(c, b, dim=0))
    rel2 = float((c - b).norm() / (b.norm() + 1e-12))
    print("%-32s engine cos %.4f rel %.4f | autocast cos %.4f rel %.4f | norm %.3e" % (n, cos, rel, cos2, rel2, float(b.norm())))
