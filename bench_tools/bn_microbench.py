"""Stand-alone timing of the BN forward / backward-apply kernels at the ResNet-18 layer shapes (batch 128)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dps_b200.ops import _lib as L
from dps_b200.models import engine as E
E._decl()
lib = L.lib()
p = lambda t: None if t is None else t.data_ptr()
for (P, Cc) in [(131072, 64), (32768, 128), (8192, 256), (2048, 512)]:
    x = torch.randn(P, Cc, device="cuda").to(torch.bfloat16); y = torch.empty_like(x); r = torch.randn_like(x)
    stats = torch.zeros(2 * Cc, device="cuda", dtype=torch.int64); sums = torch.zeros(2 * Cc, device="cuda", dtype=torch.int64)
    g, b = torch.ones(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    m, rs, rm, rv = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda"), torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
    nb = torch.zeros((), dtype=torch.int64, device="cuda"); dg, db = torch.zeros(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    st = C.c_void_p(L.stream_ptr())
    def fwd(res): lib.dps_bn_forward(p(x), p(stats), p(g), p(b), p(res), p(y), p(m), p(rs), p(rm), p(rv), p(nb), P, Cc, 1, 1, 1e-5, 0.1, st)
    def app(): lib.dps_bn_backward_apply(p(x), p(r), p(m), p(rs), p(g), p(sums), p(y), p(dg), p(db), P, Cc, st)
    for name, fn, nbytes in [("fwd", lambda: fwd(None), 2 * x.numel() * 2), ("fwd+res", lambda: fwd(r), 3 * x.numel() * 2), ("bwd_apply", app, 3 * x.numel() * 2)]:
        for _ in range(5): fn()
        gph = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(200): fn()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 200
        print(f"P={P:6d} C={Cc:3d} {name:9s} {us:6.2f} us  {nbytes / us / 1e6:6.2f} TB/s")
