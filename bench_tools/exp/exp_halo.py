"""Hardware probe: shifted-window A operand from a shared-memory halo tile (see bench_tools/exp/exp_halo.cu;
built separately into bench_tools/bin/libexp_halo.so, not part of the product library)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from dps_b200.ops import _lib as L
L.lib()
lib = C.CDLL(os.path.join(ROOT, "bench_tools", "bin", "libexp_halo.so"))
lib.dps_exp_halo_tile.restype = C.c_int
lib.dps_exp_halo_tile.argtypes = [C.c_void_p] * 3 + [C.c_int] * 7 + [C.c_void_p]
N, H, W = 2, 32, 32
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(N, H, W, 64, device="cuda", generator=g).to(torch.bfloat16)
w = (torch.randn(64, 3, 3, 64, device="cuda", generator=g) / 24).to(torch.bfloat16)
ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
for variant in (0, 1):
    for (n, h0, w0) in [(0, 0, 0), (1, 16, 24), (0, 16, 8)]:
        out = torch.zeros(128, 64, device="cuda")
        rc = lib.dps_exp_halo_tile(x.data_ptr(), w.data_ptr(), out.data_ptr(), N, H, W, n, h0, w0, variant, L.stream_ptr())
        torch.cuda.synchronize()
        want = ref[n, h0:h0 + 16, w0:w0 + 8].reshape(128, 64)
        err = float((out - want).abs().max() / want.abs().max())
        rows_ok = ((out - want).abs().amax(1) < 2e-2 * want.abs().max()).sum().item()
        print(f"variant {variant} tile ({n},{h0},{w0}): rc={rc} relerr={err:.4f} rows_ok={rows_ok}/128")
