"""Probe: what does cuBLAS (library baseline) reach on the implicit-GEMM shapes of the conv layers (as plain GEMMs)?"""
import torch
shapes = [("layer1", 131072, 64, 576), ("layer2", 32768, 128, 1152), ("layer3", 8192, 256, 2304), ("layer4", 2048, 512, 4608),
          ("wgrad4", 512, 4608, 2048), ("wgrad3", 256, 2304, 8192), ("wgrad1", 64, 576, 131072)]
for name, M, N, K in shapes:
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
    for _ in range(5):
        a @ b
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            c = a @ b
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    print(f"{name}: M{M} N{N} K{K}  {us:7.2f} us  {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s")
