"""every conv shape of the model x {fprop, dgrad, wgrad} against a plain PyTorch fp32
reference of the same op (same bf16-rounded inputs)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# (H, Cin, Cout, R, stride)  -- CIFAR ResNet-18 conv shapes, H == W
SHAPES = [
    (32, 64, 64, 3, 1),      # layer1 (x4) and the im2col'ed stem GEMM shape family
    (32, 64, 128, 3, 2),     # layer2.0.conv1
    (16, 128, 128, 3, 1),    # layer2
    (32, 64, 128, 1, 2),     # layer2 shortcut
    (16, 128, 256, 3, 2),
    (8, 256, 256, 3, 1),
    (16, 128, 256, 1, 2),
    (8, 256, 512, 3, 2),
    (4, 512, 512, 3, 1),
    (8, 256, 512, 1, 2),
]
N = 8


def _mk(H, Cin, Cout, R, stride, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(N, H, H, Cin, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(Cout, R, R, Cin, device="cuda", generator=g) / (R * R * Cin) ** 0.5).to(torch.bfloat16)
    Ho = H // stride
    dy = torch.randn(N, Ho, Ho, Cout, device="cuda", generator=g).to(torch.bfloat16)
    return x, w, dy


def _ref(x, w, dy, R, stride):
    xf = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wf = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    y = F.conv2d(xf, wf, stride=stride, padding=R // 2)
    y.backward(dy.float().permute(0, 3, 1, 2))
    return (y.detach().permute(0, 2, 3, 1), xf.grad.permute(0, 2, 3, 1), wf.grad.permute(0, 2, 3, 1))


def _relerr(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6))


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "H%d_%d->%d_k%d_s%d" % s)
def test_fprop(shape):
    from dps_b200.ops.conv import ConvPlan
    H, Cin, Cout, R, stride = shape
    x, w, dy = _mk(*shape)
    y_ref, _, _ = _ref(x, w, dy, R, stride)
    out = torch.zeros(N, H // stride, H // stride, Cout, device="cuda", dtype=torch.bfloat16)
    stats = torch.zeros(2 * Cout, device="cuda", dtype=torch.int64)     # fixed point: sum x2^32, sumsq x2^24
    ConvPlan.fprop(x, w, out, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride, stats=stats).run()
    torch.cuda.synchronize()
    assert _relerr(out, y_ref) < 1.5e-2, _relerr(out, y_ref)
    # fused BN statistics are taken over the stored bf16 values
    of = out.float().reshape(-1, Cout)
    assert torch.allclose((stats[:Cout].double() / 2 ** 32).float(), of.sum(0), rtol=2e-3, atol=2e-1)
    assert torch.allclose((stats[Cout:].double() / 2 ** 24).float(), (of * of).sum(0), rtol=2e-3, atol=2e-1)


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "H%d_%d->%d_k%d_s%d" % s)
def test_dgrad(shape):
    from dps_b200.ops.conv import ConvPlan
    H, Cin, Cout, R, stride = shape
    x, w, dy = _mk(*shape)
    _, dx_ref, _ = _ref(x, w, dy, R, stride)
    if R == 1 and stride == 2:
        # shortcut dgrad accumulates into an existing gradient at parity (0,0) and leaves the rest alone
        base = torch.randn(N, H, H, Cin, device="cuda").to(torch.bfloat16)
        dx = base.clone()
        ConvPlan.dgrad(dy, w, dx, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride, addend=dx).run()
        torch.cuda.synchronize()
        want = base.float() + dx_ref
        assert _relerr(dx, want) < 2e-2
        assert torch.equal(dx[:, 1::2], base[:, 1::2]) and torch.equal(dx[:, :, 1::2], base[:, :, 1::2])
    else:
        dx = torch.full((N, H, H, Cin), 7.0, device="cuda", dtype=torch.bfloat16)
        ConvPlan.dgrad(dy, w, dx, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride).run()
        torch.cuda.synchronize()
        assert _relerr(dx, dx_ref) < 1.5e-2, _relerr(dx, dx_ref)


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "H%d_%d->%d_k%d_s%d" % s)
def test_wgrad(shape):
    from dps_b200.ops.conv import ConvPlan
    H, Cin, Cout, R, stride = shape
    x, w, dy = _mk(*shape)
    _, _, dw_ref = _ref(x, w, dy, R, stride)
    dw = torch.zeros(Cout, R, R, Cin, device="cuda")
    ConvPlan.wgrad(x, dy, dw, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride).run()
    torch.cuda.synchronize()
    assert _relerr(dw, dw_ref) < 5e-3, _relerr(dw, dw_ref)


def test_dgrad_residual_addend_fused():
    from dps_b200.ops.conv import ConvPlan
    shape = (16, 128, 128, 3, 1)
    H, Cin, Cout, R, stride = shape
    x, w, dy = _mk(*shape)
    _, dx_ref, _ = _ref(x, w, dy, R, stride)
    res = torch.randn(N, H, H, Cin, device="cuda").to(torch.bfloat16)
    dx = torch.zeros_like(res)
    ConvPlan.dgrad(dy, w, dx, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride, addend=res).run()
    torch.cuda.synchronize()
    assert _relerr(dx, dx_ref + res.float()) < 2e-2


def test_full_batch_layer1_shape():
    from dps_b200.ops.conv import ConvPlan
    Nb, H, C = 128, 32, 64
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(Nb, H, H, C, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(C, 3, 3, C, device="cuda", generator=g) / 24).to(torch.bfloat16)
    out = torch.empty(Nb, H, H, C, device="cuda", dtype=torch.bfloat16)
    ConvPlan.fprop(x, w, out, N=Nb, H=H, W=H, Cin=C, Cout=C, R=3, stride=1).run()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    assert _relerr(out, ref) < 1.5e-2


@pytest.mark.parametrize("shape", [(32, 64, 64, 3, 1), (16, 128, 128, 3, 1), (32, 64, 128, 3, 2), (8, 256, 512, 3, 2)],
                         ids=lambda s: "H%d_%d->%d_k%d_s%d" % s)
def test_dgrad_with_fused_bn_backward_reduction(shape):
    """dgrad epilogue = dgrad (+addend) then the next BN's backward reduction: dz = dx*(y>0), sum(dz), sum(dz*xhat)."""
    from dps_b200.ops.conv import ConvPlan
    H, Cin, Cout, R, stride = shape
    x, w, dy = _mk(*shape)
    _, dx_ref, _ = _ref(x, w, dy, R, stride)
    g = torch.Generator(device="cuda").manual_seed(7)
    ybn = torch.randn(N, H, H, Cin, device="cuda", generator=g).to(torch.bfloat16)      # forward output of that BN
    xbn = torch.randn(N, H, H, Cin, device="cuda", generator=g).to(torch.bfloat16)      # its input (conv output)
    mean = torch.randn(Cin, device="cuda", generator=g) * 0.1
    rstd = torch.rand(Cin, device="cuda", generator=g) + 0.5
    sums = torch.zeros(2 * Cin, device="cuda", dtype=torch.int64)
    dz = torch.empty(N, H, H, Cin, device="cuda", dtype=torch.bfloat16)
    if stride == 2:      # shortcut gradient arrives at half resolution and lands on parity (0,0)
        half = (torch.randn(N, H // 2, H // 2, Cin, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
        plan = ConvPlan.dgrad(dy, w, dz, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride, addend=half)
        plan.addend_compact(True)
        dx_ref = dx_ref.clone()
        dx_ref[:, 0::2, 0::2] += half.float()
    else:
        plan = ConvPlan.dgrad(dy, w, dz, N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride)
    plan.fuse_bn_bwd(ybn, xbn, mean, rstd, sums).run()
    torch.cuda.synchronize()
    want_dz = dx_ref * (ybn.float() > 0)
    assert _relerr(dz, want_dz) < 2e-2
    assert torch.equal(dz.float() == 0, (dz.float() == 0) | (ybn.float() <= 0))
    dzf = dz.float().reshape(-1, Cin)
    xhat = (xbn.float().reshape(-1, Cin) - mean) * rstd
    got = sums.double() / 2 ** 40
    assert torch.allclose(got[:Cin].float(), dzf.sum(0), rtol=2e-3, atol=5e-2)
    assert torch.allclose(got[Cin:].float(), (dzf * xhat).sum(0), rtol=2e-3, atol=5e-2)


@pytest.mark.parametrize("Nb", [8, 128])
def test_halo2_layer2_kernels(monkeypatch, Nb):
    """halo2conv_kernel (DPS_HALO2=1): 3x3 stride-1, 128 reduction channels, resident 64-wide weight slices, halo tiles,
    register-reduction epilogue.  fprop + statistics, dgrad + addend, dgrad + fused BN-backward sums."""
    from dps_b200.ops.conv import ConvPlan
    monkeypatch.setenv("DPS_HALO2", "1")
    H, C, R = 16, 128, 3
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(Nb, H, H, C, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(C, R, R, C, device="cuda", generator=g) / (9 * C) ** 0.5).to(torch.bfloat16)
    dy = torch.randn(Nb, H, H, C, device="cuda", generator=g).to(torch.bfloat16)
    xf = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wf = w.float().permute(0, 3, 1, 2).contiguous()
    y = F.conv2d(xf, wf, padding=1)
    y.backward(dy.float().permute(0, 3, 1, 2))
    y_ref, dx_ref = y.detach().permute(0, 2, 3, 1), xf.grad.permute(0, 2, 3, 1)
    # fprop + BN statistics
    out = torch.zeros(Nb, H, H, C, device="cuda", dtype=torch.bfloat16)
    stats = torch.zeros(2 * C, device="cuda", dtype=torch.int64)
    plan = ConvPlan.fprop(x, w, out, N=Nb, H=H, W=H, Cin=C, Cout=C, R=R, stride=1, stats=stats)
    assert plan.grid[1] == 2 and plan.grid[0] == min(Nb * 2, 74) and plan.smem > 200 * 1024      # the halo2 path
    plan.run()
    torch.cuda.synchronize()
    assert _relerr(out, y_ref) < 1.5e-2
    of = out.float().reshape(-1, C)
    assert torch.allclose((stats[:C].double() / 2 ** 32).float(), of.sum(0), rtol=1e-3, atol=5e-2)
    assert torch.allclose((stats[C:].double() / 2 ** 24).float(), (of * of).sum(0), rtol=1e-3, atol=5e-1)
    # dgrad + residual addend
    res = torch.randn(Nb, H, H, C, device="cuda", generator=g).to(torch.bfloat16)
    dx = torch.zeros_like(res)
    plan = ConvPlan.dgrad(dy, w, dx, N=Nb, H=H, W=H, Cin=C, Cout=C, R=R, stride=1, addend=res)
    assert plan.grid[1] == 2
    plan.run()
    torch.cuda.synchronize()
    assert _relerr(dx, dx_ref + res.float()) < 2e-2
    # dgrad + fused BN-backward reduction
    ybn = torch.randn(Nb, H, H, C, device="cuda", generator=g).to(torch.bfloat16)
    xbn = torch.randn(Nb, H, H, C, device="cuda", generator=g).to(torch.bfloat16)
    mean = torch.randn(C, device="cuda", generator=g) * 0.1
    rstd = torch.rand(C, device="cuda", generator=g) + 0.5
    sums = torch.zeros(2 * C, device="cuda", dtype=torch.int64)
    dz = torch.empty(Nb, H, H, C, device="cuda", dtype=torch.bfloat16)
    ConvPlan.dgrad(dy, w, dz, N=Nb, H=H, W=H, Cin=C, Cout=C, R=R, stride=1).fuse_bn_bwd(ybn, xbn, mean, rstd, sums).run()
    torch.cuda.synchronize()
    assert _relerr(dz, dx_ref * (ybn.float() > 0)) < 2e-2
    dzf = dz.float().reshape(-1, C)
    xhat = (xbn.float().reshape(-1, C) - mean) * rstd
    got = sums.double() / 2 ** 40
    assert torch.allclose(got[:C].float(), dzf.sum(0), rtol=2e-3, atol=5e-2 * (Nb / 8) ** 0.5)
    assert torch.allclose(got[C:].float(), (dzf * xhat).sum(0), rtol=2e-3, atol=5e-2 * (Nb / 8) ** 0.5)


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "H%d_%d->%d_k%d_s%d" % s)
def test_all_passes_at_the_training_batch(shape):
    """The whole shape matrix again at N = 128, the batch the engine runs: other grids (several waves, CTA pairs on the
    layer-4 shapes, split reductions with many CTAs per output tile in wgrad) than at N = 8."""
    from dps_b200.ops.conv import ConvPlan
    H, Cin, Cout, R, stride = shape
    Nb = 128
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn(Nb, H, H, Cin, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(Cout, R, R, Cin, device="cuda", generator=g) / (R * R * Cin) ** 0.5).to(torch.bfloat16)
    Ho = H // stride
    dy = torch.randn(Nb, Ho, Ho, Cout, device="cuda", generator=g).to(torch.bfloat16)
    y_ref, dx_ref, dw_ref = _ref(x, w, dy, R, stride)
    kw = dict(N=Nb, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=stride)
    out = torch.zeros(Nb, Ho, Ho, Cout, device="cuda", dtype=torch.bfloat16)
    ConvPlan.fprop(x, w, out, **kw).run()
    dw = torch.zeros(Cout, R, R, Cin, device="cuda")
    ConvPlan.wgrad(x, dy, dw, **kw).run()
    if R == 1 and stride == 2:
        base = torch.randn(Nb, H, H, Cin, device="cuda", generator=g).to(torch.bfloat16)
        dx = base.clone()
        ConvPlan.dgrad(dy, w, dx, addend=dx, **kw).run()
        dx_want = base.float() + dx_ref
    else:
        dx = torch.zeros(Nb, H, H, Cin, device="cuda", dtype=torch.bfloat16)
        ConvPlan.dgrad(dy, w, dx, **kw).run()
        dx_want = dx_ref
    torch.cuda.synchronize()
    assert _relerr(out, y_ref) < 1.5e-2, ("fprop", _relerr(out, y_ref))
    assert _relerr(dx, dx_want) < 2e-2, ("dgrad", _relerr(dx, dx_want))
    assert _relerr(dw, dw_ref) < 5e-3, ("wgrad", _relerr(dw, dw_ref))


@pytest.mark.parametrize("shape", [(16, 128, 128, 3, 1), (8, 256, 256, 3, 1), (4, 512, 512, 3, 1)],
                         ids=lambda s: "H%d_%d->%d_k%d_s%d" % s)
def test_fused_bn_backward_mask_recomputed_from_x(shape):
    """A BatchNorm without residual input: the fused dgrad epilogue recomputes the ReLU mask from x
    (y = relu(x*sc + sh) > 0) and never reads y.  Oracle: dz = dx * (bn(x) > 0) in fp32 torch, and the same sums."""
    from dps_b200.ops.conv import ConvPlan
    H, Cin, Cout, R, stride = shape
    Nb = 32
    g = torch.Generator(device="cuda").manual_seed(5)
    w = (torch.randn(Cout, R, R, Cin, device="cuda", generator=g) / (R * R * Cin) ** 0.5).to(torch.bfloat16)
    dy = torch.randn(Nb, H, H, Cout, device="cuda", generator=g).to(torch.bfloat16)
    wf = w.float().permute(0, 3, 1, 2).contiguous()
    dx_ref = torch.nn.grad.conv2d_input((Nb, Cin, H, H), wf, dy.float().permute(0, 3, 1, 2).contiguous(),
                                        padding=1).permute(0, 2, 3, 1)
    xbn = (torch.randn(Nb, H, H, Cin, device="cuda", generator=g) * 1.3 + 0.2).to(torch.bfloat16)
    mean = xbn.float().reshape(-1, Cin).mean(0)
    rstd = (xbn.float().reshape(-1, Cin).var(0, unbiased=False) + 1e-5).rsqrt()
    gamma = torch.rand(Cin, device="cuda", generator=g) + 0.5
    gamma[::7] *= -1.0                                       # negative scales flip the mask: must be honoured
    beta = torch.randn(Cin, device="cuda", generator=g) * 0.3
    sc = gamma * rstd
    pre = xbn.float() * sc + (beta - mean * sc)
    poison = torch.full_like(xbn, float("nan"))              # y must not be read on this path
    sums = torch.zeros(2 * Cin, device="cuda", dtype=torch.int64)
    dz = torch.empty(Nb, H, H, Cin, device="cuda", dtype=torch.bfloat16)
    plan = ConvPlan.dgrad(dy, w, dz, N=Nb, H=H, W=H, Cin=Cin, Cout=Cout, R=R, stride=1)
    plan.fuse_bn_bwd(poison, xbn, mean, rstd, sums).mask_from_x(gamma, beta).run()
    torch.cuda.synchronize()
    want = dx_ref * (pre > 0)
    # elements whose pre-activation is within rounding of zero may fall on either side of the mask
    sure = pre.abs() > 1e-3
    assert _relerr(dz.float() * sure, want * sure) < 2e-2
    dzf = dz.float().reshape(-1, Cin)
    xhat = (xbn.float().reshape(-1, Cin) - mean) * rstd
    got = sums.double() / 2 ** 40
    assert torch.allclose(got[:Cin].float(), dzf.sum(0), rtol=2e-3, atol=5e-2 * (Nb / 8) ** 0.5)
    assert torch.allclose(got[Cin:].float(), (dzf * xhat).sum(0), rtol=2e-3, atol=5e-2 * (Nb / 8) ** 0.5)
