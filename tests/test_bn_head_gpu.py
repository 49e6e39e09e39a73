"""the BatchNorm, head and stem-staging kernels one by one against plain PyTorch fp32 of the same op
(reference ops: nn.BatchNorm2d + relu + residual add, server.py:25-40; avg-pool + Linear + CrossEntropyLoss,
server.py:56-57,73-75, worker.py:133; the 3x3 stem's input window, server.py:48).  Every BN width of the model,
pixel counts with odd tails."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

P_, i32, i64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float


def _lib():
    from dps_b200.ops import _lib as L
    lib = L.lib()
    lib.dps_bn_forward.restype = i32
    lib.dps_bn_forward.argtypes = [P_] * 11 + [i64, i32, i32, i32, f32, f32, P_]
    lib.dps_bn_backward.restype = i32
    lib.dps_bn_backward.argtypes = [P_] * 11 + [i64, i32, i32, P_]
    lib.dps_head_forward_backward.restype = i32
    lib.dps_head_forward_backward.argtypes = [P_] * 12 + [i32] * 5 + [P_]
    lib.dps_stem_im2col.restype = i32
    lib.dps_stem_im2col.argtypes = [P_, P_, i32, i32, i32, P_, i32, P_]
    return lib, L


def p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _fixed_stats(x):
    """What the conv epilogue hands to the BN kernel: per-channel sum x 2^32 and sum of squares x 2^24 (int64)."""
    xd = x.double()
    s = (xd.sum(0) * 2.0 ** 32).round().to(torch.int64)
    q = ((xd * xd).sum(0) * 2.0 ** 24).round().to(torch.int64)
    return torch.cat([s, q]).contiguous()


CASES = [(64, 4099), (128, 1031), (256, 515), (512, 129), (64, 131072), (512, 2048)]


@pytest.mark.parametrize("Cc,P", CASES, ids=lambda v: str(v))
@pytest.mark.parametrize("residual,relu", [(False, True), (True, True), (False, False)])
def test_bn_forward_training_vs_torch(Cc, P, residual, relu):
    lib, L = _lib()
    g = torch.Generator(device="cuda").manual_seed(Cc + P)
    x = (torch.randn(P, Cc, device="cuda", generator=g) * 1.5 + 0.3).to(torch.bfloat16)
    res = torch.randn(P, Cc, device="cuda", generator=g).to(torch.bfloat16) if residual else None
    gamma = torch.rand(Cc, device="cuda", generator=g) + 0.5
    beta = torch.randn(Cc, device="cuda", generator=g) * 0.2
    y = torch.empty_like(x)
    mean, rstd = torch.zeros(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    rm, rv = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
    nb = torch.zeros((), dtype=torch.int64, device="cuda")
    stats = _fixed_stats(x)
    rc = lib.dps_bn_forward(p(x), p(stats), p(gamma), p(beta), p(res), p(y), p(mean), p(rstd), p(rm), p(rv), p(nb), P,
                            Cc, int(relu), 1, 1e-5, 0.1, C.c_void_p(L.stream_ptr()))
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    # fp32 oracle: torch's batch_norm in training mode on the same (bf16-rounded) input
    rm_ref, rv_ref = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
    ref = F.batch_norm(x.float(), rm_ref, rv_ref, gamma, beta, training=True, momentum=0.1, eps=1e-5)
    if residual:
        ref = ref + res.float()
    if relu:
        ref = ref.relu()
    assert torch.allclose(y.float(), ref, rtol=1.6e-2, atol=1.6e-2), float((y.float() - ref).abs().max())
    assert torch.allclose(mean, x.float().mean(0), rtol=1e-4, atol=1e-5)
    assert torch.allclose(rstd, (x.float().var(0, unbiased=False) + 1e-5).rsqrt(), rtol=1e-3, atol=1e-5)
    assert torch.allclose(rm, rm_ref, rtol=1e-4, atol=1e-5) and torch.allclose(rv, rv_ref, rtol=1e-3, atol=1e-5)
    assert int(nb) == 1


@pytest.mark.parametrize("Cc,P", CASES, ids=lambda v: str(v))
def test_bn_forward_eval_uses_running_statistics(Cc, P):
    lib, L = _lib()
    g = torch.Generator(device="cuda").manual_seed(7 * Cc + P)
    x = torch.randn(P, Cc, device="cuda", generator=g).to(torch.bfloat16)
    gamma, beta = torch.rand(Cc, device="cuda", generator=g) + 0.5, torch.randn(Cc, device="cuda", generator=g)
    rm, rv = torch.randn(Cc, device="cuda", generator=g) * 0.1, torch.rand(Cc, device="cuda", generator=g) + 0.5
    rm0, rv0 = rm.clone(), rv.clone()
    y = torch.empty_like(x)
    nb = torch.zeros((), dtype=torch.int64, device="cuda")
    rc = lib.dps_bn_forward(p(x), None, p(gamma), p(beta), None, p(y), None, None, p(rm), p(rv), p(nb), P, Cc, 1, 0,
                            1e-5, 0.1, C.c_void_p(L.stream_ptr()))
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    ref = F.batch_norm(x.float(), rm0, rv0, gamma, beta, training=False, eps=1e-5).relu()
    assert torch.allclose(y.float(), ref, rtol=1.6e-2, atol=1.6e-2)
    assert torch.equal(rm, rm0) and torch.equal(rv, rv0) and int(nb) == 0


@pytest.mark.parametrize("Cc,P", CASES, ids=lambda v: str(v))
@pytest.mark.parametrize("relu", [True, False])
def test_bn_backward_vs_torch_autograd(Cc, P, relu):
    lib, L = _lib()
    g = torch.Generator(device="cuda").manual_seed(3 * Cc + P)
    x = (torch.randn(P, Cc, device="cuda", generator=g) * 1.2 - 0.2).to(torch.bfloat16)
    gamma = torch.rand(Cc, device="cuda", generator=g) + 0.5
    beta = torch.randn(Cc, device="cuda", generator=g) * 0.2
    dy = (torch.randn(P, Cc, device="cuda", generator=g) * 0.1).to(torch.bfloat16)
    xf = x.float().requires_grad_(True)
    gf, bf = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    z = F.batch_norm(xf, None, None, gf, bf, training=True, eps=1e-5)
    yf = z.relu() if relu else z
    yf.backward(dy.float())
    mean = x.float().mean(0)
    rstd = (x.float().var(0, unbiased=False) + 1e-5).rsqrt()
    y = yf.detach().to(torch.bfloat16)
    dz, dx = torch.empty_like(dy), torch.empty_like(dy)
    dgamma, dbeta = torch.zeros(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    sums = torch.zeros(2 * Cc, device="cuda", dtype=torch.int64)
    rc = lib.dps_bn_backward(p(dy), p(y), p(x), p(mean), p(rstd), p(gamma), p(dz), p(sums), p(dx), p(dgamma),
                             p(dbeta), P, Cc, int(relu), C.c_void_p(L.stream_ptr()))
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    scale = float(xf.grad.abs().max())
    assert float((dx.float() - xf.grad).abs().max()) < 2e-2 * scale + 1e-4
    # parameter gradients are fp32 sums over P rows of bf16-rounded terms
    assert torch.allclose(dgamma, gf.grad, rtol=2e-2, atol=2e-2 * float(gf.grad.abs().max()))
    assert torch.allclose(dbeta, bf.grad, rtol=2e-2, atol=2e-2 * float(bf.grad.abs().max()))
    if relu:
        mask = (y.float() > 0).float()
        assert torch.equal(dz.float(), dy.float() * mask)


@pytest.mark.parametrize("N,HW,Fdim,K", [(128, 16, 512, 100), (16, 16, 512, 100), (37, 4, 256, 10), (128, 784, 512, 100)],
                         ids=lambda v: str(v))
def test_head_forward_backward_vs_torch(N, HW, Fdim, K):
    lib, L = _lib()
    g = torch.Generator(device="cuda").manual_seed(N + HW)
    act = torch.randn(N, HW, Fdim, device="cuda", generator=g).relu().to(torch.bfloat16)
    w = (torch.randn(K, Fdim, device="cuda", generator=g) / Fdim ** 0.5).to(torch.bfloat16)
    bias = torch.randn(K, device="cuda", generator=g) * 0.1
    labels = torch.randint(0, K, (N,), device="cuda", generator=g)
    pooled = torch.zeros(N, Fdim, device="cuda")
    dlogits = torch.zeros(N, 128, device="cuda")
    loss_n = torch.zeros(N, device="cuda")
    correct_n = torch.zeros(N, dtype=torch.int32, device="cuda")
    d_act = torch.zeros_like(act)
    dw, db, out = torch.zeros(K, Fdim, device="cuda"), torch.zeros(K, device="cuda"), torch.zeros(2, device="cuda")
    rc = lib.dps_head_forward_backward(p(act), p(w), p(bias), p(labels), p(pooled), p(dlogits), p(loss_n), p(correct_n),
                                       p(d_act), p(dw), p(db), p(out), N, HW, Fdim, K, 1, C.c_void_p(L.stream_ptr()))
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    af = act.float().requires_grad_(True)
    wf, bf = w.float().requires_grad_(True), bias.clone().requires_grad_(True)
    logits = af.mean(1) @ wf.t() + bf
    loss = F.cross_entropy(logits, labels)
    loss.backward()
    assert abs(float(out[0]) - float(loss)) < 2e-4 * max(1.0, abs(float(loss)))
    assert int(out[1]) == int((logits.argmax(1) == labels).sum())
    assert torch.allclose(dw, wf.grad, rtol=1e-3, atol=1e-5 + 1e-3 * float(wf.grad.abs().max()))
    assert torch.allclose(db, bf.grad, rtol=1e-3, atol=1e-6 + 1e-3 * float(bf.grad.abs().max()))
    gs = float(af.grad.abs().max())
    assert float((d_act.float() - af.grad).abs().max()) < 1e-2 * gs + 1e-9


def test_head_eval_ignores_padding_rows():
    lib, L = _lib()
    N, HW, Fdim, K = 16, 16, 512, 100
    g = torch.Generator(device="cuda").manual_seed(5)
    act = torch.randn(N, HW, Fdim, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(K, Fdim, device="cuda", generator=g) / Fdim ** 0.5).to(torch.bfloat16)
    bias = torch.zeros(K, device="cuda")
    logits = act.float().mean(1) @ w.float().t()
    labels = logits.argmax(1).clone()
    labels[5:] = -1                                     # padded tail of an evaluation batch
    bufs = [torch.zeros(N, Fdim, device="cuda"), torch.zeros(N, 128, device="cuda"), torch.zeros(N, device="cuda"),
            torch.zeros(N, dtype=torch.int32, device="cuda"), torch.zeros_like(act), torch.zeros(K, Fdim, device="cuda"),
            torch.zeros(K, device="cuda"), torch.zeros(2, device="cuda")]
    rc = lib.dps_head_forward_backward(p(act), p(w), p(bias), p(labels), *[p(b) for b in bufs], N, HW, Fdim, K, 0,
                                       C.c_void_p(L.stream_ptr()))
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    assert int(bufs[-1][1]) == 5
    assert float(bufs[5].abs().max()) == 0.0            # eval mode leaves the gradients alone


@pytest.mark.parametrize("N,S", [(128, 32), (3, 32), (5, 16)], ids=lambda v: str(v))
def test_stem_im2col_vs_unfold(N, S):
    lib, L = _lib()
    g = torch.Generator(device="cuda").manual_seed(N)
    img = torch.randn(N, 3, S, S, device="cuda", generator=g)
    col = torch.zeros(N * S * S, 64, device="cuda", dtype=torch.bfloat16)
    scratch = torch.full((1000,), 7, device="cuda", dtype=torch.int64)     # cleared by the same launch
    rc = lib.dps_stem_im2col(p(img), p(col), N, S, S, p(scratch), 1000, C.c_void_p(L.stream_ptr()))
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    assert int(scratch.abs().max()) == 0
    # unfold orders columns (c, r, s); the kernel's K order is (r, s, c) to match the K-R-S-C weights
    u = F.unfold(img, 3, padding=1).view(N, 3, 9, S * S).permute(0, 3, 2, 1).reshape(N * S * S, 27)
    assert torch.equal(col[:, :27].float(), u.to(torch.bfloat16).float())
    assert float(col[:, 27:].abs().max()) == 0.0
