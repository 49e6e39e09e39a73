"""BatchNorm backward through the entry point that allocates nothing (dps_bn_backward_noalloc).

Who this is for: code that cannot hand the kernel a scratch buffer -- a process whose GPU memory is
entirely owned by a fixed-budget pool, a loop that cannot afford cudaMalloc's device-wide sync, a hook
that only gets its inputs and outputs.  The library then reduces into a built-in 8 KiB device array
(at most 512 channels) that every call shares and that still holds the previous call's sums, so the
entry point has to clear it itself.  This test drives the C ABI directly with ctypes and checks, on two
calls in a row with different shapes and data, that the results equal those of the existing
dps_bn_backward (which is given a fresh zeroed scratch each time) and, independently of that sibling, those
of a plain fp32 PyTorch restatement of BatchNorm + ReLU backward -- a bug shared by both entry points would
pass the first comparison but not the second.  The file name sorts last on purpose: if it fails under `-x`,
every other test has already run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

P_, i32, i64 = C.c_void_p, C.c_int, C.c_longlong


def _lib():
    from dps_b200.ops import _lib as L
    lib = L.lib()
    lib.dps_bn_backward.restype = i32
    lib.dps_bn_backward.argtypes = [P_] * 11 + [i64, i32, i32, P_]
    lib.dps_bn_backward_noalloc.restype = i32
    lib.dps_bn_backward_noalloc.argtypes = [P_] * 10 + [i64, i32, i32, P_]
    return lib, L


def _case(P, Cc, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    x = r(P, Cc).to(torch.bfloat16)
    y = r(P, Cc).to(torch.bfloat16)                 # sign pattern of the forward output (ReLU mask)
    dy = (r(P, Cc) * 0.1).to(torch.bfloat16)
    mean, rstd, gamma = r(Cc) * 0.1, r(Cc).abs() + 0.5, r(Cc).abs() + 0.5
    return x, y, dy, mean, rstd, gamma


def _run(lib, L, noalloc, x, y, dy, mean, rstd, gamma, P, Cc):
    dz, dx = torch.empty_like(dy), torch.empty_like(dy)
    dgamma, dbeta = torch.zeros(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    st = C.c_void_p(L.stream_ptr())
    p = lambda t: C.c_void_p(t.data_ptr())
    if noalloc:
        rc = lib.dps_bn_backward_noalloc(p(dy), p(y), p(x), p(mean), p(rstd), p(gamma), p(dz), p(dx), p(dgamma),
                                         p(dbeta), P, Cc, 1, st)
    else:
        sums = torch.zeros(2 * Cc, device="cuda", dtype=torch.int64)
        rc = lib.dps_bn_backward(p(dy), p(y), p(x), p(mean), p(rstd), p(gamma), p(dz), p(sums), p(dx), p(dgamma),
                                 p(dbeta), P, Cc, 1, st)
    assert rc == 0, lib.dps_last_error()
    torch.cuda.synchronize()
    return dz, dx, dgamma, dbeta


def _torch_oracle(x, y, dy, mean, rstd, gamma):
    """fp32 restatement: dz = dy * (y > 0); dbeta = sum dz; dgamma = sum dz * xhat;
    dx = gamma * rstd * (dz - mean(dz) - xhat * mean(dz * xhat))."""
    xf, dyf = x.float(), dy.float()
    dz = dyf * (y.float() > 0).float()
    xhat = (xf - mean) * rstd
    P = xf.shape[0]
    dbeta, dgamma = dz.sum(0), (dz * xhat).sum(0)
    dx = gamma * rstd * (dz - dbeta / P - xhat * dgamma / P)
    return dz, dx, dgamma, dbeta


def test_noalloc_bn_backward_matches_existing_entry_point_twice_in_a_row():
    lib, L = _lib()
    # second call: fewer channels and other data -- stale sums from the first call would show up here
    for P, Cc, seed in [(4096, 256, 1), (2048, 64, 2)]:
        case = _case(P, Cc, seed)
        want = _run(lib, L, False, *case, P, Cc)
        got = _run(lib, L, True, *case, P, Cc)
        for name, a, b in zip(("dz", "dx", "dgamma", "dbeta"), got, want):
            if a.dtype == torch.bfloat16:
                assert torch.allclose(a.float(), b.float(), rtol=1.5e-2, atol=1e-3), (name, Cc)
            else:
                assert torch.allclose(a, b, rtol=1e-5, atol=1e-6), (name, Cc)
        ref = _torch_oracle(*case)
        for name, a, b in zip(("dz", "dx", "dgamma", "dbeta"), got, ref):
            tol = 1.5e-2 * float(b.abs().max()) + 1e-3
            assert float((a.float() - b).abs().max()) <= tol, (name, Cc, float((a.float() - b).abs().max()), tol)


def test_noalloc_two_consecutive_calls_replayed_from_a_captured_graph():
    """The memset + two kernels of each call become graph nodes; two calls back to back share the built-in array, so
    the second call's memset must be ordered behind the first call's apply kernel in the captured graph as well.
    Replayed three times: every replay must reproduce the eager results of both calls."""
    lib, L = _lib()
    cases = [(4096, 128, 11), (1024, 512, 12)]
    data = [_case(P, Cc, seed) for P, Cc, seed in cases]
    eager = [_run(lib, L, True, *d, P, Cc) for d, (P, Cc, _) in zip(data, cases)]     # also resolves the symbol address
    outs = []
    for d, (P, Cc, _) in zip(data, cases):
        dy = d[2]
        outs.append((torch.empty_like(dy), torch.empty_like(dy), torch.zeros(Cc, device="cuda"),
                     torch.zeros(Cc, device="cuda")))
    p = lambda t: C.c_void_p(t.data_ptr())
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            st = C.c_void_p(L.stream_ptr())
            for (x, y, dy, mean, rstd, gamma), (dz, dx, dgamma, dbeta), (P, Cc, _) in zip(data, outs, cases):
                rc = lib.dps_bn_backward_noalloc(p(dy), p(y), p(x), p(mean), p(rstd), p(gamma), p(dz), p(dx),
                                                 p(dgamma), p(dbeta), P, Cc, 1, st)
                assert rc == 0, lib.dps_last_error()
    for _ in range(3):
        for dz, dx, dgamma, dbeta in outs:
            dz.fill_(float("nan")), dx.fill_(float("nan"))
            dgamma.zero_(), dbeta.zero_()                 # the entry point ACCUMULATES into dgamma / dbeta
        graph.replay()
        torch.cuda.synchronize()
        for got, want, (P, Cc, _) in zip(outs, eager, cases):
            for name, a, b in zip(("dz", "dx", "dgamma", "dbeta"), got, want):
                assert torch.equal(a, b), (name, Cc)      # integer fixed-point sums: bit-identical to the eager calls


def test_noalloc_rejects_channel_counts_beyond_the_built_in_scratch():
    lib, L = _lib()
    t = torch.zeros(8, device="cuda")
    p = C.c_void_p(t.data_ptr())
    rc = lib.dps_bn_backward_noalloc(p, p, p, p, p, p, p, p, p, p, 8, 1024, 1, C.c_void_p(L.stream_ptr()))
    assert rc != 0
    msg = lib.dps_last_error().decode()
    assert "512" in msg and "1024" in msg
