"""Python handles for the tcgen05 implicit-GEMM convolution plans (csrc/conv_tc.cu).

A plan binds device buffers (fixed addresses) to pre-encoded TMA descriptors; running it is one
kernel launch on the given stream.  Activations NHWC bf16, weights [Cout][R*R*Cin] bf16 (K-R-S-C),
weight gradients fp32 accumulated with atomics (caller zeroes).
Replaces nn.Conv2d forward/backward of the reference model (reference: src/parameter_server/server.py:24-48).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L

_declared = False


def _decl() -> None:
    global _declared
    if _declared:
        return
    P, i32 = C.c_void_p, C.c_int
    L.declare("dps_conv_plan_create", i32, [i32, P, P, P, P, P, i32, i32, i32, i32, i32, i32, i32, i32, i32,
                                            C.POINTER(P)])
    L.declare("dps_wgrad_plan_create", i32, [P, P, P, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32,
                                             C.POINTER(P)])
    L.declare("dps_conv_plan_run", i32, [P, P])
    L.declare("dps_conv_plan_fuse_bn_bwd", i32, [P, P, P, P, P, P])
    L.declare("dps_conv_plan_addend_compact", i32, [P, i32])
    L.declare("dps_conv_plan_fuse_bn_mask_from_x", i32, [P, P, P])
    L.declare("dps_conv_plan_destroy", i32, [P])
    L.declare("dps_conv_plan_info", i32, [P, C.POINTER(i32 * 3), C.POINTER(i32), C.POINTER(i32)])
    _declared = True


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class ConvPlan:
    FPROP, DGRAD, WGRAD = 0, 1, 2

    def __init__(self, handle: int, kind: int, keep):
        self.handle, self.kind, self._keep = handle, kind, keep
        info = (C.c_int * 3)()
        bn, sm = C.c_int(), C.c_int()
        L.check(L.lib().dps_conv_plan_info(C.c_void_p(handle), C.byref(info), C.byref(bn), C.byref(sm)), "plan_info")
        self.grid, self.block_n, self.smem = tuple(info), bn.value, sm.value
        self.desc = ""                      # human-readable shape, set by the factories (timeline / profiles)

    @classmethod
    def fprop(cls, x, w, out, *, N, H, W, Cin, Cout, R, stride, stats=None, w_cols=None, w_pitch=None):
        _decl()
        h = C.c_void_p()
        w_cols = R * R * Cin if w_cols is None else w_cols
        w_pitch = R * R * Cin if w_pitch is None else w_pitch
        L.check(L.lib().dps_conv_plan_create(0, _ptr(x), _ptr(w), _ptr(out), None, _ptr(stats), N, H, W, Cin, Cout,
                                             R, stride, w_cols, w_pitch, C.byref(h)), "conv_plan_create(fprop)")
        pl = cls(h.value, cls.FPROP, (x, w, out, stats))
        pl.desc = f"H{H} {Cin}->{Cout} k{R} s{stride}"
        return pl

    @classmethod
    def dgrad(cls, dy, w, dx, *, N, H, W, Cin, Cout, R, stride, addend=None, w_cols=None, w_pitch=None):
        """H, W are the conv INPUT (dx) spatial dims."""
        _decl()
        h = C.c_void_p()
        w_cols = R * R * Cin if w_cols is None else w_cols
        w_pitch = R * R * Cin if w_pitch is None else w_pitch
        L.check(L.lib().dps_conv_plan_create(1, _ptr(dy), _ptr(w), _ptr(dx), _ptr(addend), None, N, H, W, Cin, Cout,
                                             R, stride, w_cols, w_pitch, C.byref(h)), "conv_plan_create(dgrad)")
        pl = cls(h.value, cls.DGRAD, (dy, w, dx, addend))
        pl.desc = f"H{H} {Cin}->{Cout} k{R} s{stride}"
        return pl

    @classmethod
    def wgrad(cls, x, dy, dw, *, N, H, W, Cin, Cout, R, stride, valid_cols=None, dw_pitch=None, nsplit=0):
        _decl()
        h = C.c_void_p()
        valid_cols = Cin if valid_cols is None else valid_cols
        dw_pitch = R * R * Cin if dw_pitch is None else dw_pitch
        L.check(L.lib().dps_wgrad_plan_create(_ptr(x), _ptr(dy), _ptr(dw), N, H, W, Cin, Cout, R, stride, valid_cols,
                                              dw_pitch, nsplit, C.byref(h)), "wgrad_plan_create")
        pl = cls(h.value, cls.WGRAD, (x, dy, dw))
        pl.desc = f"H{H} {Cin}->{Cout} k{R} s{stride}"
        return pl

    def fuse_bn_bwd(self, y, x, mean, rstd, sums) -> "ConvPlan":
        """dgrad only: the epilogue also does the next BatchNorm's backward reduction (dz = dy*(y>0), sums)."""
        L.check(L.lib().dps_conv_plan_fuse_bn_bwd(C.c_void_p(self.handle), _ptr(y), _ptr(x), _ptr(mean), _ptr(rstd),
                                                  _ptr(sums)), "conv_plan_fuse_bn_bwd")
        self._keep = self._keep + (y, x, mean, rstd, sums)
        return self

    def mask_from_x(self, gamma, beta) -> "ConvPlan":
        """after fuse_bn_bwd, for a BN without residual input: the ReLU mask is recomputed from x, y is not read."""
        L.check(L.lib().dps_conv_plan_fuse_bn_mask_from_x(C.c_void_p(self.handle), _ptr(gamma), _ptr(beta)),
                "conv_plan_fuse_bn_mask_from_x")
        self._keep = self._keep + (gamma, beta)
        return self

    def addend_compact(self, on: bool = True) -> "ConvPlan":
        L.check(L.lib().dps_conv_plan_addend_compact(C.c_void_p(self.handle), int(on)), "conv_plan_addend_compact")
        return self

    def run(self, stream=None) -> None:
        L.check(L.lib().dps_conv_plan_run(C.c_void_p(self.handle), C.c_void_p(L.stream_ptr(stream))), "conv_plan_run")

    def __del__(self):
        try:
            if self.handle:
                L.lib().dps_conv_plan_destroy(C.c_void_p(self.handle))
                self.handle = 0
        except Exception:
            pass
