"""ResNet-18 training engine built from the hand-written sm_100a kernels.

This is the hot path on B200: a *static* schedule of kernel launches over pre-allocated NHWC bf16
buffers -- implicit-GEMM convs on tcgen05/TMEM fed by TMA (ops/conv.py, csrc/conv_tc.cu), fused
BatchNorm(+residual)+ReLU (csrc/bn.cu), fused avg-pool + FC + cross-entropy head (csrc/head.cu) --
captured once into a CUDA graph and replayed every step.  No autograd, no cuDNN/cuBLAS, no tracing
compiler.  Same network as ``models/resnet.py`` (reference: src/parameter_server/server.py:21-76);
weights are read from the worker's bf16 arena and gradients are accumulated into the flat fp32
gradient arena, both laid out by ``store/shardmap.py`` -- exactly what the Push/Pull kernels move.
"""
from __future__ import annotations

import os

import ctypes as C
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Tuple

import torch

from ..ops import _lib as L
from ..ops.conv import ConvPlan
from ..store.shardmap import STEM_PITCH, Layout
from .resnet import BLOCKS_PER_STAGE, STAGE_STRIDES, STAGE_WIDTHS

_declared = False


def _decl() -> None:
    global _declared
    if _declared:
        return
    P, i32, i64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float
    L.declare("dps_bn_forward", i32, [P, P, P, P, P, P, P, P, P, P, P, i64, i32, i32, i32, f32, f32, P])
    L.declare("dps_bn_backward", i32, [P, P, P, P, P, P, P, P, P, P, P, i64, i32, i32, P])
    L.declare("dps_bn_backward_apply", i32, [P, P, P, P, P, P, P, P, P, i64, i32, P])
    L.declare("dps_head_forward_backward", i32, [P] * 12 + [i32, i32, i32, i32, i32, P])
    L.declare("dps_stem_im2col", i32, [P, P, i32, i32, i32, P, i32, P])
    _declared = True


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class _Marker:
    """Schedule entries that launch nothing themselves (fork / join of a side branch, push hooks)."""

    def __call__(self) -> None:
        pass


class _GroupDone(_Marker):
    """Marker in the backward schedule: every weight gradient of layer group `k` has been issued."""

    def __init__(self, engine: "ResNetEngine", k: int):
        self.engine, self.k = engine, k

    def __call__(self) -> None:
        if self.engine.on_group_done is not None:
            self.engine.on_group_done(self.k)


class _Fork(_Marker):
    """The entries up to the matching _Join that are tagged for the branch stream run beside the main chain."""


class _Join(_Marker):
    pass


@dataclass
class _BN:
    name: str                 # state_dict prefix, e.g. 'layer1.0.bn1'
    C: int
    P: int
    gamma: torch.Tensor
    beta: torch.Tensor
    dgamma: torch.Tensor
    dbeta: torch.Tensor
    stats: torch.Tensor       # [2C] fwd sums (zeroed per step)
    bsums: torch.Tensor       # [2C] bwd sums (zeroed per step)
    mean: torch.Tensor
    rstd: torch.Tensor
    running_mean: torch.Tensor
    running_var: torch.Tensor
    num_batches: torch.Tensor


class ResNetEngine:
    BN_EPS = 1e-5
    BN_MOMENTUM = 0.1

    def __init__(self, layout: Layout, wbf16: torch.Tensor, small: torch.Tensor, grad: torch.Tensor,
                 batch: int, image_size: int = 32, num_classes: int = 100, device: Optional[torch.device] = None):
        _decl()
        self.lib = L.lib()
        self.layout, self.wbf16, self.small, self.grad = layout, wbf16, small, grad
        self.N, self.S, self.K = batch, image_size, num_classes
        self.dev = device or wbf16.device
        self.bns: List[_BN] = []
        self.fwd: List[Callable[[], None]] = []
        self.bwd: List[Callable[[], None]] = []
        self.evalf: List[Callable[[], None]] = []
        self.n_launch_fwd = self.n_launch_bwd = 0
        self._keep: List[object] = []
        self._scratch_need = 0
        self._scratch_slices: List[Tuple[int, int]] = []
        self._build()
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.seg_wait: Optional[Callable[[int], None]] = None   # set by the worker: wait for a pulled layer group
        # set by the worker: weight gradients (which nothing in the backward chain consumes) run on `side_stream`,
        # and `on_group_done(k)` is called there once every weight gradient of layer group k has been issued
        self.side_stream: Optional[torch.cuda.Stream] = None
        self.branch_stream: Optional[torch.cuda.Stream] = None
        self.on_group_done: Optional[Callable[[int], None]] = None

    # ------------------------------------------------------------------ helpers
    def _bf(self, *shape) -> torch.Tensor:
        return torch.zeros(*shape, dtype=torch.bfloat16, device=self.dev)

    def _w(self, name: str) -> torch.Tensor:
        e = self.layout.entries[name]
        return self.wbf16[e.offset: e.offset + e.span]

    def _g(self, name: str) -> torch.Tensor:
        e = self.layout.entries[name]
        return self.grad[e.offset: e.offset + e.span]

    def _sm(self, name: str) -> torch.Tensor:
        e = self.layout.entries[name]
        return self.small[e.offset: e.offset + e.span]

    def _stream(self) -> C.c_void_p:
        return C.c_void_p(L.stream_ptr())

    def _make_bn(self, name: str, Cc: int, P: int) -> _BN:
        f = lambda n: torch.zeros(n, dtype=torch.float32, device=self.dev)
        off = self._scratch_need
        self._scratch_need += 4 * Cc
        bn = _BN(name, Cc, P, self._sm(name + ".weight"), self._sm(name + ".bias"), self._g(name + ".weight"),
                 self._g(name + ".bias"), None, None, f(Cc), f(Cc), f(Cc), torch.ones(Cc, device=self.dev),
                 torch.zeros((), dtype=torch.int64, device=self.dev))
        self._scratch_slices.append((off, Cc))
        self.bns.append(bn)
        return bn

    # ------------------------------------------------------------------ op emitters
    def _bn_fwd(self, bn: _BN, x, y, residual, relu, training):
        L.check(self.lib.dps_bn_forward(_p(x), _p(bn.stats), _p(bn.gamma), _p(bn.beta), _p(residual), _p(y),
                                        _p(bn.mean), _p(bn.rstd), _p(bn.running_mean), _p(bn.running_var),
                                        _p(bn.num_batches), bn.P, bn.C, int(relu), int(training), self.BN_EPS,
                                        self.BN_MOMENTUM, self._stream()), "bn_forward")

    def _bn_bwd(self, bn: _BN, dy, y, x, dz, dx, relu):
        L.check(self.lib.dps_bn_backward(_p(dy), _p(y), _p(x), _p(bn.mean), _p(bn.rstd), _p(bn.gamma), _p(dz),
                                         _p(bn.bsums), _p(dx), _p(bn.dgamma), _p(bn.dbeta), bn.P, bn.C, int(relu),
                                         self._stream()), "bn_backward")

    def _bn_bwd_apply(self, bn: _BN, dz, x, dx):
        L.check(self.lib.dps_bn_backward_apply(_p(dz), _p(x), _p(bn.mean), _p(bn.rstd), _p(bn.gamma), _p(bn.bsums),
                                               _p(dx), _p(bn.dgamma), _p(bn.dbeta), bn.P, bn.C, self._stream()),
                "bn_backward_apply")

    # ------------------------------------------------------------------ graph construction
    def _conv_bn_unit(self, x, prefix_conv, prefix_bn, *, H, Cin, Cout, R, stride, relu, residual=None, stem=False):
        """conv -> BN(+residual)(+ReLU).  Returns dict with buffers and registers fwd/eval launches."""
        Ho = H // stride
        P = self.N * Ho * Ho
        c = self._bf(self.N, Ho, Ho, Cout)
        a = self._bf(self.N, Ho, Ho, Cout)
        bn = self._make_bn(prefix_bn, Cout, P)
        unit = dict(x=x, c=c, a=a, bn=bn, H=H, Cin=Cin, Cout=Cout, R=R, stride=stride, relu=relu, stem=stem,
                    wname=prefix_conv + ".weight", residual=residual)
        return unit

    def _finalize_unit_fwd(self, u, join_before_bn: bool = False):
        """Create plans (needs bn.stats allocated) and append forward / eval launches."""
        bn = u["bn"]
        kw = dict(N=self.N, H=u["H"], W=u["H"], Cin=u["Cin"], Cout=u["Cout"], R=u["R"], stride=u["stride"])
        if u["stem"]:
            kw.update(w_cols=STEM_PITCH, w_pitch=STEM_PITCH)
        w = self._w(u["wname"])
        plan = ConvPlan.fprop(u["x"], w, u["c"], stats=bn.stats, **kw)
        plan_eval = ConvPlan.fprop(u["x"], w, u["c"], stats=None, **kw)
        self._keep += [plan, plan_eval]
        u["fprop"] = plan
        res = u["residual"]
        self.fwd.append(plan.run)
        if join_before_bn:
            self.fwd.append(_Join())
        self.fwd.append(lambda bn=bn, u=u, res=res: self._bn_fwd(bn, u["c"], u["a"], res, u["relu"], True))
        self.evalf.append(plan_eval.run)
        self.evalf.append(lambda bn=bn, u=u, res=res: self._bn_fwd(bn, u["c"], u["a"], res, u["relu"], False))

    def _build(self) -> None:
        N, S = self.N, self.S
        dev = self.dev
        self.images = torch.zeros(N, 3, S, S, dtype=torch.float32, device=dev)
        self.labels = torch.zeros(N, dtype=torch.int64, device=dev)
        self.col = self._bf(N * S * S, 64)
        units: List[dict] = []
        # stem as a K=64 GEMM over im2col rows
        stem = self._conv_bn_unit(self.col.view(N, S, S, 64), "conv1", "bn1", H=S, Cin=64, Cout=STAGE_WIDTHS[0], R=1,
                                  stride=1, relu=True, stem=True)
        units.append(stem)
        x, H, Cin = stem["a"], S, STAGE_WIDTHS[0]
        blocks = []
        for li, (Cout, s0) in enumerate(zip(STAGE_WIDTHS, STAGE_STRIDES), start=1):
            for b in range(BLOCKS_PER_STAGE):
                stride = s0 if b == 0 else 1
                pre = f"layer{li}.{b}"
                u1 = self._conv_bn_unit(x, pre + ".conv1", pre + ".bn1", H=H, Cin=Cin, Cout=Cout, R=3, stride=stride,
                                        relu=True)
                Ho = H // stride
                us = None
                if stride != 1 or Cin != Cout:
                    us = self._conv_bn_unit(x, pre + ".shortcut.0", pre + ".shortcut.1", H=H, Cin=Cin, Cout=Cout, R=1,
                                            stride=stride, relu=False)
                res = us["a"] if us is not None else x
                u2 = self._conv_bn_unit(u1["a"], pre + ".conv2", pre + ".bn2", H=Ho, Cin=Cout, Cout=Cout, R=3, stride=1,
                                        relu=True, residual=res)
                # forward order: conv1/bn1, shortcut (so its output exists before bn2), conv2/bn2
                units += [u1] + ([us] if us is not None else []) + [u2]
                blocks.append(dict(x=x, u1=u1, u2=u2, us=us, H=H, Ho=Ho, Cin=Cin, Cout=Cout, stride=stride))
                x, H, Cin = u2["a"], Ho, Cout
        self.units, self.blocks = units, blocks
        self.feat, self.featH, self.F = x, H, Cin

        # per-step zeroed scratch: [fwd stats | bwd sums] for every BN (64-bit fixed point: order-independent atomics)
        self.scratch = torch.zeros(self._scratch_need, dtype=torch.int64, device=dev)
        for bn, (off, Cc) in zip(self.bns, self._scratch_slices):
            bn.stats = self.scratch[off: off + 2 * Cc]
            bn.bsums = self.scratch[off + 2 * Cc: off + 4 * Cc]

        # ---------------- forward schedule ----------------
        # first launch of a step: input staging for the stem, which also clears the per-step statistics scratch
        im2col = lambda: L.check(self.lib.dps_stem_im2col(_p(self.images), _p(self.col), N, S, S, _p(self.scratch),
                                                          self.scratch.numel(), self._stream()), "stem_im2col")
        self.fwd.append(im2col)
        self.evalf.append(im2col)
        # layer groups of the segmented pull: [small params + stem + layer1 | layer2 | layer3 | layer4 | fc]
        self.segment_boundaries = [0] + [self.layout.entries[n].offset for n in
                                         ("layer2.0.conv1.weight", "layer3.0.conv1.weight", "layer4.0.conv1.weight",
                                          "fc.weight")] + [self.layout.total]
        first_of_group = {id(stem): 0}
        for bi, blk in enumerate(blocks):
            if bi > 0 and bi % BLOCKS_PER_STAGE == 0:
                first_of_group[id(blk["u1"])] = bi // BLOCKS_PER_STAGE
        # A projection shortcut (1x1 conv + BN) depends only on the block input: it runs on the branch stream beside
        # conv1 / bn1 / conv2 of its block and is joined in front of bn2, which adds it as the residual.
        self._branch_ids = set()
        shortcut_of = {id(b["u1"]): b["us"] for b in blocks if b["us"] is not None}
        joins_before_bn = {id(b["u2"]) for b in blocks if b["us"] is not None}
        for u in units:
            if any(u is s for s in shortcut_of.values()):
                continue                                        # emitted with its block's conv1 below
            if id(u) in first_of_group:
                k = first_of_group[id(u)]
                self.fwd.append(lambda k=k: self.seg_wait(k) if self.seg_wait is not None else None)
            if id(u) in shortcut_of:
                self.fwd.append(_Fork())
                n0 = len(self.fwd)
                self._finalize_unit_fwd(shortcut_of[id(u)])
                self._branch_ids |= {id(f) for f in self.fwd[n0:]}
            self._finalize_unit_fwd(u, join_before_bn=id(u) in joins_before_bn)
        self.fwd.append(lambda: self.seg_wait(4) if self.seg_wait is not None else None)

        # head
        HW = self.featH * self.featH
        self.pooled = torch.zeros(N, self.F, device=dev)
        self.dlogits = torch.zeros(N, 128, device=dev)
        self.loss_n = torch.zeros(N, device=dev)
        self.correct_n = torch.zeros(N, dtype=torch.int32, device=dev)
        self.out = torch.zeros(2, device=dev)             # [mean loss, #correct]
        self.d_feat = self._bf(N, self.featH, self.featH, self.F)
        fcw, fcb = self._w("fc.weight"), self._sm("fc.bias")
        gw, gb = self._g("fc.weight"), self._g("fc.bias")

        def head(training: bool):
            L.check(self.lib.dps_head_forward_backward(
                _p(self.feat), _p(fcw), _p(fcb), _p(self.labels), _p(self.pooled), _p(self.dlogits), _p(self.loss_n),
                _p(self.correct_n), _p(self.d_feat), _p(gw), _p(gb), _p(self.out), N, HW, self.F, self.K,
                int(training), self._stream()), "head")
        self.fwd.append(lambda: head(True))
        self.evalf.append(lambda: head(False))

        # ---------------- backward schedule ----------------
        # Every dgrad epilogue also performs the backward *reduction* of the BatchNorm that consumes its output
        # (masked gradient dz + per-channel sums), so most BNs only need the apply pass.  Stand-alone reductions
        # remain for the last block's bn2 (fed by the head) and the three projection-shortcut BNs.
        self.n_bn_full = self.n_bn_apply = 0

        def bn_full(bn, dy, y, xx, dz, dx, relu):
            self.bwd.append(lambda: self._bn_bwd(bn, dy, y, xx, dz, dx, relu))
            self.n_bn_full += 1

        def bn_apply(bn, dz, xx, dx):
            self.bwd.append(lambda: self._bn_bwd_apply(bn, dz, xx, dx))
            self.n_bn_apply += 1

        def fuse(plan, unit):          # unit: the conv+BN unit whose BN consumes this dgrad's output
            b = unit["bn"]
            plan.fuse_bn_bwd(unit["a"], unit["c"], b.mean, b.rstd, b.bsums)
            if unit["residual"] is None and unit["relu"] and os.environ.get("DPS_MASK_X", "1") != "0":
                # bn1 of a block / the stem: the mask is recomputed from x, y is not read
                plan.mask_from_x(b.gamma, b.beta)
            return plan

        self._side_plans = set()                 # weight-gradient plans: leaves of the backward graph
        nseg = len(self.segment_boundaries) - 1
        self.bwd.append(_GroupDone(self, nseg - 1))             # fc gradients come from the head kernel
        g_out, g_out_fused = self.d_feat, False
        for bi in range(len(blocks) - 1, -1, -1):
            blk = blocks[bi]
            u1, u2, us = blk["u1"], blk["u2"], blk["us"]
            Cin, Cout, H, Ho, stride = blk["Cin"], blk["Cout"], blk["H"], blk["Ho"], blk["stride"]
            prev_unit = blocks[bi - 1]["u2"] if bi > 0 else stem          # producer of this block's input
            g_c2 = self._bf(N, Ho, Ho, Cout)
            g_a1 = self._bf(N, Ho, Ho, Cout)
            g_c1 = self._bf(N, Ho, Ho, Cout)
            g_x = self._bf(N, H, H, Cin)
            dz2 = g_out                                        # masked gradient lives in g_out
            if g_out_fused:
                bn_apply(u2["bn"], g_out, u2["c"], g_c2)
            else:
                bn_full(u2["bn"], g_out, u2["a"], u2["c"], g_out, g_c2, True)
            if us is not None:
                # projection shortcut: BN backward, weight gradient and the half-resolution input gradient depend only
                # on the masked block gradient dz2 -> branch stream, joined in front of conv1's stride-2 dgrad
                g_cs = self._bf(N, Ho, Ho, Cout)
                g_s = self._bf(N, Ho, Ho, Cin)
                self.bwd.append(_Fork())
                n0 = len(self.bwd)
                bn_full(us["bn"], dz2, None, us["c"], None, g_cs, False)
                kws = dict(N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=1, stride=stride)
                p_ws = ConvPlan.wgrad(blk["x"], g_cs, self._g(us["wname"]), **kws)
                p_ds = ConvPlan.dgrad(g_cs, self._w(us["wname"]), g_s, N=N, H=Ho, W=Ho, Cin=Cin, Cout=Cout, R=1,
                                      stride=1)
                self.bwd += [p_ws.run, p_ds.run]
                self._branch_ids |= {id(f) for f in self.bwd[n0:]}
                self._keep += [p_ws, p_ds]
            kw2 = dict(N=N, H=Ho, W=Ho, Cin=Cout, Cout=Cout, R=3, stride=1)
            p_w2 = ConvPlan.wgrad(u1["a"], g_c2, self._g(u2["wname"]), **kw2)
            p_d2 = fuse(ConvPlan.dgrad(g_c2, self._w(u2["wname"]), g_a1, **kw2), u1)
            self.bwd += [p_w2.run, p_d2.run]
            self._side_plans |= {id(p_w2)}
            bn_apply(u1["bn"], g_a1, u1["c"], g_c1)
            kw1 = dict(N=N, H=H, W=H, Cin=Cin, Cout=Cout, R=3, stride=stride)
            p_w1 = ConvPlan.wgrad(blk["x"], g_c1, self._g(u1["wname"]), **kw1)
            self.bwd.append(p_w1.run)
            self._side_plans |= {id(p_w1)}
            self._keep += [p_w2, p_d2, p_w1]
            if us is None:
                p_d1 = fuse(ConvPlan.dgrad(g_c1, self._w(u1["wname"]), g_x, addend=dz2, **kw1), prev_unit)
                self.bwd.append(p_d1.run)
                self._keep.append(p_d1)
            else:
                # the shortcut's gradient (branch above) is at half resolution: the stride-2 dgrad of conv1 adds it at
                # the pixels of parity (0,0)
                p_d1 = ConvPlan.dgrad(g_c1, self._w(u1["wname"]), g_x, addend=g_s, **kw1).addend_compact(True)
                fuse(p_d1, prev_unit)
                self.bwd += [_Join(), p_d1.run]
                self._keep += [p_d1]
            g_out, g_out_fused = g_x, True
            if bi > 0 and bi % BLOCKS_PER_STAGE == 0:            # first block of layer (bi/2 + 1): group complete
                self.bwd.append(_GroupDone(self, bi // BLOCKS_PER_STAGE))
        # stem
        g_c0 = self._bf(N, S, S, STAGE_WIDTHS[0])
        bn_apply(stem["bn"], g_out, stem["c"], g_c0)
        p_w0 = ConvPlan.wgrad(self.col.view(N, S, S, 64), g_c0, self._g("conv1.weight"), N=N, H=S, W=S, Cin=64,
                              Cout=STAGE_WIDTHS[0], R=1, stride=1, valid_cols=STEM_PITCH, dw_pitch=STEM_PITCH)
        self.bwd.append(p_w0.run)
        self._side_plans |= {id(p_w0)}
        self._keep.append(p_w0)
        self.bwd.append(_GroupDone(self, 0))
        # launches per step (bn_backward is two kernels)
        self.kernels_per_step = self._count_kernels()

    def _count_kernels(self) -> int:
        n_bn = len(self.bns)
        n_conv = len(self.units)
        fwd = 1 + n_conv + n_bn + 1                          # im2col (+ scratch clear), convs, bn, head
        n_dgrad = n_conv - 1
        bwd = 2 * self.n_bn_full + self.n_bn_apply + n_conv + n_dgrad     # (group markers launch nothing themselves)
        return fwd + bwd

    # ------------------------------------------------------------------ execution
    def load_batch(self, images: torch.Tensor, labels: torch.Tensor) -> None:
        self.images.copy_(images, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)

    def _on_side(self, f) -> bool:
        return isinstance(f, _GroupDone) or id(getattr(f, "__self__", None)) in self._side_plans

    def _run(self, sched) -> None:
        """Run a schedule.  Without side streams everything goes to the current stream in list order (a valid serial
        order).  With them: entries of a _Fork.._Join section that belong to the branch run on `branch_stream`
        (projection shortcuts, beside the block's main path); weight gradients and the per-group push hooks run on
        `side_stream`, forked right after the kernel that produced their input and joined at the end -- nothing in
        the dgrad / BatchNorm chain depends on them, so they fill the bubbles between the kernels of that chain."""
        W, X = self.side_stream, self.branch_stream
        cur = torch.cuda.current_stream()
        forked = used_w = False
        for f in sched:
            if isinstance(f, _Fork):
                if X is not None:
                    X.wait_stream(cur)
                    forked = True
            elif isinstance(f, _Join):
                if forked:
                    cur.wait_stream(X)
                    forked = False
            elif X is not None and id(f) in self._branch_ids:
                if W is not None and self._on_side(f):           # the shortcut's weight gradient: input comes from X
                    W.wait_stream(X)
                    used_w = True
                    with torch.cuda.stream(W):
                        f()
                else:
                    with torch.cuda.stream(X):
                        f()
            elif W is not None and self._on_side(f):
                W.wait_stream(cur)
                used_w = True
                with torch.cuda.stream(W):
                    f()
            else:
                f()
        if used_w:                       # (a stream that took no work must not be joined inside a graph capture)
            cur.wait_stream(W)

    def backward(self) -> None:
        self._run(self.bwd)

    def forward_backward(self) -> None:
        self._run(self.fwd)
        self._run(self.bwd)

    def forward_eval(self) -> None:
        for f in self.evalf:
            f()

    def capture(self, extra: Optional[Callable[[], None]] = None, warmup: int = 1,
                pre: Optional[Callable[[], None]] = None, inputs=None,
                post: Optional[Callable[[int], None]] = None) -> None:
        """Capture [pre] + forward + backward + [extra] (e.g. pull fork / push + join) into a CUDA graph.

        `inputs` = [(images, labels), ...]: capture ONE GRAPH PER INPUT BUFFER PAIR (`self.graphs[k]` reads pair k;
        `post(k)` is captured at the end of graph k, e.g. the D2H copy of the loss).  Replaying graph k while the next
        batch is copied into pair k^1 removes the device-to-device hand-over of the batch from the step."""
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            saved, self.seg_wait = self.seg_wait, None       # the warm-up pass is compute only: no pull waits ...
            hook, self.on_group_done = self.on_group_done, None      # ... and no SEND of the overlapped push (it would
            for _ in range(warmup):                                  # leave the arrival counters ahead of the rounds)
                self.forward_backward()
            self.seg_wait, self.on_group_done = saved, hook
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.dev)
        own = (self.images, self.labels)
        self.graphs = []
        for k, (x, y) in enumerate(inputs if inputs else [own]):
            self.images, self.labels = x, y                  # the schedule reads these attributes when it launches
            g = torch.cuda.CUDAGraph()
            # 'relaxed': other threads (the loader's pinned-memory allocations) may call the runtime while this
            # thread captures; nothing inside the capture allocates.
            with torch.cuda.graph(g, capture_error_mode="relaxed"):
                if pre is not None:
                    pre()
                self.forward_backward()
                if extra is not None:
                    extra()
                if post is not None:
                    post(k)
            self.graphs.append(g)
        self.images, self.labels = own                       # eager / evaluation paths keep the engine's own buffers
        self.graph = self.graphs[0]

    @torch.no_grad()
    def eval_batch(self, images: torch.Tensor, labels: torch.Tensor) -> int:
        """Forward in eval mode (running BN statistics); returns the number of correct top-1 predictions.
        Labels < 0 mark padding rows and never count as correct."""
        self.load_batch(images, labels)
        self.forward_eval()
        return int(self.out[1].item())

    def step(self) -> None:
        if self.graph is not None:
            self.graph.replay()
        else:
            self.forward_backward()

    # ------------------------------------------------------------------ state
    def bn_buffers(self) -> Dict[str, torch.Tensor]:
        out = {}
        for bn in self.bns:
            out[bn.name + ".running_mean"] = bn.running_mean
            out[bn.name + ".running_var"] = bn.running_var
            out[bn.name + ".num_batches_tracked"] = bn.num_batches
        return out

    def load_bn_buffers(self, sd) -> None:
        for k, v in self.bn_buffers().items():
            if k in sd:
                v.copy_(torch.as_tensor(sd[k]).to(v.device, v.dtype))

    def activation_bytes(self) -> int:
        return sum(u["c"].numel() * 2 + u["a"].numel() * 2 for u in self.units)
