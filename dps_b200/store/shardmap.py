"""Arena layout and shard map of the parameter store.

The reference keeps parameters as a dict of 122 numpy arrays on one server (reference:
src/parameter_server/server.py:93-98).  Here the 62 trainable tensors live in ONE flat arena
(same layout for the fp32 master, the fp32 gradient arena and every worker's bf16 weight
arena) so that push/pull/optimizer are element-wise kernels over contiguous memory:

    [ small region : BN gamma/beta in forward order, fc.bias ]   (also delivered in fp32)
    [ conv weights in forward order, physical layout K,R,S,C  ]   (what the implicit-GEMM TMA maps read)
    [ fc.weight ]

* conv weights are stored K-R-S-C (channels-last); the state_dict view is a strided
  [K,C,R,S] tensor over the same memory, so checkpoints stay reference-compatible with no copy.
* the 3-channel stem weight rows are padded from 27 to 32 elements (TMA needs 16-byte row pitch).
* the arena is sharded block-cyclically: global block k (``block_elems``) -> owner k % W, local
  block k // W.  A pull therefore streams the arena front-to-back from all owners at once, which is
  what lets it overlap the next forward pass layer by layer.

BN running statistics (60 buffers) are worker-local state and are not part of the arena
(policy in Config.bn_buffer_policy).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Mapping, Tuple

import torch

BLOCK_ELEMS = 2048
TENSOR_ALIGN = 128
STEM_PITCH = 32


@dataclass
class ParamEntry:
    name: str
    offset: int
    shape: Tuple[int, ...]          # state_dict shape
    kind: str                        # 'vec' | 'conv' | 'stem' | 'fc'
    span: int                        # elements reserved in the arena (>= numel)

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    def strides(self) -> Tuple[int, ...]:
        if self.kind == "conv":
            k, c, r, s = self.shape
            return (r * s * c, 1, s * c, c)
        if self.kind == "stem":
            k, c, r, s = self.shape
            return (STEM_PITCH, 1, s * c, c)
        if self.kind == "fc":
            return (self.shape[1], 1)
        return (1,)


def _align(x: int, a: int) -> int:
    return (x + a - 1) // a * a


@dataclass
class Layout:
    entries: Dict[str, ParamEntry]
    order: List[str]
    n_small: int
    total: int                      # arena elements, multiple of world*block_elems
    world: int
    block_elems: int = BLOCK_ELEMS

    @property
    def shard_elems(self) -> int:
        return self.total // self.world

    # ---- views ------------------------------------------------------------------------
    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        e = self.entries[name]
        return flat.as_strided(e.shape, e.strides(), e.offset)

    def views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {n: self.view(flat, n) for n in self.order}

    def flatten_into(self, flat: torch.Tensor, tensors: Mapping[str, torch.Tensor]) -> None:
        with torch.no_grad():
            for n in self.order:
                if n in tensors:
                    self.view(flat, n).copy_(tensors[n].to(flat.dtype))

    def unflatten(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {n: self.view(flat, n).contiguous().clone() for n in self.order}

    # ---- shards -----------------------------------------------------------------------
    def to_shards(self, flat: torch.Tensor) -> torch.Tensor:
        """[total] -> [W, E] (owner-major) following the block-cyclic map."""
        nb = self.total // (self.world * self.block_elems)
        return flat.view(nb, self.world, self.block_elems).permute(1, 0, 2).reshape(self.world, -1)

    def from_shards(self, shards: torch.Tensor) -> torch.Tensor:
        nb = self.total // (self.world * self.block_elems)
        return shards.view(self.world, nb, self.block_elems).permute(1, 0, 2).reshape(-1)

    def owner_of(self, global_index: int) -> Tuple[int, int]:
        k, r = divmod(global_index, self.block_elems)
        return k % self.world, (k // self.world) * self.block_elems + r

    def describe(self) -> List[dict]:
        return [{"name": n, "offset": self.entries[n].offset, "shape": list(self.entries[n].shape),
                 "kind": self.entries[n].kind} for n in self.order]


def forward_ordered_names(model: torch.nn.Module) -> Tuple[List[str], List[str]]:
    """(small parameter names, big parameter names) in forward order."""
    small, big = [], []
    for name, p in model.named_parameters():
        (small if p.dim() == 1 else big).append(name)
    return small, big


def build_layout(model: torch.nn.Module, world: int, block_elems: int = BLOCK_ELEMS) -> Layout:
    small, big = forward_ordered_names(model)
    params = dict(model.named_parameters())
    entries: Dict[str, ParamEntry] = {}
    order: List[str] = []
    off = 0
    for n in small:
        p = params[n]
        span = _align(p.numel(), 64)
        entries[n] = ParamEntry(n, off, tuple(p.shape), "vec", span)
        order.append(n)
        off += span
    n_small = _align(off, block_elems)
    off = n_small
    for n in big:
        p = params[n]
        if p.dim() == 4:
            k, c, r, s = p.shape
            if c * r * s * 2 % 16 != 0:            # stem: 27-element rows -> padded pitch
                kind, span = "stem", k * STEM_PITCH
                assert c * r * s <= STEM_PITCH
            else:
                kind, span = "conv", p.numel()
        else:
            kind, span = "fc", p.numel()
        entries[n] = ParamEntry(n, off, tuple(p.shape), kind, _align(span, TENSOR_ALIGN))
        order.append(n)
        off += entries[n].span
    total = _align(off, world * block_elems)
    return Layout(entries, order, n_small, total, world, block_elems)
