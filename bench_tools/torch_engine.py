"""BASELINE (not product code): ATen/cuDNN execution of the same network over the same arenas, injected with
``Job(cfg, engine_factory=TorchEngine)``.  It is the apples-to-apples compute baseline for the
native engine (same Push/Pull kernels, library convolutions instead of the tcgen05 ones).  Parameters
are fp32 views refreshed from the pulled bf16 arena; ``.grad`` tensors are strided views INTO the flat
fp32 gradient arena, so autograd accumulates straight into what the Push kernel sends.
"""
from __future__ import annotations

import copy
from typing import Dict, Optional

import torch
from torch import nn

from dps_b200.store.shardmap import Layout


class TorchEngine:
    def __init__(self, layout: Layout, wbf16: torch.Tensor, small: torch.Tensor, grad: torch.Tensor,
                 model_cpu: nn.Module, batch: int, image_size: int = 32, device: Optional[torch.device] = None,
                 autocast: bool = True):
        self.layout, self.wbf16, self.small, self.grad = layout, wbf16, small, grad
        self.dev = device or wbf16.device
        self.model = copy.deepcopy(model_cpu).to(self.dev).to(memory_format=torch.channels_last)
        self.autocast = autocast
        self.wf32 = torch.zeros(layout.total, dtype=torch.float32, device=self.dev)
        self.params: Dict[str, nn.Parameter] = dict(self.model.named_parameters())
        for n, p in self.params.items():
            p.data = layout.view(self.wf32, n)
            p.grad = layout.view(self.grad, n)
        self.images = torch.zeros(batch, 3, image_size, image_size, device=self.dev)
        self.labels = torch.zeros(batch, dtype=torch.int64, device=self.dev)
        self.out = torch.zeros(2, device=self.dev)
        self.graph = None
        self.kernels_per_step = 0

    def _refresh(self) -> None:
        self.wf32.copy_(self.wbf16)
        self.wf32[: self.layout.n_small].copy_(self.small)

    def load_batch(self, images, labels) -> None:
        self.images.copy_(images, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)

    def forward_backward(self) -> None:
        self._refresh()
        self.model.train()
        x = self.images.contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.autocast):
            logits = self.model(x)
        loss = nn.functional.cross_entropy(logits.float(), self.labels)
        loss.backward()
        self.out[0] = loss.detach()
        self.out[1] = (logits.argmax(1) == self.labels).sum()

    def capture(self, extra=None, warmup: int = 1) -> None:      # no graph on this path
        self.graph = None

    @torch.no_grad()
    def eval_batch(self, images, labels) -> int:
        self.load_batch(images, labels)
        self._refresh()
        self.model.eval()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.autocast):
            logits = self.model(self.images.contiguous(memory_format=torch.channels_last))
        return int((logits.argmax(1) == self.labels).sum())

    def bn_buffers(self) -> Dict[str, torch.Tensor]:
        return dict(self.model.named_buffers())

    def load_bn_buffers(self, sd) -> None:
        for k, v in self.bn_buffers().items():
            if k in sd:
                v.copy_(torch.as_tensor(sd[k]).to(v.device, v.dtype))
