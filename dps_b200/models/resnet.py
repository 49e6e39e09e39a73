"""CIFAR-style ResNet-18 as a plain ``torch.nn.Module``.

This is the *portable* definition of the model: it runs on CPU (gRPC path,
oracle for kernel tests) and on CUDA through ATen/cuDNN (``engine='torch'``,
the DDP baseline).  The hot path on B200 does not use it -- see
``models/engine.py`` for the sm_100a kernel engine -- but both share the exact
parameter names so checkpoints and the parameter store are interchangeable.

Architecture (reference: src/parameter_server/server.py:21-76, identical copies in
src/workers/worker.py:21-76 and baseline/baseline_training.py:37-95):
3x3 stem (3->64, stride 1, no bias) + BN + ReLU, **no max-pool**, four stages of two
residual blocks with widths 64/128/256/512 and first-block strides 1/2/2/2, global
average pool, Linear(512, num_classes).  ``state_dict()`` has 122 entries
(62 trainable tensors = 11 220 132 parameters, 60 BN buffers).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
from torch import nn

STAGE_WIDTHS: Tuple[int, ...] = (64, 128, 256, 512)
STAGE_STRIDES: Tuple[int, ...] = (1, 2, 2, 2)
BLOCKS_PER_STAGE = 2


def _conv3x3(cin: int, cout: int, stride: int) -> nn.Conv2d:
    return nn.Conv2d(cin, cout, kernel_size=3, stride=stride, padding=1, bias=False)


class BasicBlock(nn.Module):
    """Two 3x3 conv+BN; projection shortcut (1x1 conv + BN) when shape changes."""

    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        self.conv1 = _conv3x3(cin, cout, stride)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = _conv3x3(cout, cout, 1)
        self.bn2 = nn.BatchNorm2d(cout)
        proj: List[nn.Module] = []
        if stride != 1 or cin != cout:
            proj = [nn.Conv2d(cin, cout, kernel_size=1, stride=stride, bias=False),
                    nn.BatchNorm2d(cout)]
        self.shortcut = nn.Sequential(*proj)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = torch.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return torch.relu(y + self.shortcut(x))


class ResNet18(nn.Module):
    def __init__(self, num_classes: int = 100,
                 widths: Sequence[int] = STAGE_WIDTHS,
                 strides: Sequence[int] = STAGE_STRIDES):
        super().__init__()
        self.conv1 = _conv3x3(3, widths[0], 1)
        self.bn1 = nn.BatchNorm2d(widths[0])
        cin = widths[0]
        for i, (w, s) in enumerate(zip(widths, strides), start=1):
            blocks = []
            for b in range(BLOCKS_PER_STAGE):
                blocks.append(BasicBlock(cin, w, s if b == 0 else 1))
                cin = w
            setattr(self, f"layer{i}", nn.Sequential(*blocks))
        self.avg_pool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(cin, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = torch.relu(self.bn1(self.conv1(x)))
        for i in range(1, 5):
            x = getattr(self, f"layer{i}")(x)
        return self.fc(torch.flatten(self.avg_pool(x), 1))


def build_model(num_classes: int = 100, seed: int | None = None) -> ResNet18:
    """Seedable construction (the reference has no seed anywhere)."""
    if seed is None:
        return ResNet18(num_classes)
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        return ResNet18(num_classes)


def trainable_names(model: nn.Module) -> List[str]:
    return [n for n, _ in model.named_parameters()]


def buffer_names(model: nn.Module) -> List[str]:
    return [n for n, _ in model.named_buffers()]
