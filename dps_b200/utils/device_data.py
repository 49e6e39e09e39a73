"""Device-resident dataset + loader: the whole input pipeline of a step is ONE kernel launch on the GPU.

Role: the reference's per-worker `Subset` + `DataLoader(shuffle=True, num_workers=2)` with the torchvision transform
stack (reference: src/workers/worker.py:140-197).  B200 design: the uint8 dataset (150 MB for CIFAR-100) is uploaded
once and stays in HBM; every batch is produced by `dps_augment_batch` (csrc/data.cu: gather by shuffled index +
RandomCrop(pad 4) + RandomHorizontalFlip + ToTensor + Normalize) straight into device memory, so a step needs no host
transform work and no host->device copy.  Same partition semantics as the host loader (contiguous block per worker,
the last worker takes the remainder; full test set everywhere).

The crop/flip randomness is counter-based (see the kernel); `host_reference_batch` restates it on the CPU for tests.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

from ..ops import _lib as L
from .data import CIFAR100_MEAN, CIFAR100_STD, ArrayDataset

_M64 = (1 << 64) - 1
_declared = False


def _declare() -> None:
    global _declared
    if not _declared:
        P = C.c_void_p
        L.declare("dps_augment_batch", C.c_int, [P, P, P, P, P, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64,
                                                C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), P])
        _declared = True


def splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _M64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def aug_params(seed: int, counter: int, pad: int = 4) -> Tuple[int, int, int]:
    """(dy, dx, flip) of the sample with global counter `counter` -- same arithmetic as the kernel."""
    z = splitmix64((seed * 0x100000001B3 + counter) & _M64)
    return z % (2 * pad + 1), (z >> 16) % (2 * pad + 1), (z >> 40) & 1


def host_reference_batch(images_u8: np.ndarray, labels: np.ndarray, idx, seed: int, counter0: int, train: bool,
                         pad: int = 4) -> Tuple[torch.Tensor, torch.Tensor]:
    """Plain fp32 restatement of dps_augment_batch (pad -> crop -> flip -> /255 -> normalise)."""
    mean = torch.tensor(CIFAR100_MEAN).view(3, 1, 1)
    std = torch.tensor(CIFAR100_STD).view(3, 1, 1)
    out = []
    for b, s in enumerate(idx):
        x = torch.from_numpy(images_u8[int(s)]).permute(2, 0, 1).float() / 255.0
        S = x.shape[-1]
        if train:
            dy, dx, flip = aug_params(seed, counter0 + b, pad)
            xp = torch.nn.functional.pad(x, (pad, pad, pad, pad))
            x = xp[:, dy:dy + S, dx:dx + S]
            if flip:
                x = x.flip(-1)
        out.append((x - mean) / std)
    return torch.stack(out), torch.from_numpy(labels[np.asarray(idx, dtype=np.int64)])


class DeviceLoader:
    """Iterates (images fp32 [B,3,S,S], labels int64 [B]) DEVICE tensors over samples [start, end) of `ds`.

    The yielded tensors are slots of a small ring: consume a batch (e.g. `Worker.train_step`) before asking for the
    fourth one after it.  Every batch has exactly `batch_size` rows while training (the ragged tail is dropped, as the
    static-shape engine requires); in eval mode the tail batch is returned short."""

    RING = 4

    def __init__(self, ds: ArrayDataset, start: int, end: int, batch_size: int, *, train: bool, seed: int = 0,
                 device: Optional[torch.device] = None, pad: int = 4, resident=None):
        _declare()
        self.device = torch.device(device if device is not None else "cuda")
        self.start, self.end, self.batch_size, self.train, self.seed, self.pad = start, end, batch_size, train, seed, pad
        if resident is None:
            resident = (torch.from_numpy(np.ascontiguousarray(ds.images)).to(self.device),
                        torch.from_numpy(np.ascontiguousarray(ds.labels).astype(np.int64)).to(self.device))
        self.images_u8, self.labels_dev = resident
        n, S = self.images_u8.shape[0], self.images_u8.shape[1]
        assert self.images_u8.dtype == torch.uint8 and self.images_u8.shape[1:] == (S, S, 3) and end <= n
        self.S = S
        self.gen = torch.Generator(device="cpu").manual_seed(seed)
        self._x = [torch.empty(batch_size, 3, S, S, dtype=torch.float32, device=self.device) for _ in range(self.RING)]
        self._y = [torch.empty(batch_size, dtype=torch.int64, device=self.device) for _ in range(self.RING)]
        self._mean = (C.c_float * 3)(*CIFAR100_MEAN)
        self._std = (C.c_float * 3)(*CIFAR100_STD)
        self._slot = 0
        self.counter = 0                     # global sample counter: advances over epochs, never repeats
        self.last_order: Optional[torch.Tensor] = None

    @property
    def resident(self):
        return self.images_u8, self.labels_dev

    def __len__(self) -> int:
        n = self.end - self.start
        return n // self.batch_size if self.train else (n + self.batch_size - 1) // self.batch_size

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        n = self.end - self.start
        if n <= 0:
            return
        order = (torch.randperm(n, generator=self.gen) if self.train else torch.arange(n)) + self.start
        self.last_order = order
        order_dev = order.to(self.device, non_blocking=False)
        lib = L.lib()
        B = self.batch_size
        for i in range(len(self)):
            k = self._slot % self.RING
            self._slot += 1
            valid = min(B, n - i * B)
            L.check(lib.dps_augment_batch(self.images_u8.data_ptr(), self.labels_dev.data_ptr(),
                                          order_dev.data_ptr() + 8 * i * B, self._x[k].data_ptr(),
                                          self._y[k].data_ptr(), B, self.S, valid, self.seed, self.counter,
                                          1 if self.train else 0, self.pad, self._mean, self._std, L.stream_ptr()),
                    "dps_augment_batch")
            self.counter += B
            if valid == B:
                yield self._x[k], self._y[k]
            else:
                yield self._x[k][:valid], self._y[k][:valid]
