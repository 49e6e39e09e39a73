"""Data: per-worker disjoint contiguous partitions, synthetic CIFAR-100-shaped source,
optional real CIFAR-100, batched augmentation, evaluation helper.

Partitioning follows the reference exactly (reference: src/workers/worker.py:166-179):
worker i takes ``[i*floor(N/W), (i+1)*floor(N/W))`` and the last worker also takes the
remainder.  Transforms match worker.py:145-155 (RandomCrop(32, pad 4), horizontal
flip, normalise with the CIFAR-100 mean/std) but are applied to whole batches as
tensor ops instead of per-sample PIL transforms in DataLoader workers.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterator, Optional, Tuple

import os
import time

import numpy as np
import torch

CIFAR100_MEAN = (0.5071, 0.4867, 0.4408)
CIFAR100_STD = (0.2675, 0.2565, 0.2761)


def partition_range(worker_id: int, total_workers: int, n_samples: int) -> Tuple[int, int]:
    """[start, end) of worker ``worker_id``; ids >= W get an empty range instead of the
    reference's IndexError."""
    if worker_id < 0 or worker_id >= total_workers:
        return (n_samples, n_samples)
    per = n_samples // total_workers
    start = worker_id * per
    end = n_samples if worker_id == total_workers - 1 else start + per
    return (start, end)


@dataclass
class ArrayDataset:
    """uint8 NHWC images + int64 labels held in host memory."""
    images: np.ndarray   # [N, H, W, 3] uint8
    labels: np.ndarray   # [N] int64

    def __len__(self) -> int:
        return int(self.labels.shape[0])


def synthetic_dataset(n: int, num_classes: int = 100, image_size: int = 32, seed: int = 0,
                      learnable: bool = True) -> ArrayDataset:
    """Deterministic CIFAR-100-shaped data.  With ``learnable`` each class has a fixed
    low-frequency template plus noise, so accuracy moves when training works.

    Large sets are generated once per machine and cached as .npz (``DPS_DATA_CACHE`` = directory, or ``0`` to
    disable): with one process per GPU, eight ranks generating the same 150 MB at start-up was the slowest part of a
    short job.  One rank generates (exclusive lock file), the others wait for the file."""
    cache = os.environ.get("DPS_DATA_CACHE", os.path.join(
        os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "build", "data_cache"))
    if n < 10000 or cache == "0":
        return _synthetic_dataset(n, num_classes, image_size, seed, learnable)
    path = os.path.join(cache, f"synth_n{n}_k{num_classes}_s{image_size}_seed{seed}_{int(learnable)}.npz")
    lock = path + ".lock"
    try:
        os.makedirs(cache, exist_ok=True)
        deadline = time.time() + 120.0
        while True:
            if os.path.exists(path):
                with np.load(path) as z:
                    return ArrayDataset(z["images"], z["labels"])
            try:
                fd = os.open(lock, os.O_CREAT | os.O_EXCL | os.O_WRONLY)
            except FileExistsError:
                if time.time() > deadline:
                    break                              # the generating rank died: fall through and generate locally
                time.sleep(0.05)
                continue
            try:
                ds = _synthetic_dataset(n, num_classes, image_size, seed, learnable)
                tmp = f"{path}.{os.getpid()}.tmp.npz"
                np.savez(tmp, images=ds.images, labels=ds.labels)
                os.replace(tmp, path)
                return ds
            finally:
                os.close(fd)
                os.unlink(lock)
    except OSError:
        pass
    return _synthetic_dataset(n, num_classes, image_size, seed, learnable)


def _synthetic_dataset(n: int, num_classes: int, image_size: int, seed: int, learnable: bool) -> ArrayDataset:
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, num_classes, size=n, dtype=np.int64)
    if learnable:
        trng = np.random.default_rng(12345)     # templates do not depend on the split seed
        coarse = trng.normal(0.0, 1.0, size=(num_classes, 4, 4, 3)).astype(np.float32)
        rep = image_size // 4
        templates = np.repeat(np.repeat(coarse, rep, axis=1), rep, axis=2)
        templates = templates[:, :image_size, :image_size]
        noise = rng.normal(0.0, 0.5, size=(n, image_size, image_size, 3)).astype(np.float32)
        x = templates[labels] * 0.5 + noise
        images = np.clip(x * 48.0 + 120.0, 0, 255).astype(np.uint8)
    else:
        images = rng.integers(0, 256, size=(n, image_size, image_size, 3), dtype=np.uint8)
    return ArrayDataset(images, labels)


def cifar100_dataset(root: str, train: bool) -> ArrayDataset:
    """Real CIFAR-100 through torchvision if the files are already on disk (no network)."""
    import torchvision
    ds = torchvision.datasets.CIFAR100(root=root, train=train, download=False)
    return ArrayDataset(np.asarray(ds.data, dtype=np.uint8), np.asarray(ds.targets, dtype=np.int64))


def load_dataset(kind: str, train: bool, n: int, num_classes: int, image_size: int, seed: int,
                 root: str = "./data") -> ArrayDataset:
    if kind == "cifar100":
        return cifar100_dataset(root, train)
    return synthetic_dataset(n, num_classes, image_size, seed=seed + (0 if train else 7919))


class BatchLoader:
    """Iterates a [start, end) slice of an ArrayDataset in shuffled mini-batches.

    Yields (images fp32 NCHW normalised, labels int64) host tensors; with
    ``pin=True`` they are written into reusable pinned buffers (the device path
    copies them with one async H2D per step).
    """

    def __init__(self, ds: ArrayDataset, start: int, end: int, batch_size: int, *, train: bool,
                 seed: int = 0, drop_last: bool = False, pin: bool = False, prefetch: int = 0):
        self.ds, self.start, self.end = ds, start, end
        self.prefetch = prefetch            # > 0: batches are produced by a background thread, this many ahead
        self.batch_size, self.train, self.drop_last = batch_size, train, drop_last
        self.gen = torch.Generator().manual_seed(seed)
        self.mean = torch.tensor(CIFAR100_MEAN).view(1, 3, 1, 1)
        self.std = torch.tensor(CIFAR100_STD).view(1, 3, 1, 1)
        self.pin = pin and torch.cuda.is_available()

    def __len__(self) -> int:
        n = self.end - self.start
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def _augment(self, x: torch.Tensor) -> torch.Tensor:
        # RandomCrop(size, padding=4) + RandomHorizontalFlip, whole batch at once
        b, _, h, w = x.shape
        padded = torch.nn.functional.pad(x, (4, 4, 4, 4))
        dy = torch.randint(0, 9, (b,), generator=self.gen)
        dx = torch.randint(0, 9, (b,), generator=self.gen)
        rows = (dy.view(b, 1) + torch.arange(h).view(1, h))            # [b,h]
        cols = (dx.view(b, 1) + torch.arange(w).view(1, w))            # [b,w]
        flip = torch.rand(b, generator=self.gen) < 0.5
        cols = torch.where(flip.view(b, 1), cols.flip(1), cols)
        bi = torch.arange(b).view(b, 1, 1)
        out = padded.permute(0, 2, 3, 1)[bi, rows.view(b, h, 1), cols.view(b, 1, w)]  # [b,h,w,c]
        return out.permute(0, 3, 1, 2)

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        if self.prefetch <= 0:
            yield from self._batches()
            return
        # The reference uses DataLoader(num_workers=2) worker processes (worker.py:181-187); one producer thread is
        # enough here: the tensor ops release the GIL and a step consumes a batch every ~1.5 ms.
        import queue
        import threading
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def produce():
            try:
                for b in self._batches():
                    while not stop.is_set():
                        try:
                            q.put(b, timeout=0.1)
                            break
                        except queue.Full:
                            continue
                    if stop.is_set():
                        return
                q.put(None)
            except BaseException as e:               # surface loader errors in the consumer
                q.put(e)

        th = threading.Thread(target=produce, daemon=True, name="dps-loader")
        th.start()
        try:
            while True:
                b = q.get()
                if b is None:
                    break
                if isinstance(b, BaseException):
                    raise b
                yield b
        finally:
            stop.set()

    def _batches(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        n = self.end - self.start
        if n <= 0:
            return
        order = (torch.randperm(n, generator=self.gen) if self.train else torch.arange(n)) + self.start
        for i in range(len(self)):
            idx = order[i * self.batch_size:(i + 1) * self.batch_size].numpy()
            x = torch.from_numpy(self.ds.images[idx]).permute(0, 3, 1, 2).float().div_(255.0)
            if self.train:
                x = self._augment(x)
            x = ((x - self.mean) / self.std).contiguous()
            y = torch.from_numpy(self.ds.labels[idx])
            if self.pin:
                x, y = x.pin_memory(), y.pin_memory()
            yield x, y


def make_loaders(cfg, worker_id: int, total_workers: int, pin: bool = False, device=None):
    """(train_loader over this worker's partition, test_loader over the full test set).
    Every worker sees the whole test set, as in the reference (worker.py:189-194);
    use ``evaluate(..., shard=(rank, world))`` to split the work."""
    train = load_dataset(cfg.dataset, True, cfg.train_samples, cfg.num_classes, cfg.image_size,
                         cfg.seed, cfg.data_root)
    test = load_dataset(cfg.dataset, False, cfg.test_samples, cfg.num_classes, cfg.image_size,
                        cfg.seed, cfg.data_root)
    s, e = partition_range(worker_id, total_workers, len(train))
    if device is not None and getattr(cfg, "data_on_device", True) and torch.device(device).type == "cuda":
        # B200 path: dataset resident in HBM, one augmentation kernel per batch, no host work per step
        from .device_data import DeviceLoader
        tl = DeviceLoader(train, s, e, cfg.batch_size, train=True, seed=cfg.seed * 1000 + worker_id, device=device)
        vl = DeviceLoader(test, 0, len(test), cfg.batch_size, train=False, seed=0, device=device)
        return tl, vl, (s, e)
    tl = BatchLoader(train, s, e, cfg.batch_size, train=True, seed=cfg.seed * 1000 + worker_id, pin=pin,
                     prefetch=4 if pin else 0)
    vl = BatchLoader(test, 0, len(test), cfg.batch_size, train=False, seed=0, pin=pin, prefetch=4 if pin else 0)
    return tl, vl, (s, e)


@torch.no_grad()
def evaluate(model: torch.nn.Module, loader: BatchLoader, device: Optional[torch.device] = None,
             max_batches: Optional[int] = None) -> float:
    """Top-1 accuracy in percent (reference: worker.py:313-331)."""
    was_training = model.training
    model.eval()
    correct = total = 0
    for i, (x, y) in enumerate(loader):
        if max_batches is not None and i >= max_batches:
            break
        if device is not None:
            x, y = x.to(device), y.to(device)
        pred = model(x).argmax(1)
        correct += int((pred == y).sum())
        total += int(y.numel())
    model.train(was_training)
    return 100.0 * correct / max(1, total)
