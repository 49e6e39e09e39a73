"""Launcher that runs an UNMODIFIED reference script (server.py / worker.py from baseline/_ref) with
torchvision's CIFAR-100 replaced by an in-memory synthetic dataset of the same shape -- the reference
insists on download=True (reference: src/workers/worker.py:158-164) and there is no network.

    python bench_tools/ref_launch.py <train_samples> <test_samples> <script.py> [script args...]

Nothing under baseline/_ref is edited; only the dataset class the script imports is substituted.
"""
import runpy
import sys

import numpy as np
import torch
import torchvision
from PIL import Image


class SyntheticCIFAR100(torch.utils.data.Dataset):
    train_samples = 50000
    test_samples = 10000

    def __init__(self, root="./data", train=True, download=False, transform=None, target_transform=None):
        n = self.train_samples if train else self.test_samples
        rng = np.random.default_rng(0 if train else 1)
        self.data = rng.integers(0, 256, size=(n, 32, 32, 3), dtype=np.uint8)
        self.targets = rng.integers(0, 100, size=n).tolist()
        self.transform, self.target_transform = transform, target_transform

    def __len__(self):
        return len(self.targets)

    def __getitem__(self, i):
        img, tgt = Image.fromarray(self.data[i]), self.targets[i]
        if self.transform is not None:
            img = self.transform(img)
        if self.target_transform is not None:
            tgt = self.target_transform(tgt)
        return img, tgt


def main():
    SyntheticCIFAR100.train_samples = int(sys.argv[1])
    SyntheticCIFAR100.test_samples = int(sys.argv[2])
    script = sys.argv[3]
    torchvision.datasets.CIFAR100 = SyntheticCIFAR100
    sys.argv = [script] + sys.argv[4:]
    runpy.run_path(script, run_name="__main__")


if __name__ == "__main__":
    main()
