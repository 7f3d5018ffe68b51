"""Synthetic stand-in for torchvision.datasets.CIFAR10 used by the reference baseline harness
(importable at module level so DataLoader workers can unpickle it).  Random uint8 [n, hw, hw, 3]
images -> PIL -> the reference's OWN transforms (ref:data.py:11-19)."""
import os

import numpy as np
import torch.utils.data
from PIL import Image


class FakeCIFAR10(torch.utils.data.Dataset):
    def __init__(self, root=None, train=True, transform=None, target_transform=None, download=False):
        n = int(os.environ.get('REF_TRAIN_SIZE' if train else 'REF_TEST_SIZE', '512' if train else '256'))
        hw = int(os.environ.get('REF_IMAGE_SIZE', '32'))
        rng = np.random.default_rng(0 if train else 1)
        self.data = rng.integers(0, 256, size=(n, hw, hw, 3), dtype=np.uint8)
        self.targets = rng.integers(0, 10, size=n).tolist()
        self.transform = transform

    def __len__(self):
        return len(self.targets)

    def __getitem__(self, i):
        img = Image.fromarray(self.data[i])
        if self.transform is not None:
            img = self.transform(img)
        return img, self.targets[i]
