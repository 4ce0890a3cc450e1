"""Drop-in for the dataset -> tensor stage of fall_2025/TENT_transform.py (:11-31, :56-59) and
fall_2025/TTA_transforms.py:10-36: severity-major slicing of a CIFAR-C .npy file and ToTensor+Normalize, batched."""
from __future__ import annotations

import os

import numpy as np
import torch

from . import batch as B

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)


class _Compose:
    """Stands in for T.Compose([T.ToTensor(), T.Normalize(mean, std)]): PIL / HWC uint8 -> CHW f32 tensor."""

    def __init__(self, mean=CIFAR_MEAN, std=CIFAR_STD):
        self.mean, self.std = mean, std

    def __call__(self, img):
        from ._pil import to_dev
        return B.to_tensor(to_dev(img), self.mean, self.std)[0]

    def batch(self, x_u8: torch.Tensor) -> torch.Tensor:
        return B.to_tensor(x_u8, self.mean, self.std)


transform = _Compose()


class CIFARC_Dataset(torch.utils.data.Dataset):
    """CIFARC_Dataset(root, corruption, severity, transform=None): data[(s-1)*10000 : s*10000]."""

    def __init__(self, root, corruption, severity, transform=None):
        self.transform = transform
        all_images = np.load(os.path.join(root, f"{corruption}.npy"), mmap_mode="r")
        all_labels = np.load(os.path.join(root, "labels.npy"))
        start_idx, end_idx = (severity - 1) * 10000, severity * 10000
        self.data = np.ascontiguousarray(all_images[start_idx:end_idx])
        self.labels = all_labels[start_idx:end_idx]

    def __len__(self):
        return len(self.data)

    def __getitem__(self, idx):
        img, label = self.data[idx], self.labels[idx]
        if self.transform is not None:
            return self.transform(img), label
        from PIL import Image
        return Image.fromarray(img), label

    def device_batches(self, batch_size=200, transform=transform):
        """Whole-severity fast path: pinned host staging -> device -> one ToTensor+Normalize launch per batch."""
        pinned = torch.from_numpy(self.data).pin_memory()
        for i in range(0, len(self.data), batch_size):
            xb = pinned[i:i + batch_size].cuda(non_blocking=True)
            yield transform.batch(xb), torch.as_tensor(self.labels[i:i + batch_size])
