"""Drop-in for the entropy gate of fall_2025/Initial_Experiments.py: compute_shannon_entropy (:95-113, input already
in [0,1]) and the per-image decision of its evaluation loop (:147-157: denormalise, clamp, entropy > threshold)."""
from __future__ import annotations

import torch

from . import batch as B

ENTROPY_THRESHOLD = 7.46
CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)


def compute_shannon_entropy(image_tensor: torch.Tensor) -> float:
    """compute_shannon_entropy(image_tensor): (C,H,W) tensor in [0,1] -> Shannon entropy (bits) as a float."""
    t = image_tensor if image_tensor.is_cuda else image_tensor.cuda()
    return float(B.shannon_entropy(t.unsqueeze(0))[0])


def shannon_entropy_batch(images: torch.Tensor, normalised: bool = True) -> torch.Tensor:
    """(n,3,h,w) f32 CUDA tensor -> f64[n]; normalised=True applies the loop's denormalise + clamp first."""
    return B.shannon_entropy(images, CIFAR_MEAN, CIFAR_STD) if normalised else B.shannon_entropy(images)


def adaptive_mask(images: torch.Tensor, threshold: float = ENTROPY_THRESHOLD) -> torch.Tensor:
    """Which images of a normalised batch get the adaptive transforms (:157)."""
    return shannon_entropy_batch(images) > threshold
