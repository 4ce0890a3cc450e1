"""Drop-in for the image-entropy gate of fall_2025/TENT_Pipelines.py: compute_shannon_entropy (:99-121) and the
per-batch gate of evaluate_with_tent_conditional (:257-274).  TENT itself (model adaptation) is out of scope."""
from __future__ import annotations

import torch

from . import batch as B

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)


def shannon_entropy_batch(images: torch.Tensor) -> torch.Tensor:
    """(n,3,h,w) normalised f32 CUDA tensor -> f64[n] entropies, one launch, no host round trip."""
    return B.shannon_entropy(images, CIFAR_MEAN, CIFAR_STD)


def compute_shannon_entropy(image_tensor: torch.Tensor) -> float:
    """compute_shannon_entropy(image_tensor): (C,H,W) normalised tensor -> Shannon entropy (bits) as a float."""
    t = image_tensor if image_tensor.is_cuda else image_tensor.cuda()
    return float(shannon_entropy_batch(t.unsqueeze(0))[0])


def high_entropy_gate(images: torch.Tensor, entropy_threshold: float = 7.16):
    """The decision of evaluate_with_tent_conditional (:262-274) for one batch: (mask of images whose entropy exceeds
    the threshold, number of them, adapt = more than half of the batch is high-entropy)."""
    mask = shannon_entropy_batch(images) > entropy_threshold
    num = int(mask.sum())
    return mask, num, num > images.size(0) / 2
