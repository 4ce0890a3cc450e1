"""Drop-in for fall_2025/TTA_transforms.py (:10-36, :52-165): CIFARC_Dataset, base_transform and the three 4-view
TTA strategies.  Each view callable maps a PIL image to a normalised (3,32,32) f32 CUDA tensor like the reference's
T.Compose pipelines (same torch RNG consumption); `strategy_views_batch` produces all four views of a whole batch in
one fused launch."""
from __future__ import annotations

import numpy as np
import torch
import torchvision.transforms as T

from . import batch as B
from . import views as V
from ._pil import to_dev
from .TENT_transform import CIFARC_Dataset  # noqa: F401  (same class in both reference files)

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)

corruptions = ['defocus_blur', 'glass_blur', 'motion_blur', 'zoom_blur', 'snow', 'frost', 'fog', 'brightness',
               'contrast', 'elastic_transform', 'pixelate', 'jpeg_compression']


class _View:
    """One T.Compose([... , ToTensor, Normalize]) of the reference, as a list of (kind, arg) steps."""

    def __init__(self, *steps):
        self.steps = steps

    def draw(self, w=32, h=32):
        """Consume the torch RNG exactly like the torchvision transforms and return the drawn parameters."""
        params = []
        for kind, arg in self.steps:
            if kind == "flip":
                torch.rand(1)  # RandomHorizontalFlip(p=1.0) still draws
                params.append(None)
            elif kind == "crop":
                i = int(torch.randint(0, h + 2 * arg - 32 + 1, size=(1,)).item())
                j = int(torch.randint(0, w + 2 * arg - 32 + 1, size=(1,)).item())
                params.append((i, j))
            elif kind in ("brightness", "contrast"):
                lo, hi = max(0.0, 1 - arg), 1 + arg
                torch.randperm(4)  # ColorJitter.get_params draws the op order first
                params.append(float(torch.empty(1).uniform_(lo, hi)))
            else:
                params.append(None)
        return params

    def apply_u8(self, x, params):
        for (kind, arg), p in zip(self.steps, params):
            if kind == "flip":
                x = B.hflip(x)
            elif kind == "crop":
                x = B.crop_reflect(x, p[0], p[1], arg, 32)
            elif kind == "resize_cc":
                x = B.resize_center_crop(x, arg, 32)
            elif kind == "brightness":
                x = B.brightness(x, p)
            elif kind == "contrast":
                x = B.enhance_contrast(x, p)
        return x

    def __call__(self, img):
        x = to_dev(img)
        params = self.draw(x.shape[2], x.shape[1])
        return B.to_tensor(self.apply_u8(x, params), CIFAR_MEAN, CIFAR_STD)[0]


base_transform = _View()
tta_transforms_gentle = [_View(), _View(("flip", None)), _View(("crop", 4)), _View(("flip", None), ("crop", 4))]
tta_transforms_multiscale = [_View(), _View(("resize_cc", 36)), _View(("resize_cc", 40)), _View(("flip", None))]
tta_transforms_color = [_View(), _View(("brightness", 0.2)), _View(("contrast", 0.2)), _View(("flip", None))]

corruption_to_strategy = {
    'defocus_blur': 'multiscale', 'glass_blur': 'gentle', 'motion_blur': 'multiscale', 'zoom_blur': 'multiscale',
    'snow': 'color', 'frost': 'color', 'fog': 'color', 'brightness': 'color', 'contrast': 'color',
    'elastic_transform': 'gentle', 'pixelate': 'multiscale', 'jpeg_compression': 'gentle',
}


def get_tta_transforms(corruption_type):
    strategy = corruption_to_strategy.get(corruption_type, 'gentle')
    if strategy == 'multiscale':
        return tta_transforms_multiscale
    elif strategy == 'color':
        return tta_transforms_color
    return tta_transforms_gentle


def sample_strategy_draws(strategy: str, n: int):
    """Per-image draws of one strategy's four views in the reference's order -> dict of arrays."""
    views = {"gentle": tta_transforms_gentle, "multiscale": tta_transforms_multiscale,
             "color": tta_transforms_color}[strategy]
    crop = np.zeros((n, 2, 2), np.int32)
    jit = np.ones((n, 2), np.float64)
    for k in range(n):
        ps = [v.draw() for v in views]
        if strategy == "gentle":
            crop[k, 0] = ps[2][0]
            crop[k, 1] = ps[3][1]
        elif strategy == "color":
            jit[k, 0] = ps[1][0]
            jit[k, 1] = ps[2][0]
    return {"crop_ij": crop, "jitter": jit}


def strategy_views_batch(x_u8: torch.Tensor, strategy: str, draws=None, out=None) -> torch.Tensor:
    """(n,32,32,3) uint8 -> (n,4,3,32,32) normalised f32: the four views of a strategy in ONE fused launch."""
    n = x_u8.shape[0]
    if draws is None:
        draws = sample_strategy_draws(strategy, n)
    return V.tta_strategy(x_u8, strategy, draws.get("crop_ij"), draws.get("jitter"), CIFAR_MEAN, CIFAR_STD, out=out)
