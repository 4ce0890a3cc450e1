"""Drop-ins for the batched float-tensor corruptions (pipenline/angellic.py:34-46,
fall_2025/angellic_with_corrupt.py:51-62)."""
from __future__ import annotations

import torch

from . import _lib as L
from . import batch as B


def add_gaussian_noise(images, mean=0.0, std=0.1):
    """(images + (randn_like(images)*std + mean)).clamp(0, 1) — same torch RNG draw as the reference."""
    noise = torch.randn_like(images)
    return B.float_corrupt(images, L.FC_GAUSSIAN_NOISE, std, mean, noise=noise, clamp=True)


def add_brightness(images, factor=0.3):
    return B.float_corrupt(images, L.FC_BRIGHTNESS, factor, clamp=True)


def add_contrast(images, factor=1.5):
    return B.float_corrupt(images, L.FC_CONTRAST, factor, clamp=True)


def add_gaussian_noise_unclamped(img, std=0.05):
    """fall_2025/angellic_with_corrupt.py:51-52: img + randn_like(img)*std."""
    noise = torch.randn_like(img)
    return B.float_corrupt(img, L.FC_ADD_NOISE, std, noise=noise, clamp=False)


def gaussian_blur_img(img, kernel_size=3):
    """TF.gaussian_blur(img, kernel_size) — fall_2025/angellic_with_corrupt.py:54-56."""
    return B.gaussian_blur_f32(img, kernel_size)
