"""Drop-in for fall_2025/transformations_code (:39-66): hflip ("vert_flip"), rand_crop, zoom, perspective warp."""
from __future__ import annotations

import numpy as np
import torch
from PIL import Image

from . import batch as B
from ._pil import to_dev, to_pil
from .transformation import apply_scale  # noqa: F401  (the reference file re-defines the same helpers)
from .transformation import (apply_rotation, apply_contrast, apply_shear, apply_blur, apply_brightness,  # noqa: F401
                             apply_gaussian_noise, apply_translation)


def vert_flip(img):
    """Image.FLIP_LEFT_RIGHT (the reference's name is a misnomer)."""
    return to_pil(B.hflip(to_dev(img)))


def rand_crop(img):
    """Random crop with 0.78 scale factor, then resize((32, 32)) (BICUBIC); draws x then y from np.random."""
    w, h = img.size
    cs = int(0.78 * w)
    x, y = np.random.randint(0, w - cs + 1), np.random.randint(0, h - cs + 1)
    return to_pil(B.crop_resize(to_dev(img), x, y, cs, 32))


def apply_random_zoom(img, scale_factor):
    return apply_scale(img, scale_factor)


def apply_perspective_warp(img, distortion_scale=0.2):
    """ToTensor -> RandomPerspective(distortion_scale, p=1.0) -> ToPILImage.  The corner draws come from
    torchvision's own get_params (same torch RNG consumption as the reference)."""
    import torchvision.transforms as T
    w, h = img.size
    torch.rand(1)  # RandomPerspective.forward draws torch.rand(1) for the p test before get_params
    start, end = T.RandomPerspective.get_params(w, h, distortion_scale)
    return to_pil(B.perspective_tensor(to_dev(img), B.perspective_coeffs(start, end), fill=0.0))
