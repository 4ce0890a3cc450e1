"""Drop-in for the view-generation part of fall_2025/tta_initial_claudeFixed.py (:25-78, :218-230): tensor<->PIL
helpers, the four augmentations and the 16-view expansion (AUG_COUNTS = [1,5,5,5])."""
from __future__ import annotations

import numpy as np
import torch
from PIL import Image

from . import batch as B
from . import views as V
from ._pil import to_dev, to_pil

CIFAR100_MEAN = (0.5071, 0.4867, 0.4408)
CIFAR100_STD = (0.2675, 0.2565, 0.2761)


def _cuda(t):
    return t if t.is_cuda else t.cuda()


def denormalize_tensor(tensor, mean=CIFAR100_MEAN, std=CIFAR100_STD):
    return B.channel_affine(_cuda(tensor), mean, std, True)


def normalize_tensor(tensor, mean=CIFAR100_MEAN, std=CIFAR100_STD):
    return B.channel_affine(_cuda(tensor), mean, std, False)


def tensor_to_pil(tensor):
    """normalised (3,h,w) f32 -> PIL: x*std+mean, clamp[0,1], *255 truncation."""
    t = tensor if tensor.is_cuda else tensor.cuda()
    return to_pil(B.to_uint8(t.unsqueeze(0), CIFAR100_MEAN, CIFAR100_STD, clamp=True))


def pil_to_tensor(pil_image):
    return B.to_tensor(to_dev(pil_image), CIFAR100_MEAN, CIFAR100_STD)[0]


def horiz_flip(image):
    return to_pil(B.hflip(to_dev(image)))


def rand_crop(image):
    w, h = image.size
    crop_size = int(w * 0.78)
    x = np.random.randint(0, w - crop_size + 1)
    y = np.random.randint(0, h - crop_size + 1)
    return to_pil(B.crop_resize(to_dev(image), x, y, crop_size, 32))


def rand_brightness(image):
    factor = np.random.uniform(0.5, 1.5)
    return to_pil(B.brightness(to_dev(image), factor))


def rand_contrast(image):
    factor = np.random.uniform(0.8, 1.2)
    return to_pil(B.enhance_contrast(to_dev(image), factor))


AUG_TYPES = [horiz_flip, rand_crop, rand_brightness, rand_contrast]
AUG_COUNTS = [1, 5, 5, 5]


def sample_view_draws(n: int, w: int = 32, h: int = 32):
    """np.random draws of the 16-view expansion for n images, in the reference's order (per image: 5x(x,y),
    5x brightness, 5x contrast)."""
    cs = int(w * 0.78)
    crop = np.empty((n, 5, 2), np.int32)
    br = np.empty((n, 5), np.float64)
    ct = np.empty((n, 5), np.float64)
    for i in range(n):
        for k in range(5):
            crop[i, k, 0] = np.random.randint(0, w - cs + 1)
            crop[i, k, 1] = np.random.randint(0, h - cs + 1)
        for k in range(5):
            br[i, k] = np.random.uniform(0.5, 1.5)
        for k in range(5):
            ct[i, k] = np.random.uniform(0.8, 1.2)
    return crop, br, ct


def views16_batch(tn: torch.Tensor, crop_xy=None, brightness=None, contrast=None, out=None) -> torch.Tensor:
    """(n,3,32,32) normalised f32 -> (n,16,3,32,32) normalised f32 in ONE fused launch: tensor_to_pil, then
    [flip, 5x crop(24)+BICUBIC resize, 5x Brightness, 5x Contrast], each followed by pil_to_tensor."""
    n = tn.shape[0]
    if crop_xy is None:
        crop_xy, brightness, contrast = sample_view_draws(n, tn.shape[3], tn.shape[2])
    return V.tta16(tn, np.asarray(crop_xy).reshape(n, 5, 2), np.asarray(brightness).reshape(n, 5),
                   np.asarray(contrast).reshape(n, 5), CIFAR100_MEAN, CIFAR100_STD, out=out)


def generate_views(tensor):
    """tta_initial_claudeFixed.py:218-230 for one normalised image: list of 16 normalised tensors."""
    t = tensor if tensor.is_cuda else tensor.cuda()
    return list(views16_batch(t.unsqueeze(0))[0])
