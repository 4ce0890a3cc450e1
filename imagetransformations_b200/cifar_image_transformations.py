"""Drop-in for the transformation stage of pipenline/cifar_image_transformations.py (:17-152): tensor<->PIL helpers,
TransformationPool (the reference's own CIFAR-C-style generators + enhancers) and Individual.apply_transformations.
Draw order from the global `random` / `np.random` generators is the reference's."""
from __future__ import annotations

import random

import numpy as np
import torch
from PIL import Image

from . import _lib as L
from . import batch as B
from ._pil import to_dev, to_pil

CIFAR10_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR10_STD = (0.2023, 0.1994, 0.2010)

GAUSSIAN_LEVELS = [0.08, 0.12, 0.18, 0.26, 0.38]
IMPULSE_LEVELS = [0.03, 0.06, 0.09, 0.17, 0.27]
SHOT_LEVELS = [60, 25, 12, 5, 3]
DEFOCUS_LEVELS = [3, 4, 6, 8, 10]


def _cuda(t):
    return t if t.is_cuda else t.cuda()


def denormalize_tensor(tensor, mean=CIFAR10_MEAN, std=CIFAR10_STD):
    return B.channel_affine(_cuda(tensor), mean, std, True)


def normalize_tensor(tensor, mean=CIFAR10_MEAN, std=CIFAR10_STD):
    return B.channel_affine(_cuda(tensor), mean, std, False)


def tensor_to_pil(tensor):
    t = tensor if tensor.is_cuda else tensor.cuda()
    return to_pil(B.to_uint8(t.unsqueeze(0), CIFAR10_MEAN, CIFAR10_STD, clamp=True))


def pil_to_tensor(pil_image):
    return B.to_tensor(to_dev(pil_image), CIFAR10_MEAN, CIFAR10_STD)[0]


def _dev(image):
    """PIL image or (n,h,w,3) uint8 CUDA batch -> (batch tensor, was_pil)."""
    if isinstance(image, torch.Tensor):
        return image, False
    return to_dev(image), True


def _ret(x, was_pil):
    return to_pil(x) if was_pil else x


class TransformationPool:
    """Every op accepts a PIL image (reference signature) or a (n,h,w,3) uint8 CUDA batch (batched form; random
    draws are then made per image in batch order)."""

    def gaussian_noise(image, severity=None):
        if severity is None:
            severity = random.choice([1, 2, 3, 4, 5])
        x, pil = _dev(image)
        noise = np.random.normal(0, GAUSSIAN_LEVELS[severity - 1] * 255, tuple(x.shape))
        return _ret(B.noise_injected(x, L.NOISE_GAUSS_F64, noise), pil)

    def impulse_noise(image, severity=None):
        if severity is None:
            severity = random.choice([1, 2, 3, 4, 5])
        x, pil = _dev(image)
        mask = np.random.random(tuple(x.shape[:3]))
        return _ret(B.noise_injected(x, L.NOISE_IMPULSE, mask, IMPULSE_LEVELS[severity - 1]), pil)

    def shot_noise(image, severity=None):
        if severity is None:
            severity = random.choice([1, 2, 3, 4, 5])
        x, pil = _dev(image)
        lam = SHOT_LEVELS[severity - 1]
        # the Poisson rates are data dependent: same f32 expression as the reference, same generator call
        scaled = x.cpu().numpy().astype(np.float32) / 255.0 * lam
        counts = np.random.poisson(scaled)
        return _ret(B.noise_injected(x, L.NOISE_SHOT, counts, lam), pil)

    def defocus_blur(image, severity=None):
        if severity is None:
            severity = random.choice([1, 2, 3, 4, 5])
        x, pil = _dev(image)
        return _ret(B.pil_gaussian_blur(x, DEFOCUS_LEVELS[severity - 1]), pil)

    def enhance_contrast(image, factor=None):
        if factor is None:
            factor = random.uniform(0.5, 2.0)
        x, pil = _dev(image)
        return _ret(B.enhance_contrast(x, factor), pil)

    def enhance_brightness(image, factor=None):
        if factor is None:
            factor = random.uniform(0.5, 2.0)
        x, pil = _dev(image)
        return _ret(B.brightness(x, factor), pil)

    def enhance_sharpness(image, factor=None):
        if factor is None:
            factor = random.uniform(0.5, 3.0)
        x, pil = _dev(image)
        return _ret(B.enhance_sharpness(x, factor), pil)

    def enhance_color(image, factor=None):
        if factor is None:
            factor = random.uniform(0.5, 2.0)
        x, pil = _dev(image)
        return _ret(B.enhance_color(x, factor), pil)

    def motion_blur(image, size=None):
        if size is None:
            size = random.choice([5, 7, 9, 11])
        x, pil = _dev(image)
        return _ret(B.box_filter_h(x, size), pil)

    def histogram_equalization(image):
        x, pil = _dev(image)
        return _ret(B.histogram_equalization(x), pil)


class Individual:
    def __init__(self, transformations=None):
        self.transformations = transformations if transformations else []
        self.fitness = 0.0
        self.uncertainty_score = float('inf')
        self.transformed_image = None
        self.prediction = None
        self.confidence = 0.0

    def apply_transformations(self, original_image):
        """(3,h,w) normalised f32 -> same (or a (n,3,h,w) batch): tensor_to_pil, the chain, pil_to_tensor — the
        image stays on the device as uint8 between ops."""
        t = original_image if original_image.is_cuda else original_image.cuda()
        single = t.dim() == 3
        x = B.to_uint8(t.unsqueeze(0) if single else t, CIFAR10_MEAN, CIFAR10_STD, clamp=True)
        for transform_name, transform_params in self.transformations:
            transform_func = getattr(TransformationPool, transform_name)
            x = transform_func(x, **transform_params) if transform_params else transform_func(x)
        out = B.to_tensor(x, CIFAR10_MEAN, CIFAR10_STD)
        self.transformed_image = out[0] if single else out
        return self.transformed_image
