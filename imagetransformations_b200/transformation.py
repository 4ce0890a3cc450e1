"""Drop-in for the reference's transformation.py (:173-325): same names, arguments, defaults, return types and
RNG draw order; the arithmetic runs in libimgtx on the GPU.  `*_batch` variants take/return (n,h,w,c) uint8 CUDA
tensors and are what a throughput user calls.
"""
from __future__ import annotations

import math
import os
import random
from typing import List, Tuple

import numpy as np
import torch
from PIL import Image

from . import _lib as L
from . import batch as B
from ._pil import to_dev, to_pil, batch_to_dev, batch_to_pil


# ---- batched forms -------------------------------------------------------------------------------
apply_scale_batch = B.apply_scale
apply_blur_batch = B.apply_blur
apply_shear_batch = B.apply_shear


def apply_rotation_batch(x, angle, out=None):
    """img.rotate(-angle, fillcolor=0) — transformation.py:198-201."""
    return B.rotate(x, -np.asarray(angle, dtype=np.float64), fill=0, out=out)


def apply_contrast_batch(x, contrast_amount, out=None):
    return B.convert_scale_abs(x, contrast_amount, out=out)


def apply_brightness_batch(x, brightness_factor, out=None):
    f = brightness_factor
    f = (1.0 + f) if np.isscalar(f) else (1.0 + np.asarray(f, dtype=np.float64))
    return B.brightness(x, f, out=out)


def apply_translation_batch(x, tx, ty, out=None):
    return B.translate(x, tx, ty, out=out)


def apply_gaussian_noise_batch(x, noise_std, noise=None, seed=None, out=None):
    """noise given (f32, already scaled by noise_std*255): parity mode. Otherwise on-chip Philox draws."""
    if noise is not None:
        return B.noise_injected(x, L.NOISE_GAUSS_F32, noise, out=out)
    seed = random.getrandbits(63) if seed is None else seed
    return B.noise_philox(x, L.NOISE_GAUSS_F32, float(noise_std) * 255.0, seed, out=out)


# ---- reference signatures ------------------------------------------------------------------------
def apply_scale(img: Image.Image, scale_factor: float) -> Image.Image:
    return to_pil(B.apply_scale(to_dev(img), scale_factor))


def apply_rotation(img: Image.Image, angle: float) -> Image.Image:
    return to_pil(apply_rotation_batch(to_dev(img), angle))


def apply_contrast(img: Image.Image, contrast_amount: float) -> Image.Image:
    return to_pil(B.convert_scale_abs(to_dev(img), contrast_amount))


def apply_shear(img: Image.Image, shear_factor: float) -> Image.Image:
    return to_pil(B.apply_shear(to_dev(img), shear_factor))


def apply_blur(img: Image.Image, blur_radius: float) -> Image.Image:
    if blur_radius == 0:
        return img  # the reference returns the same object (transformation.py:245-246)
    return to_pil(B.apply_blur(to_dev(img), blur_radius))


def apply_brightness(img: Image.Image, brightness_factor: float) -> Image.Image:
    return to_pil(B.brightness(to_dev(img), 1.0 + brightness_factor))


def apply_gaussian_noise(img: Image.Image, noise_std: float) -> Image.Image:
    x = to_dev(img)
    # same draw as the reference (global numpy RNG, transformation.py:274)
    noise = np.random.normal(0, noise_std * 255, tuple(x.shape[1:])).astype(np.float32)
    return to_pil(B.noise_injected(x, L.NOISE_GAUSS_F32, noise[None]))


def apply_translation(img: Image.Image, tx: float, ty: float) -> Image.Image:
    return to_pil(B.translate(to_dev(img), int(tx), int(ty)))


def apply_camera_distance(img: Image.Image, distance_factor: float) -> Image.Image:
    neutral_distance = 2.75
    return apply_scale(img, neutral_distance / distance_factor)


def apply_xy_translation_3d(img: Image.Image, tx: float, ty: float) -> Image.Image:
    width, height = img.size
    return apply_translation(img, int(tx * width), int(ty * height))


def apply_rotation_3d(img: Image.Image, angle: float) -> Image.Image:
    return apply_rotation(img, angle)


TRANSFORMATIONS_2D = {
    'scale': {'min': 0.9, 'max': 1.4, 'step': 0.1},
    'rotation': {'min': -22.5, 'max': 22.5, 'step': 2.5},
    'lighten_darken': {'min': -0.05, 'max': 0.05, 'step': 0.01},
    'gaussian_noise': {'min': 0.0, 'max': 0.1, 'step': 0.01},
    'translation': {'min': -50, 'max': 50, 'step': 5},
    'contrast': {'min': 0, 'max': 1, 'step': 0.1},
    'blur': {'min': 0, 'max': 5, 'step': 0.5},
    'shear': {'min': 0, 'max': 1, 'step': 0.1},
}

_BATCH_FN = {
    'scale': lambda x, v, d: B.apply_scale(x, v),
    'rotation': lambda x, v, d: apply_rotation_batch(x, v),
    'lighten_darken': lambda x, v, d: B.brightness(x, 1.0 + v),
    'gaussian_noise': lambda x, v, d: B.noise_injected(x, L.NOISE_GAUSS_F32, d),
    'translation': lambda x, v, d: B.translate(x, int(v[0]), int(v[1])),
    'contrast': lambda x, v, d: B.convert_scale_abs(x, v),
    'blur': lambda x, v, d: B.apply_blur(x, v),
    'shear': lambda x, v, d: B.apply_shear(x, v),
}


def possible_values(params):
    n = int((params['max'] - params['min']) / params['step']) + 1
    return [params['min'] + j * params['step'] for j in range(n)]


def apply_all_transformations(images, output_dir: str = None, transformations=None) -> list:
    """transformation.py:92-170: for every (img, path) apply each of the 8 transformation types with a value drawn
    by `random.choice` from the type's grid (same draw order).  Images sharing (type, value, size) are transformed
    in one batched launch.  JPEG files are written only when output_dir is given."""
    transformations = TRANSFORMATIONS_2D if transformations is None else transformations
    plan = []  # (image index, type, value, draws, filename)
    for i, (img, path) in enumerate(images):
        name, _ = os.path.splitext(os.path.basename(path))
        for ttype, params in transformations.items():
            vals = possible_values(params)
            draws = None
            if ttype == 'translation':
                tx = random.choice(vals)
                ty = random.choice(vals)
                value = (tx, ty)
                fname = f"{name}_{ttype}_{tx}_{ty}_corrupted.jpg"
            else:
                value = random.choice(vals)
                fname = f"{name}_{ttype}_{value}_corrupted.jpg"
            if ttype == 'gaussian_noise':
                w, h = img.size
                draws = np.random.normal(0, value * 255, (h, w, 3)).astype(np.float32)
            plan.append((i, ttype, value, draws, fname))
    groups = {}
    for k, (i, ttype, value, draws, fname) in enumerate(plan):
        groups.setdefault((ttype, value, images[i][0].size), []).append(k)
    results = [None] * len(plan)
    for (ttype, value, size), ks in groups.items():
        if ttype == 'blur' and value == 0:
            for k in ks:
                results[k] = images[plan[k][0]][0]
            continue
        x = batch_to_dev([images[plan[k][0]][0] for k in ks])
        d = np.stack([plan[k][3] for k in ks]) if ttype == 'gaussian_noise' else None
        for k, p in zip(ks, batch_to_pil(_BATCH_FN[ttype](x, value, d))):
            results[k] = p
    if output_dir is not None:
        os.makedirs(output_dir, exist_ok=True)
        for (i, ttype, value, draws, fname), p in zip(plan, results):
            p.save(os.path.join(output_dir, fname))
    return results
