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


# ------------------------------------------------------------------------------------------------
# The population in one launch (imgtx_pool_chain_f32)
# ------------------------------------------------------------------------------------------------
POOL_KINDS = {"gaussian_noise": 0, "impulse_noise": 1, "shot_noise": 2, "defocus_blur": 3, "enhance_contrast": 4,
              "enhance_brightness": 5, "enhance_sharpness": 6, "enhance_color": 7, "motion_blur": 8,
              "histogram_equalization": 9}
POOL_OP_DTYPE = np.dtype([("kind", "<i4"), ("iparam", "<i4"), ("fparam", "<f4"), ("draw", "<i4")])
assert POOL_OP_DTYPE.itemsize == 16


def encode_chain(transformations) -> np.ndarray:
    """[(name, params), ...] -> POOL_OP_DTYPE records.  Parameters left out are drawn from the global `random` module exactly
    as the op itself would draw them (severity: random.choice([1..5]); factor: random.uniform(0.5, 2.0), sharpness
    random.uniform(0.5, 3.0); size: random.choice([5, 7, 9, 11])), in chain order.  Noise ops get draw = -1 (on chip)."""
    rec = np.zeros(len(transformations), POOL_OP_DTYPE)
    for i, (name, params) in enumerate(transformations):
        params = params or {}
        k = POOL_KINDS[name]
        rec[i]["kind"], rec[i]["draw"] = k, -1
        if k <= 3:
            sev = params.get("severity")
            rec[i]["iparam"] = random.choice([1, 2, 3, 4, 5]) if sev is None else int(sev)
        elif k <= 7:
            f = params.get("factor")
            rec[i]["fparam"] = (random.uniform(0.5, 3.0) if k == 6 else random.uniform(0.5, 2.0)) if f is None else float(f)
        elif k == 8:
            sz = params.get("size")
            rec[i]["iparam"] = random.choice([5, 7, 9, 11]) if sz is None else int(sz)
    return rec


def pool_chain_batch(x: torch.Tensor, ops: np.ndarray, seed: int = 0, draws=None, shared_input: bool = False,
                     mean=CIFAR10_MEAN, std=CIFAR10_STD) -> torch.Tensor:
    """x: (n,3,h,w) normalised f32 CUDA batch, or one (3,h,w) image shared by all n individuals (shared_input);
    ops: (n, max_ops) POOL_OP_DTYPE (kind -1 ends a list).  draws: {"gauss": (k,h,w,3) f64, "impulse": (k,h,w) f64,
    "shot": (k,h,w,3) int64} blocks that ops[..]["draw"] index; without them the noise is drawn on chip from `seed`."""
    x = B._check_f32(x)
    ops = np.ascontiguousarray(ops)
    if ops.dtype != POOL_OP_DTYPE or ops.ndim != 2:
        raise ValueError("ops must be an (n, max_ops) array of POOL_OP_DTYPE records")
    n, max_ops = ops.shape
    if shared_input:
        if x.dim() != 3:
            raise ValueError("shared_input takes one (3,h,w) image")
        stride = 0
    else:
        if x.dim() != 4 or x.shape[0] != n:
            raise ValueError("one input image per individual")
        stride = 3 * x.shape[-2] * x.shape[-1]
    if x.shape[-3] != 3:
        raise ValueError("RGB images only")
    h, w = int(x.shape[-2]), int(x.shape[-1])
    d = draws or {}
    def f64(a):
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(x.device)
    g = f64(d["gauss"]) if "gauss" in d else None
    m = f64(d["impulse"]) if "impulse" in d else None
    s = torch.from_numpy(np.ascontiguousarray(d["shot"], dtype=np.int64)).to(x.device) if "shot" in d else None
    for name, arr, per in (("gauss", g, h * w * 3), ("impulse", m, h * w), ("shot", s, h * w * 3)):
        sel = ops["draw"][(ops["kind"] == POOL_KINDS[{"gauss": "gaussian_noise", "impulse": "impulse_noise", "shot": "shot_noise"}[name]])]
        if len(sel) and sel.max() >= 0 and (arr is None or (int(sel.max()) + 1) * per > arr.numel()):
            raise ValueError(f"ops refer to {name} draw blocks that were not given")
    out = torch.empty((n, 3, h, w), dtype=torch.float32, device=x.device)
    with torch.cuda.device_of(x):
        od = torch.from_numpy(ops.view(np.uint8).reshape(n, -1)).to(x.device)
        mk, mp = B._fptr(list(mean))
        sk, sp = B._fptr(list(std))
        L.check(L.load().imgtx_pool_chain_f32(x.data_ptr(), stride, out.data_ptr(), n, h, w, od.data_ptr(), max_ops, mp, sp,
                                              g.data_ptr() if g is not None else None,
                                              m.data_ptr() if m is not None else None,
                                              s.data_ptr() if s is not None else None,
                                              int(seed) & 0xFFFFFFFFFFFFFFFF, B._stream()))
    return out


def apply_population(individuals, original_image: torch.Tensor, seed: int = 0) -> torch.Tensor:
    """Every Individual of a population applied to the same (3,h,w) normalised image in ONE launch (the loop of
    evolutionary_search, pipenline/cifar_image_transformations.py:236-260, calls apply_transformations once per individual:
    6+ launches, three numpy draws and one device->host round trip each).  Op parameters left open are drawn per
    individual as the ops would draw them; the noise is drawn on chip (seed).  Sets transformed_image on each individual."""
    t = _cuda(original_image)
    chains = [encode_chain(ind.transformations) for ind in individuals]
    max_ops = max(1, max(len(c) for c in chains))
    ops = np.zeros((len(chains), max_ops), POOL_OP_DTYPE)
    ops["kind"] = -1
    for i, c in enumerate(chains):
        ops[i, :len(c)] = c
    out = pool_chain_batch(t, ops, seed=seed, shared_input=True)
    for i, ind in enumerate(individuals):
        ind.transformed_image = out[i]
    return out
