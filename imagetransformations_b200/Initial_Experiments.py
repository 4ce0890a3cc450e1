"""Drop-in for fall_2025/Initial_Experiments.py: the entropy gate (compute_shannon_entropy :95-113 and the per-image decision
of the evaluation loop :147-157) and the ADAPTIVE TRANSFORMS it triggers (:88-92, applied at :161-167):

    adaptive_transforms = T.Compose([T.RandomHorizontalFlip(p=0.5),
                                     T.RandomAffine(degrees=15, translate=(0.1, 0.1), scale=(0.9, 1.1)),
                                     T.RandomPerspective(distortion_scale=0.2, p=0.5)])

on PIL images (NEAREST affine, BILINEAR perspective, fill 0).  `adaptive_transforms(img)` keeps the reference's call shape (PIL in,
PIL out) and consumes the global torch RNG exactly as the Compose does (flip: torch.rand(1); RandomAffine.get_params: angle,
tx, ty, scale; perspective: torch.rand(1), then get_params when it fires), so a script that seeds torch gets the same draws.
The pixels come from ONE fused launch of libimgtx (imgtx_adaptive_chain_u8); `adaptive_transforms_batch` runs a whole
device batch from recorded draws, and `adapt_batch` is the loop body of :147-167 for a normalised batch.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib as L
from . import batch as B

ENTROPY_THRESHOLD = 7.46
CIFAR_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR_STD = (0.2023, 0.1994, 0.2010)

ADAPTIVE_DTYPE = np.dtype([("flags", "<i4"), ("a", "<i4", (6,)), ("pad", "<i4"), ("persp", "<f8", (8,))])
assert ADAPTIVE_DTYPE.itemsize == 96


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


# ------------------------------------------------------------------------------------------------
# draws -> per-image records
# ------------------------------------------------------------------------------------------------
def draw_adaptive_params(w: int, h: int) -> dict:
    """One image's draws, consuming the global torch RNG in the Compose's order (:88-92)."""
    import torchvision.transforms as T
    flip = bool(torch.rand(1) < 0.5)
    angle, (tx, ty), scale, _ = T.RandomAffine.get_params([-15.0, 15.0], (0.1, 0.1), (0.9, 1.1), None, [w, h])
    fired = bool(torch.rand(1) < 0.5)
    start, end = T.RandomPerspective.get_params(w, h, 0.2) if fired else (None, None)
    return {"flip": flip, "angle": float(angle), "translate": (int(tx), int(ty)), "scale": float(scale), "persp": fired,
            "start": start, "end": end}


def _fix(v: float) -> int:
    return int(math.floor(v * 65536.0 + 0.5))


def adaptive_record(d: dict, w: int, h: int) -> np.ndarray:
    """The imgtx_adaptive_t of one image: torchvision's inverse affine matrix about the image centre (F.affine on PIL:
    center = [w/2, h/2]) FIX'ed like libImaging's affine_fixed, and _get_perspective_coeffs of the corner draws."""
    from torchvision.transforms import functional as TF
    rec = np.zeros((), ADAPTIVE_DTYPE)
    m = TF._get_inverse_affine_matrix([w * 0.5, h * 0.5], d["angle"], [float(d["translate"][0]), float(d["translate"][1])],
                                      d["scale"], [0.0, 0.0])
    flags = (1 if d["flip"] else 0) | 2
    if m[1] == 0.0 and m[3] == 0.0:
        raise NotImplementedError("axis-aligned affine (libImaging ImagingScaleAffine path): use the unfused ops")
    rec["a"] = [_fix(m[0]), _fix(m[1]), _fix(m[2] + m[0] * 0.5 + m[1] * 0.5), _fix(m[3]), _fix(m[4]),
                _fix(m[5] + m[3] * 0.5 + m[4] * 0.5)]
    if d["persp"]:
        flags |= 4
        rec["persp"] = TF._get_perspective_coeffs(d["start"], d["end"])
    rec["flags"] = flags
    return rec


def adaptive_transforms_batch(x: torch.Tensor, draws, out=None) -> torch.Tensor:
    """x: (n,h,w,3) uint8 CUDA; draws: a list of n dicts from draw_adaptive_params (or an ADAPTIVE_DTYPE array)."""
    x = B._check_u8(x)
    n, h, w, c = x.shape
    recs = draws if isinstance(draws, np.ndarray) else np.stack([adaptive_record(d, w, h) for d in draws])
    if recs.dtype != ADAPTIVE_DTYPE or recs.shape != (n,):
        raise ValueError("one imgtx_adaptive_t record per image")
    out = B._check_out(out, x)
    with torch.cuda.device_of(x):
        dev = torch.from_numpy(np.ascontiguousarray(recs).view(np.uint8).reshape(n, -1)).to(x.device)
        L.check(L.load().imgtx_adaptive_chain_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, dev.data_ptr(), B._stream()))
    return out


def adaptive_transforms(img):
    """adaptive_transforms(img_pil) -> PIL image (:164), same torch RNG consumption as the reference's Compose."""
    from ._pil import to_dev, to_pil
    x = to_dev(img)
    d = draw_adaptive_params(x.shape[2], x.shape[1])
    return to_pil(adaptive_transforms_batch(x, [d]))


def adapt_batch(images: torch.Tensor, threshold: float = ENTROPY_THRESHOLD) -> torch.Tensor:
    """Loop body of :147-167 for a normalised (n,3,h,w) CUDA batch: images whose entropy exceeds the threshold go through
    ToPILImage(denormalised, clamped) -> adaptive_transforms -> ToTensor + Normalize; the others pass unchanged.  Draws are
    taken per selected image in batch order, as the reference's per-image loop does."""
    mask = adaptive_mask(images, threshold)
    idx = torch.nonzero(mask).flatten()
    if idx.numel() == 0:
        return images.clone()
    sel = images.index_select(0, idx)
    u8 = B.to_uint8(sel, CIFAR_MEAN, CIFAR_STD, clamp=True)
    n, h, w, _ = u8.shape
    y = adaptive_transforms_batch(u8, [draw_adaptive_params(w, h) for _ in range(n)])
    out = images.clone()
    out.index_copy_(0, idx, B.to_tensor(y, CIFAR_MEAN, CIFAR_STD))
    return out
