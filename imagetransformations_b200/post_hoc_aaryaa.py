"""Drop-in for the augmentation side of fall_2025/post_hoc_aaryaa.py: TestTimeAugmentation (:24-51) on f32 tensors and
the mean over the T+1 logit vectors (:317-342; also fall_2025/TTA_transforms.py:208-209).  The calibrators and the
model forward stay out of scope."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import batch as B

CHAIN_DTYPE = np.dtype([("flip", "<i4"), ("perspective", "<i4"), ("brightness", "<f4"), ("reserved", "<f4"),
                        ("rot", "<f4", 6), ("aff", "<f4", 6), ("persp", "<f4", 8)])
assert CHAIN_DTYPE.itemsize == 96


def draw_chain(w: int, h: int) -> dict:
    """Consume the global torch RNG exactly as the reference's Compose does for one image and return the draws:
    flip, angle, scale, brightness, perspective (startpoints, endpoints) or None."""
    import torchvision.transforms as T
    flip = bool(torch.rand(1) < 0.5)                                     # RandomHorizontalFlip(p=0.5)
    angle = T.RandomRotation.get_params([-10.0, 10.0])                   # RandomRotation(degrees=10)
    _, _, scale, _ = T.RandomAffine.get_params((0.0, 0.0), None, (1.0, 1.1), None, [w, h])
    _, b, _, _, _ = T.ColorJitter.get_params((0.9, 1.1), None, None, None)   # randperm(4), then the factor
    persp = None
    if torch.rand(1) < 0.5:                                              # RandomPerspective(0.2, p=0.5)
        persp = T.RandomPerspective.get_params(w, h, 0.2)
    return {"flip": flip, "angle": float(angle), "scale": float(scale), "brightness": float(b), "perspective": persp}


def chain_record(d: dict) -> np.ndarray:
    """Draws -> the imgtx_tta_chain_t record (torchvision's own host parameter math, all in f64, cast to f32 last)."""
    from torchvision.transforms.functional import _get_inverse_affine_matrix
    r = np.zeros((), CHAIN_DTYPE)
    r["flip"] = int(d["flip"])
    r["brightness"] = d["brightness"]
    r["rot"] = _get_inverse_affine_matrix([0.0, 0.0], -d["angle"], [0.0, 0.0], 1.0, [0.0, 0.0])
    r["aff"] = _get_inverse_affine_matrix([0.0, 0.0], 0.0, [0.0, 0.0], d["scale"], [0.0, 0.0])
    if d["perspective"] is not None:
        r["perspective"] = 1
        r["persp"] = B.perspective_coeffs(*d["perspective"])
    return r


def tta_chain_batch(x: torch.Tensor, records: np.ndarray, views: int = 1, out=None) -> torch.Tensor:
    """x (n,c,h,w) f32 CUDA, records CHAIN_DTYPE[n*views] -> (n*views,c,h,w): one launch for every view of every image."""
    x = B._check_f32(x)
    n, c, h, w = x.shape
    rec = np.ascontiguousarray(records, CHAIN_DTYPE).reshape(-1)
    if len(rec) != n * views:
        raise ValueError("need one chain record per output image")
    rd = torch.from_numpy(rec.view(np.uint8).reshape(-1).copy()).to(x.device)
    out = torch.empty((n * views, c, h, w), dtype=torch.float32, device=x.device) if out is None else out
    L.check(L.load().imgtx_tensor_tta_f32(x.data_ptr(), out.data_ptr(), n, views, c, h, w, rd.data_ptr(), B._stream()))
    return out


def mean_views(stack: torch.Tensor) -> torch.Tensor:
    """(V, ...) f32 CUDA -> mean over V in the reference's accumulation order (sequential adds, one division)."""
    s = B._check_f32(stack)
    out = torch.empty(s.shape[1:], dtype=torch.float32, device=s.device)
    L.check(L.load().imgtx_mean_views_f32(s.data_ptr(), out.data_ptr(), out.numel(), s.shape[0], B._stream()))
    return out


class TestTimeAugmentation:
    """TestTimeAugmentation(): `tta(image)` applies one random augmentation to a (C,H,W) f32 tensor; same global
    torch RNG consumption as the reference.  `views(x, T)` draws T chains per image and runs them in one launch."""
    __test__ = False

    def __call__(self, image: torch.Tensor) -> torch.Tensor:
        t = image if image.is_cuda else image.cuda()
        rec = chain_record(draw_chain(t.shape[-1], t.shape[-2]))
        return tta_chain_batch(t.unsqueeze(0), rec.reshape(1))[0]

    def views(self, x: torch.Tensor, T: int) -> torch.Tensor:
        n, c, h, w = x.shape
        recs = np.stack([chain_record(draw_chain(w, h)) for _ in range(n * T)])
        return tta_chain_batch(x, recs, views=T).view(n, T, c, h, w)
