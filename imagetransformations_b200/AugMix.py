"""Drop-in for fall_2025/AugMix.py (:17-62).  `augmix(image_tensor, severity, width, depth)` keeps the reference's
signature and RNG draw order (np.random.dirichlet / beta / randint, random.choice) and runs the whole width x depth
chain, the mix and the final blend as ONE fused kernel launch; `augmix_batch` does the same for a whole batch with the
draws supplied (or sampled per image in reference order)."""
from __future__ import annotations

import math
import random

import numpy as np
import torch

from . import _lib as L
from . import batch as B
from ._pil import to_dev, to_pil

CIFAR10_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR10_STD = (0.2023, 0.1994, 0.2010)
ALPHA = 1.0

AUG_OP_DTYPE = np.dtype([("kind", "<i4"), ("a", "<i4", (6,)), ("param", "<i4")])
assert AUG_OP_DTYPE.itemsize == 32


def _cuda(t):
    return t if t.is_cuda else t.cuda()


def normalize_tensor(tensor):
    return B.channel_affine(_cuda(tensor), CIFAR10_MEAN, CIFAR10_STD, False)


def denormalize_tensor(tensor):
    return B.channel_affine(_cuda(tensor), CIFAR10_MEAN, CIFAR10_STD, True)


# ---- the eight PIL ops (single-image drop-ins) -----------------------------------------------------
def rotate(img, severity):
    return to_pil(B.rotate(to_dev(img), severity * random.choice([-1, 1])))


def posterize(img, severity):
    return to_pil(B.posterize(to_dev(img), int(severity)))


def shear_x(img, severity):
    return to_pil(B.affine_nearest(to_dev(img), (1, severity * 0.3, 0, 0, 1, 0)))


def shear_y(img, severity):
    return to_pil(B.affine_nearest(to_dev(img), (1, 0, 0, severity * 0.3, 1, 0)))


def translate_x(img, severity):
    return to_pil(B.affine_nearest(to_dev(img), (1, 0, severity * 2, 0, 1, 0)))


def translate_y(img, severity):
    return to_pil(B.affine_nearest(to_dev(img), (1, 0, 0, 0, 1, severity * 2)))


def equalize(img, _):
    return to_pil(B.equalize(to_dev(img)))


def solarize(img, severity):
    return to_pil(B.solarize(to_dev(img), int(severity * 20)))


AUG_OPS = [rotate, posterize, shear_x, shear_y, translate_x, translate_y, equalize, solarize]


# ---- op compilation (host, doubles -> 16.16 ints exactly as Pillow's C does) ------------------------------
def _fix(v: float) -> int:
    return int(math.floor(v * 65536.0 + 0.5))


def _affine_fixed(m, w, h):
    """Pillow affine (NEAREST): 16.16 coefficients of the fixed-point path, or of the scale/translate path."""
    if m[1] == 0 and m[3] == 0:
        if m[0] == 1 and m[4] == 1:
            return [65536, 0, int(math.floor(m[2] + 0.5)) * 65536, 0, 65536, int(math.floor(m[5] + 0.5)) * 65536]
        raise NotImplementedError("pure-scale affine is not an AugMix op")
    return [_fix(m[0]), _fix(m[1]), _fix(m[2] + m[0] * 0.5 + m[1] * 0.5),
            _fix(m[3]), _fix(m[4]), _fix(m[5] + m[3] * 0.5 + m[4] * 0.5)]


def compile_op(op: int, severity, sign: int, w: int, h: int):
    """(AUG_OPS index, severity, rotate sign) -> one imgtx_augmix_op_t record."""
    rec = np.zeros((), AUG_OP_DTYPE)
    if op < 0:
        return rec
    if op == 0:
        ang = (severity * sign) % 360.0
        if ang == 0:
            return rec
        rec["kind"] = L.AUG_GEOM
        if ang == 180:
            rec["a"] = [-65536, 0, (w - 1) * 65536, 0, -65536, (h - 1) * 65536]
        elif ang == 90 and w == h:
            rec["a"] = [0, -65536, (w - 1) * 65536, 65536, 0, 0]
        elif ang == 270 and w == h:
            rec["a"] = [0, 65536, 0, -65536, 0, (h - 1) * 65536]
        else:
            rec["a"] = _affine_fixed(B.pil_rotate_matrix(ang, w, h), w, h)
    elif op == 1:
        rec["kind"] = L.AUG_POSTERIZE
        rec["param"] = ~((1 << (8 - int(severity))) - 1) & 0xFF
    elif op in (2, 3, 4, 5):
        rec["kind"] = L.AUG_GEOM
        m = {2: (1, severity * 0.3, 0, 0, 1, 0), 3: (1, 0, 0, severity * 0.3, 1, 0),
             4: (1, 0, severity * 2, 0, 1, 0), 5: (1, 0, 0, 0, 1, severity * 2)}[op]
        rec["a"] = _affine_fixed([float(v) for v in m], w, h)
    elif op == 6:
        rec["kind"] = L.AUG_EQUALIZE
    elif op == 7:
        rec["kind"] = L.AUG_SOLARIZE
        rec["param"] = int(severity * 20)
    else:
        raise ValueError(op)
    return rec


def sample_draws(width=3, depth=-1):
    """The draws one augmix() call consumes, in the reference's order."""
    ws = np.random.dirichlet([ALPHA] * width)
    m = np.random.beta(ALPHA, ALPHA)
    dmax = depth if depth > 0 else 3
    ops = np.full((width, dmax), -1, np.int32)
    signs = np.zeros((width, dmax), np.int32)
    for i in range(width):
        d = depth if depth > 0 else np.random.randint(1, 4)
        for j in range(d):
            op = random.choice(range(len(AUG_OPS)))
            ops[i, j] = op
            if op == 0:
                signs[i, j] = random.choice([-1, 1])
    return ws, m, ops, signs


def augmix_batch(x: torch.Tensor, severity, ws, m, ops, signs, out=None) -> torch.Tensor:
    """x (n,3,h,w) f32 CUDA.  ws (n,width) f64, m (n,) f64, ops/signs (n,width,depth) int (op < 0: chain ended)."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] == 3):
        raise TypeError("expected a CUDA float32 tensor of shape (n, 3, h, w)")
    x = x.contiguous()
    n, _, h, w = x.shape
    ws = np.asarray(ws, np.float64).reshape(n, -1)
    m = np.asarray(m, np.float64).reshape(n)
    ops = np.asarray(ops).reshape(n, ws.shape[1], -1)
    signs = np.asarray(signs).reshape(ops.shape)
    width, depth = ops.shape[1], ops.shape[2]
    sev = np.broadcast_to(np.asarray(severity, np.float64), (n,))
    recs = np.zeros((n, width, depth), AUG_OP_DTYPE)
    cache = {}
    for i in range(n):
        for a in range(width):
            for b in range(depth):
                key = (int(ops[i, a, b]), float(sev[i]), int(signs[i, a, b]))
                if key not in cache:
                    cache[key] = compile_op(key[0], sev[i] if sev[i] != int(sev[i]) else int(sev[i]), key[2], w, h)
                recs[i, a, b] = cache[key]
    wts = np.concatenate([ws.astype(np.float32), m.astype(np.float32)[:, None],
                          (1.0 - m).astype(np.float32)[:, None]], axis=1)
    rd = torch.from_numpy(recs.view(np.uint8).reshape(-1).copy()).to(x.device)
    wd = torch.from_numpy(np.ascontiguousarray(wts)).to(x.device)
    out = torch.empty_like(x) if out is None else out
    L.check(L.load().imgtx_augmix_f32(x.data_ptr(), out.data_ptr(), n, h, w, width, depth, rd.data_ptr(),
                                      wd.data_ptr(), B._stream()))
    return out


def augmix(image_tensor, severity=3, width=3, depth=-1):
    """augmix(image_tensor (3,h,w) f32, severity, width, depth) -> (3,h,w) f32 on the input's device."""
    ws, m, ops, signs = sample_draws(width, depth)
    x = image_tensor if image_tensor.is_cuda else image_tensor.cuda()
    out = augmix_batch(x.unsqueeze(0), severity, ws[None], [m], ops[None], signs[None])[0]
    return out if image_tensor.is_cuda else out.to(image_tensor.device)
