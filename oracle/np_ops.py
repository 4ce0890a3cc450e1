"""CPU oracle: numpy restatement of the reference's image-transformation hot path.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; the product (imagetransformations_b200/) must never route through it.

The reference (aaryaamoharir/ImageTransformations) delegates all arithmetic to third-party libraries that
are NOT vendored in /root/reference and whose versions it does not pin (requirements.txt:1-4):
Pillow (12.2.0 here), OpenCV (4.13.0), torchvision (0.26.0), torch (2.11.0), numpy (2.3.5).  Each function
below restates the published algorithm of the library routine the cited reference line calls, batched over
a leading image axis, integer-exact where the library is integer.  Parity is pinned de facto: every function
here is checked (tests/test_oracle_golden.py) against golden vectors produced by running the reference's
own functions in the build container (oracle/make_golden.py -> tests/golden/*.npz).

Array conventions: uint8 images are (n, h, w, c); float tensors are (n, c, h, w) float32.
"""
from __future__ import annotations

import math

import numpy as np

f32 = np.float32

# =================================================================================================
# A.4 / A.5 / A.4b  pointwise maps
# =================================================================================================


def pil_blend_lut(degenerate: int, factor: float) -> np.ndarray:
    """256-entry table of Pillow ImagingBlend(degenerate, p, factor) (libImaging/Blend.c).

    reference: ImageEnhance.*.enhance -> Image.blend; transformation.py:261-269 (Brightness),
    fall_2025/tta_initial_claudeFixed.py:64-72, pipenline/cifar_image_transformations.py:81-106.
    """
    p = np.arange(256, dtype=np.int32)
    f = f32(factor)
    if f == 0:
        return np.full(256, degenerate, np.uint8)
    if f == 1:
        return p.astype(np.uint8)
    t = (f32(degenerate) + (f * (p - degenerate).astype(f32)).astype(f32)).astype(f32)
    if 0 <= f <= 1:
        return t.astype(np.int32).astype(np.uint8)
    out = np.where(t <= 0, 0, np.where(t >= 255, 255, np.trunc(np.clip(t, 0, 255)))).astype(np.uint8)
    return out


def enhance_brightness(img: np.ndarray, factor) -> np.ndarray:
    """ImageEnhance.Brightness(img).enhance(factor); factor scalar or per-image (n,)."""
    factor = np.broadcast_to(np.asarray(factor, np.float64), (img.shape[0],))
    out = np.empty_like(img)
    for i in range(img.shape[0]):
        out[i] = pil_blend_lut(0, factor[i])[img[i]]
    return out


def convert_scale_abs(img: np.ndarray, alpha) -> np.ndarray:
    """cv2.convertScaleAbs(img, alpha=alpha, beta=0) — transformation.py:203-210.
    f32 product, round-half-even, saturate."""
    alpha = np.broadcast_to(np.asarray(alpha, np.float64), (img.shape[0],)).astype(f32)
    a = alpha.reshape((-1,) + (1,) * (img.ndim - 1))
    v = np.abs(img.astype(f32) * a).astype(f32)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


def posterize(img: np.ndarray, bits: int) -> np.ndarray:
    """ImageOps.posterize — fall_2025/AugMix.py:31."""
    mask = ~((1 << (8 - int(bits))) - 1) & 0xFF
    return img & np.uint8(mask)


def solarize(img: np.ndarray, threshold: int) -> np.ndarray:
    """ImageOps.solarize — fall_2025/AugMix.py:37."""
    return np.where(img < threshold, img, 255 - img).astype(np.uint8)


def equalize_lut(hist: np.ndarray) -> np.ndarray:
    """ImageOps.equalize LUT for one channel histogram (256,) — fall_2025/AugMix.py:36."""
    nz = hist[hist > 0]
    ident = np.arange(256, dtype=np.uint8)
    if len(nz) <= 1:
        return ident
    step = (int(hist.sum()) - int(nz[-1])) // 255
    if step == 0:
        return ident
    lut = np.empty(256, np.uint8)
    n = step // 2
    for i in range(256):
        lut[i] = min(n // step, 255)
        n += int(hist[i])
    return lut


def equalize(img: np.ndarray) -> np.ndarray:
    out = np.empty_like(img)
    for i in range(img.shape[0]):
        for ch in range(img.shape[3]):
            h = np.bincount(img[i, :, :, ch].ravel(), minlength=256)
            out[i, :, :, ch] = equalize_lut(h)[img[i, :, :, ch]]
    return out


def luma(img: np.ndarray) -> np.ndarray:
    """Pillow RGB->L: (19595 R + 38470 G + 7471 B + 0x8000) >> 16."""
    r = img[..., 0].astype(np.int64)
    g = img[..., 1].astype(np.int64)
    b = img[..., 2].astype(np.int64)
    return ((19595 * r + 38470 * g + 7471 * b + 0x8000) >> 16).astype(np.uint8)


def contrast_degenerate(img: np.ndarray) -> np.ndarray:
    """int(ImageStat.Stat(L).mean[0] + 0.5) per image."""
    L = luma(img).reshape(img.shape[0], -1).astype(np.int64)
    s = L.sum(axis=1)
    return np.array([int(si / L.shape[1] + 0.5) for si in s], dtype=np.int32)


def enhance_contrast(img: np.ndarray, factor) -> np.ndarray:
    """ImageEnhance.Contrast(img).enhance(factor) — tta_initial_claudeFixed.py:69-72."""
    factor = np.broadcast_to(np.asarray(factor, np.float64), (img.shape[0],))
    d = contrast_degenerate(img)
    out = np.empty_like(img)
    for i in range(img.shape[0]):
        out[i] = pil_blend_lut(int(d[i]), factor[i])[img[i]]
    return out


def _blend_arrays(deg: np.ndarray, img: np.ndarray, factor: float) -> np.ndarray:
    f = f32(factor)
    if f == 0:
        return deg.copy()
    if f == 1:
        return img.copy()
    d = deg.astype(np.int32)
    p = img.astype(np.int32)
    t = (d.astype(f32) + (f * (p - d).astype(f32)).astype(f32)).astype(f32)
    if 0 <= f <= 1:
        return t.astype(np.int32).astype(np.uint8)
    return np.where(t <= 0, 0, np.where(t >= 255, 255, np.trunc(np.clip(t, 0, 255)))).astype(np.uint8)


def enhance_color(img: np.ndarray, factor) -> np.ndarray:
    """ImageEnhance.Color: degenerate = L replicated to RGB — cifar_image_transformations.py:102-106."""
    factor = np.broadcast_to(np.asarray(factor, np.float64), (img.shape[0],))
    L = luma(img)
    out = np.empty_like(img)
    for i in range(img.shape[0]):
        deg = np.repeat(L[i][..., None], 3, axis=2)
        out[i] = _blend_arrays(deg, img[i], factor[i])
    return out


def smooth3x3(img: np.ndarray) -> np.ndarray:
    """ImageFilter.SMOOTH: [1,1,1;1,5,1;1,1,1]/13, f32 accumulate, +0.5 floor, border copied."""
    n, h, w, c = img.shape
    out = img.copy()
    if h < 3 or w < 3:
        return out
    k = np.array([[1, 1, 1], [1, 5, 1], [1, 1, 1]], f32) / f32(13)
    # Pillow stores the kernel flipped vertically, accumulates bottom row first; the kernel is symmetric.
    acc = np.zeros((n, h - 2, w - 2, c), f32)
    src = img.astype(f32)
    # libImaging Filter.c (ImagingFilter3x3) sums in0 (y-1), in1 (y), in2 (y+1) with kernel rows [0..2],[3..5],[6..8]
    # ss = 0.5; ss += MAC3(in1..) ; ss += MAC3(in_1) ; ss += MAC3(in0)
    def mac(row, kr):
        return (src[:, row[0]:row[1], 0:w - 2] * kr[0] + src[:, row[0]:row[1], 1:w - 1] * kr[1]
                + src[:, row[0]:row[1], 2:w] * kr[2]).astype(f32)
    acc = f32(0.5) + np.zeros_like(acc)
    acc = (acc + mac((2, h), k[0])).astype(f32)
    acc = (acc + mac((1, h - 1), k[1])).astype(f32)
    acc = (acc + mac((0, h - 2), k[2])).astype(f32)
    out[:, 1:h - 1, 1:w - 1] = np.clip(np.floor(acc), 0, 255).astype(np.uint8)
    return out


def enhance_sharpness(img: np.ndarray, factor) -> np.ndarray:
    """ImageEnhance.Sharpness: degenerate = SMOOTH filter — cifar_image_transformations.py:95-99."""
    factor = np.broadcast_to(np.asarray(factor, np.float64), (img.shape[0],))
    deg = smooth3x3(img)
    out = np.empty_like(img)
    for i in range(img.shape[0]):
        out[i] = _blend_arrays(deg[i], img[i], factor[i])
    return out


# =================================================================================================
# A.1  Pillow NEAREST affine (Image.rotate / Image.transform(AFFINE))
# =================================================================================================


def _fix(v: float) -> int:
    return int(math.floor(v * 65536.0 + 0.5))


def rotate_matrix(angle: float, w: int, h: int):
    """The 6 coefficients Image.rotate(angle) hands to the C transform (Image.py rotate, expand=False)."""
    angle = angle % 360.0
    a = -math.radians(angle)
    m = [round(math.cos(a), 15), round(math.sin(a), 15), 0.0,
         round(-math.sin(a), 15), round(math.cos(a), 15), 0.0]
    cx, cy = w / 2.0, h / 2.0

    def tr(x, y):
        return m[0] * x + m[1] * y + m[2], m[3] * x + m[4] * y + m[5]

    m[2], m[5] = tr(-cx, -cy)
    m[2] += cx
    m[5] += cy
    return m


def affine_nearest_coeffs(m):
    """FIX'ed 16.16 coefficients (a0,a1,a2,a3,a4,a5) of libImaging Geometry.c affine_fixed."""
    return [_fix(m[0]), _fix(m[1]), _fix(m[2] + m[0] * 0.5 + m[1] * 0.5),
            _fix(m[3]), _fix(m[4]), _fix(m[5] + m[3] * 0.5 + m[4] * 0.5)]


def affine_nearest(img: np.ndarray, m, out_size=None, fill=0) -> np.ndarray:
    """Image.transform(size, AFFINE, m) with NEAREST resampling. m: 6 doubles (uniform) or (n,6)."""
    n, h, w, c = img.shape
    ow, oh = out_size if out_size is not None else (w, h)
    m = np.asarray(m, np.float64)
    if m.ndim == 1:
        m = np.broadcast_to(m, (n, 6))
    out = np.empty((n, oh, ow, c), np.uint8)
    fillv = np.broadcast_to(np.asarray(fill, np.uint8), (c,))
    ys, xs = np.mgrid[0:oh, 0:ow]
    for i in range(n):
        mi = m[i]
        if mi[1] == 0 and mi[3] == 0:
            # scale/translate path (affine_transform non-fixed): repeated double addition
            xo = mi[2] + mi[0] * 0.5
            xin = np.empty(ow, np.int64)
            for x in range(ow):
                xin[x] = math.floor(xo)  # COORD(v) = v < 0 ? -1 : (int)v  — same as floor for v>=0 / marks <0
                if xo < 0:
                    xin[x] = -1
                xo += mi[0]
            yo = mi[5] + mi[4] * 0.5
            yin = np.empty(oh, np.int64)
            for y in range(oh):
                yin[y] = math.floor(yo)
                if yo < 0:
                    yin[y] = -1
                yo += mi[4]
            XI = np.broadcast_to(xin[None, :], (oh, ow))
            YI = np.broadcast_to(yin[:, None], (oh, ow))
        else:
            a = affine_nearest_coeffs(mi)
            XI = (a[2] + ys * a[1] + xs * a[0]) >> 16
            YI = (a[5] + ys * a[4] + xs * a[3]) >> 16
        inside = (XI >= 0) & (XI < w) & (YI >= 0) & (YI < h)
        o = np.empty((oh, ow, c), np.uint8)
        o[:] = fillv
        o[inside] = img[i][YI[inside], XI[inside]]
        out[i] = o
    return out


def pil_rotate_nearest(img: np.ndarray, angle, fill=0) -> np.ndarray:
    """img.rotate(angle, fillcolor=fill) (NEAREST, expand=False) — transformation.py:198-201 calls it
    with -angle; fall_2025/AugMix.py:30 with +-severity."""
    n, h, w, c = img.shape
    angle = np.broadcast_to(np.asarray(angle, np.float64), (n,))
    out = np.empty_like(img)
    for i in range(n):
        a = angle[i] % 360.0
        if a == 0:
            out[i] = img[i]
        elif a == 180:
            out[i] = img[i][::-1, ::-1]
        elif a == 90 and h == w:
            out[i] = np.rot90(img[i], 1)
        elif a == 270 and h == w:
            out[i] = np.rot90(img[i], 3)
        else:
            out[i] = affine_nearest(img[i:i + 1], rotate_matrix(a, w, h), fill=fill)[0]
    return out


def translate(img: np.ndarray, tx, ty) -> np.ndarray:
    """apply_translation — transformation.py:284-307: integer shift, zero fill."""
    n, h, w, c = img.shape
    tx = np.broadcast_to(np.asarray(tx), (n,))
    ty = np.broadcast_to(np.asarray(ty), (n,))
    out = np.zeros_like(img)
    for i in range(n):
        px, py = int(tx[i]), int(ty[i])
        x0, y0 = max(0, -px), max(0, -py)
        x1, y1 = min(w, w - px), min(h, h - py)
        if x0 < x1 and y0 < y1:
            out[i, max(0, py):max(0, py) + (y1 - y0), max(0, px):max(0, px) + (x1 - x0)] = img[i, y0:y1, x0:x1]
    return out


def hflip(img: np.ndarray) -> np.ndarray:
    return img[:, :, ::-1].copy()


# =================================================================================================
# A.7  ToTensor / Normalize / tensor_to_pil
# =================================================================================================


def to_tensor(img: np.ndarray) -> np.ndarray:
    """T.ToTensor(): HWC u8 -> CHW f32 u/255 (true division)."""
    return (img.transpose(0, 3, 1, 2).astype(f32) / f32(255)).astype(f32)


def normalize(t: np.ndarray, mean, std) -> np.ndarray:
    m = np.asarray(mean, f32).reshape(1, -1, 1, 1)
    s = np.asarray(std, f32).reshape(1, -1, 1, 1)
    return ((t - m).astype(f32) / s).astype(f32)


def to_tensor_normalize(img, mean, std):
    return normalize(to_tensor(img), mean, std)


def denormalize(t: np.ndarray, mean, std) -> np.ndarray:
    m = np.asarray(mean, f32).reshape(1, -1, 1, 1)
    s = np.asarray(std, f32).reshape(1, -1, 1, 1)
    return ((t * s).astype(f32) + m).astype(f32)


def to_pil_image(t: np.ndarray) -> np.ndarray:
    """TF.to_pil_image on float CHW: x.mul(255).byte() — truncation with modulo-256 wrap (A.9)."""
    v = (t.astype(f32) * f32(255)).astype(f32)
    q = np.trunc(v).astype(np.int64) & 255
    return q.astype(np.uint8).transpose(0, 2, 3, 1).copy()


def tensor_to_pil(t: np.ndarray, mean, std) -> np.ndarray:
    """denormalize -> clamp[0,1] -> ToPILImage — tta_initial_claudeFixed.py:37-41,
    pipenline/cifar_image_transformations.py:27-30."""
    d = np.clip(denormalize(t, mean, std), f32(0), f32(1)).astype(f32)
    return to_pil_image(d)


# =================================================================================================
# A.12  sign-step family
# =================================================================================================


def torch_sign(g: np.ndarray) -> np.ndarray:
    return (g > 0).astype(f32) - (g < 0).astype(f32)


def _clamp_nan(x, lo, hi):
    # torch.clamp / max / min propagate NaN from x
    with np.errstate(invalid="ignore"):
        y = np.where(x < lo, lo, x)
        y = np.where(y > hi, hi, y)
    return y.astype(f32)


def fgsm_step(x, g, eps, direction=+1.0, lo=None, hi=None):
    """x + direction*eps*sign(g), optionally clamped to [lo,hi] (scalars or per-channel (c,)).
    attacks/fgsm.py:190-192, attacks/fgsm/cifar-10/fgsm_cifar10.py:25-28, reverse_fgsm_cifar10.py:74-80."""
    e = f32(eps)
    pert = (e * torch_sign(g)).astype(f32)
    out = (x + pert).astype(f32) if direction > 0 else (x - pert).astype(f32)
    if lo is not None:
        lo = np.asarray(lo, f32)
        hi = np.asarray(hi, f32)
        if lo.ndim == 1:
            lo = lo.reshape(1, -1, 1, 1)
            hi = hi.reshape(1, -1, 1, 1)
        out = _clamp_nan(out, lo, hi)
    return out


def pgd_step(x, x0, g, alpha, eps, lo, hi):
    """pgd_cifar10.py:37-45 / pgd_cifar10_corrupted.py:72-83: step, project to eps-ball, clamp."""
    xa = (x + (f32(alpha) * torch_sign(g)).astype(f32)).astype(f32)
    eta = _clamp_nan((xa - x0).astype(f32), f32(-eps), f32(eps))
    out = (x0 + eta).astype(f32)
    lo = np.asarray(lo, f32)
    hi = np.asarray(hi, f32)
    if lo.ndim == 1:
        lo = lo.reshape(1, -1, 1, 1)
        hi = hi.reshape(1, -1, 1, 1)
    return _clamp_nan(out, lo, hi)
