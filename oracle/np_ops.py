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
# SURVEY 8f-3  per-image Shannon-entropy gate
# =================================================================================================


def shannon_entropy(t: np.ndarray, mean=None, std=None):
    """compute_shannon_entropy for a batch (n,c,h,w) f32 -> (entropy f64[n], counts i64[n,256]).

    fall_2025/TENT_Pipelines.py:99-121 (mean/std given: x*std+mean per channel in f32 with the Python scalars cast to
    f32, then clamp to [0,1]) and fall_2025/Initial_Experiments.py:95-113 (mean is None: values used as they are).
    np.histogram(bins=256, range=(0,1)) on f32 data: values outside [0,1] (and NaN) are not counted, the bin is
    min(floor(256*v), 255) (256*v is exact in f32; the edges k/256 are exact, so numpy's edge fix-ups never fire);
    density=True only rescales.  scipy.stats.entropy(hist[hist>0], base=2) = -sum(p ln p)/ln 2 with p = count/N in f64.
    An image with no countable value gives NaN (0/0), as the reference does."""
    v = np.asarray(t, f32)
    n = v.shape[0]
    if mean is not None:
        m = np.asarray(mean, f32).reshape(1, -1, 1, 1)
        s = np.asarray(std, f32).reshape(1, -1, 1, 1)
        v = (v * s).astype(f32)
        v = (v + m).astype(f32)
        v = np.where(np.isnan(v), v, np.minimum(np.maximum(v, f32(0)), f32(1)))
    v = v.reshape(n, int(np.prod(v.shape[1:])))
    ent = np.empty(n, np.float64)
    counts = np.zeros((n, 256), np.int64)
    for i in range(n):
        r = v[i]
        r = r[(r >= 0) & (r <= 1)]
        idx = np.minimum(np.floor(r * f32(256)).astype(np.int64), 255)
        counts[i] = np.bincount(idx, minlength=256)
        tot = counts[i].sum()
        with np.errstate(invalid="ignore", divide="ignore"):
            p = counts[i][counts[i] > 0] / np.float64(tot)
            ent[i] = -(p * np.log(p)).sum() / np.log(2.0) if tot > 0 else np.nan
    return ent, counts


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


def patch_step(p, g, eps):
    """fall_2025/angellic_patches_fgsm.py:67-69: p -= eps*sign(g); clamp_(0, 1) (f32)."""
    return fgsm_step(p, g, eps, -1.0, 0.0, 1.0)


def odin_perturb(x, g, eps):
    """fall_2025/generalizedOdin2.py:47: x - eps*sign(-g) (f32, no clamp)."""
    s = torch_sign(-np.asarray(g, np.float32))
    return (np.asarray(x, np.float32) - (np.float32(eps) * s).astype(np.float32)).astype(np.float32)


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


# =================================================================================================
# A.3  Pillow ImagingResample (Image.resize): two passes, Q22 coefficients, uint8 intermediate
# =================================================================================================
BOX, BILINEAR, BICUBIC, LANCZOS = 0, 1, 2, 3
_SUPPORT = {BOX: 0.5, BILINEAR: 1.0, BICUBIC: 2.0, LANCZOS: 3.0}


def _filter_value(filt: int, x: float) -> float:
    if filt == BOX:
        return 1.0 if -0.5 < x <= 0.5 else 0.0
    if filt == BILINEAR:
        x = abs(x)
        return 1.0 - x if x < 1.0 else 0.0
    if filt == BICUBIC:
        a = -0.5
        x = abs(x)
        if x < 1.0:
            return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
        if x < 2.0:
            return (((x - 5) * x + 8) * x - 4) * a
        return 0.0
    if filt == LANCZOS:
        def sinc(v):
            if v == 0.0:
                return 1.0
            v = v * math.pi
            return math.sin(v) / v
        if -3.0 <= x < 3.0:
            return sinc(x) * sinc(x / 3)
        return 0.0
    raise ValueError(filt)


def resample_coeffs(in_size: int, out_size: int, filt: int, box=None):
    """libImaging/Resample.c precompute_coeffs + normalize_coeffs_8bpc.
    Returns (xmin[out], n[out], kk[out, ksize] int32 Q22)."""
    in0, in1 = (0.0, float(in_size)) if box is None else box
    scale = filterscale = (in1 - in0) / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = _SUPPORT[filt] * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    xmins = np.zeros(out_size, np.int32)
    ns = np.zeros(out_size, np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = in0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [_filter_value(filt, (x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        for x in range(xmax):
            if ww != 0.0:
                w[x] /= ww
        xmins[xx] = xmin
        ns[xx] = xmax
        for x in range(xmax):
            v = w[x]
            kk[xx, x] = int(-0.5 + v * (1 << 22)) if v < 0 else int(0.5 + v * (1 << 22))
    return xmins, ns, kk


def _resample_axis(img: np.ndarray, out_size: int, filt: int, axis: int) -> np.ndarray:
    """One pass along `axis` (1 = vertical/H, 2 = horizontal/W) of an (n,h,w,c) uint8 batch."""
    in_size = img.shape[axis]
    xmins, ns, kk = resample_coeffs(in_size, out_size, filt)
    shp = list(img.shape)
    shp[axis] = out_size
    out = np.empty(shp, np.uint8)
    src = img.astype(np.int64)
    for xx in range(out_size):
        x0, n = int(xmins[xx]), int(ns[xx])
        k = kk[xx, :n].astype(np.int64)
        if axis == 2:
            acc = (src[:, :, x0:x0 + n, :] * k[None, None, :, None]).sum(axis=2)
            out[:, :, xx, :] = np.clip((acc + (1 << 21)) >> 22, 0, 255)
        else:
            acc = (src[:, x0:x0 + n, :, :] * k[None, :, None, None]).sum(axis=1)
            out[:, xx, :, :] = np.clip((acc + (1 << 21)) >> 22, 0, 255)
    return out


def pil_resize(img: np.ndarray, size, filt: int = BICUBIC) -> np.ndarray:
    """Image.resize((ow, oh), filt): horizontal pass first, each pass skipped if the size is unchanged."""
    ow, oh = size
    out = img
    if ow != img.shape[2]:
        out = _resample_axis(out, ow, filt, 2)
    if oh != img.shape[1]:
        out = _resample_axis(out, oh, filt, 1)
    return out.copy()


def apply_scale(img: np.ndarray, s: float) -> np.ndarray:
    """transformation.py:173-196: LANCZOS resize to int(w*s) then centre crop / zero-pad paste."""
    n, h, w, c = img.shape
    nw, nh = int(w * s), int(h * s)
    sc = pil_resize(img, (nw, nh), LANCZOS)
    if s > 1.0:
        l, t = (nw - w) // 2, (nh - h) // 2
        return sc[:, t:t + h, l:l + w].copy()
    if s < 1.0:
        out = np.zeros_like(img)
        px, py = (w - nw) // 2, (h - nh) // 2
        out[:, py:py + nh, px:px + nw] = sc
        return out
    return sc


def crop_resize(img: np.ndarray, x, y, cs: int, out_size: int = 32, filt: int = BICUBIC) -> np.ndarray:
    """rand_crop — tta_initial_claudeFixed.py:56-62: crop (x,y,x+cs,y+cs) then resize((32,32)) (BICUBIC default)."""
    n = img.shape[0]
    x = np.broadcast_to(np.asarray(x), (n,))
    y = np.broadcast_to(np.asarray(y), (n,))
    crops = np.stack([img[i, y[i]:y[i] + cs, x[i]:x[i] + cs] for i in range(n)])
    return pil_resize(crops, (out_size, out_size), filt)


def resize_center_crop(img: np.ndarray, size: int, crop: int) -> np.ndarray:
    """T.Resize(size) + T.CenterCrop(crop) on PIL (BILINEAR) — TTA_transforms.py:104-113."""
    n, h, w, c = img.shape
    if w <= h:
        ow, oh = size, int(size * h / w)
    else:
        oh, ow = size, int(size * w / h)
    r = pil_resize(img, (ow, oh), BILINEAR)
    top = int(round((oh - crop) / 2.0))
    left = int(round((ow - crop) / 2.0))
    return r[:, top:top + crop, left:left + crop].copy()


# =================================================================================================
# A.2  Pillow generic transform with BILINEAR / BICUBIC filters (f64)
# =================================================================================================


def _bicubic_w(v1, v2, v3, v4, d):
    p1 = v2
    p2 = -v1 + v3
    p3 = 2 * (v1 - v2) + v3 - v4
    p4 = -v1 + v2 - v3 + v4
    return p1 + d * (p2 + d * (p3 + d * p4))


def pil_transform(img: np.ndarray, coeffs, out_size, mode="affine", resample="bicubic", fill=0) -> np.ndarray:
    """Image.transform(size, AFFINE|PERSPECTIVE, coeffs, resample=BILINEAR|BICUBIC, fillcolor=fill).
    libImaging/Geometry.c ImagingGenericTransform + bilinear_filter8 / bicubic_filter8.
    reference: transformation.py:212-226 (apply_shear), fall_2025/Initial_Experiments.py:88-92."""
    n, h, w, c = img.shape
    ow, oh = out_size
    a = [float(v) for v in coeffs]
    out = np.empty((n, oh, ow, c), np.uint8)
    out[:] = np.broadcast_to(np.asarray(fill, np.uint8), (c,))
    ys, xs = np.mgrid[0:oh, 0:ow]
    xin = xs + 0.5
    yin = ys + 0.5
    if mode == "affine":
        sx = a[0] * xin + a[1] * yin + a[2]
        sy = a[3] * xin + a[4] * yin + a[5]
    else:
        den = a[6] * xin + a[7] * yin + 1
        sx = (a[0] * xin + a[1] * yin + a[2]) / den
        sy = (a[3] * xin + a[4] * yin + a[5]) / den
    valid = ~((sx < 0.0) | (sx >= w) | (sy < 0.0) | (sy >= h))
    sx = sx - 0.5
    sy = sy - 0.5
    x0 = np.floor(sx).astype(np.int64)
    y0 = np.floor(sy).astype(np.int64)
    dx = sx - x0
    dy = sy - y0
    src = img.astype(np.float64)

    def px(yy, xx):
        return src[:, yy, np.clip(xx, 0, w - 1), :]  # (n, oh, ow, c)

    if resample == "bilinear":
        def row(yy):
            yyc = np.clip(yy, 0, h - 1)
            l = src[:, yyc, np.clip(x0, 0, w - 1), :]
            r = src[:, yyc, np.clip(x0 + 1, 0, w - 1), :]
            return l + (r - l) * dx[None, :, :, None]
        v1 = row(y0)
        ok2 = ((y0 + 1) >= 0) & ((y0 + 1) < h)
        v2 = np.where(ok2[None, :, :, None], row(y0 + 1), v1)
        v = v1 + (v2 - v1) * dy[None, :, :, None]
        res = v.astype(np.int64).astype(np.uint8)  # (UINT8) v
    else:
        def row(yy):
            yyc = np.clip(yy, 0, h - 1)
            t = [src[:, yyc, np.clip(x0 + k, 0, w - 1), :] for k in (-1, 0, 1, 2)]
            return _bicubic_w(t[0], t[1], t[2], t[3], dx[None, :, :, None])
        # rows y0-1 .. y0+2; a row outside the image reuses the previous row's value (v1 always computed
        # from clamp of y0-1? no: libImaging uses in = im->image8[YCLIP(im, y)] for first row only)
        v1 = row(y0 - 1)
        def ok(k):
            yy = y0 + k
            return ((yy >= 0) & (yy < h))[None, :, :, None]
        v2 = np.where(ok(0), row(y0), v1)
        v3 = np.where(ok(1), row(y0 + 1), v2)
        v4 = np.where(ok(2), row(y0 + 2), v3)
        v = _bicubic_w(v1, v2, v3, v4, dy[None, :, :, None])
        res = np.where(v <= 0.0, 0, np.where(v >= 255.0, 255, np.trunc(np.clip(v, 0, 255)))).astype(np.uint8)
    m = valid[None, :, :, None]
    return np.where(m, res, out)


def apply_shear(img: np.ndarray, sf: float) -> np.ndarray:
    """transformation.py:212-226."""
    n, h, w, c = img.shape
    shift = int(math.ceil(sf * h))
    return pil_transform(img, (1, sf, -shift if sf > 0 else 0, 0, 1, 0), (w + shift, h), "affine", "bicubic", 255)


# =================================================================================================
# A.6  cv2.GaussianBlur on uint8 (Q8 separable, REFLECT_101)
# =================================================================================================


def cv_gaussian_kernel_q8(ksize: int, sigma: float) -> np.ndarray:
    """OpenCV getGaussianKernelBitExact for CV_8U: 64-bit kernel quantised to Q8 with error diffusion."""
    n = ksize
    x = np.arange(n, dtype=np.float64) - (n - 1) * 0.5
    scale2x = -0.5 / (sigma * sigma)
    g = np.exp(scale2x * x * x)
    g = g / g.sum()
    k = np.zeros(n, np.int64)
    err = 0.0
    for i in range(n // 2):
        adj = g[i] * 256 + err
        v = float(np.rint(adj))
        err = adj - v
        k[i] = k[n - 1 - i] = int(v)
    k[n // 2] = 256 - (k.sum())
    return k


def _reflect101_idx(idx, n):
    idx = np.asarray(idx)
    if n == 1:
        return np.zeros_like(idx)
    period = 2 * n - 2
    m = np.mod(idx, period)
    return np.where(m >= n, period - m, m)


def cv_gaussian_blur(img: np.ndarray, ksize: int, sigma: float) -> np.ndarray:
    """cv2.GaussianBlur(img, (ksize,ksize), sigma) for uint8: fixed-point path (ufixedpoint16)."""
    n, h, w, c = img.shape
    k = cv_gaussian_kernel_q8(ksize, sigma)
    r = ksize // 2
    xi = _reflect101_idx(np.arange(-r, w + r), w)
    yi = _reflect101_idx(np.arange(-r, h + r), h)
    src = img.astype(np.int64)
    hp = np.zeros((n, h, w, c), np.int64)
    for i in range(ksize):
        hp += k[i] * src[:, :, xi[i:i + w], :]
    hp = np.minimum(hp, 65535)
    vp = np.zeros((n, h, w, c), np.int64)
    for i in range(ksize):
        vp += k[i] * hp[:, yi[i:i + h], :, :]
    return np.clip((vp + (1 << 15)) >> 16, 0, 255).astype(np.uint8)


def apply_blur(img: np.ndarray, r: float) -> np.ndarray:
    """transformation.py:228-257."""
    if r == 0:
        return img.copy()
    ks = int(r * 6)
    if ks % 2 == 0:
        ks += 1
    if ks < 3:
        ks = 3
    return cv_gaussian_blur(img, ks, r)


# =================================================================================================
# A.6b  Pillow ImageFilter.GaussianBlur (3 extended box passes per axis)
# =================================================================================================


def pil_gaussian_box_radius(radius: float) -> float:
    """libImaging/BoxBlur.c _gaussian_blur_radius with n=3 passes, in C float/double mix."""
    sigma2 = f32(radius * radius / 3)  # float sigma2 = radius*radius/passes
    L = f32(math.sqrt(12.0 * float(sigma2) + 1.0))
    l = f32(math.floor((float(L) - 1.0) / 2.0))
    a = f32((2 * l + 1) * (l * (l + 1) - 3 * sigma2))
    a = f32(a / f32(6 * (sigma2 - (l + 1) * (l + 1))))
    return float(f32(l + a))


def _box_pass_x(img: np.ndarray, R: float) -> np.ndarray:
    """One ImagingLineBoxBlur8 pass along axis 2 (x) with edge replication."""
    n, h, w, c = img.shape
    ri = int(R)
    ww = int(np.uint32(f32(1 << 24) / f32(f32(R) * 2 + 1)))
    fw = ((1 << 24) - (ri * 2 + 1) * ww) // 2
    src = img.astype(np.int64)
    idx = np.arange(w)
    acc = np.zeros((n, h, w, c), np.int64)
    for i in range(-ri, ri + 1):
        acc += src[:, :, np.clip(idx + i, 0, w - 1), :]
    far = src[:, :, np.clip(idx - ri - 1, 0, w - 1), :] + src[:, :, np.clip(idx + ri + 1, 0, w - 1), :]
    v = (acc * ww + far * fw) & 0xFFFFFFFF
    v = ((v + (1 << 23)) & 0xFFFFFFFF) >> 24
    return v.astype(np.uint8)


def pil_gaussian_blur(img: np.ndarray, radius: float) -> np.ndarray:
    """img.filter(ImageFilter.GaussianBlur(radius)) — pipenline/cifar_image_transformations.py:72-77."""
    R = pil_gaussian_box_radius(radius)
    out = img
    for _ in range(3):
        out = _box_pass_x(out, R)
    out = out.transpose(0, 2, 1, 3)
    for _ in range(3):
        out = _box_pass_x(out, R)
    return out.transpose(0, 2, 1, 3).copy()


# =================================================================================================
# A.6c  motion_blur (cv2.filter2D horizontal box) and A.14 histogram_equalization
# =================================================================================================


def motion_blur_box(img: np.ndarray, size: int) -> np.ndarray:
    """cv2.filter2D(img, -1, K) with K = size x size, middle row 1/size
    (pipenline/cifar_image_transformations.py:109-119): horizontal box, REFLECT_101, f32 accumulate,
    round-half-even."""
    n, h, w, c = img.shape
    r = (size - 1) // 2
    xi = _reflect101_idx(np.arange(-r, w + r), w)
    kv = f32(np.float64(1.0) / size)
    src = img.astype(f32)
    acc = np.zeros((n, h, w, c), f32)
    for i in range(size):
        acc = (acc + (src[:, :, xi[i:i + w], :] * kv).astype(f32)).astype(f32)
    return np.clip(np.rint(acc), 0, 255).astype(np.uint8)


def _D14(x):
    return (x + 8192) >> 14


def histogram_equalization(img: np.ndarray) -> np.ndarray:
    """RGB->YUV (OpenCV Q14), equalizeHist(Y), YUV->RGB — cifar_image_transformations.py:122-129."""
    r = img[..., 0].astype(np.int64)
    g = img[..., 1].astype(np.int64)
    b = img[..., 2].astype(np.int64)
    Y = _D14(4899 * r + 9617 * g + 1868 * b)
    U = np.clip(_D14((b - Y) * 8061 + (128 << 14)), 0, 255)
    V = np.clip(_D14((r - Y) * 14369 + (128 << 14)), 0, 255)
    Y = np.clip(Y, 0, 255)
    out = np.empty_like(img)
    for i in range(img.shape[0]):
        y = Y[i]
        hist = np.bincount(y.ravel(), minlength=256)
        total = y.size
        nz = np.nonzero(hist)[0]
        first = int(nz[0])
        if hist[first] == total:
            yeq = np.full_like(y, first)
        else:
            scale = f32(255.0) / f32(total - hist[first])
            lut = np.zeros(256, np.int64)
            s = 0
            for j in range(first + 1, 256):
                s += int(hist[j])
                lut[j] = min(max(int(np.rint(f32(s) * scale)), 0), 255)
            yeq = lut[y]
        u = U[i] - 128
        v = V[i] - 128
        R = yeq + _D14(v * 18678)
        G = yeq + _D14(v * (-9519) + u * (-6472))
        Bc = yeq + _D14(u * 33292)
        out[i, ..., 0] = np.clip(R, 0, 255)
        out[i, ..., 1] = np.clip(G, 0, 255)
        out[i, ..., 2] = np.clip(Bc, 0, 255)
    return out


# =================================================================================================
# A.8  noise (draws injected)
# =================================================================================================


def gaussian_noise_f32(img: np.ndarray, noise_f32: np.ndarray) -> np.ndarray:
    """transformation.apply_gaussian_noise (:272-281): noise already scaled (std*255) and cast to f32."""
    v = (img.astype(f32) + noise_f32.astype(f32)).astype(f32)
    return np.clip(v, 0, 255).astype(np.uint8)


def gaussian_noise_f64(img: np.ndarray, noise_f64: np.ndarray) -> np.ndarray:
    """TransformationPool.gaussian_noise (:39-48): f32 image + f64 noise (already scaled) summed in f64."""
    v = img.astype(f32).astype(np.float64) + noise_f64
    return np.clip(v, 0, 255).astype(np.uint8)


def impulse_noise(img: np.ndarray, mask: np.ndarray, prob: float) -> np.ndarray:
    """TransformationPool.impulse_noise (:50-59): mask (n,h,w) uniform f64 shared by channels."""
    out = img.copy()
    out[mask < prob / 2] = 0
    out[mask > 1 - prob / 2] = 255
    return out


def shot_noise_scaled(img: np.ndarray, lam: float) -> np.ndarray:
    """The Poisson rate array the reference hands to np.random.poisson (f32)."""
    return (img.astype(f32) / f32(255.0) * f32(lam)).astype(f32)


def shot_noise(img: np.ndarray, counts: np.ndarray, lam: float) -> np.ndarray:
    """TransformationPool.shot_noise (:61-70) with the Poisson draws injected: k/lam*255.0 in f64."""
    v = counts.astype(np.float64) / lam * 255.0
    return np.clip(v, 0, 255).astype(np.uint8)
