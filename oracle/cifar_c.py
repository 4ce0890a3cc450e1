"""CPU oracle for the CIFAR-10-C-style 15-corruption x 5-severity suite.

TEST INFRASTRUCTURE ONLY (see oracle/np_ops.py header).

PARITY STATUS.  The reference only *names* the 15 corruption types
(attacks/fgsm/cifar-10/reverse_fgsm_cifar10_corrupted.py:18-23) and loads Hendrycks & Dietterich's pre-made .npy
files; it contains code for four of them only, with ImageNet-C constants (pipenline/cifar_image_transformations.py:
39-77, restated and pinned in oracle/np_ops.py / np_chains.py).  The canonical generator (make_cifar_c.py of
hendrycks/robustness, not vendored, not pinned; depends on skimage and Wand/ImageMagick, neither installed) is
restated here from its published algorithm.  => "parity unpinned" for this file, with three partial pins, all in
tests/test_oracle_cifar_c.py:
  (1) every function with a library counterpart available in the container (cv2.filter2D, scipy.ndimage
      gaussian_filter / zoom / map_coordinates, cv2.warpAffine, PIL BOX resize, PIL/libjpeg-turbo JPEG round trip)
      is checked against that library call (<= 1 LSB; pixelate and jpeg_compression bit-exact);
  (2) the deterministic types are checked against the reference's real CIFAR-10-C fixtures
      (CIFAR-10-C/store/*.png, all severity 1) with the clean-image estimate of SURVEY App. B, at the noise floor;
  (3) the random draws are explicit inputs, so statistics of the on-chip (Philox) mode are checked separately.

ARITHMETIC.  Each function states its arithmetic explicitly (op order, f32 vs f64) so that the CUDA kernels can be
bit-identical to it.  Pointwise types whose canonical code is float64 (noise, brightness, contrast, fog, frost) are
f64 here; defocus / glass / the elastic displacement field follow the library routines' own f64 summation order (bit-identical
to cv2.filter2D / scipy.ndimage.gaussian_filter in the tests); zoom_blur is specified as ONE merged sparse linear operator
per severity applied in f32 (the canonical per-zoom loop, restated as zoom_blur_lib and bit-identical to scipy.ndimage.zoom,
is kept as its cross-check, <= 1 LSB); motion_blur and snow (ImageMagick in the canonical code) are f32 restatements.  Output everywhere: np.uint8(v) truncation of a value in [0, 255].

Array conventions: x is (n, 32, 32, 3) uint8 (any h, w where the canonical code allows); draws are per image.
"""
from __future__ import annotations

import math

import numpy as np

from . import np_ops as O

f32 = np.float32
f64 = np.float64

CORRUPTIONS = ["gaussian_noise", "shot_noise", "impulse_noise", "defocus_blur", "glass_blur", "motion_blur",
               "zoom_blur", "snow", "frost", "fog", "brightness", "contrast", "elastic_transform", "pixelate",
               "jpeg_compression"]

# canonical CIFAR-10-C severity tables ([recalled])
PARAMS = {
    "gaussian_noise": [0.04, 0.06, 0.08, 0.09, 0.10],
    "shot_noise": [500, 250, 100, 75, 50],
    "impulse_noise": [0.01, 0.02, 0.03, 0.05, 0.07],
    "speckle_noise": [0.06, 0.1, 0.12, 0.16, 0.2],
    "defocus_blur": [(0.3, 0.4), (0.4, 0.5), (0.5, 0.6), (1, 0.2), (1.5, 0.1)],
    "glass_blur": [(0.05, 1, 1), (0.25, 1, 1), (0.4, 1, 1), (0.25, 1, 2), (0.4, 1, 2)],
    "motion_blur": [(10, 1), (10, 1.5), (10, 2), (10, 2.5), (12, 3)],
    "zoom_blur": [(1, 1.06), (1, 1.11), (1, 1.16), (1, 1.21), (1, 1.26)],  # np.arange(lo, hi, 0.01)
    "snow": [(0.1, 0.2, 1, 0.6, 8, 3, 0.95), (0.1, 0.2, 1, 0.5, 10, 4, 0.9), (0.15, 0.3, 1.75, 0.55, 10, 4, 0.9),
             (0.25, 0.3, 2.25, 0.6, 12, 6, 0.85), (0.3, 0.3, 1.25, 0.65, 14, 12, 0.8)],
    "frost": [(1, 0.2), (1, 0.3), (0.9, 0.4), (0.85, 0.4), (0.75, 0.45)],
    "fog": [(0.2, 3), (0.5, 3), (0.75, 2.5), (1, 2), (1.5, 1.75)],
    "brightness": [0.05, 0.1, 0.15, 0.2, 0.3],
    "contrast": [0.75, 0.5, 0.4, 0.3, 0.15],
    "elastic_transform": [(0, 0, 0.08), (0.05, 0.2, 0.07), (0.08, 0.06, 0.06), (0.1, 0.04, 0.05), (0.1, 0.03, 0.03)],
    "pixelate": [0.95, 0.9, 0.85, 0.75, 0.65],
    "jpeg_compression": [80, 65, 58, 50, 40],
}


def _u8(v):
    return np.asarray(v).astype(np.int64).astype(np.uint8)  # values already clipped to [0,255]: truncation


def _reflect101(i, n):
    return O._reflect101_idx(i, n)


# =================================================================================================
# noise (float64, pointwise)
# =================================================================================================
def gaussian_noise(x, noise):
    """clip(x/255 + N(0,c), 0, 1)*255; `noise` = np.random.normal(size=x.shape, scale=c) (f64, already scaled)."""
    return _u8(np.clip(x / 255.0 + noise, 0, 1) * 255)


def shot_rate(x, c):
    return x / 255.0 * c


def shot_noise(x, counts, c):
    """clip(Poisson(x/255*c)/c, 0, 1)*255; `counts` = np.random.poisson(shot_rate(x, c))."""
    return _u8(np.clip(counts / f64(c), 0, 1) * 255)


def impulse_noise(x, u_flip, u_salt, c):
    """skimage.util.random_noise(mode='s&p', amount=c): per ELEMENT, flipped = u_flip <= c, salted = u_salt <= 0.5;
    flipped&salted -> 1, flipped&~salted -> 0; then clip*255 (identity on untouched elements)."""
    out = x / 255.0
    flipped = u_flip <= c
    salted = u_salt <= 0.5
    out = np.where(flipped & salted, 1.0, out)
    out = np.where(flipped & ~salted, 0.0, out)
    return _u8(np.clip(out, 0, 1) * 255)


def speckle_noise(x, noise):
    """clip(x + x*N(0,c), 0, 1)*255 on x/255 (extra type present in CIFAR-10-C/store)."""
    v = x / 255.0
    return _u8(np.clip(v + v * noise, 0, 1) * 255)


# =================================================================================================
# defocus_blur: disk kernel (anti-aliased with a 3x3 Gaussian) applied per channel, REFLECT_101
# =================================================================================================
def gaussian_kernel3_f32(sigma):
    """cv2.getGaussianKernel(3, sigma, CV_32F): exp in double, normalised in double, stored as f32."""
    cf = np.exp(-0.5 / (sigma * sigma) * np.array([1.0, 0.0, 1.0]))
    return (cf / cf.sum()).astype(f32)


def disk_kernel(radius, alias_blur):
    """make_cifar_c.disk(): (X^2+Y^2 <= r^2) on the 17x17 grid L=-8..8, normalised (f32), then
    cv2.GaussianBlur(ksize=(3,3), sigmaX=alias_blur) (separable f32, symmetric form, REFLECT_101)."""
    if radius <= 8:
        L = np.arange(-8, 8 + 1)
    else:
        raise NotImplementedError
    X, Y = np.meshgrid(L, L)
    d = np.array((X ** 2 + Y ** 2) <= radius ** 2, dtype=f32)
    d /= np.sum(d)
    k = gaussian_kernel3_f32(alias_blur)

    def sym3(a, axis):
        n = a.shape[axis]
        idx = _reflect101(np.arange(-1, n + 1), n)
        lo = np.take(a, idx[0:n], axis)
        hi = np.take(a, idx[2:n + 2], axis)
        return (a * k[1] + (lo + hi).astype(f32) * k[0]).astype(f32)

    return sym3(sym3(d, 1), 0)


def kernel_taps(kern):
    """Non-zero taps of a small correlation kernel as [(dy, dx, w)] in row-major order."""
    c = kern.shape[0] // 2
    ys, xs = np.nonzero(kern)
    return [(int(y) - c, int(x) - c, f32(kern[y, x])) for y, x in zip(ys, xs)]


def filter2d_reflect_f64(src, taps):
    """Correlation of (n,h,w,c) f64 with sparse taps (f32 weights widened), BORDER_REFLECT_101, f64 accumulation in
    row-major tap order — bit-identical to cv2.filter2D on a float64 image for these kernels (tests)."""
    n, h, w, c = src.shape
    acc = np.zeros_like(src)
    for dy, dx, wt in taps:
        yi = _reflect101(np.arange(h) + dy, h)
        xi = _reflect101(np.arange(w) + dx, w)
        acc = acc + src[:, yi][:, :, xi] * f64(wt)
    return acc


def defocus_blur(x, radius, alias):
    """cv2.filter2D(x/255, -1, disk(radius, alias)) per channel; clip*255.  f64."""
    acc = filter2d_reflect_f64(x / 255.0, kernel_taps(disk_kernel(radius, alias)))
    return _u8(np.clip(acc, 0, 1) * 255)


# =================================================================================================
# glass_blur: Gaussian -> uint8 -> sequential local pixel swaps -> Gaussian
# =================================================================================================
def gaussian_weights(sigma, truncate):
    """scipy.ndimage._gaussian_kernel1d(sigma, 0, int(truncate*sigma+0.5)), f64 (index 0 = centre tap)."""
    lw = int(truncate * float(sigma) + 0.5)
    full = np.exp(-0.5 / (sigma * sigma) * np.arange(-lw, lw + 1, dtype=f64) ** 2)
    full = full / full.sum()
    return full[lw:]


def sym_filter(a, wts, axis, mode):
    """scipy correlate1d with a symmetric kernel: tmp = x[l]*w0; for k = lw..1: tmp += (x[l-k] + x[l+k])*w_k.
    mode 'nearest' (edge replicate) or 'reflect' (half-sample symmetric).  f64; bit-identical to
    scipy.ndimage.gaussian_filter (tests)."""
    n = a.shape[axis]
    lw = len(wts) - 1
    base = np.arange(n)

    def idx(i):
        if mode == "nearest":
            return np.clip(i, 0, n - 1)
        period = 2 * n
        m = np.mod(i, period)
        return np.where(m >= n, period - 1 - m, m)

    acc = a * wts[0]
    for k in range(lw, 0, -1):
        acc = acc + (np.take(a, idx(base - k), axis) + np.take(a, idx(base + k), axis)) * wts[k]
    return acc


def gaussian_filter_sym(a, sigma, mode, truncate, axes=(1, 2)):
    """skimage.filters.gaussian / scipy gaussian_filter over the spatial axes (first axis first); sigma == 0 is the
    identity (scipy skips axes with sigma <= 1e-15)."""
    if sigma <= 1e-15:
        return a.copy()
    wts = gaussian_weights(sigma, truncate)
    for ax in axes:
        a = sym_filter(a, wts, ax, mode)
    return a


def glass_swaps(u, deltas):
    """make_cifar_c.glass_blur inner loops.  u (n,32,32,3) uint8, modified copy returned.
    deltas (n, iters, 30, 30, 2) in {-1, 0}: np.random.randint(-1, 1, size=(2,)) -> (dx, dy) for h=31..2, w=31..2."""
    u = u.copy()
    n, H, W, _ = u.shape
    for i in range(n):
        for it in range(deltas.shape[1]):
            for a, h in enumerate(range(H - 1, 1, -1)):
                for b, w in enumerate(range(W - 1, 1, -1)):
                    dx, dy = int(deltas[i, it, a, b, 0]), int(deltas[i, it, a, b, 1])
                    hp, wp = h + dy, w + dx
                    t = u[i, h, w].copy()
                    u[i, h, w] = u[i, hp, wp]
                    u[i, hp, wp] = t
    return u


def glass_blur(x, sigma, iters, deltas):
    g1 = gaussian_filter_sym(x / 255.0, sigma, "nearest", 4.0)
    u = _u8(g1 * 255)
    u = glass_swaps(u, deltas[:, :iters])
    g2 = gaussian_filter_sym(u / 255.0, sigma, "nearest", 4.0)
    return _u8(np.clip(g2, 0, 1) * 255)


# =================================================================================================
# motion_blur: ImageMagick MotionBlurImage restated (one-sided Gaussian line kernel, edge virtual pixels)
# =================================================================================================
def motion_kernel(radius, sigma, angle_deg):
    """(weights f32 [width], offsets int [width,2] (dx, dy)).  width = 2*ceil(radius)+1;
    k[i] = exp(-i^2/(2 sigma^2))/(sqrt(2 pi) sigma), normalised; point = width*(sin a, cos a);
    offset[i] = (ceil(i*point.y/hypot - 0.5), ceil(i*point.x/hypot - 0.5))."""
    width = int(2.0 * math.ceil(radius) + 1.0)
    i = np.arange(width, dtype=f64)
    k = np.exp(-(i * i) / (2.0 * sigma * sigma)) / (math.sqrt(2.0 * math.pi) * sigma)
    k = k / k.sum()
    a = math.radians(angle_deg)
    px, py = width * math.sin(a), width * math.cos(a)
    hyp = math.hypot(px, py)
    off = np.stack([np.ceil(i * py / hyp - 0.5), np.ceil(i * px / hyp - 0.5)], axis=1).astype(np.int64)
    return k.astype(f32), off


def line_blur_f32(src, k, off):
    """sum_i k[i] * src[clamp(y+dy_i), clamp(x+dx_i)] in f32, tap order i = 0..width-1.  src (h,w,c) f32."""
    h, w = src.shape[:2]
    acc = np.zeros_like(src)
    for wt, (dx, dy) in zip(k, off):
        yi = np.clip(np.arange(h) + dy, 0, h - 1)
        xi = np.clip(np.arange(w) + dx, 0, w - 1)
        acc = (acc + (src[yi][:, xi] * wt).astype(f32)).astype(f32)
    return acc


def motion_blur(x, radius, sigma, angles):
    """angles (n,) degrees ~ U(-45, 45).  Result rounded to the nearest integer like an 8-bit ImageMagick write."""
    out = np.empty_like(x)
    for i in range(x.shape[0]):
        k, off = motion_kernel(radius, sigma, float(angles[i]))
        acc = line_blur_f32(x[i].astype(f32), k, off)
        out[i] = _u8(np.clip(np.floor((acc + f32(0.5)).astype(f32)), 0, 255))
    return out


# =================================================================================================
# zoom_blur: mean of the image and its centre zooms (order-1 spline, "align corners" mapping of scipy.ndimage.zoom)
# =================================================================================================
def zoom_factors(severity):
    lo, hi = PARAMS["zoom_blur"][severity - 1]
    return np.arange(lo, hi, 0.01)


def zoom_axis_table(size, z):
    """Per output index of clipped_zoom(img, z) along one axis: (i0, i1, t) into the ORIGINAL axis: crop of
    ch = ceil(size/z) at top = (size-ch)//2, zoom to round(ch*z) samples with coordinate o*(ch-1)/(out-1), centre
    trim to `size`."""
    ch = int(np.ceil(size / z))
    top = (size - ch) // 2
    osz = int(round(ch * z))
    trim = (osz - size) // 2
    o = np.arange(size) + trim
    if osz > 1:
        cc = o.astype(f64) * (f64(ch - 1) / f64(osz - 1))
    else:
        cc = np.zeros(size)
    i0 = np.floor(cc).astype(np.int64)
    t = (cc - i0).astype(f32)
    i0 = np.clip(i0, 0, ch - 1)
    i1 = np.minimum(i0 + 1, ch - 1)
    return i0 + top, i1 + top, t


def clipped_zoom_f32(src, z):
    """src (n,h,w,c) f32.  v = ((p00*(1-ty))*(1-tx) + (p01*(1-ty))*tx) + (p10*ty)*(1-tx) + (p11*ty)*tx, f32,
    accumulated left to right."""
    n, h, w, c = src.shape
    y0, y1, ty = zoom_axis_table(h, z)
    x0, x1, tx = zoom_axis_table(w, z)
    ty = ty[None, :, None, None]
    tx = tx[None, None, :, None]
    uy = (f32(1) - ty).astype(f32)
    ux = (f32(1) - tx).astype(f32)
    p00 = src[:, y0][:, :, x0]
    p01 = src[:, y0][:, :, x1]
    p10 = src[:, y1][:, :, x0]
    p11 = src[:, y1][:, :, x1]
    acc = ((p00 * uy).astype(f32) * ux).astype(f32)
    acc = (acc + ((p01 * uy).astype(f32) * tx).astype(f32)).astype(f32)
    acc = (acc + ((p10 * ty).astype(f32) * ux).astype(f32)).astype(f32)
    acc = (acc + ((p11 * ty).astype(f32) * tx).astype(f32)).astype(f32)
    return acc


def clipped_zoom_lib(src, z):
    """scipy's own arithmetic: interpolation weights and sum in f64 over the f32 image, result cast to f32 —
    bit-identical to scipy.ndimage.zoom(order=1) (tests).  src (n,h,w,c) f32."""
    n, h, w, c = src.shape

    def tab(size):
        ch = int(np.ceil(size / z))
        top = (size - ch) // 2
        osz = int(round(ch * z))
        trim = (osz - size) // 2
        cc = (np.arange(size) + trim).astype(f64) * (f64(ch - 1) / f64(osz - 1)) if osz > 1 else np.zeros(size)
        i0 = np.floor(cc).astype(np.int64)
        t = cc - i0
        i0 = np.clip(i0, 0, ch - 1)
        return i0 + top, np.minimum(i0 + 1, ch - 1) + top, t

    y0, y1, ty = tab(h)
    x0, x1, tx = tab(w)
    ty = ty[None, :, None, None]
    tx = tx[None, None, :, None]
    s = src.astype(f64)
    acc = (s[:, y0][:, :, x0] * (1 - ty)) * (1 - tx)
    acc = acc + (s[:, y0][:, :, x1] * (1 - ty)) * tx
    acc = acc + (s[:, y1][:, :, x0] * ty) * (1 - tx)
    acc = acc + (s[:, y1][:, :, x1] * ty) * tx
    return acc.astype(f32)


def zoom_blur_lib(x, severity):
    """The canonical evaluation order: out += clipped_zoom(x, z) per z in f32; (x + out)/(n+1); clip*255."""
    src = (x.astype(f32) / f32(255)).astype(f32)
    zs = zoom_factors(severity)
    out = np.zeros_like(src)
    for z in zs:
        out = (out + clipped_zoom_lib(src, float(z))).astype(f32)
    v = ((src + out).astype(f32) / f32(len(zs) + 1)).astype(f32)
    return _u8((np.clip(v, 0, 1) * f32(255)).astype(f32))


def zoom_operator(h, w, severity):
    """zoom_blur is one fixed linear map per severity: out[p] = sum_q W[p,q] src[q] with
    W = (I + sum_z Z_z)/(n+1).  Returns CSR (indptr int32 [h*w+1], indices int32, weights f32) with the taps of each
    output pixel merged (weights summed in f64) and sorted by source index."""
    zs = zoom_factors(severity)
    W = np.zeros((h * w, h * w), f64)
    rows = np.arange(h * w)
    W[rows, rows] += 1.0
    for z in zs:
        y0, y1, ty = zoom_axis_table(h, float(z))
        x0, x1, tx = zoom_axis_table(w, float(z))
        ty, tx = ty.astype(f64), tx.astype(f64)
        for ya, wy in ((y0, 1 - ty), (y1, ty)):
            for xa, wx in ((x0, 1 - tx), (x1, tx)):
                q = (ya[:, None] * w + xa[None, :]).reshape(-1)
                np.add.at(W, (rows, q), (wy[:, None] * wx[None, :]).reshape(-1))
    W /= (len(zs) + 1)
    indptr = [0]
    indices, weights = [], []
    for p in range(h * w):
        nz = np.nonzero(W[p])[0]
        indices.append(nz)
        weights.append(W[p, nz])
        indptr.append(indptr[-1] + len(nz))
    return (np.array(indptr, np.int32), np.concatenate(indices).astype(np.int32),
            np.concatenate(weights).astype(f32))


def zoom_operator_sparse(h, w, severity):
    """zoom_operator without the dense (h*w)^2 matrix (20 GB at 224x224): the same entries in the same visiting order,
    equal (row, column) pairs summed in f64 in that order.  tests/test_oracle_cifar_c.py checks it against the dense
    builder bit for bit at small sizes."""
    zs = zoom_factors(severity)
    P = h * w
    rows = np.arange(P, dtype=np.int64)
    ent = [(rows, rows, np.ones(P, f64))]
    for z in zs:
        y0, y1, ty = zoom_axis_table(h, float(z))
        x0, x1, tx = zoom_axis_table(w, float(z))
        ty, tx = ty.astype(f64), tx.astype(f64)
        for ya, wy in ((y0, 1 - ty), (y1, ty)):
            for xa, wx in ((x0, 1 - tx), (x1, tx)):
                ent.append((rows, (ya[:, None] * w + xa[None, :]).reshape(-1).astype(np.int64),
                            (wy[:, None] * wx[None, :]).reshape(-1)))
    r = np.concatenate([e[0] for e in ent])
    q = np.concatenate([e[1] for e in ent])
    v = np.concatenate([e[2] for e in ent])
    order = np.lexsort((np.arange(len(r)), q, r))        # by (row, column), ties in visiting order
    r, q, v = r[order], q[order], v[order]
    new = np.ones(len(r), bool)
    new[1:] = (r[1:] != r[:-1]) | (q[1:] != q[:-1])
    gid = np.cumsum(new) - 1
    pos = np.arange(len(r)) - np.nonzero(new)[0][gid]
    tot = np.zeros(int(gid[-1]) + 1, f64)
    for k in range(int(pos.max()) + 1):
        m = pos == k
        tot[gid[m]] = tot[gid[m]] + v[m]
    tot = tot / (len(zs) + 1)
    ur, uq = r[new], q[new]
    nzm = tot != 0
    ur, uq, tot = ur[nzm], uq[nzm], tot[nzm]
    indptr = np.zeros(P + 1, np.int64)
    np.add.at(indptr, ur + 1, 1)
    return np.cumsum(indptr).astype(np.int32), uq.astype(np.int32), tot.astype(f32)


def apply_csr_f32(src, indptr, indices, weights):
    """out[:, p, :] = sum_k weights[k] * src[:, indices[k], :] for k in row p, f32, in stored order.
    src (n, h*w, c) f32."""
    n, P, c = src.shape
    out = np.zeros_like(src)
    counts = np.diff(indptr)
    for j in range(int(counts.max())):
        rows = np.nonzero(counts > j)[0]
        k = indptr[rows] + j
        out[:, rows] = (out[:, rows] + (src[:, indices[k]] * weights[k][None, :, None]).astype(f32)).astype(f32)
    return out


_zoom_ops = {}


def zoom_blur(x, severity):
    """Fast evaluation used as the CUDA kernel's specification: the merged sparse operator applied in f32."""
    n, h, w, c = x.shape
    key = (h, w, severity)
    if key not in _zoom_ops:
        _zoom_ops[key] = zoom_operator(h, w, severity) if h * w <= 4096 else zoom_operator_sparse(h, w, severity)
    src = (x.astype(f32) / f32(255)).astype(f32).reshape(n, h * w, c)
    v = apply_csr_f32(src, *_zoom_ops[key]).reshape(n, h, w, c)
    return _u8((np.clip(v, 0, 1) * f32(255)).astype(f32))


# =================================================================================================
# snow
# =================================================================================================
def snow(x, severity, layer_noise, angles):
    """layer_noise (n,h,w) f64 = np.random.normal(size, loc=c0, scale=c1); angles (n,) ~ U(-135,-45)."""
    c = PARAMS["snow"][severity - 1]
    n, h, w, _ = x.shape
    out = np.empty_like(x)
    xf = (x.astype(f32) / f32(255)).astype(f32)
    for i in range(n):
        layer = clipped_zoom_f32(layer_noise[i].astype(f32)[None, :, :, None], float(c[2]))[0, :, :, 0]
        layer = np.where(layer < f32(c[3]), f32(0), layer)
        lq = _u8((np.clip(layer, 0, 1) * f32(255)).astype(f32))                      # 8-bit grayscale PNG
        k, off = motion_kernel(c[4], c[5], float(angles[i]))
        acc = line_blur_f32(lq.astype(f32)[:, :, None], k, off)[:, :, 0]
        lb = _u8(np.clip(np.floor((acc + f32(0.5)).astype(f32)), 0, 255))          # 8-bit blob back
        sl = (lb.astype(f32) / f32(255)).astype(f32)[:, :, None]
        xi = xf[i]
        gray = ((xi[:, :, 0] * f32(0.299)).astype(f32) + (xi[:, :, 1] * f32(0.587)).astype(f32)).astype(f32)
        gray = (gray + (xi[:, :, 2] * f32(0.114)).astype(f32)).astype(f32)
        whit = ((gray * f32(1.5)).astype(f32) + f32(0.5)).astype(f32)[:, :, None]
        mix = ((f32(c[6]) * xi).astype(f32) + (f32(1 - c[6]) * np.maximum(xi, whit)).astype(f32)).astype(f32)
        v = ((mix + sl).astype(f32) + sl[::-1, ::-1]).astype(f32)
        out[i] = _u8((np.clip(v, 0, 1) * f32(255)).astype(f32))
    return out


# =================================================================================================
# frost / fog / brightness / contrast (float64, pointwise + per-image statistics)
# =================================================================================================
def frost(x, frost_crop, severity):
    """clip(c0*x + c1*frost_crop, 0, 255); frost_crop (n,h,w,3) uint8 is the random crop of a frost texture (the
    canonical textures are not available: any texture may be injected)."""
    c0, c1 = PARAMS["frost"][severity - 1]
    return _u8(np.clip(c0 * x.astype(f64) + c1 * frost_crop.astype(f64), 0, 255))


def fog(x, plasma, severity):
    """x/255 + c0*plasma[:h,:w]; renormalise by max/(max+c0); plasma (n,h,w) f64 from plasma_fractal(wibbledecay=c1)."""
    c0 = PARAMS["fog"][severity - 1][0]
    v = x / 255.0
    mx = v.reshape(v.shape[0], -1).max(axis=1).reshape(-1, 1, 1, 1)
    v = v + c0 * plasma[..., None]
    return _u8(np.clip(v * mx / (mx + c0), 0, 1) * 255)


def plasma_fractal(mapsize, wibbledecay, rng_uniform):
    """make_cifar_c.plasma_fractal (diamond-square); rng_uniform(lo, hi, shape) supplies the draws in call order."""
    maparray = np.empty((mapsize, mapsize), dtype=np.float64)
    maparray[0, 0] = 0
    stepsize = mapsize
    wibble = 100.0

    def wibbledmean(array):
        return array / 4 + wibble * rng_uniform(-wibble, wibble, array.shape)

    while stepsize >= 2:
        cornerref = maparray[0:mapsize:stepsize, 0:mapsize:stepsize]
        squareaccum = cornerref + np.roll(cornerref, shift=-1, axis=0)
        squareaccum += np.roll(squareaccum, shift=-1, axis=1)
        maparray[stepsize // 2:mapsize:stepsize, stepsize // 2:mapsize:stepsize] = wibbledmean(squareaccum)
        mapsize_ = maparray.shape[0]
        drgrid = maparray[stepsize // 2:mapsize_:stepsize, stepsize // 2:mapsize_:stepsize]
        ulgrid = maparray[0:mapsize_:stepsize, 0:mapsize_:stepsize]
        ldrsum = drgrid + np.roll(drgrid, 1, axis=0)
        lulsum = ulgrid + np.roll(ulgrid, -1, axis=1)
        ltsum = ldrsum + lulsum
        maparray[0:mapsize_:stepsize, stepsize // 2:mapsize_:stepsize] = wibbledmean(ltsum)
        tdrsum = drgrid + np.roll(drgrid, 1, axis=1)
        tulsum = ulgrid + np.roll(ulgrid, -1, axis=0)
        ttsum = tdrsum + tulsum
        maparray[stepsize // 2:mapsize_:stepsize, 0:mapsize_:stepsize] = wibbledmean(ttsum)
        stepsize //= 2
        wibble /= wibbledecay
    maparray -= maparray.min()
    return maparray / maparray.max()


def brightness(x, c):
    """skimage rgb2hsv -> V = clip(V + c) -> hsv2rgb -> clip*255, all f64 (formulas of skimage.color.colorconv)."""
    arr = x / 255.0
    v = arr.max(-1)
    delta = arr.max(-1) - arr.min(-1)
    with np.errstate(invalid="ignore", divide="ignore"):
        s = delta / v
        s = np.where(delta == 0.0, 0.0, s)
        r, g, b = arr[..., 0], arr[..., 1], arr[..., 2]
        hr = (g - b) / delta
        hg = 2.0 + (b - r) / delta
        hb = 4.0 + (r - g) / delta
    h = np.zeros_like(v)
    h = np.where(r == v, hr, h)     # skimage assigns red, then green, then blue: later matches win
    h = np.where(g == v, hg, h)
    h = np.where(b == v, hb, h)
    h = (h / 6.0) % 1.0
    h = np.where(delta == 0.0, 0.0, h)
    v = np.clip(v + c, 0, 1)
    hi = np.floor(h * 6)
    f = h * 6 - hi
    p = v * (1 - s)
    q = v * (1 - f * s)
    t = v * (1 - (1 - f) * s)
    hi = hi.astype(np.uint8) % 6
    out = np.choose(hi[..., None] * 0 + hi[..., None],
                    [np.stack((v, t, p), -1), np.stack((q, v, p), -1), np.stack((p, v, t), -1),
                     np.stack((p, q, v), -1), np.stack((t, p, v), -1), np.stack((v, p, q), -1)])
    return _u8(np.clip(out, 0, 1) * 255)


def contrast(x, c):
    """(x - mean)*c + mean with the per-image per-channel mean; mean = (sum of bytes / (h*w)) / 255 in f64."""
    v = x / 255.0
    n, h, w, ch = x.shape
    s = x.reshape(n, -1, ch).astype(np.int64).sum(axis=1)
    means = (s / f64(h * w) / 255.0).reshape(n, 1, 1, ch)
    return _u8(np.clip((v - means) * c + means, 0, 1) * 255)


# =================================================================================================
# elastic_transform
# =================================================================================================
def affine_from_points(pts1, pts2):
    """cv2.getAffineTransform: the 2x3 matrix M with M @ [x,y,1] = pts2 for the three pts1 (f64 solve)."""
    A = np.zeros((6, 6))
    b = np.zeros(6)
    for i in range(3):
        A[i, 0:3] = [pts1[i][0], pts1[i][1], 1]
        A[i + 3, 3:6] = [pts1[i][0], pts1[i][1], 1]
        b[i] = pts2[i][0]
        b[i + 3] = pts2[i][1]
    return np.linalg.solve(A, b).reshape(2, 3)


def warp_affine_f32(src, M):
    """cv2.warpAffine(src f32 (h,w,c), M, (w,h), INTER_LINEAR, BORDER_REFLECT_101): inverse map in f64, coordinates in
    1/1024 fixed point rounded to 1/32 sub-pixels, f32 bilinear table weights."""
    h, w, c = src.shape
    M = np.array(M, f64).copy()
    D = M[0, 0] * M[1, 1] - M[0, 1] * M[1, 0]
    D = 1.0 / D if D != 0 else 0.0
    A11, A22 = M[1, 1] * D, M[0, 0] * D
    M[0, 0] = A11
    M[0, 1] *= -D
    M[1, 0] *= -D
    M[1, 1] = A22
    b1 = -M[0, 0] * M[0, 2] - M[0, 1] * M[1, 2]
    b2 = -M[1, 0] * M[0, 2] - M[1, 1] * M[1, 2]
    M[0, 2], M[1, 2] = b1, b2
    xs = np.arange(w)
    ys = np.arange(h)
    adelta = np.rint(M[0, 0] * xs * 1024).astype(np.int64)
    bdelta = np.rint(M[1, 0] * xs * 1024).astype(np.int64)
    X0 = np.rint((M[0, 1] * ys + M[0, 2]) * 1024).astype(np.int64) + 16
    Y0 = np.rint((M[1, 1] * ys + M[1, 2]) * 1024).astype(np.int64) + 16
    X = (X0[:, None] + adelta[None, :]) >> 5
    Y = (Y0[:, None] + bdelta[None, :]) >> 5
    sx, sy = X >> 5, Y >> 5
    fx = ((X & 31).astype(f32) * f32(1.0 / 32)).astype(f32)
    fy = ((Y & 31).astype(f32) * f32(1.0 / 32)).astype(f32)
    wx0, wy0 = (f32(1) - fx).astype(f32), (f32(1) - fy).astype(f32)

    def tap(yy, xx):
        return src[_reflect101(yy, h), _reflect101(xx, w)]

    w00 = (wy0 * wx0).astype(f32)[..., None]
    w01 = (wy0 * fx).astype(f32)[..., None]
    w10 = (fy * wx0).astype(f32)[..., None]
    w11 = (fy * fx).astype(f32)[..., None]
    acc = (tap(sy, sx) * w00).astype(f32)
    acc = (acc + (tap(sy, sx + 1) * w01).astype(f32)).astype(f32)
    acc = (acc + (tap(sy + 1, sx) * w10).astype(f32)).astype(f32)
    acc = (acc + (tap(sy + 1, sx + 1) * w11).astype(f32)).astype(f32)
    return acc


def map_coordinates_reflect_f32(img, yy, xx):
    """scipy.ndimage.map_coordinates(order=1, mode='reflect') at (yy, xx) f32 coordinates, per channel: weights and
    sum in f64 over the f32 image, cast to f32.  img (h,w,c)."""
    h, w, c = img.shape

    def refl(i, n):
        period = 2 * n
        m = np.mod(i, period)
        return np.where(m >= n, period - 1 - m, m)

    yy = yy.astype(f64)
    xx = xx.astype(f64)
    y0 = np.floor(yy)
    x0 = np.floor(xx)
    ty = (yy - y0)[..., None]
    tx = (xx - x0)[..., None]
    y0 = y0.astype(np.int64)
    x0 = x0.astype(np.int64)
    s = img.astype(f64)
    acc = (s[refl(y0, h), refl(x0, w)] * (1 - ty)) * (1 - tx)
    acc = acc + (s[refl(y0, h), refl(x0 + 1, w)] * (1 - ty)) * tx
    acc = acc + (s[refl(y0 + 1, h), refl(x0, w)] * ty) * (1 - tx)
    acc = acc + (s[refl(y0 + 1, h), refl(x0 + 1, w)] * ty) * tx
    return acc.astype(f32)


def elastic_points(size=32):
    cs = np.float32((size, size)) // 2
    sq = min((size, size)) // 3
    return np.float32([cs + sq, [cs[0] + sq, cs[1] - sq], cs - sq])


def elastic_transform(x, severity, jitter, ux, uy):
    """jitter (n,3,2) f32 = np.random.uniform(-c2, c2, (3,2)).astype(f32); ux, uy (n,h,w) f64 = uniform(-1,1) fields
    (dx from ux, dy from uy; drawn dx first)."""
    n, h, w, _ = x.shape
    c = PARAMS["elastic_transform"][severity - 1]
    alpha, sigma = c[0] * h, c[1] * h
    out = np.empty_like(x)
    pts1 = elastic_points(h)
    ys, xs = np.mgrid[0:h, 0:w]
    for i in range(n):
        img = (x[i].astype(f32) / f32(255)).astype(f32)
        M = affine_from_points(pts1, pts1 + jitter[i].astype(f32))
        img = warp_affine_f32(img, M)
        dx = (gaussian_filter_sym(ux[i][None], sigma, "reflect", 3.0)[0] * alpha).astype(f32)
        dy = (gaussian_filter_sym(uy[i][None], sigma, "reflect", 3.0)[0] * alpha).astype(f32)
        v = map_coordinates_reflect_f32(img, (ys.astype(f32) + dy).astype(f32), (xs.astype(f32) + dx).astype(f32))
        out[i] = _u8((np.clip(v, 0, 1) * f32(255)).astype(f32))
    return out


# =================================================================================================
# pixelate (Pillow BOX resize down and up: exact) and jpeg_compression (libjpeg baseline round trip: exact)
# =================================================================================================
def pixelate(x, c):
    n, h, w, _ = x.shape
    small = O.pil_resize(x, (int(w * c), int(h * c)), O.BOX)
    return O.pil_resize(small, (w, h), O.BOX)


_JPEG_LUMA = np.array([16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56,
                       14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113,
                       92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99]).reshape(8, 8)
_JPEG_CHROMA = np.array([17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99,
                         99, 47, 66, 99, 99, 99, 99, 99, 99] + [99] * 32).reshape(8, 8)


def jpeg_qtables(quality):
    """jpeg_set_quality(quality, force_baseline=TRUE)."""
    q = max(1, min(100, int(quality)))
    scale = 5000 // q if q < 50 else 200 - q * 2
    out = []
    for base in (_JPEG_LUMA, _JPEG_CHROMA):
        t = (base.astype(np.int64) * scale + 50) // 100
        out.append(np.clip(t, 1, 255))
    return out


_FIX = lambda v: int(v * (1 << 13) + 0.5)  # noqa: E731  CONST_BITS = 13
F_0_298631336, F_0_390180644, F_0_541196100, F_0_765366865 = _FIX(0.298631336), _FIX(0.390180644), _FIX(0.541196100), _FIX(0.765366865)
F_0_899976223, F_1_175875602, F_1_501321110, F_1_847759065 = _FIX(0.899976223), _FIX(1.175875602), _FIX(1.501321110), _FIX(1.847759065)
F_1_961570560, F_2_053119869, F_2_562915447, F_3_072711026 = _FIX(1.961570560), _FIX(2.053119869), _FIX(2.562915447), _FIX(3.072711026)


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _fdct_1d(d, shift_even, descale_odd):
    """One pass of jfdctint.c (jpeg_fdct_islow) over the last axis of d (..., 8) int64."""
    tmp0, tmp7 = d[..., 0] + d[..., 7], d[..., 0] - d[..., 7]
    tmp1, tmp6 = d[..., 1] + d[..., 6], d[..., 1] - d[..., 6]
    tmp2, tmp5 = d[..., 2] + d[..., 5], d[..., 2] - d[..., 5]
    tmp3, tmp4 = d[..., 3] + d[..., 4], d[..., 3] - d[..., 4]
    tmp10, tmp13 = tmp0 + tmp3, tmp0 - tmp3
    tmp11, tmp12 = tmp1 + tmp2, tmp1 - tmp2
    o = np.empty_like(d)
    o[..., 0] = shift_even(tmp10 + tmp11)
    o[..., 4] = shift_even(tmp10 - tmp11)
    z1 = (tmp12 + tmp13) * F_0_541196100
    o[..., 2] = descale_odd(z1 + tmp13 * F_0_765366865)
    o[..., 6] = descale_odd(z1 + tmp12 * (-F_1_847759065))
    z1, z2, z3, z4 = tmp4 + tmp7, tmp5 + tmp6, tmp4 + tmp6, tmp5 + tmp7
    z5 = (z3 + z4) * F_1_175875602
    t4, t5, t6, t7 = tmp4 * F_0_298631336, tmp5 * F_2_053119869, tmp6 * F_3_072711026, tmp7 * F_1_501321110
    z1, z2 = z1 * (-F_0_899976223), z2 * (-F_2_562915447)
    z3, z4 = z3 * (-F_1_961570560) + z5, z4 * (-F_0_390180644) + z5
    o[..., 7] = descale_odd(t4 + z1 + z3)
    o[..., 5] = descale_odd(t5 + z2 + z4)
    o[..., 3] = descale_odd(t6 + z2 + z3)
    o[..., 1] = descale_odd(t7 + z1 + z4)
    return o


def jpeg_fdct_islow(block):
    """block (..., 8, 8) samples - 128 -> coefficients scaled by 8 (PASS1_BITS = 2)."""
    d = block.astype(np.int64)
    d = _fdct_1d(d, lambda v: v << 2, lambda v: _descale(v, 13 - 2))                      # rows
    d = np.swapaxes(d, -1, -2)
    d = _fdct_1d(d, lambda v: _descale(v, 2), lambda v: _descale(v, 13 + 2))              # columns
    return np.swapaxes(d, -1, -2)


def _idct_1d(d, descale):
    """One pass of jidctint.c (jpeg_idct_islow) over the last axis."""
    z2, z3 = d[..., 2], d[..., 6]
    z1 = (z2 + z3) * F_0_541196100
    tmp2 = z1 + z3 * (-F_1_847759065)
    tmp3 = z1 + z2 * F_0_765366865
    z2, z3 = d[..., 0], d[..., 4]
    tmp0, tmp1 = (z2 + z3) << 13, (z2 - z3) << 13
    tmp10, tmp13 = tmp0 + tmp3, tmp0 - tmp3
    tmp11, tmp12 = tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = d[..., 7], d[..., 5], d[..., 3], d[..., 1]
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * F_1_175875602
    t0, t1, t2, t3 = t0 * F_0_298631336, t1 * F_2_053119869, t2 * F_3_072711026, t3 * F_1_501321110
    z1, z2 = z1 * (-F_0_899976223), z2 * (-F_2_562915447)
    z3, z4 = z3 * (-F_1_961570560) + z5, z4 * (-F_0_390180644) + z5
    t0, t1, t2, t3 = t0 + z1 + z3, t1 + z2 + z4, t2 + z2 + z3, t3 + z1 + z4
    o = np.empty_like(d)
    o[..., 0], o[..., 7] = descale(tmp10 + t3), descale(tmp10 - t3)
    o[..., 1], o[..., 6] = descale(tmp11 + t2), descale(tmp11 - t2)
    o[..., 2], o[..., 5] = descale(tmp12 + t1), descale(tmp12 - t1)
    o[..., 3], o[..., 4] = descale(tmp13 + t0), descale(tmp13 - t0)
    return o


def jpeg_idct_islow(coef):
    """dequantised coefficients (..., 8, 8) -> samples 0..255 (range-limited, +128)."""
    d = coef.astype(np.int64)
    d = np.swapaxes(_idct_1d(np.swapaxes(d, -1, -2), lambda v: _descale(v, 13 - 2)), -1, -2)   # columns first
    d = _idct_1d(d, lambda v: _descale(v, 13 + 2 + 3))                                          # rows
    return np.clip(d + 128, 0, 255)


def _blocks(plane):
    n, h, w = plane.shape
    return plane.reshape(n, h // 8, 8, w // 8, 8).transpose(0, 1, 3, 2, 4)


def _unblocks(b):
    n, by, bx, _, _ = b.shape
    return b.transpose(0, 1, 3, 2, 4).reshape(n, by * 8, bx * 8)


def _jpeg_plane(plane, qt):
    coef = jpeg_fdct_islow(_blocks(plane.astype(np.int64) - 128))
    qv = (qt.astype(np.int64) << 3)
    a = np.abs(coef)
    quant = np.sign(coef) * ((a + (qv >> 1)) // qv)
    return _unblocks(jpeg_idct_islow(quant * qt))


def jpeg_compression(x, quality):
    """PIL save(JPEG, quality) + reload, for sizes that are multiples of 16: RGB->YCbCr (jccolor.c), 4:2:0
    h2v2_downsample, islow FDCT, quantise, dequantise, islow IDCT, h2v2 fancy upsample, YCbCr->RGB (jdcolor.c)."""
    n, h, w, _ = x.shape
    assert h % 16 == 0 and w % 16 == 0, "edge padding of partial MCUs is not restated"
    r, g, b = (x[..., i].astype(np.int64) for i in range(3))
    half, cb = 1 << 15, 128 << 16
    fx = lambda v: int(v * 65536 + 0.5)  # noqa: E731
    Y = (fx(0.29900) * r + fx(0.58700) * g + fx(0.11400) * b + half) >> 16
    Cb = (-fx(0.16874) * r - fx(0.33126) * g + fx(0.50000) * b + cb + half - 1) >> 16
    Cr = (fx(0.50000) * r - fx(0.41869) * g - fx(0.08131) * b + cb + half - 1) >> 16

    def down(p):
        q = p.reshape(n, h // 2, 2, w // 2, 2)
        bias = np.tile(np.array([1, 2]), w // 4)[None, None, :]
        return (q[:, :, 0, :, 0] + q[:, :, 0, :, 1] + q[:, :, 1, :, 0] + q[:, :, 1, :, 1] + bias) >> 2

    ql, qc = jpeg_qtables(quality)
    Y2 = _jpeg_plane(Y, ql)
    Cb2 = _jpeg_plane(down(Cb), qc)
    Cr2 = _jpeg_plane(down(Cr), qc)

    def up(p):
        """h2v2_fancy_upsample (jdsample.c): vertical 3:1 blend to two rows, then horizontal 3:1 with +8 / +7 bias."""
        hh, ww = p.shape[1], p.shape[2]
        above = np.concatenate([p[:, :1], p[:, :-1]], axis=1)
        below = np.concatenate([p[:, 1:], p[:, -1:]], axis=1)
        rows = np.empty((n, hh * 2, ww), np.int64)
        rows[:, 0::2] = p * 3 + above
        rows[:, 1::2] = p * 3 + below
        out = np.empty((n, hh * 2, ww * 2), np.int64)
        left = np.concatenate([rows[:, :, :1], rows[:, :, :-1]], axis=2)
        right = np.concatenate([rows[:, :, 1:], rows[:, :, -1:]], axis=2)
        out[:, :, 0::2] = (rows * 3 + left + 8) >> 4
        out[:, :, 1::2] = (rows * 3 + right + 7) >> 4
        out[:, :, 0] = (rows[:, :, 0] * 4 + 8) >> 4
        out[:, :, -1] = (rows[:, :, -1] * 4 + 7) >> 4
        return out

    Cbu, Cru = up(Cb2) - 128, up(Cr2) - 128
    R = Y2 + ((fx(1.40200) * Cru + half) >> 16)
    G = Y2 + ((-fx(0.34414) * Cbu - fx(0.71414) * Cru + half) >> 16)
    B = Y2 + ((fx(1.77200) * Cbu + half) >> 16)
    return np.stack([np.clip(R, 0, 255), np.clip(G, 0, 255), np.clip(B, 0, 255)], axis=-1).astype(np.uint8)


# =================================================================================================
# dispatcher
# =================================================================================================
def draw(name, severity, x, rng):
    """Draw the random inputs of one corruption for batch x from numpy RandomState `rng`, in the canonical order."""
    n, h, w, c = x.shape
    p = PARAMS[name][severity - 1]
    if name == "gaussian_noise":
        return {"noise": rng.normal(size=x.shape, scale=p)}
    if name == "shot_noise":
        return {"counts": rng.poisson(shot_rate(x, p))}
    if name == "impulse_noise":
        return {"u_flip": rng.random_sample(x.shape), "u_salt": rng.random_sample(x.shape)}
    if name == "glass_blur":
        return {"deltas": rng.randint(-p[1], p[1], size=(n, p[2], h - 2, w - 2, 2)).astype(np.int8)}
    if name == "motion_blur":
        return {"angles": rng.uniform(-45, 45, n)}
    if name == "snow":
        return {"layer": rng.normal(size=(n, h, w), loc=p[0], scale=p[1]), "angles": rng.uniform(-135, -45, n)}
    if name == "frost":
        return {"frost": rng.randint(0, 256, size=x.shape).astype(np.uint8)}
    if name == "fog":
        # the diamond-square map needs a power-of-two side: sizes in between draw the next power of two and keep the top-left
        # h x w corner, as the canonical ImageNet-C generator does (mapsize 256, [:224, :224]); 32 x 32 is unchanged
        ms = 1 << int(np.ceil(np.log2(max(h, w))))
        return {"plasma": np.stack([plasma_fractal(ms, p[1], lambda lo, hi, s: rng.uniform(lo, hi, s))[:h, :w]
                                    for _ in range(n)])}
    if name == "elastic_transform":
        c2 = p[2] * h
        return {"jitter": rng.uniform(-c2, c2, size=(n, 3, 2)).astype(f32), "ux": rng.uniform(-1, 1, (n, h, w)),
                "uy": rng.uniform(-1, 1, (n, h, w))}
    return {}


def corrupt(name, severity, x, d=None):
    p = PARAMS[name][severity - 1]
    d = d or {}
    if name == "gaussian_noise":
        return gaussian_noise(x, d["noise"])
    if name == "shot_noise":
        return shot_noise(x, d["counts"], p)
    if name == "impulse_noise":
        return impulse_noise(x, d["u_flip"], d["u_salt"], p)
    if name == "defocus_blur":
        return defocus_blur(x, p[0], p[1])
    if name == "glass_blur":
        return glass_blur(x, p[0], p[2], d["deltas"])
    if name == "motion_blur":
        return motion_blur(x, p[0], p[1], d["angles"])
    if name == "zoom_blur":
        return zoom_blur(x, severity)
    if name == "snow":
        return snow(x, severity, d["layer"], d["angles"])
    if name == "frost":
        return frost(x, d["frost"], severity)
    if name == "fog":
        return fog(x, d["plasma"], severity)
    if name == "brightness":
        return brightness(x, p)
    if name == "contrast":
        return contrast(x, p)
    if name == "elastic_transform":
        return elastic_transform(x, severity, d["jitter"], d["ux"], d["uy"])
    if name == "pixelate":
        return pixelate(x, p)
    if name == "jpeg_compression":
        return jpeg_compression(x, p)
    raise KeyError(name)
