"""CPU oracle, part 2: the reference's multi-op chains restated on top of oracle/np_ops.py.

TEST INFRASTRUCTURE ONLY (see oracle/np_ops.py header).  Each function cites the reference lines it
follows; random draws are explicit arguments.  Pinned by tests/test_oracle_golden.py
against tests/golden/*.npz (made from the live reference by oracle/make_golden.py).
"""
from __future__ import annotations

import numpy as np

from . import np_ops as O

f32 = np.float32

CIFAR10_MEAN, CIFAR10_STD = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)
CIFAR100_MEAN, CIFAR100_STD = (0.5071, 0.4867, 0.4408), (0.2675, 0.2565, 0.2761)

# ------------------------------------------------------------------------------------------------
# AugMix — fall_2025/AugMix.py:30-62
# ------------------------------------------------------------------------------------------------
AUG_OP_NAMES = ["rotate", "posterize", "shear_x", "shear_y", "translate_x", "translate_y", "equalize", "solarize"]


def augmix_op(u: np.ndarray, op: int, severity, sign: int = 1) -> np.ndarray:
    """One AUG_OPS[op](pil, severity) on a (n,h,w,3) uint8 batch — AugMix.py:30-37."""
    if op == 0:
        return O.pil_rotate_nearest(u, severity * sign)
    if op == 1:
        return O.posterize(u, int(severity))
    if op == 2:
        return O.affine_nearest(u, (1, severity * 0.3, 0, 0, 1, 0))
    if op == 3:
        return O.affine_nearest(u, (1, 0, 0, severity * 0.3, 1, 0))
    if op == 4:
        return O.affine_nearest(u, (1, 0, severity * 2, 0, 1, 0))
    if op == 5:
        return O.affine_nearest(u, (1, 0, 0, 0, 1, severity * 2))
    if op == 6:
        return O.equalize(u)
    if op == 7:
        return O.solarize(u, int(severity * 20))
    raise ValueError(op)


def augmix(x: np.ndarray, severity, ws, m, ops, signs) -> np.ndarray:
    """augmix(image_tensor, severity, width, depth) — AugMix.py:45-62 with the draws injected.
    x: (c,h,w) f32; ws (width,) f64; m f64; ops/signs (width, maxdepth) int, op < 0 = chain ended."""
    x = np.asarray(x, f32)
    mix = np.zeros_like(x)
    for i in range(len(ws)):
        u = O.to_pil_image(x[None])  # TF.to_pil_image: mul(255).byte() with mod-256 wrap
        aug = x
        for j in range(ops.shape[1]):
            if ops[i, j] < 0:
                break
            u = augmix_op(u, int(ops[i, j]), severity, int(signs[i, j]))
            aug = O.to_tensor(u)[0]
            u = O.to_pil_image(aug[None])  # identity on u/255 values
        mix = (mix + (f32(ws[i]) * aug).astype(f32)).astype(f32)
    return ((f32(1 - m) * x).astype(f32) + (f32(m) * mix).astype(f32)).astype(f32)


# ------------------------------------------------------------------------------------------------
# Adaptive-TTA 16 views — fall_2025/tta_initial_claudeFixed.py:37-78,218-230
# ------------------------------------------------------------------------------------------------
def tta_views16(tn: np.ndarray, crop_xy, brightness, contrast, mean=CIFAR100_MEAN, std=CIFAR100_STD) -> np.ndarray:
    """tn (3,32,32) normalised f32 -> (16,3,32,32): tensor_to_pil, then [1 flip, 5 rand_crop, 5 rand_brightness,
    5 rand_contrast] (AUG_COUNTS), each followed by pil_to_tensor."""
    u = O.tensor_to_pil(tn[None], mean, std)
    views = [O.hflip(u)]
    w = u.shape[2]
    cs = int(w * 0.78)
    for x, y in np.asarray(crop_xy):
        views.append(O.crop_resize(u, int(x), int(y), cs, 32, O.BICUBIC))
    for f in brightness:
        views.append(O.enhance_brightness(u, float(f)))
    for f in contrast:
        views.append(O.enhance_contrast(u, float(f)))
    return O.to_tensor_normalize(np.concatenate(views), mean, std)


# ------------------------------------------------------------------------------------------------
# TTA strategies — fall_2025/TTA_transforms.py:69-138
# ------------------------------------------------------------------------------------------------
def crop_reflect(u: np.ndarray, i: int, j: int, pad: int = 4, size: int = 32) -> np.ndarray:
    """T.RandomCrop(size, padding=pad, padding_mode='reflect') with drawn offsets (i row, j col)."""
    p = np.pad(u, ((0, 0), (pad, pad), (pad, pad), (0, 0)), mode="reflect")
    return p[:, i:i + size, j:j + size].copy()


def tta_strategy(u: np.ndarray, strategy: str, crop_ij=None, jitter=None, mean=CIFAR10_MEAN, std=CIFAR10_STD):
    """u (1,32,32,3) uint8 -> (4,3,32,32) f32: the four views of tta_transforms_<strategy>."""
    if strategy == "gentle":
        (i0, j0), (i1, j1) = np.asarray(crop_ij)
        views = [u, O.hflip(u), crop_reflect(u, int(i0), int(j0)), crop_reflect(O.hflip(u), int(i1), int(j1))]
    elif strategy == "multiscale":
        views = [u, O.resize_center_crop(u, 36, 32), O.resize_center_crop(u, 40, 32), O.hflip(u)]
    elif strategy == "color":
        jb, jc = float(jitter[0][0]), float(jitter[1][1])
        views = [u, O.enhance_brightness(u, jb), O.enhance_contrast(u, jc), O.hflip(u)]
    else:
        raise ValueError(strategy)
    return O.to_tensor_normalize(np.concatenate(views), mean, std)


# ------------------------------------------------------------------------------------------------
# TransformationPool chains — pipenline/cifar_image_transformations.py:39-152
# ------------------------------------------------------------------------------------------------
POOL_GAUSS = [0.08, 0.12, 0.18, 0.26, 0.38]
POOL_IMPULSE = [0.03, 0.06, 0.09, 0.17, 0.27]
POOL_SHOT = [60, 25, 12, 5, 3]
POOL_DEFOCUS = [3, 4, 6, 8, 10]


def pool_op(u: np.ndarray, name: str, params: dict, draws: dict | None = None) -> np.ndarray:
    """TransformationPool.<name>(pil, **params) on a uint8 batch; draws carries the injected RNG output."""
    d = draws or {}
    if name == "gaussian_noise":
        return O.gaussian_noise_f64(u, d["noise"].reshape(u.shape))
    if name == "impulse_noise":
        return O.impulse_noise(u, d["mask"].reshape(u.shape[:3]), POOL_IMPULSE[params["severity"] - 1])
    if name == "shot_noise":
        return O.shot_noise(u, d["counts"].reshape(u.shape), POOL_SHOT[params["severity"] - 1])
    if name == "defocus_blur":
        return O.pil_gaussian_blur(u, POOL_DEFOCUS[params["severity"] - 1])
    if name == "enhance_contrast":
        return O.enhance_contrast(u, params["factor"])
    if name == "enhance_brightness":
        return O.enhance_brightness(u, params["factor"])
    if name == "enhance_sharpness":
        return O.enhance_sharpness(u, params["factor"])
    if name == "enhance_color":
        return O.enhance_color(u, params["factor"])
    if name == "motion_blur":
        return O.motion_blur_box(u, params["size"])
    if name == "histogram_equalization":
        return O.histogram_equalization(u)
    raise ValueError(name)


def individual_apply(tn: np.ndarray, chain, draws=None, mean=CIFAR10_MEAN, std=CIFAR10_STD) -> np.ndarray:
    """Individual(chain).apply_transformations(tn) — cifar_image_transformations.py:141-152. tn (3,h,w) f32."""
    u = O.tensor_to_pil(tn[None], mean, std)
    for name, params in chain:
        u = pool_op(u, name, params or {}, draws)
    return O.to_tensor_normalize(u, mean, std)[0]


# ------------------------------------------------------------------------------------------------
# torchvision tensor backend pieces
# ------------------------------------------------------------------------------------------------
def perspective_coeffs(startpoints, endpoints) -> np.ndarray:
    """torchvision.transforms.functional._get_perspective_coeffs: least-squares solve in f64, 8 coefficients
    mapping OUTPUT pixel -> INPUT pixel."""
    a = np.zeros((8, 8), np.float64)
    for i, (p1, p2) in enumerate(zip(endpoints, startpoints)):
        a[2 * i] = [p1[0], p1[1], 1, 0, 0, 0, -p2[0] * p1[0], -p2[0] * p1[1]]
        a[2 * i + 1] = [0, 0, 0, p1[0], p1[1], 1, -p2[1] * p1[0], -p2[1] * p1[1]]
    b = np.asarray(startpoints, np.float64).reshape(8)
    return np.linalg.lstsq(a, b, rcond=None)[0]


def _fma(a, b, c):
    """f32 fused multiply-add through f64 (the product of two f32 is exact in f64)."""
    return (np.asarray(a).astype(np.float64) * np.asarray(b).astype(np.float64) + np.asarray(c).astype(np.float64)).astype(f32)


def _unnormalize(g, size):
    """grid_sample's align_corners=False coordinate map as ATen's CPU kernel evaluates it: fma(g + 1, size/2, -0.5)."""
    return _fma((g + f32(1)).astype(f32), f32(size / 2.0), f32(-0.5))


def perspective_tensor_bilinear(t: np.ndarray, coeffs, fill=0.0) -> np.ndarray:
    """TF.perspective(tensor, ..., BILINEAR, fill) — _perspective_grid + grid_sample(zeros, align_corners=False)
    + mask blend.  t (n,c,h,w) f32 -> same.  All f32, in the evaluation order of the CPU reference (bmm: x*a rounded,
    fma(y, b, .), + c; unnormalise by fma; taps accumulated nw, ne, sw, se by fma): measured against
    torchvision 0.26 / torch 2.11 this leaves <= 7e-7 absolute on a handful of elements per thousand images."""
    n, c, h, w = t.shape
    co = [f32(v) for v in coeffs]
    xs = np.linspace(0.5, w - 0.5, w, dtype=np.float64).astype(f32)
    ys = np.linspace(0.5, h - 0.5, h, dtype=np.float64).astype(f32)
    X, Y = np.meshgrid(xs, ys)
    t1x = [f32(co[0] / f32(0.5 * w)), f32(co[1] / f32(0.5 * w)), f32(co[2] / f32(0.5 * w))]
    t1y = [f32(co[3] / f32(0.5 * h)), f32(co[4] / f32(0.5 * h)), f32(co[5] / f32(0.5 * h))]

    def dot(r):
        return (_fma(Y, r[1], (X * r[0]).astype(f32)) + r[2]).astype(f32)

    g1x, g1y, g2 = dot(t1x), dot(t1y), dot([co[6], co[7], f32(1)])
    gx = ((g1x / g2).astype(f32) - f32(1)).astype(f32)
    gy = ((g1y / g2).astype(f32) - f32(1)).astype(f32)
    ix, iy = _unnormalize(gx, w), _unnormalize(gy, h)
    x0 = np.floor(ix)
    y0 = np.floor(iy)
    wx1 = (ix - x0).astype(f32)
    wy1 = (iy - y0).astype(f32)
    wx0 = (f32(1) - wx1).astype(f32)
    wy0 = (f32(1) - wy1).astype(f32)
    x0 = x0.astype(np.int64)
    y0 = y0.astype(np.int64)
    src = np.concatenate([t, np.ones((n, 1, h, w), f32)], axis=1)

    def tap(yy, xx):
        ok = (xx >= 0) & (xx < w) & (yy >= 0) & (yy < h)
        v = src[:, :, np.clip(yy, 0, h - 1), np.clip(xx, 0, w - 1)]
        return np.where(ok[None, None], v, f32(0))

    out = (tap(y0, x0) * (wx0 * wy0).astype(f32)).astype(f32)
    out = _fma(tap(y0, x0 + 1), (wx1 * wy0).astype(f32), out)
    out = _fma(tap(y0 + 1, x0), (wx0 * wy1).astype(f32), out)
    out = _fma(tap(y0 + 1, x0 + 1), (wx1 * wy1).astype(f32), out)
    img, mask = out[:, :c], out[:, c:]
    return (img * mask + (f32(1) - mask) * f32(fill)).astype(f32)


def perspective_warp_u8(u: np.ndarray, startpoints, endpoints) -> np.ndarray:
    """apply_perspective_warp — fall_2025/transformations_code:50-66: ToTensor, RandomPerspective(p=1) on the
    tensor (bilinear, fill 0), ToPILImage (mul(255).byte())."""
    t = O.to_tensor(u)
    o = perspective_tensor_bilinear(t, perspective_coeffs(startpoints, endpoints), 0.0)
    return O.to_pil_image(o)


def affine_nearest_tensor(t: np.ndarray, matrix) -> np.ndarray:
    """F.rotate / F.affine on an f32 (c,h,w) tensor with NEAREST and fill 0 — torchvision _functional_tensor.py
    _gen_affine_grid + grid_sample(nearest, zeros, align_corners=False).  The grid's 3-term dot products follow
    torch.bmm's f32 evaluation on CPU (x*a rounded, fma(y, b, .), + c — checked bit-exact against
    _gen_affine_grid), so the picks are torch's."""
    c, h, w = t.shape
    th = np.asarray(matrix, np.float64).astype(f32).reshape(2, 3)
    rx = (th[0] / f32(0.5 * w)).astype(f32)
    ry = (th[1] / f32(0.5 * h)).astype(f32)
    X, Y = np.meshgrid((f32(-w * 0.5 + 0.5) + np.arange(w, dtype=f32)).astype(f32),
                       (f32(-h * 0.5 + 0.5) + np.arange(h, dtype=f32)).astype(f32))

    def dot(r):
        return (_fma(Y, r[1], (X * r[0]).astype(f32)) + r[2]).astype(f32)

    gx, gy = dot(rx), dot(ry)
    ix, iy = _unnormalize(gx, w), _unnormalize(gy, h)
    px, py = np.rint(ix), np.rint(iy)
    ok = (px >= 0) & (px < w) & (py >= 0) & (py < h)
    sx = np.where(ok, px, 0).astype(np.int64)
    sy = np.where(ok, py, 0).astype(np.int64)
    return np.where(ok[None], t[:, sy, sx], f32(0)).astype(f32)


def tensor_tta_chain(t: np.ndarray, flip: bool, rot_matrix, aff_matrix, brightness: float, persp_coeffs=None) -> np.ndarray:
    """One TestTimeAugmentation call — fall_2025/post_hoc_aaryaa.py:24-51 — on an f32 (c,h,w) tensor with the draws
    given: hflip, F.rotate (NEAREST), F.affine (NEAREST, scale only), adjust_brightness = (b*x + (1-b)*0).clamp(0,1),
    F.perspective (BILINEAR, fill 0) when it fired."""
    v = np.asarray(t, f32)
    if flip:
        v = v[:, :, ::-1]
    v = affine_nearest_tensor(v, rot_matrix)
    v = affine_nearest_tensor(v, aff_matrix)
    v = ((f32(brightness) * v).astype(f32) + f32(0)).astype(f32)
    v = np.where(np.isnan(v), v, np.minimum(np.maximum(v, f32(0)), f32(1))).astype(f32)
    if persp_coeffs is not None:
        v = perspective_tensor_bilinear(v[None], persp_coeffs, 0.0)[0]
    return v


def mean_views(stack: np.ndarray) -> np.ndarray:
    """l_hat = (0 + L0 + L1 + ...)/(T+1) — post_hoc_aaryaa.py:334-342; torch.stack(...).mean(0) — TTA_transforms.py:208."""
    s = np.zeros(stack.shape[1:], f32)
    for v in stack:
        s = (s + v.astype(f32)).astype(f32)
    return (s / f32(stack.shape[0])).astype(f32)


def gaussian_blur3_f32(t: np.ndarray, kernel_size: int = 3) -> np.ndarray:
    """TF.gaussian_blur(img, kernel_size) with the default sigma — fall_2025/angellic_with_corrupt.py:54-56.
    t (n,c,h,w) f32; reflect padding; depthwise conv with the outer-product kernel."""
    k = kernel_size
    sigma = 0.3 * ((k - 1) * 0.5 - 1) + 0.8
    half = (k - 1) * 0.5
    xs = np.linspace(-half, half, k, dtype=np.float64).astype(f32)
    pdf = np.exp((f32(-0.5) * ((xs / f32(sigma)).astype(f32) ** 2).astype(f32)).astype(f32)).astype(f32)
    k1 = (pdf / pdf.sum(dtype=f32)).astype(f32)
    k2 = (k1[:, None] * k1[None, :]).astype(f32)
    r = k // 2
    p = np.pad(t, ((0, 0), (0, 0), (r, r), (r, r)), mode="reflect")
    n, c, h, w = t.shape
    acc = np.zeros_like(t)
    for dy in range(k):
        for dx in range(k):
            acc = (acc + (p[:, :, dy:dy + h, dx:dx + w] * k2[dy, dx]).astype(f32)).astype(f32)
    return acc


def float_corrupt(x, mode: str, a=0.0, b=0.0, r=None):
    """pipenline/angellic.py:34-46, fall_2025/angellic_with_corrupt.py:51-52 (f32, torch op order)."""
    x = np.asarray(x, f32)
    if mode == "gaussian_noise":   # (x + (randn*std + mean)).clamp(0,1)
        v = (x + ((r * f32(a)).astype(f32) + f32(b)).astype(f32)).astype(f32)
        return np.clip(v, 0, 1).astype(f32)
    if mode == "brightness":       # (x + f).clamp(0,1)
        return np.clip((x + f32(a)).astype(f32), 0, 1).astype(f32)
    if mode == "contrast":         # ((x - .5)*f + .5).clamp(0,1)
        return np.clip((((x - f32(0.5)).astype(f32) * f32(a)).astype(f32) + f32(0.5)).astype(f32), 0, 1).astype(f32)
    if mode == "add_noise":        # x + randn*std
        return (x + (r * f32(a)).astype(f32)).astype(f32)
    raise ValueError(mode)


# =================================================================================================
# H26  adaptive_transforms of fall_2025/Initial_Experiments.py:88-92 on PIL images
# =================================================================================================
def adaptive_chain(u: np.ndarray, flip: bool, affine_matrix, persp_coeffs=None) -> np.ndarray:
    """RandomHorizontalFlip -> RandomAffine -> RandomPerspective on one (h,w,3) uint8 image: FLIP_LEFT_RIGHT, then
    Image.transform(AFFINE, m, NEAREST, fillcolor 0) with m = torchvision's _get_inverse_affine_matrix about the centre, then
    (if it fired) Image.transform(PERSPECTIVE, coeffs, BILINEAR, fillcolor 0)."""
    x = u[None]
    if flip:
        x = O.hflip(x)
    x = O.affine_nearest(x, np.asarray(affine_matrix, np.float64))
    if persp_coeffs is not None:
        x = O.pil_transform(x, np.asarray(persp_coeffs, np.float64), (u.shape[1], u.shape[0]), mode="perspective",
                            resample="bilinear", fill=0)
    return x[0]

