"""Batched device API: torch CUDA tensors in, torch CUDA tensors out, every op one call into libimgtx.

uint8 image batches are (n, h, w, c) contiguous; float batches are (n, c, h, w) float32 contiguous.
PyTorch is used for device memory and streams only; all arithmetic is in the CUDA library.
Per-image parameters may be Python scalars (uniform over the batch), sequences / numpy arrays or
tensors of length n.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib as L

GEOM_DTYPE = np.dtype([("kind", "<i4"), ("dx", "<i4"), ("dy", "<i4"), ("pad", "<i4"), ("m", "<f8", (6,))])
assert GEOM_DTYPE.itemsize == C.sizeof(L.GeomRecord) == 64


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _check_u8(x: torch.Tensor) -> torch.Tensor:
    if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.uint8 and x.dim() == 4):
        raise TypeError("expected a CUDA uint8 tensor of shape (n, h, w, c)")
    return x.contiguous()


def _check_out(out: Optional[torch.Tensor], like: torch.Tensor, shape=None, dtype=None) -> torch.Tensor:
    """A caller-supplied `out` is written through its data pointer as a dense tensor: insist that it is one."""
    shape = tuple(like.shape) if shape is None else tuple(shape)
    dtype = like.dtype if dtype is None else dtype
    if out is None:
        return torch.empty(shape, dtype=dtype, device=like.device)
    if not (isinstance(out, torch.Tensor) and out.device == like.device and out.dtype == dtype and
            tuple(out.shape) == shape and out.is_contiguous()):
        raise ValueError(f"out must be a contiguous {dtype} tensor of shape {shape} on {like.device}")
    return out


def _check_f32(x: torch.Tensor) -> torch.Tensor:
    if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.float32):
        raise TypeError("expected a CUDA float32 tensor")
    return x.contiguous()


def _per_image_f32(param, n: int, device) -> Optional[torch.Tensor]:
    """None for a scalar, else a device f32 tensor of length n."""
    if isinstance(param, torch.Tensor):
        if param.numel() == 1 and param.dim() == 0:
            return None
        t = param.to(device=device, dtype=torch.float32).contiguous()
    elif np.isscalar(param):
        return None
    else:
        t = torch.from_numpy(np.asarray(param, dtype=np.float64).astype(np.float32)).to(device)
    if t.numel() != n:
        raise ValueError(f"per-image parameter has {t.numel()} entries for {n} images")
    return t


def _fptr(a: Optional[Sequence[float]]):
    if a is None:
        return None, None
    arr = (C.c_float * len(a))(*[float(v) for v in a])
    return arr, C.cast(arr, C.c_void_p)


# ------------------------------------------------------------------------------------------------
# pointwise
# ------------------------------------------------------------------------------------------------
def pointwise(x: torch.Tensor, op: int, param, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    x = _check_u8(x)
    n = x.shape[0]
    out = torch.empty_like(x) if out is None else out
    per = _per_image_f32(param, n, x.device)
    scalar = 0.0 if per is not None else float(param)
    image_bytes = x[0].numel() if n else 0
    L.check(L.load().imgtx_pointwise_u8(x.data_ptr(), out.data_ptr(), n, image_bytes, op, scalar,
                                        per.data_ptr() if per is not None else None, _stream()))
    return out


def brightness(x, factor, out=None):
    """ImageEnhance.Brightness(img).enhance(factor) on a batch."""
    return pointwise(x, L.PW_BRIGHTNESS, factor, out)


def convert_scale_abs(x, alpha, out=None):
    """cv2.convertScaleAbs(img, alpha=alpha, beta=0) on a batch."""
    return pointwise(x, L.PW_SCALE_ABS, alpha, out)


def posterize(x, bits, out=None):
    return pointwise(x, L.PW_POSTERIZE, bits, out)


def solarize(x, threshold, out=None):
    return pointwise(x, L.PW_SOLARIZE, threshold, out)


# ------------------------------------------------------------------------------------------------
# per-image statistics + map
# ------------------------------------------------------------------------------------------------
def enhance(x: torch.Tensor, kind: int, factor=1.0, out=None) -> torch.Tensor:
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    per = _per_image_f32(factor, n, x.device)
    scalar = 0.0 if per is not None else float(factor)
    L.check(L.load().imgtx_enhance_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, kind, scalar,
                                      per.data_ptr() if per is not None else None, _stream()))
    return out


def enhance_contrast(x, factor, out=None):
    """ImageEnhance.Contrast(img).enhance(factor)."""
    return enhance(x, L.ENH_CONTRAST, factor, out)


def enhance_color(x, factor, out=None):
    return enhance(x, L.ENH_COLOR, factor, out)


def enhance_sharpness(x, factor, out=None):
    return enhance(x, L.ENH_SHARPNESS, factor, out)


def equalize(x, out=None):
    """ImageOps.equalize."""
    return enhance(x, L.ENH_EQUALIZE, 1.0, out)


def histogram_equalization(x, out=None):
    """cv2 YUV equalizeHist round trip (TransformationPool.histogram_equalization)."""
    return enhance(x, L.ENH_HISTEQ_YUV, 1.0, out)


# ------------------------------------------------------------------------------------------------
# geometry
# ------------------------------------------------------------------------------------------------
def geom_records(n: int) -> np.ndarray:
    return np.zeros(n, dtype=GEOM_DTYPE)


def geometry(x: torch.Tensor, rec, out_hw=None, fill=None, out=None) -> torch.Tensor:
    """rec: one GEOM_DTYPE record (uniform) or an array of n records (per image), or a device uint8
    tensor holding n packed records."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    oh, ow = out_hw if out_hw is not None else (h, w)
    out = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device) if out is None else out
    uni = None
    per_ptr = None
    keep = None
    if isinstance(rec, torch.Tensor):
        if rec.numel() != n * 64:
            raise ValueError("packed geometry tensor has the wrong size")
        keep = rec
        per_ptr = rec.data_ptr()
    else:
        rec = np.asarray(rec, dtype=GEOM_DTYPE).reshape(-1)
        if rec.size == 1:
            uni = L.GeomRecord.from_buffer_copy(rec.tobytes())
        else:
            if rec.size != n:
                raise ValueError(f"{rec.size} geometry records for {n} images")
            keep = torch.from_numpy(rec.view(np.uint8).copy()).to(x.device)
            per_ptr = keep.data_ptr()
    fill_arr = None
    if fill is not None:
        fv = [int(v) for v in (fill if isinstance(fill, (tuple, list)) else [fill] * c)]
        fv = (fv + [0] * 4)[:4]
        fill_arr = (C.c_uint8 * 4)(*fv)
    L.check(L.load().imgtx_geometry_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, oh, ow,
                                       C.byref(uni) if uni is not None else None, per_ptr,
                                       C.cast(fill_arr, C.c_void_p) if fill_arr is not None else None,
                                       _stream()))
    return out


def pil_rotate_matrix(angle: float, w: int, h: int):
    """The coefficient list PIL.Image.rotate(angle, expand=False) builds in Python before calling C."""
    angle = angle % 360.0
    a = -math.radians(angle)
    m = [round(math.cos(a), 15), round(math.sin(a), 15), 0.0,
         round(-math.sin(a), 15), round(math.cos(a), 15), 0.0]
    cx, cy = w / 2.0, h / 2.0
    m[2] = m[0] * (-cx) + m[1] * (-cy) + m[2] + cx
    m[5] = m[3] * (-cx) + m[4] * (-cy) + m[5] + cy
    return m


def rotate_records(angles, w: int, h: int) -> np.ndarray:
    """Geometry records reproducing img.rotate(angle) (counter-clockwise degrees, NEAREST)."""
    angles = np.atleast_1d(np.asarray(angles, dtype=np.float64))
    rec = geom_records(len(angles))
    for i, ang in enumerate(angles):
        a = float(ang) % 360.0
        if a == 0:
            rec[i]["kind"] = L.GEOM_IDENTITY
        elif a == 180:
            rec[i]["kind"] = L.GEOM_ROT180
        elif a in (90.0, 270.0) and w == h:
            rec[i]["kind"] = L.GEOM_ROT90 if a == 90 else L.GEOM_ROT270
        else:
            rec[i]["kind"] = L.GEOM_AFFINE
            rec[i]["m"] = pil_rotate_matrix(a, w, h)
    return rec


def rotate(x, angle, fill=0, out=None):
    """PIL img.rotate(angle, fillcolor=fill): angle scalar or per image."""
    n, h, w, c = x.shape
    rec = rotate_records(angle, w, h)
    return geometry(x, rec, fill=fill, out=out)


def affine_nearest(x, m, out_hw=None, fill=0, out=None):
    """PIL img.transform(size, AFFINE, m) with NEAREST resampling; m is (6,) or (n, 6)."""
    m = np.asarray(m, dtype=np.float64)
    rec = geom_records(1 if m.ndim == 1 else m.shape[0])
    rec["kind"] = L.GEOM_AFFINE
    rec["m"] = m.reshape(-1, 6)
    return geometry(x, rec, out_hw=out_hw, fill=fill, out=out)


def translate(x, tx, ty, out=None):
    """reference apply_translation: integer shift with zero fill (int() truncation of tx, ty)."""
    tx = np.atleast_1d(np.asarray(tx))
    ty = np.atleast_1d(np.asarray(ty))
    k = max(len(tx), len(ty))
    rec = geom_records(k)
    rec["kind"] = L.GEOM_SHIFT
    rec["dx"] = np.trunc(np.broadcast_to(tx, (k,)).astype(np.float64)).astype(np.int32)
    rec["dy"] = np.trunc(np.broadcast_to(ty, (k,)).astype(np.float64)).astype(np.int32)
    return geometry(x, rec, out=out)


def hflip(x, out=None):
    rec = geom_records(1)
    rec["kind"] = L.GEOM_FLIP_LR
    return geometry(x, rec, out=out)


def crop_reflect(x, i, j, pad: int, size: int, out=None):
    """T.RandomCrop(size, padding=pad, padding_mode='reflect') with the drawn offsets (i: row, j: col)."""
    i = np.atleast_1d(np.asarray(i))
    j = np.atleast_1d(np.asarray(j))
    k = max(len(i), len(j))
    rec = geom_records(k)
    rec["kind"] = L.GEOM_CROP_REFLECT
    rec["dy"] = np.broadcast_to(i, (k,))
    rec["dx"] = np.broadcast_to(j, (k,))
    rec["pad"] = pad
    return geometry(x, rec, out_hw=(size, size), out=out)


# ------------------------------------------------------------------------------------------------
# Pillow resample
# ------------------------------------------------------------------------------------------------
_table_cache = {}


def resample_table_host(in_size: int, out_size: int, filt: int) -> np.ndarray:
    """Pillow's Q22 coefficient table for one axis, built by the library's host helper."""
    lib = L.load()
    words = lib.imgtx_resample_table_words(in_size, out_size, filt)
    if words < 0:
        raise ValueError("bad resample table request")
    tab = np.empty(words, dtype=np.int32)
    L.check(lib.imgtx_resample_table_host(in_size, out_size, filt, tab.ctypes.data))
    return tab


def _resample_table(in_size: int, out_size: int, filt: int, device) -> torch.Tensor:
    key = (in_size, out_size, filt, str(device))
    t = _table_cache.get(key)
    if t is None:
        t = torch.from_numpy(resample_table_host(in_size, out_size, filt)).to(device)
        _table_cache[key] = t
    return t


def resample(x: torch.Tensor, resized_wh, filt: int, crop_box=None, crop_xy=None, out_hw=None, place_xy=(0, 0),
             out=None) -> torch.Tensor:
    """canvas(out_hw) <- resize(crop(x, box), resized_wh, filt) placed at place_xy (negative = crop).

    crop_box = (x, y, w, h) uniform; crop_xy = (n, 2) per-image origins (x, y) overriding the box origin."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    bx, by, bw, bh = crop_box if crop_box is not None else (0, 0, w, h)
    rw, rh = resized_wh
    oh, ow = out_hw if out_hw is not None else (rh, rw)
    out = _check_out(out, x, (n, oh, ow, c))
    ht = _resample_table(bw, rw, filt, x.device) if rw != bw else None
    vt = _resample_table(bh, rh, filt, x.device) if rh != bh else None
    cxy = None
    if crop_xy is not None:
        if not isinstance(crop_xy, torch.Tensor):   # host origins can be validated for free (device ones are clamped in the kernel)
            a = np.asarray(crop_xy, dtype=np.int64).reshape(n, 2)
            if n and (a.min() < 0 or a[:, 0].max() > w - bw or a[:, 1].max() > h - bh):
                raise ValueError("crop_xy places the crop box outside the image")
        cxy = crop_xy if isinstance(crop_xy, torch.Tensor) else torch.from_numpy(
            np.ascontiguousarray(np.asarray(crop_xy, dtype=np.int32).reshape(n, 2)))
        cxy = cxy.to(device=x.device, dtype=torch.int32).contiguous()
    L.check(L.load().imgtx_resample_u8(
        x.data_ptr(), out.data_ptr(), n, h, w, c, int(bx), int(by), int(bw), int(bh),
        cxy.data_ptr() if cxy is not None else None, int(rw), int(rh),
        ht.data_ptr() if ht is not None else None, ht.numel() if ht is not None else 0,
        vt.data_ptr() if vt is not None else None, vt.numel() if vt is not None else 0,
        oh, ow, int(place_xy[0]), int(place_xy[1]), _stream()))
    return out


def apply_scale(x: torch.Tensor, s: float, out=None) -> torch.Tensor:
    """reference apply_scale: LANCZOS resize to int(w*s) x int(h*s), centre crop (s>1) or zero-pad paste (s<1)."""
    n, h, w, c = x.shape
    nw, nh = int(w * s), int(h * s)
    if s > 1.0:
        place = (-((nw - w) // 2), -((nh - h) // 2))
        return resample(x, (nw, nh), L.FILTER_LANCZOS, out_hw=(h, w), place_xy=place, out=out)
    if s < 1.0:
        place = ((w - nw) // 2, (h - nh) // 2)
        return resample(x, (nw, nh), L.FILTER_LANCZOS, out_hw=(h, w), place_xy=place, out=out)
    return resample(x, (nw, nh), L.FILTER_LANCZOS, out=out)


def crop_resize(x: torch.Tensor, cx, cy, cs: int, out_size: int = 32, filt: int = None, out=None) -> torch.Tensor:
    """rand_crop: crop (cx, cy, cx+cs, cy+cs) per image, then resize((out_size, out_size)) (BICUBIC)."""
    n = x.shape[0]
    filt = L.FILTER_BICUBIC if filt is None else filt
    cxy = np.stack([np.broadcast_to(np.asarray(cx), (n,)), np.broadcast_to(np.asarray(cy), (n,))], axis=1)
    return resample(x, (out_size, out_size), filt, crop_box=(0, 0, cs, cs), crop_xy=cxy, out=out)


def resize_center_crop(x: torch.Tensor, size: int, crop: int, out=None) -> torch.Tensor:
    """T.Resize(size) (BILINEAR, smaller edge) + T.CenterCrop(crop) on PIL images."""
    n, h, w, c = x.shape
    if w <= h:
        ow, oh = size, int(size * h / w)
    else:
        oh, ow = size, int(size * w / h)
    top = int(round((oh - crop) / 2.0))
    left = int(round((ow - crop) / 2.0))
    return resample(x, (ow, oh), L.FILTER_BILINEAR, out_hw=(crop, crop), place_xy=(-left, -top), out=out)


# ------------------------------------------------------------------------------------------------
# stencils
# ------------------------------------------------------------------------------------------------
def cv_gaussian_blur(x: torch.Tensor, ksize: int, sigma: float, out=None, literal: bool = False) -> torch.Tensor:
    """cv2.GaussianBlur(img, (ksize, ksize), sigma) on uint8.  `literal` (tests): the plain per-tap loop instead of the
    packed dp4a form — both are the same integer arithmetic."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    L.check(L.load().imgtx_cv_gaussian_blur_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, int(ksize), float(sigma),
                                               L.CC_REPLAY_ALL if literal else 0, _stream()))
    return out


def apply_blur(x: torch.Tensor, blur_radius: float, out=None) -> torch.Tensor:
    """reference apply_blur: ksize = odd(int(6 r)) (min 3), sigma = r; r == 0 is the identity."""
    if blur_radius == 0:
        return pointwise(x, L.PW_IDENTITY, 0.0, out)
    ksize = int(blur_radius * 6)
    if ksize % 2 == 0:
        ksize += 1
    if ksize < 3:
        ksize = 3
    return cv_gaussian_blur(x, ksize, blur_radius, out)


def pil_gaussian_blur(x: torch.Tensor, radius: float, out=None) -> torch.Tensor:
    """img.filter(ImageFilter.GaussianBlur(radius))."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    L.check(L.load().imgtx_pil_gaussian_blur_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, float(radius), _stream()))
    return out


def box_filter_h(x: torch.Tensor, size: int, out=None) -> torch.Tensor:
    """TransformationPool.motion_blur: cv2.filter2D with a horizontal 1/size box."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    L.check(L.load().imgtx_box_filter_h_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, int(size), _stream()))
    return out


# ------------------------------------------------------------------------------------------------
# noise
# ------------------------------------------------------------------------------------------------
_NOISE_DRAW_DTYPE = {L.NOISE_GAUSS_F32: torch.float32, L.NOISE_GAUSS_F64: torch.float64,
                     L.NOISE_IMPULSE: torch.float64, L.NOISE_SHOT: torch.int64}


def noise_injected(x: torch.Tensor, kind: int, draws, param: float = 0.0, out=None) -> torch.Tensor:
    """Parity mode: `draws` are the reference's RNG draws (numpy array or tensor, see include/imgtx.h)."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    dt = _NOISE_DRAW_DTYPE[kind]
    d = draws if isinstance(draws, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(draws))
    d = d.to(device=x.device, dtype=dt).contiguous()
    want = n * h * w * (1 if kind == L.NOISE_IMPULSE else c)
    if d.numel() != want:
        raise ValueError(f"expected {want} draws, got {d.numel()}")
    L.check(L.load().imgtx_noise_injected_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, kind, d.data_ptr(), float(param),
                                             _stream()))
    return out


def noise_philox(x: torch.Tensor, kind: int, param: float, seed: int, out=None) -> torch.Tensor:
    """Throughput mode: draws generated on chip (Philox4x32-10).  Gaussian noise on bytes is "add floor(std Z), clip" — one
    discrete law — so it is sampled in the output domain through a Walker alias table (one Philox word, one lookup per
    element; transformation.py:272-281) whenever the table exists (std <= 39); Box-Muller otherwise."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    if kind in (L.NOISE_GAUSS_F32, L.NOISE_GAUSS_F64) and param > 0 and (n * h * w * c) % 4 == 0:
        from .cifar_c import noise_alias_table
        tab = noise_alias_table(L.CC_GAUSSIAN, float(param) / 255.0, x.device)
        if tab is not None:
            with torch.cuda.device_of(x):
                L.check(L.load().imgtx_cc_noise_onchip_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, L.CC_GAUSSIAN,
                                                          tab.data_ptr(), tab.numel(), int(seed) & 0xFFFFFFFFFFFFFFFF, 0, None,
                                                          _stream()))
            return out
    L.check(L.load().imgtx_noise_philox_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, kind, float(param),
                                           int(seed) & 0xFFFFFFFFFFFFFFFF, _stream()))
    return out


# ------------------------------------------------------------------------------------------------
# tensors
# ------------------------------------------------------------------------------------------------
def to_tensor(x: torch.Tensor, mean=None, std=None, out=None) -> torch.Tensor:
    """ToTensor() [+ Normalize(mean, std)]: (n,h,w,c) u8 -> (n,c,h,w) f32."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device) if out is None else out
    mk, mp = _fptr(mean)
    sk, sp = _fptr(std)
    L.check(L.load().imgtx_to_tensor_f32(x.data_ptr(), out.data_ptr(), n, h, w, c, mp, sp, _stream()))
    return out


def to_uint8(t: torch.Tensor, mean=None, std=None, clamp=False, out=None) -> torch.Tensor:
    """tensor_to_pil (mean/std given, clamp=True) or TF.to_pil_image (plain): (n,c,h,w) f32 -> (n,h,w,c) u8."""
    t = _check_f32(t)
    n, c, h, w = t.shape
    out = torch.empty((n, h, w, c), dtype=torch.uint8, device=t.device) if out is None else out
    flags = (L.TOU8_DENORMALIZE if mean is not None else 0) | (L.TOU8_CLAMP01 if clamp else 0)
    mk, mp = _fptr(mean)
    sk, sp = _fptr(std)
    L.check(L.load().imgtx_to_uint8_f32(t.data_ptr(), out.data_ptr(), n, h, w, c, mp, sp, flags, _stream()))
    return out


# ------------------------------------------------------------------------------------------------
# sign-step family
# ------------------------------------------------------------------------------------------------
def _bounds(lo, hi, c):
    if lo is None:
        return L.CLAMP_NONE, None, None, None, None
    lo_l = [float(v) for v in np.atleast_1d(np.asarray(lo, dtype=np.float64).reshape(-1))]
    hi_l = [float(v) for v in np.atleast_1d(np.asarray(hi, dtype=np.float64).reshape(-1))]
    if len(lo_l) == 1 and len(hi_l) == 1:
        mode = L.CLAMP_SCALAR
    else:
        if len(lo_l) != c or len(hi_l) != c:
            raise ValueError("per-channel bounds must have one entry per channel")
        mode = L.CLAMP_PER_CHANNEL
    lk, lp = _fptr(lo_l)
    hk, hp = _fptr(hi_l)
    return mode, lk, lp, hk, hp


def shannon_entropy(x: torch.Tensor, mean=None, std=None, return_counts: bool = False):
    """Per-image Shannon entropy (bits) of the 256-bin histogram of all c*h*w values: (n,c,h,w) f32 -> f64[n] on the
    device (and the exact u32 counts (n,256) when return_counts).  mean/std given: x is normalised and is mapped
    x*std+mean, clamp[0,1] first (fall_2025/TENT_Pipelines.py:99-121); else values outside [0,1] are not counted
    (fall_2025/Initial_Experiments.py:95-113)."""
    x = _check_f32(x)
    if x.dim() == 3:
        x = x.unsqueeze(0)
    if x.dim() != 4:
        raise ValueError("expected a (n,c,h,w) or (c,h,w) tensor")
    n, c, h, w = x.shape
    ent = torch.empty(n, dtype=torch.float64, device=x.device)
    counts = torch.empty((n, 256), dtype=torch.int32, device=x.device) if return_counts else None
    mk, mp = _fptr(mean)
    sk, sp = _fptr(std)
    L.check(L.load().imgtx_shannon_entropy_f32(x.data_ptr(), n, c, h, w, mp, sp, ent.data_ptr(),
                                               counts.data_ptr() if counts is not None else None, _stream()))
    return (ent, counts) if return_counts else ent


def sign_step(x: torch.Tensor, g: torch.Tensor, eps, direction: float = 1.0, lo=None, hi=None,
              return_pert: bool = False, out: Optional[torch.Tensor] = None):
    """x + direction*eps*sign(g) (optionally clamped). eps scalar -> tensor like x; eps sequence of E
    values -> (E, *x.shape) from ONE pass over x and g.  `out` (scalar eps only) may be x itself: every element is read
    before it is written, by the same thread."""
    x = _check_f32(x)
    g = _check_f32(g)
    if x.shape != g.shape:
        raise ValueError("x and g must have the same shape")
    shp = x.shape
    if x.dim() == 4:
        n, c, hw = shp[0], shp[1], shp[2] * shp[3]
    elif x.dim() == 3:
        n, c, hw = 1, shp[0], shp[1] * shp[2]
    else:
        n, c, hw = 1, 1, x.numel()
    many = not np.isscalar(eps)
    eps_l = [float(e) for e in (eps if many else [eps])]
    E = len(eps_l)
    if out is not None:
        if many or not (out.is_cuda and out.dtype == torch.float32 and out.shape == x.shape and out.is_contiguous()):
            raise ValueError("out must be a contiguous float32 CUDA tensor shaped like x (scalar eps only)")
        out = out.view((1,) + tuple(shp))
    else:
        out = torch.empty((E,) + tuple(shp), dtype=torch.float32, device=x.device)
    pert = torch.empty_like(out) if return_pert else None
    mode, lk, lp, hk, hp = _bounds(lo, hi, c)
    ek, ep = _fptr(eps_l)
    L.check(L.load().imgtx_sign_step_f32(x.data_ptr(), g.data_ptr(), out.data_ptr(),
                                         pert.data_ptr() if pert is not None else None, n, c, hw, ep, E,
                                         float(direction), mode, lp, hp, _stream()))
    if not many:
        out = out[0]
        pert = pert[0] if pert is not None else None
    return (out, pert) if return_pert else out


def sign_step_post(x: torch.Tensor, g: torch.Tensor, eps, lut, mean, std, direction: float = 1.0, lo=None, hi=None):
    """BASELINE config 5 in one kernel: sign step (+clamp) -> tensor_to_pil(mean, std) -> 256-entry uint8 map `lut` ->
    pil_to_tensor(mean, std), for every eps of the list: (E, n, c, h, w) f32 from ONE pass over x and g."""
    x, g = _check_f32(x), _check_f32(g)
    if x.shape != g.shape or x.dim() != 4:
        raise ValueError("x and g must be (n, c, h, w) tensors of the same shape")
    n, c, hw = x.shape[0], x.shape[1], x.shape[2] * x.shape[3]
    eps_l = [float(e) for e in (eps if not np.isscalar(eps) else [eps])]
    out = torch.empty((len(eps_l),) + tuple(x.shape), dtype=torch.float32, device=x.device)
    mode, lk, lp, hk, hp = _bounds(lo, hi, c)
    ek, ep = _fptr(eps_l)
    mk, mp = _fptr(mean)
    sk, sp = _fptr(std)
    lut = np.ascontiguousarray(np.asarray(lut, np.uint8).reshape(256))
    L.check(L.load().imgtx_sign_step_post_f32(x.data_ptr(), g.data_ptr(), out.data_ptr(), n, c, hw, ep, len(eps_l),
                                              float(direction), mode, lp, hp, mp, sp, lut.ctypes.data, _stream()))
    return out


def pgd_step(x, x0, g, alpha: float, eps: float, lo=None, hi=None):
    x, x0, g = _check_f32(x), _check_f32(x0), _check_f32(g)
    shp = x.shape
    if x.dim() == 4:
        n, c, hw = shp[0], shp[1], shp[2] * shp[3]
    else:
        n, c, hw = 1, 1, x.numel()
    out = torch.empty_like(x)
    mode, lk, lp, hk, hp = _bounds(lo, hi, c)
    L.check(L.load().imgtx_pgd_step_f32(x.data_ptr(), x0.data_ptr(), g.data_ptr(), out.data_ptr(), n, c, hw,
                                        float(alpha), float(eps), mode, lp, hp, _stream()))
    return out


def channel_affine(x: torch.Tensor, mean, std, denormalize: bool) -> torch.Tensor:
    """normalize_tensor / denormalize_tensor of the reference on a (c,h,w) or (n,c,h,w) f32 CUDA tensor:
    (x - mean) / std, or x * std + mean when denormalize (pipenline/cifar_image_transformations.py:19-25)."""
    x = _check_f32(x)
    if x.dim() not in (3, 4):
        raise ValueError("expected a (c,h,w) or (n,c,h,w) tensor")
    c, h, w = x.shape[-3:]
    n = x.shape[0] if x.dim() == 4 else 1
    if len(mean) != c or len(std) != c:
        raise ValueError("one mean / std per channel")
    out = torch.empty_like(x)
    mk, mp = _fptr(list(mean))
    sk, sp = _fptr(list(std))
    with torch.cuda.device_of(x):
        L.check(L.load().imgtx_channel_affine_f32(x.data_ptr(), out.data_ptr(), n, c, h * w, mp, sp, int(denormalize),
                                                  _stream()))
    return out


def float_corrupt(x, mode: int, a: float = 0.0, b: float = 0.0, noise=None, clamp: bool = True):
    x = _check_f32(x)
    out = torch.empty_like(x)
    if noise is not None:
        noise = _check_f32(noise)
    L.check(L.load().imgtx_float_corrupt_f32(x.data_ptr(), noise.data_ptr() if noise is not None else None,
                                             out.data_ptr(), x.numel(), mode, float(a), float(b), int(clamp),
                                             _stream()))
    return out


# ------------------------------------------------------------------------------------------------
# interpolating warps
# ------------------------------------------------------------------------------------------------
def pil_transform(x: torch.Tensor, coeffs, out_hw=None, perspective: bool = False, bicubic: bool = False, fill=0,
                  out=None) -> torch.Tensor:
    """PIL img.transform(size, AFFINE|PERSPECTIVE, coeffs, resample=BILINEAR|BICUBIC, fillcolor=fill).
    coeffs: (6|8,) uniform or (n, 6|8) per image (doubles)."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    oh, ow = out_hw if out_hw is not None else (h, w)
    out = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device) if out is None else out
    k = 8 if perspective else 6
    co = np.asarray(coeffs, dtype=np.float64)
    host = dev = None
    if co.ndim == 1:
        host = np.zeros(8, np.float64)
        host[:k] = co[:k]
    else:
        full = np.zeros((n, 8), np.float64)
        full[:, :k] = co.reshape(n, -1)[:, :k]
        dev = torch.from_numpy(full).to(x.device)
    fv = [int(v) for v in (fill if isinstance(fill, (tuple, list)) else [fill] * c)]
    fill_arr = (C.c_uint8 * 4)(*((fv + [0] * 4)[:4]))
    L.check(L.load().imgtx_pil_transform_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, oh, ow, int(perspective),
                                            int(bicubic), host.ctypes.data if host is not None else None,
                                            dev.data_ptr() if dev is not None else None,
                                            C.cast(fill_arr, C.c_void_p), _stream()))
    return out


def apply_shear(x: torch.Tensor, sf: float, out=None) -> torch.Tensor:
    """reference apply_shear: AFFINE (1, sf, -shift, 0, 1, 0), BICUBIC, white fill, width + ceil(sf*h)."""
    n, h, w, c = x.shape
    shift = int(math.ceil(sf * h))
    return pil_transform(x, (1, sf, -shift if sf > 0 else 0, 0, 1, 0), out_hw=(h, w + shift), bicubic=True,
                         fill=255, out=out)


def perspective_coeffs(startpoints, endpoints) -> np.ndarray:
    """torchvision _get_perspective_coeffs (least squares in f64): output pixel -> input pixel."""
    a = np.zeros((8, 8), np.float64)
    for i, (p1, p2) in enumerate(zip(endpoints, startpoints)):
        a[2 * i] = [p1[0], p1[1], 1, 0, 0, 0, -p2[0] * p1[0], -p2[0] * p1[1]]
        a[2 * i + 1] = [0, 0, 0, p1[0], p1[1], 1, -p2[1] * p1[0], -p2[1] * p1[1]]
    b = np.asarray(startpoints, np.float64).reshape(8)
    return np.linalg.lstsq(a, b, rcond=None)[0]


def perspective_tensor(x: torch.Tensor, coeffs, fill: float = 0.0, out=None) -> torch.Tensor:
    """ToTensor -> TF.perspective(tensor, BILINEAR, fill) -> ToPILImage on a u8 batch; coeffs (n,8) or (8,)."""
    x = _check_u8(x)
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    co = np.broadcast_to(np.asarray(coeffs, np.float64).astype(np.float32).reshape(-1, 8), (n, 8))
    cd = torch.from_numpy(np.array(co, dtype=np.float32, order="C")).to(x.device)
    L.check(L.load().imgtx_perspective_tensor_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, cd.data_ptr(), float(fill),
                                                 _stream()))
    return out


def gaussian_blur_f32(t: torch.Tensor, kernel_size: int = 3, sigma: float = None) -> torch.Tensor:
    """TF.gaussian_blur(t, kernel_size[, sigma]) on f32 (..., h, w)."""
    t = _check_f32(t)
    if sigma is None:
        sigma = 0.3 * ((kernel_size - 1) * 0.5 - 1) + 0.8
    h, w = t.shape[-2:]
    out = torch.empty_like(t)
    L.check(L.load().imgtx_gaussian_blur_f32(t.data_ptr(), out.data_ptr(), t.numel() // (h * w), h, w, int(kernel_size),
                                             float(sigma), _stream()))
    return out
