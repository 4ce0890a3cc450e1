"""Fused TTA view expansion (host side of imgtx_views_f32): builds the uniform view list, the per-image parameter
tensor and the Pillow Q22 resample tables, and launches ONE kernel that writes every view of every image."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L
from . import batch as B

VIEW_DTYPE = np.dtype([("kind", "<i4"), ("table_h", "<i4"), ("table_v", "<i4"), ("i0", "<i4"), ("i1", "<i4"),
                       ("reserved", "<i4", (3,))])
assert VIEW_DTYPE.itemsize == 32

_blob_cache = {}


def _table_blob(specs, device):
    """specs: list of (in_size, out_size, filter) -> (device blob, host offsets int32[n+1])."""
    key = (tuple(specs), str(device))
    if key not in _blob_cache:
        tabs = [B.resample_table_host(a, b, f) for a, b, f in specs]
        offs = np.zeros(len(tabs) + 1, np.int32)
        for i, t in enumerate(tabs):
            offs[i + 1] = offs[i] + len(t)
        blob = np.concatenate(tabs) if tabs else np.zeros(1, np.int32)
        _blob_cache[key] = (torch.from_numpy(blob).to(device), offs)
    return _blob_cache[key]


def expand(x: torch.Tensor, views: np.ndarray, params: np.ndarray, table_specs, out_mean, out_std, in_mean=None,
           in_std=None, out=None) -> torch.Tensor:
    """x: (n,h,w,3) uint8 or (n,3,h,w) normalised f32 (CUDA).  views: VIEW_DTYPE[V].  params: (n,V,2) f32."""
    is_f32 = x.dtype == torch.float32
    if not x.is_cuda or x.dim() != 4 or x.dtype not in (torch.float32, torch.uint8):
        raise TypeError("expected a CUDA uint8 (n,h,w,3) or float32 (n,3,h,w) tensor")
    x = x.contiguous()
    if is_f32:
        n, c, h, w = x.shape
    else:
        n, h, w, c = x.shape
    if c != 3:
        raise ValueError("views need RGB images")
    V = len(views)
    out = torch.empty((n, V, 3, h, w), dtype=torch.float32, device=x.device) if out is None else out
    pd = params if isinstance(params, torch.Tensor) else torch.from_numpy(
        np.ascontiguousarray(np.asarray(params, np.float64).astype(np.float32).reshape(n, V, 2)))
    pd = pd.to(device=x.device, dtype=torch.float32).contiguous()
    blob, offs = _table_blob(table_specs, x.device)
    vk = np.ascontiguousarray(views)
    km, kmp = B._fptr(in_mean)
    ks, ksp = B._fptr(in_std)
    om, omp = B._fptr(out_mean)
    os_, osp = B._fptr(out_std)
    L.check(L.load().imgtx_views_f32(x.data_ptr(), int(is_f32), out.data_ptr(), n, h, w, V, vk.ctypes.data,
                                     pd.data_ptr(), blob.data_ptr(), offs.ctypes.data, len(table_specs), kmp, ksp,
                                     omp, osp, B._stream()))
    return out


def tta16_draws(n: int, crop_xy, brightness, contrast, device) -> torch.Tensor:
    """The (n,16,2) f32 per-view parameter block of tta16 on the device: upload the draws once, reuse across calls."""
    params = np.zeros((n, 16, 2), np.float32)
    params[:, 1:6, :] = np.asarray(crop_xy, np.float64).reshape(n, 5, 2)
    params[:, 6:11, 0] = np.asarray(brightness, np.float64).reshape(n, 5)
    params[:, 11:16, 0] = np.asarray(contrast, np.float64).reshape(n, 5)
    return torch.from_numpy(params).to(device)


def tta16(tn: torch.Tensor, crop_xy, brightness, contrast, mean, std, out=None, draws=None) -> torch.Tensor:
    """The 16-view expansion of tta_initial_claudeFixed.py (AUG_COUNTS [1,5,5,5]) for normalised f32 input (or a
    uint8 batch, which skips the tensor_to_pil prologue).  draws: the device block of tta16_draws (then crop_xy,
    brightness and contrast are ignored and nothing is copied from the host)."""
    n = tn.shape[0]
    h, w = (tn.shape[2], tn.shape[3]) if tn.dtype == torch.float32 else (tn.shape[1], tn.shape[2])
    cs = int(w * 0.78)
    views = np.zeros(16, VIEW_DTYPE)
    views["kind"][0] = L.VIEW_FLIP
    views["kind"][1:6] = L.VIEW_CROP_RESIZE
    views["i0"][1:6] = cs
    views["table_h"][1:6] = 0
    views["table_v"][1:6] = 1
    views["kind"][6:11] = L.VIEW_BRIGHTNESS
    views["kind"][11:16] = L.VIEW_CONTRAST
    if draws is not None:
        if not (isinstance(draws, torch.Tensor) and draws.is_cuda and draws.dtype == torch.float32
                and tuple(draws.shape) == (n, 16, 2)):
            raise TypeError("draws must be the CUDA float32 (n,16,2) block of tta16_draws")
        params = draws
    else:
        params = np.zeros((n, 16, 2), np.float64)
        params[:, 1:6, :] = np.asarray(crop_xy).reshape(n, 5, 2)
        params[:, 6:11, 0] = np.asarray(brightness).reshape(n, 5)
        params[:, 11:16, 0] = np.asarray(contrast).reshape(n, 5)
    specs = [(cs, w, L.FILTER_BICUBIC), (cs, h, L.FILTER_BICUBIC)]
    f32_in = tn.dtype == torch.float32
    return expand(tn, views, params, specs, mean, std, mean if f32_in else None, std if f32_in else None, out=out)


def tta_strategy(x_u8: torch.Tensor, strategy: str, crop_ij, jitter, mean, std, out=None) -> torch.Tensor:
    """The four views of TTA_transforms.tta_transforms_<strategy> for a uint8 batch."""
    n, h, w, _ = x_u8.shape
    views = np.zeros(4, VIEW_DTYPE)
    params = np.zeros((n, 4, 2), np.float64)
    specs = []
    if strategy == "gentle":
        views["kind"] = [L.VIEW_IDENTITY, L.VIEW_FLIP, L.VIEW_CROP_REFLECT, L.VIEW_FLIP_CROP_REFLECT]
        views["i0"][2:4] = 4
        params[:, 2:4, :] = np.asarray(crop_ij).reshape(n, 2, 2)
    elif strategy == "multiscale":
        views["kind"] = [L.VIEW_IDENTITY, L.VIEW_RESIZE_CENTER_CROP, L.VIEW_RESIZE_CENTER_CROP, L.VIEW_FLIP]
        for k, size in ((1, 36), (2, 40)):
            ow, oh = (size, int(size * h / w)) if w <= h else (int(size * w / h), size)
            views["table_h"][k] = len(specs)
            specs.append((w, ow, L.FILTER_BILINEAR))
            views["table_v"][k] = len(specs)
            specs.append((h, oh, L.FILTER_BILINEAR))
            views["i0"][k] = int(round((ow - w) / 2.0))
            views["i1"][k] = int(round((oh - h) / 2.0))
    elif strategy == "color":
        views["kind"] = [L.VIEW_IDENTITY, L.VIEW_BRIGHTNESS, L.VIEW_CONTRAST, L.VIEW_FLIP]
        j = np.asarray(jitter, np.float64).reshape(n, 2)
        params[:, 1, 0] = j[:, 0]
        params[:, 2, 0] = j[:, 1]
    else:
        raise ValueError(strategy)
    return expand(x_u8, views, params, specs, mean, std, out=out)
