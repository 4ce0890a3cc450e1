"""CIFAR-10-C-style corruption generation: 15 types x 5 severities, batched on the GPU.

The reference names the types (attacks/fgsm/cifar-10/reverse_fgsm_cifar10_corrupted.py:18-23) and consumes
(50000,32,32,3) uint8 severity-major .npy files (fall_2025/TTA_transforms.py:16-26, transformation.py:35-41); its own
generators exist for four types with ImageNet-C constants (pipenline/cifar_image_transformations.py:39-77 — those are the
`table="reference"` variants here and the TransformationPool drop-ins).  `table="cifar"` (default) is the canonical
CIFAR-10-C recipe.  Severity tables are data; small host-side tables (Gaussian taps, the disk kernel, the
merged zoom operator, JPEG quantisers) are built here with numpy and handed to libimgtx.

Random inputs: pass `draws` (a dict of arrays, see `DRAW_KEYS`) to inject them (parity mode), or leave it None to have
them drawn on chip from Philox keyed by (seed, image index) (throughput mode).
"""
from __future__ import annotations

import ctypes as C
import functools
import math
import os
from typing import Optional

import numpy as np
import torch

from . import _lib as L
from . import batch as B

CORRUPTIONS = ["gaussian_noise", "shot_noise", "impulse_noise", "defocus_blur", "glass_blur", "motion_blur",
               "zoom_blur", "snow", "frost", "fog", "brightness", "contrast", "elastic_transform", "pixelate",
               "jpeg_compression"]
EXTRA = ["speckle_noise"]

CIFAR_TABLE = {
    "gaussian_noise": [0.04, 0.06, 0.08, 0.09, 0.10],
    "shot_noise": [500, 250, 100, 75, 50],
    "impulse_noise": [0.01, 0.02, 0.03, 0.05, 0.07],
    "speckle_noise": [0.06, 0.1, 0.12, 0.16, 0.2],
    "defocus_blur": [(0.3, 0.4), (0.4, 0.5), (0.5, 0.6), (1, 0.2), (1.5, 0.1)],
    "glass_blur": [(0.05, 1, 1), (0.25, 1, 1), (0.4, 1, 1), (0.25, 1, 2), (0.4, 1, 2)],
    "motion_blur": [(10, 1), (10, 1.5), (10, 2), (10, 2.5), (12, 3)],
    "zoom_blur": [(1, 1.06), (1, 1.11), (1, 1.16), (1, 1.21), (1, 1.26)],
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
# the reference's own generators (ImageNet-C constants) — pipenline/cifar_image_transformations.py:39-77
REFERENCE_TABLE = {
    "gaussian_noise": [0.08, 0.12, 0.18, 0.26, 0.38],
    "impulse_noise": [0.03, 0.06, 0.09, 0.17, 0.27],
    "shot_noise": [60, 25, 12, 5, 3],
    "defocus_blur": [3, 4, 6, 8, 10],
}
DRAW_KEYS = {
    "gaussian_noise": ("noise",), "speckle_noise": ("noise",), "shot_noise": ("counts",),
    "impulse_noise": ("u_flip", "u_salt"), "glass_blur": ("deltas",), "motion_blur": ("angles",),
    "snow": ("layer", "angles"), "frost": ("frost",), "fog": ("plasma",), "elastic_transform": ("jitter", "ux", "uy"),
}


# ------------------------------------------------------------------------------------------------
# host-side tables (numpy, float64 unless stated)
# ------------------------------------------------------------------------------------------------
def _reflect101(i, n):
    if n == 1:
        return np.zeros_like(i)
    period = 2 * n - 2
    m = np.mod(i, period)
    return np.where(m >= n, period - m, m)


@functools.lru_cache(maxsize=None)
def disk_taps(radius: float, alias_blur: float):
    """make_cifar_c.disk(radius, alias_blur) as sparse taps: the 17x17 indicator disk, normalised in f32, blurred with
    OpenCV's 3x3 f32 Gaussian (symmetric separable form, REFLECT_101).  Returns ((ntaps,2) int32 dy/dx, (ntaps,) f64)."""
    f32 = np.float32
    Lr = np.arange(-8, 8 + 1)
    X, Y = np.meshgrid(Lr, Lr)
    d = np.array((X ** 2 + Y ** 2) <= radius ** 2, dtype=f32)
    d /= np.sum(d)
    cf = np.exp(-0.5 / (alias_blur * alias_blur) * np.array([1.0, 0.0, 1.0]))
    k = (cf / cf.sum()).astype(f32)

    def sym3(a, axis):
        n = a.shape[axis]
        idx = _reflect101(np.arange(-1, n + 1), n)
        lo = np.take(a, idx[0:n], axis)
        hi = np.take(a, idx[2:n + 2], axis)
        return (a * k[1] + (lo + hi).astype(f32) * k[0]).astype(f32)

    kern = sym3(sym3(d, 1), 0)
    ys, xs = np.nonzero(kern)
    dydx = np.stack([ys - 8, xs - 8], axis=1).astype(np.int32)
    return np.ascontiguousarray(dydx), kern[ys, xs].astype(np.float64)


@functools.lru_cache(maxsize=None)
def gaussian_taps(sigma: float, truncate: float) -> np.ndarray:
    """scipy.ndimage._gaussian_kernel1d: centre tap first, radius int(truncate*sigma + 0.5)."""
    lw = int(truncate * float(sigma) + 0.5)
    full = np.exp(-0.5 / (sigma * sigma) * np.arange(-lw, lw + 1, dtype=np.float64) ** 2)
    full = full / full.sum()
    return np.ascontiguousarray(full[lw:])


@functools.lru_cache(maxsize=None)
def motion_kernel(radius: float, sigma: float) -> np.ndarray:
    """ImageMagick's one-sided Gaussian line kernel (f32), width 2*ceil(radius)+1."""
    width = int(2.0 * math.ceil(radius) + 1.0)
    i = np.arange(width, dtype=np.float64)
    k = np.exp(-(i * i) / (2.0 * sigma * sigma)) / (math.sqrt(2.0 * math.pi) * sigma)
    return (k / k.sum()).astype(np.float32)


def motion_offsets(width: int, angles_deg) -> np.ndarray:
    """(n, width, 2) int32 (dx, dy) pixel offsets of the line taps for each angle."""
    a = np.radians(np.asarray(angles_deg, np.float64)).reshape(-1, 1)
    px, py = width * np.sin(a), width * np.cos(a)
    hyp = np.hypot(px, py)
    i = np.arange(width, dtype=np.float64)[None, :]
    off = np.stack([np.ceil(i * py / hyp - 0.5), np.ceil(i * px / hyp - 0.5)], axis=2)
    return np.ascontiguousarray(off.astype(np.int32))


@functools.lru_cache(maxsize=None)
def zoom_axis_table(size: int, z: float):
    """clipped_zoom along one axis: source indices (i0, i1) and fraction t (f32) per output sample
    (scipy.ndimage.zoom order=1 coordinate o*(in-1)/(out-1), centre crop before, centre trim after)."""
    ch = int(np.ceil(size / z))
    top = (size - ch) // 2
    osz = int(round(ch * z))
    trim = (osz - size) // 2
    o = np.arange(size) + trim
    cc = o.astype(np.float64) * (np.float64(ch - 1) / np.float64(osz - 1)) if osz > 1 else np.zeros(size)
    i0 = np.floor(cc).astype(np.int64)
    t = (cc - i0).astype(np.float32)
    i0 = np.clip(i0, 0, ch - 1)
    i1 = np.minimum(i0 + 1, ch - 1)
    return i0 + top, i1 + top, t


_zoom_cache = {}


_zoom_level_cache = {}


def zoom_level_tables(h: int, w: int, severity: int, device):
    """Per-level, per-axis tables of zoom_blur on `device`: (yidx (nz,2,h) int32, yt (nz,h) f32, xidx (nz,2,w), xt (nz,w), nz)."""
    key = (h, w, severity, str(device))
    if key not in _zoom_level_cache:
        lo, hi = CIFAR_TABLE["zoom_blur"][severity - 1]
        zs = np.arange(lo, hi, 0.01)
        ya = [zoom_axis_table(h, float(z)) for z in zs]
        xa = [zoom_axis_table(w, float(z)) for z in zs]
        yi = np.stack([np.stack([a[0], a[1]]) for a in ya]).astype(np.int32)
        xi = np.stack([np.stack([a[0], a[1]]) for a in xa]).astype(np.int32)
        yt = np.stack([a[2] for a in ya]).astype(np.float32)
        xt = np.stack([a[2] for a in xa]).astype(np.float32)
        _zoom_level_cache[key] = tuple(torch.from_numpy(np.ascontiguousarray(a)).to(device) for a in (yi, yt, xi, xt)) + (len(zs),)
    return _zoom_level_cache[key]


def zoom_operator(h: int, w: int, severity: int, device):
    """zoom_blur as one sparse linear operator W = (I + sum_z Zoom_z)/(n+1) in CSR form on `device` (taps of a row merged
    in f64, sorted by source index, weights f32)."""
    key = (h, w, severity, str(device))
    if key in _zoom_cache:
        return _zoom_cache[key]
    lo, hi = CIFAR_TABLE["zoom_blur"][severity - 1]
    zs = np.arange(lo, hi, 0.01)
    P = h * w
    # entries (row, column, weight) in the order the canonical loop visits them: the identity, then per zoom factor the four
    # bilinear taps; equal (row, column) pairs are then summed in float64 IN THAT ORDER (a stable sort keeps it), which is
    # what a dense accumulation does, without the P x P matrix (20 GB at 224 x 224)
    rows = np.arange(P, dtype=np.int64)
    r_l, c_l, w_l = [rows], [rows], [np.ones(P, np.float64)]
    for z in zs:
        y0, y1, ty = zoom_axis_table(h, float(z))
        x0, x1, tx = zoom_axis_table(w, float(z))
        ty, tx = ty.astype(np.float64), tx.astype(np.float64)
        for ya, wy in ((y0, 1 - ty), (y1, ty)):
            for xa, wx in ((x0, 1 - tx), (x1, tx)):
                r_l.append(rows)
                c_l.append((ya[:, None] * w + xa[None, :]).reshape(-1).astype(np.int64))
                w_l.append((wy[:, None] * wx[None, :]).reshape(-1))
    ekey = np.concatenate(r_l) * P + np.concatenate(c_l)
    wts = np.concatenate(w_l)
    order = np.argsort(ekey, kind="stable")
    ekey, wts = ekey[order], wts[order]
    first = np.concatenate([[True], ekey[1:] != ekey[:-1]])
    group = np.cumsum(first) - 1
    start = np.nonzero(first)[0]
    rank = np.arange(len(ekey)) - start[group]
    acc = np.zeros(len(start), np.float64)
    for r in range(int(rank.max()) + 1):
        sel = rank == r
        acc[group[sel]] += wts[sel]
    acc /= (len(zs) + 1)
    keep = acc != 0.0
    uekey, acc = ekey[start][keep], acc[keep]
    nzr, nzc = uekey // P, uekey % P
    indptr = np.zeros(P + 1, np.int64)
    np.add.at(indptr, nzr + 1, 1)
    indptr = np.cumsum(indptr).astype(np.int32)
    wf32 = acc.astype(np.float32)
    op = (torch.from_numpy(indptr).to(device), torch.from_numpy(nzc.astype(np.int32)).to(device),
          torch.from_numpy(wf32).to(device), int(np.diff(indptr).max()))
    _zoom_cache[key] = op
    _zoom_host_csr[key] = (indptr, nzc.astype(np.int64), acc)
    return op


_zoom_host_csr = {}
_zoom_mma_cache = {}
OP32_MAGIC = 0x4F503332   # 'OP32'


def op32_mma_tables(indptr, cols, weights):
    """A sparse row-stochastic operator with non-negative weights on 32x32 images (zoom_blur's merged operator) compiled for
    imgtx_cc_op32_mma_u8 (csrc/cifar_c_zoom.cu): 64 tiles of 4x4 output pixels, each contracting over the aligned 4-pixel
    groups ("quads") of the source rows its taps touch, 8 quads = one m16n8k32 integer-MMA step, at most 3 steps per tile.
    Weights are Q16 (rounded to nearest, the largest of a row absorbs the rounding so a row sums to 2^16 and flat regions
    stay exact), split into a low and a high byte limb.  Returns (int32 table words, total steps) or None when a tile needs
    more than 3 steps."""
    indptr, cols, weights = np.asarray(indptr, np.int64), np.asarray(cols, np.int64), np.asarray(weights, np.float64)
    if len(indptr) != 1025 or (weights < 0).any():
        return None
    meta = np.zeros((64, 4), np.int32)
    bias = np.zeros((64, 8), np.uint32)
    qtab = np.zeros((64, 3, 4, 2), np.int32)
    hi = np.zeros((64, 3, 32, 4), np.uint32)
    lo_steps = []
    lane = np.arange(32)
    gid, tig = lane >> 2, lane & 3
    m = np.arange(16)

    def tile_pixels(blk):
        ty, tx = divmod(blk, 8)
        return (ty * 4 + ((m % 8) >> 1)) * 32 + tx * 4 + ((m % 8) & 1) * 2 + (m >> 3)   # rows m and m + 8 are neighbours

    def tile_quads(blk):
        return np.unique(np.concatenate([cols[indptr[q]:indptr[q + 1]] for q in tile_pixels(blk)]) // 4)
    # warp w of the kernel owns table slots 4w .. 4w+3: deal the pixel blocks out by step count so the warps carry equal work
    need = [(len(tile_quads(blk)) + 7) // 8 for blk in range(64)]
    order = sorted(range(64), key=lambda blk: -need[blk])
    slot_of = {}
    for i, blk in enumerate(order):
        rnd, k = divmod(i, 16)
        wv = k if rnd % 2 == 0 else 15 - k
        slot_of[wv * 4 + rnd] = blk
    for t in range(64):
        blk = slot_of[t]
        pix = tile_pixels(blk)
        quads = tile_quads(blk)
        ns = (len(quads) + 7) // 8
        if ns > 3 or ns < 1:
            return None
        # Slot order.  The kernel's byte planes have a row pitch of 10 words and an image pitch = 4 (mod 32) words, so the bank
        # of a B-fragment load is 4 * image + 2 * row + column (mod 32): a half-step (the 4 quads the lanes tig = 0..3 load
        # together, for 8 images) is conflict-free iff its quads differ in (2 * row + column) mod 4.  Deal the quads to the
        # half-steps accordingly; padding slots take one of the quads 0..3 (row 0) with a residue the half-step lacks.
        G = 2 * ns
        members, used = [[] for _ in range(G)], [set() for _ in range(G)]
        res = ((quads >> 3) * 2 + (quads & 7)) % 4
        for r in sorted(range(4), key=lambda r: -int((res == r).sum())):
            for qd in quads[res == r]:
                free = [g_ for g_ in range(G) if len(members[g_]) < 4]
                good = [g_ for g_ in free if r not in used[g_]]
                g_ = min(good or free, key=lambda g_: len(members[g_]))
                members[g_].append(int(qd))
                used[g_].add(r)
        qpos = np.full(256, -1, np.int64)
        qpad = np.zeros(ns * 8, np.int64)
        for g_ in range(G):
            for j, qd in enumerate(members[g_]):
                qpos[qd] = g_ * 4 + j
                qpad[g_ * 4 + j] = qd
            for j in range(len(members[g_]), 4):
                lack = [r for r in range(4) if r not in used[g_]]
                r = lack[0] if lack else 0
                used[g_].add(r)
                qpad[g_ * 4 + j] = r          # a zero-weight column: any readable quad does
        W = np.zeros((16, 96), np.int64)
        rb = np.full(16, 32, np.int64)       # rounding bias per row: 2^-11 LSB, so an exactly integral sum floors to itself
        for r, q in enumerate(pix):
            cs, ws = cols[indptr[q]:indptr[q + 1]], weights[indptr[q]:indptr[q + 1]]
            if len(cs) == 0:
                continue
            wq = np.rint(ws * 65536.0).astype(np.int64)
            wq[np.argmax(wq)] += int(round(ws.sum() * 65536.0)) - wq.sum()
            if wq.max() > 65535:              # a weight of 1.0: 65535 x + 256 still floors to x for every byte x
                wq = np.minimum(wq, 65535)
                rb[r] = 256
            W[r, qpos[cs // 4] * 4 + cs % 4] = wq
        meta[t, 0], meta[t, 1], meta[t, 2] = ns, len(lo_steps), (blk // 8) * 128 + (blk % 8) * 4
        bias[t] = (rb[:8] | (rb[8:] << 16)).astype(np.uint32)
        for sidx in range(ns):
            qa, qb = qpad[sidx * 8 + np.arange(4)], qpad[sidx * 8 + 4 + np.arange(4)]
            qtab[t, sidx, :, 0] = ((qa >> 3) * 10 + (qa & 7)) * 4      # byte offset in a plane of row pitch 10 words
            qtab[t, sidx, :, 1] = ((qb >> 3) * 10 + (qb & 7)) * 4
            fr = np.zeros((2, 32, 4), np.uint32)
            for reg in range(4):
                row = gid + (reg & 1) * 8
                kb = sidx * 32 + (reg >> 1) * 16 + tig * 4
                for i in range(4):
                    w_ = W[row, kb + i]
                    fr[0, :, reg] |= ((w_ & 255) << (8 * i)).astype(np.uint32)
                    fr[1, :, reg] |= ((w_ >> 8) << (8 * i)).astype(np.uint32)
            lo_steps.append(fr[0])
            hi[t, sidx] = fr[1]
    lo = np.stack(lo_steps)
    head = np.zeros(16, np.int32)
    parts = [head, meta.reshape(-1), bias.view(np.int32).reshape(-1), qtab.reshape(-1), lo.view(np.int32).reshape(-1),
             hi.view(np.int32).reshape(-1)]
    offs = np.cumsum([0] + [len(a) for a in parts])
    head[:8] = [OP32_MAGIC, len(lo_steps), offs[1], offs[2], offs[3], offs[4], offs[5], offs[6]]
    return np.concatenate(parts).astype(np.int32), len(lo_steps)


def zoom_mma_tables(severity: int, device):
    """zoom_blur's merged operator at 32x32 compiled for the integer tensor cores, on `device` (None if it does not fit)."""
    key = (32, 32, severity, str(device))
    if key not in _zoom_mma_cache:
        zoom_operator(32, 32, severity, device)
        r = op32_mma_tables(*_zoom_host_csr[key])
        _zoom_mma_cache[key] = None if r is None else (torch.from_numpy(r[0]).to(device), r[1])
    return _zoom_mma_cache[key]


_JPEG_LUMA = np.array([16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56,
                       14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113,
                       92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99], np.int64)
_JPEG_CHROMA = np.array([17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99,
                         99, 47, 66, 99, 99, 99, 99, 99, 99] + [99] * 32, np.int64)


@functools.lru_cache(maxsize=None)
def jpeg_qtables(quality: int) -> np.ndarray:
    """libjpeg jpeg_set_quality(quality, force_baseline): (2, 64) int32, natural order."""
    q = max(1, min(100, int(quality)))
    scale = 5000 // q if q < 50 else 200 - q * 2
    return np.stack([np.clip((b * scale + 50) // 100, 1, 255) for b in (_JPEG_LUMA, _JPEG_CHROMA)]).astype(np.int32)


def affine_from_jitter(jitter: np.ndarray, size: int) -> np.ndarray:
    """cv2.getAffineTransform(pts1, pts1 + jitter) for each image: (n, 6) f64 row-major 2x3 matrices."""
    cs = np.float32((size, size)) // 2
    sq = min((size, size)) // 3
    pts1 = np.float32([cs + sq, [cs[0] + sq, cs[1] - sq], cs - sq])
    out = np.empty((len(jitter), 6), np.float64)
    A = np.zeros((6, 6))
    for i in range(3):
        A[i, 0:3] = [pts1[i][0], pts1[i][1], 1]
        A[i + 3, 3:6] = [pts1[i][0], pts1[i][1], 1]
    for k in range(len(jitter)):
        pts2 = pts1 + np.asarray(jitter[k], np.float32)
        out[k] = np.linalg.solve(A, np.concatenate([pts2[:, 0], pts2[:, 1]]).astype(np.float64))
    return out


_alias_cache = {}


def noise_alias_table_host(kind: int, p0: float):
    """Walker alias table (int32 words, layout in csrc/cifar_c_fast.cuh) of the output-domain law of on-chip gaussian_noise
    (p0 = sigma) or shot_noise (p0 = c); None when the parameter is outside the tabulated range."""
    lib = L.load()
    words = lib.imgtx_cc_noise_table_words(kind, float(p0))
    if words <= 0:
        return None
    tab = np.empty(words, np.int32)
    L.check(lib.imgtx_cc_noise_table_host(kind, float(p0), tab.ctypes.data, words))
    return tab


def noise_alias_table(kind: int, p0: float, device):
    key = (kind, float(p0), str(device))
    if key not in _alias_cache:
        tab = noise_alias_table_host(kind, p0)
        _alias_cache[key] = None if tab is None else torch.from_numpy(tab).to(device)
    return _alias_cache[key]


_frost_cache = {}


def procedural_frost_bank(device, n_tex: int = 5, size: int = 96, seed: int = 1234) -> torch.Tensor:
    """(n_tex, size, size, 3) uint8 stand-in for the canonical frost photographs (which are not available): smoothed
    multi-octave noise with a cold tint.  Documented substitute; any bank of textures can be passed instead."""
    key = (str(device), n_tex, size, seed)
    if key not in _frost_cache:
        rng = np.random.default_rng(seed)
        bank = np.empty((n_tex, size, size, 3), np.uint8)
        for t in range(n_tex):
            acc = np.zeros((size, size))
            for octave in (4, 8, 16, 32):
                g = rng.random((size // octave + 2, size // octave + 2))
                acc += np.kron(g, np.ones((octave, octave)))[:size, :size] * octave
            acc = (acc - acc.min()) / (acc.max() - acc.min())
            crystals = (rng.random((size, size)) > 0.93) * rng.random((size, size))
            v = np.clip(0.55 * acc + 0.6 * crystals, 0, 1)
            bank[t] = np.stack([v * 215, v * 235, v * 255], axis=-1).astype(np.uint8)
        _frost_cache[key] = torch.from_numpy(bank).to(device)
    return _frost_cache[key]


# ------------------------------------------------------------------------------------------------
# one corruption
# ------------------------------------------------------------------------------------------------
@functools.lru_cache(maxsize=None)
def _snow_zoom_tables(h: int, w: int, z: float):
    y0, y1, ty = zoom_axis_table(h, z)
    x0, x1, tx = zoom_axis_table(w, z)
    return (np.ascontiguousarray(np.concatenate([y0, y1, x0, x1]).astype(np.int32)),
            np.ascontiguousarray(np.concatenate([ty, tx]).astype(np.float32)))


MODES = {"exact": 0, "fast": L.CC_FAST, "replay": L.CC_REPLAY_ALL}


def _dev(a, dtype, device):
    if a is None:
        return None
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    return t.to(device=device, dtype=dtype).contiguous()


def _ptr(t):
    return t.data_ptr() if t is not None else None


def corrupt(x: torch.Tensor, name: str, severity: int, draws: Optional[dict] = None, seed: int = 0, table: str = "cifar",
            out: Optional[torch.Tensor] = None, frost_bank: Optional[torch.Tensor] = None,
            mode: Optional[str] = None, metrics: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x: (n,h,w,3) uint8 CUDA -> same shape.  `draws` injects the random inputs (DRAW_KEYS[name]); None = on-chip RNG.

    mode: "exact"  the float64-specified arithmetic of oracle/cifar_c.py, bit for bit (default when draws are injected);
          "fast"   float32 / fixed-point throughput kernels, every byte within 1 LSB of "exact" on the same draws
                   (default when the draws are made on chip: no bit-wise oracle exists for those);
          "replay" (tests) "exact" with every element forced through its literal replay.
    metrics: optional (n, 2) int32/uint32 CUDA tensor receiving the pass's per-image metric (byte sum, position-weighted sum
          mod 2^32) — written by the kernel's epilogue, ready to be all-gathered (shard.py)."""
    x = B._check_u8(x)
    with torch.cuda.device_of(x):
        return _corrupt(x, name, severity, draws, seed, table, out, frost_bank, mode, metrics)


def _corrupt(x, name, severity, draws, seed, table, out, frost_bank, mode, metrics=None):
    n, h, w, c = x.shape
    if not 1 <= severity <= 5:
        raise ValueError("severity must be 1..5")
    d = draws or {}
    if mode is None:
        mode = "exact" if draws is not None else "fast"   # an (even empty) draws dict asks for the parity path
    if mode not in MODES:
        raise ValueError("mode must be 'exact', 'fast' or 'replay'")
    flags = MODES[mode]
    dev = x.device
    lib = L.load()
    st = B._stream()
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    out = B._check_out(out, x)
    mp = None
    if metrics is not None:
        if not (metrics.is_cuda and metrics.dtype in (torch.int32, torch.uint32) and tuple(metrics.shape) == (n, 2) and
                metrics.is_contiguous()):
            raise ValueError("metrics must be a contiguous (n, 2) int32 CUDA tensor")
        mp = metrics.data_ptr()
    if table == "reference":
        res = _corrupt_reference(x, name, severity, d, seed, out)
        if mp is not None:
            L.check(lib.imgtx_image_checksums_u8(res.data_ptr(), n, h * w * c, mp, st))
        return res
    if table != "cifar":
        raise ValueError("table must be 'cifar' or 'reference'")
    p = CIFAR_TABLE[name][severity - 1]
    xp, op = x.data_ptr(), out.data_ptr()

    def point(kind, p0, p1=0.0, a=None, b=None, tex=(0, 0, 0)):
        L.check(lib.imgtx_cc_pointwise_u8(xp, op, n, h, w, c, kind, float(p0), float(p1), _ptr(a), _ptr(b), seed, tex[0],
                                          tex[1], tex[2], flags, mp, st))

    # NB every device staging tensor is bound to a local name until the call returns: a temporary would hand its block
    # back to the caching allocator before the next one is staged
    if name in ("gaussian_noise", "shot_noise") and not d and (h * w * c) % 16 == 0:
        kind = L.CC_GAUSSIAN if name == "gaussian_noise" else L.CC_SHOT
        if kind == L.CC_SHOT and h * w * c <= 16384:
            # small images: the flat table (one shared-memory lookup per element); the library answers "unsupported" when
            # the batch cannot be staged beside it (unaligned views) and the row-record table below takes over
            flat = noise_alias_table(L.CC_SHOT_FLAT, float(p), dev)
            if flat is not None:
                rc = lib.imgtx_cc_noise_onchip_u8(xp, op, n, h, w, c, L.CC_SHOT_FLAT, flat.data_ptr(), flat.numel(), seed, flags, mp, st)
                if rc != L.ERR_UNSUPPORTED:
                    L.check(rc)
                    return out
        tab = noise_alias_table(kind, float(p), dev)
        if tab is not None:
            L.check(lib.imgtx_cc_noise_onchip_u8(xp, op, n, h, w, c, kind, tab.data_ptr(), tab.numel(), seed, flags, mp, st))
            return out
    if name in ("gaussian_noise", "speckle_noise"):
        point(L.CC_GAUSSIAN if name == "gaussian_noise" else L.CC_SPECKLE, p, a=_dev(d.get("noise"), torch.float64, dev))
    elif name == "shot_noise":
        point(L.CC_SHOT, p, a=_dev(d.get("counts"), torch.int64, dev))
    elif name == "impulse_noise":
        point(L.CC_IMPULSE, p, a=_dev(d.get("u_flip"), torch.float64, dev), b=_dev(d.get("u_salt"), torch.float64, dev))
    elif name == "brightness":
        point(L.CC_BRIGHTNESS, p)
    elif name == "contrast":
        point(L.CC_CONTRAST, p)
    elif name == "fog":
        point(L.CC_FOG, p[0], p[1], a=_dev(d.get("plasma"), torch.float64, dev))
    elif name == "frost":
        if "frost" in d:
            point(L.CC_FROST, p[0], p[1], a=_dev(d["frost"], torch.uint8, dev))
        else:
            # the stand-in textures must be larger than the image: 96 x 96 for CIFAR, twice the image side above that
            bank = (procedural_frost_bank(dev, size=96 if max(h, w) < 96 else 2 * max(h, w)) if frost_bank is None
                    else frost_bank)
            point(L.CC_FROST, p[0], p[1], a=bank, tex=(bank.shape[0], bank.shape[1], bank.shape[2]))
    elif name == "defocus_blur":
        dydx, wts = disk_taps(p[0], p[1])
        L.check(lib.imgtx_cc_defocus_u8(xp, op, n, h, w, c, len(wts), dydx.ctypes.data, wts.ctypes.data, flags, mp, st))
    elif name == "zoom_blur" and (flags & L.CC_FAST) and h * w > 1024:
        yi, yt, xi, xt, nz = zoom_level_tables(h, w, severity, dev)
        L.check(lib.imgtx_cc_zoom_levels_u8(xp, op, n, h, w, c, nz, yi.data_ptr(), yt.data_ptr(), xi.data_ptr(), xt.data_ptr(),
                                            mp, st))
    elif name == "zoom_blur" and (flags & L.CC_FAST) and h == 32 and w == 32 and zoom_mma_tables(severity, dev) is not None:
        tab, steps = zoom_mma_tables(severity, dev)
        L.check(lib.imgtx_cc_op32_mma_u8(xp, op, n, tab.data_ptr(), tab.numel(), steps, mp, st))
    elif name == "zoom_blur":
        ip, idx, wt, max_row = zoom_operator(h, w, severity, dev)
        L.check(lib.imgtx_cc_csr_ell_u8(xp, op, n, h, w, c, ip.data_ptr(), idx.data_ptr(), wt.data_ptr(), idx.numel(),
                                        max_row, flags, mp, st))
    elif name == "glass_blur":
        wts = gaussian_taps(p[0], 4.0)
        deltas = _dev(d.get("deltas"), torch.int8, dev)
        if deltas is not None and deltas.numel() != n * p[2] * (h - 2) * (w - 2) * 2:
            raise ValueError("deltas must have shape (n, iters, h-2, w-2, 2)")
        L.check(lib.imgtx_cc_glass_u8(xp, op, n, h, w, c, wts.ctypes.data, len(wts) - 1, p[2], _ptr(deltas), seed, flags, mp, st))
    elif name == "motion_blur":
        k = motion_kernel(p[0], p[1])
        off = _dev(motion_offsets(len(k), d["angles"]), torch.int32, dev) if "angles" in d else None
        L.check(lib.imgtx_cc_motion_u8(xp, op, n, h, w, c, len(k), k.ctypes.data, _ptr(off), -45.0, 45.0, seed, flags, mp, st))
    elif name == "snow":
        k = motion_kernel(p[4], p[5])
        off = _dev(motion_offsets(len(k), d["angles"]), torch.int32, dev) if "angles" in d else None
        zi, zt = _snow_zoom_tables(h, w, float(p[2]))
        layer = _dev(d.get("layer"), torch.float64, dev)
        L.check(lib.imgtx_cc_snow_u8(xp, op, n, h, w, c, float(np.float32(p[3])), float(np.float32(p[6])),
                                     float(np.float32(1 - p[6])), zi.ctypes.data, zt.ctypes.data, len(k), k.ctypes.data,
                                     _ptr(off), _ptr(layer), float(p[0]), float(p[1]), -135.0, -45.0, seed, flags, mp, st))
    elif name == "elastic_transform":
        alpha, sigma, c2 = p[0] * h, p[1] * h, p[2] * h
        if sigma > 1e-15:
            wts = gaussian_taps(sigma, 3.0)
            lw = len(wts) - 1
        else:
            wts, lw = np.zeros(1), -1
        M = _dev(affine_from_jitter(d["jitter"], h), torch.float64, dev) if "jitter" in d else None
        ux, uy = _dev(d.get("ux"), torch.float64, dev), _dev(d.get("uy"), torch.float64, dev)
        L.check(lib.imgtx_cc_elastic_u8(xp, op, n, h, w, c, float(alpha), wts.ctypes.data, lw, _ptr(M), _ptr(ux), _ptr(uy),
                                        float(c2), seed, flags, mp, st))
    elif name == "pixelate":
        sw, sh = int(w * p), int(h * p)
        if sw == w or sh == h:     # Pillow skips a pass whose size does not change: the general path (two resample launches)
            small = B.resample(x, (sw, sh), L.FILTER_BOX)
            B.resample(small, (w, h), L.FILTER_BOX, out=out)
            if mp is not None:
                L.check(lib.imgtx_image_checksums_u8(op, n, h * w * c, mp, st))
        else:
            rc = lib.imgtx_box_pixelate_u8(xp, op, n, h, w, c, sw, sh, flags, mp, st)   # gather forms; 1 = not applicable
            if rc == 1 and h * w > 96 * 96:     # the fused chain's shared-memory planes do not fit: two resample launches
                small = B.resample(x, (sw, sh), L.FILTER_BOX)
                B.resample(small, (w, h), L.FILTER_BOX, out=out)
                rc = lib.imgtx_image_checksums_u8(op, n, h * w * c, mp, st) if mp is not None else 0
            elif rc == 1:
                t = [B._resample_table(a, b, L.FILTER_BOX, dev) for a, b in ((w, sw), (h, sh), (sw, w), (sh, h))]
                rc = lib.imgtx_resize_chain_u8(xp, op, n, h, w, c, sw, sh, t[0].data_ptr(), t[0].numel(), t[1].data_ptr(),
                                               t[1].numel(), t[2].data_ptr(), t[2].numel(), t[3].data_ptr(), t[3].numel(), st)
                if rc == 0 and mp is not None:
                    rc = lib.imgtx_image_checksums_u8(op, n, h * w * c, mp, st)
            L.check(rc)
    elif name == "jpeg_compression":
        qt = jpeg_qtables(p)
        L.check(lib.imgtx_cc_jpeg_u8(xp, op, n, h, w, c, qt.ctypes.data, flags, mp, st))
    else:
        raise KeyError(name)
    return out


def _corrupt_reference(x, name, severity, d, seed, out):
    """The reference's own four generators (TransformationPool, ImageNet-C constants)."""
    p = REFERENCE_TABLE[name][severity - 1]
    if name == "gaussian_noise":
        if "noise" in d:
            return B.noise_injected(x, L.NOISE_GAUSS_F64, d["noise"], out=out)
        return B.noise_philox(x, L.NOISE_GAUSS_F64, p * 255.0, seed, out=out)
    if name == "impulse_noise":
        if "mask" in d:
            return B.noise_injected(x, L.NOISE_IMPULSE, d["mask"], p, out=out)
        return B.noise_philox(x, L.NOISE_IMPULSE, p, seed, out=out)
    if name == "shot_noise":
        if "counts" in d:
            return B.noise_injected(x, L.NOISE_SHOT, d["counts"], p, out=out)
        return B.noise_philox(x, L.NOISE_SHOT, p, seed, out=out)
    if name == "defocus_blur":
        return B.pil_gaussian_blur(x, p, out=out)
    raise KeyError(f"the reference has no generator for {name}")


# ------------------------------------------------------------------------------------------------
# the suite
# ------------------------------------------------------------------------------------------------
def suite_passes(corruptions=None, severities=(1, 2, 3, 4, 5)):
    return [(c, s) for c in (corruptions or CORRUPTIONS) for s in severities]


def generate_suite(x: torch.Tensor, corruptions=None, severities=(1, 2, 3, 4, 5), seed: int = 0, out=None, sink=None,
                   mode: Optional[str] = None, metrics: Optional[torch.Tensor] = None, graph: bool = False, lanes: int = 1):
    """Run every (corruption, severity) pass over the device-resident batch x.  Each pass is one kernel launch.  `out`:
    optional (len(passes), n, h, w, 3) uint8 CUDA tensor receiving every pass; otherwise `sink(name, severity, tensor)` is
    called per pass with a reused buffer (e.g. to checksum or stage to the host).  `metrics`: optional (len(passes), n, 2)
    int32 tensor receiving the per-image metric of every pass.  graph=True returns a SuiteGraph: the whole step captured once
    in a CUDA graph and replayed with one host call (the seeds are kernel arguments, so a replay re-draws the same noise;
    capture again for new seeds); lanes: see SuiteGraph."""
    if graph:
        return SuiteGraph(x, corruptions, severities, seed, out, mode, metrics, lanes)
    passes = suite_passes(corruptions, severities)
    buf = None if out is not None else torch.empty_like(x)
    for k, (name, sev) in enumerate(passes):
        dst = out[k] if out is not None else buf
        corrupt(x, name, sev, seed=seed * 1000 + k, out=dst, mode=mode, metrics=None if metrics is None else metrics[k])
        if sink is not None:
            sink(name, sev, dst)
    return out if out is not None else buf


class SuiteGraph:
    """The 75-pass step as one CUDA graph.  The first (eager) run fills the host-side table caches and device-side constant
    tables; the capture then records only kernel launches (the library allocates nothing and never synchronises).

    lanes > 1: the passes read the same batch and write their own outputs, so nothing orders them but the output buffers.
    Pass k is captured on stream k % lanes (and writes out[k % len(out)], len(out) a multiple of lanes, so two lanes never
    share a buffer): the graph becomes `lanes` independent chains and a kernel's tail — its last, partly idle wave of CTAs
    and the launch gap behind it — is filled by the other chain's next kernel."""

    def __init__(self, x, corruptions=None, severities=(1, 2, 3, 4, 5), seed=0, out=None, mode=None, metrics=None, lanes=1):
        self.passes = suite_passes(corruptions, severities)
        self.x = x
        self.lanes = max(1, int(lanes))
        self.out = out if out is not None else [torch.empty_like(x) for _ in range(2 * ((self.lanes + 1) // 2))]
        if self.lanes > 1 and not isinstance(self.out, torch.Tensor) and len(self.out) % self.lanes:
            raise ValueError("SuiteGraph: the number of rotating output buffers must be a multiple of lanes")
        self.metrics = metrics
        self._args = (seed, mode)
        self._run()                       # warm: caches, cudaFuncSetAttribute
        torch.cuda.synchronize()
        self._side = [torch.cuda.Stream() for _ in range(self.lanes - 1)]
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._run(self._side)

    def _run(self, side=()):
        seed, mode = self._args
        main = torch.cuda.current_stream()
        streams = [main] + list(side)
        if side:
            fork = torch.cuda.Event()
            fork.record(main)
            for st in side:
                st.wait_event(fork)
        for k, (name, sev) in enumerate(self.passes):
            dst = self.out[k] if isinstance(self.out, torch.Tensor) else self.out[k % len(self.out)]
            with torch.cuda.stream(streams[k % len(streams)]):
                corrupt(self.x, name, sev, seed=seed * 1000 + k, out=dst, mode=mode,
                        metrics=None if self.metrics is None else self.metrics[k])
        for st in side:
            join = torch.cuda.Event()
            join.record(st)
            main.wait_event(join)

    def replay(self):
        self.graph.replay()

    @property
    def launches(self):
        return len(self.passes)


def write_cifar_c(root: str, x_u8: np.ndarray, labels: np.ndarray, corruptions=None, seed: int = 0, batch: int = 10000,
                  mode: Optional[str] = None):
    """Write Hendrycks-layout files: <root>/<corruption>.npy of shape (5*N, h, w, 3), severity-major, plus labels.npy
    (the layout CIFARC_Dataset slices: fall_2025/TTA_transforms.py:16-26)."""
    os.makedirs(root, exist_ok=True)
    n = len(x_u8)
    np.save(os.path.join(root, "labels.npy"), np.tile(np.asarray(labels), 5))
    pinned = torch.from_numpy(np.ascontiguousarray(x_u8)).pin_memory()
    for ci, name in enumerate(corruptions or CORRUPTIONS):
        dst = np.lib.format.open_memmap(os.path.join(root, f"{name}.npy"), mode="w+", dtype=np.uint8,
                                        shape=(5 * n,) + tuple(x_u8.shape[1:]))
        for sev in range(1, 6):
            for i in range(0, n, batch):
                xb = pinned[i:i + batch].cuda(non_blocking=True)
                yb = corrupt(xb, name, sev, seed=(seed * 1000 + ci * 5 + sev) * 65536 + i // batch, mode=mode)
                dst[(sev - 1) * n + i:(sev - 1) * n + i + len(xb)] = yb.cpu().numpy()
        dst.flush()
