"""Freeze golden vectors from the LIVE reference (TEST INFRASTRUCTURE; run in the build container only).

    python -m oracle.make_golden          # writes tests/golden/*.npz

Every case runs one of the reference's own hot-path functions (loaded from /root/reference by
oracle/ref_loader.py, with the installed Pillow / OpenCV / torchvision / torch / numpy doing the arithmetic
exactly as they do for the reference) on seeded inputs, captures the RNG draws the function consumed
 and stores inputs, draws and outputs.  tests/test_oracle_golden.py pins oracle/np_ops.py to
these files on CPU; the `-m gpu` tests pin the CUDA library to the same files on the B200.

A golden file is an .npz with one JSON manifest (key "manifest") listing cases
    {"fn": name, "args": {...}, "in": array-key, "draws": {name: array-key}, "out": array-key}
plus the arrays.  Library versions used are recorded in the manifest header.
"""
from __future__ import annotations

import glob
import io
import json
import os
import random
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_loader as R  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


class Book:
    def __init__(self, name):
        self.name = name
        self.arrays = {}
        self.cases = []

    def put(self, key, arr):
        if key not in self.arrays:
            self.arrays[key] = np.ascontiguousarray(arr)
        return key

    def case(self, fn, inp, out, args=None, draws=None, note=None):
        i = len(self.cases)
        c = {"fn": fn, "args": args or {}, "in": inp, "out": self.put(f"o{i}", out), "draws": {}}
        for k, v in (draws or {}).items():
            c["draws"][k] = v if isinstance(v, str) else self.put(f"d{i}_{k}", v)
        if note:
            c["note"] = note
        self.cases.append(c)

    def save(self):
        import PIL, cv2, torch, torchvision, scipy
        header = {"versions": {"pillow": PIL.__version__, "opencv": cv2.__version__, "torch": torch.__version__,
                               "torchvision": torchvision.__version__, "numpy": np.__version__,
                               "scipy": scipy.__version__},
                  "cases": self.cases}
        os.makedirs(GOLDEN_DIR, exist_ok=True)
        path = os.path.join(GOLDEN_DIR, self.name + ".npz")
        np.savez_compressed(path, manifest=np.frombuffer(json.dumps(header).encode(), dtype=np.uint8), **self.arrays)
        print(f"{path}: {len(self.cases)} cases, {os.path.getsize(path) / 1e6:.2f} MB")


# ------------------------------------------------------------------------------------------------
# inputs
# ------------------------------------------------------------------------------------------------
def make_inputs():
    from PIL import Image
    rng = np.random.default_rng(1234)
    pngs = sorted(glob.glob(os.path.join(R.REF_ROOT, "CIFAR-10-C", "store", "*.png")))
    store = np.stack([np.array(Image.open(p).convert("RGB")) for p in pngs])  # (95,32,32,3)
    x32 = np.concatenate([rng.integers(0, 256, (4, 32, 32, 3), dtype=np.uint8), store[[0, 22, 47, 61]]])
    tiles = store[:49].reshape(7, 7, 32, 32, 3).transpose(0, 2, 1, 3, 4).reshape(1, 224, 224, 3)
    xodd = rng.integers(0, 256, (2, 37, 53, 3), dtype=np.uint8)
    # degenerate images for the histogram paths: constant, two-level, all-but-one
    flat = np.full((1, 32, 32, 3), 77, np.uint8)
    two = np.where(rng.random((1, 32, 32, 3)) < 0.5, 10, 200).astype(np.uint8)
    one_off = np.full((1, 32, 32, 3), 5, np.uint8)
    one_off[0, 3, 4] = (250, 9, 5)
    xedge = np.concatenate([flat, two, one_off])
    return {"x32": x32, "x224": np.ascontiguousarray(tiles), "xodd": xodd, "xedge": xedge, "store": store}


def pil_list(x):
    from PIL import Image
    return [Image.fromarray(a) for a in x]


def unpil(lst):
    return np.stack([np.array(p) for p in lst])


def grid(lo, hi, step):
    n = int((hi - lo) / step) + 1
    return [lo + j * step for j in range(n)]


# ------------------------------------------------------------------------------------------------
# transformation.py + transformations_code
# ------------------------------------------------------------------------------------------------
def gen_transformation(X):
    b = Book("transformation")
    for k in ("x32", "x224", "xodd"):
        b.put(k, X[k])
    t = R.transformation()

    def run(fn, key, **args):
        f = getattr(t, fn)
        out = unpil([f(p, *args.values()) for p in pil_list(X[key])])
        b.case(fn, key, out, args=args)

    for s in grid(0.9, 1.4, 0.1):
        run("apply_scale", "x32", scale_factor=s)
    for s in (0.9, 1.2000000000000002):
        run("apply_scale", "x224", scale_factor=s)
    for s in (0.9, 1.3):
        run("apply_scale", "xodd", scale_factor=s)
    for a in grid(-22.5, 22.5, 2.5) + [90.0, 180.0, -90.0, 33.3]:
        run("apply_rotation", "x32", angle=a)
    run("apply_rotation", "x224", angle=12.5)
    for a in (-7.5, 0.0, 90.0):
        run("apply_rotation", "xodd", angle=a)
    for c in grid(0, 1, 0.1):
        run("apply_contrast", "x32", contrast_amount=c)
    run("apply_contrast", "x224", contrast_amount=0.7000000000000001)
    for sf in (0.0, 0.1, 0.5, 1.0):
        run("apply_shear", "x32", shear_factor=sf)
    run("apply_shear", "x224", shear_factor=0.30000000000000004)
    run("apply_shear", "xodd", shear_factor=0.2)
    for r in grid(0, 5, 0.5):
        run("apply_blur", "x32", blur_radius=r)
    run("apply_blur", "x224", blur_radius=2.0)
    run("apply_blur", "xodd", blur_radius=1.5)
    for f in grid(-0.05, 0.05, 0.01):
        run("apply_brightness", "x32", brightness_factor=f)
    run("apply_brightness", "x224", brightness_factor=0.03)
    for tx, ty in ((5, -10), (0, 0), (-50, 20), (35, 0), (31, -31), (-5.9, 7.9)):
        run("apply_translation", "x32", tx=tx, ty=ty)
    run("apply_translation", "xodd", tx=3, ty=-7)
    run("apply_camera_distance", "x32", distance_factor=2.5)
    run("apply_xy_translation_3d", "x32", tx=0.1, ty=-0.2)
    run("apply_rotation_3d", "x32", angle=-15.0)

    # noise: np.random.normal(0, std*255, shape).astype(f32) is the only draw (transformation.py:274)
    for seed, std in ((11, 0.0), (12, 0.05), (13, 0.1)):
        outs, draws = [], []
        for i, p in enumerate(pil_list(X["x32"][:3])):
            np.random.seed(seed * 100 + i)
            outs.append(np.array(t.apply_gaussian_noise(p, std)))
            np.random.seed(seed * 100 + i)
            draws.append(np.random.normal(0, std * 255, (32, 32, 3)).astype(np.float32))
        b.put("x32n", X["x32"][:3])
        b.case("apply_gaussian_noise", "x32n", np.stack(outs), args={"noise_std": std}, draws={"noise": np.stack(draws)})

    tc = R.transformations_code()
    b.case("vert_flip", "x32", unpil([tc.vert_flip(p) for p in pil_list(X["x32"])]))
    b.case("vert_flip", "xodd", unpil([tc.vert_flip(p) for p in pil_list(X["xodd"])]))
    outs, xy = [], []
    for i, p in enumerate(pil_list(X["x32"])):
        np.random.seed(500 + i)
        outs.append(np.array(tc.rand_crop(p)))
        np.random.seed(500 + i)
        xy.append((np.random.randint(0, 32 - 24 + 1), np.random.randint(0, 32 - 24 + 1)))
    b.case("rand_crop", "x32", np.stack(outs), draws={"xy": np.array(xy, np.int32)})
    for s in (1.0, 1.05, 1.1):
        b.case("apply_random_zoom", "x32", unpil([tc.apply_random_zoom(p, s) for p in pil_list(X["x32"])]),
               args={"scale_factor": s})

    # apply_perspective_warp: RandomPerspective(p=1) on a tensor; capture get_params' corner points
    import torch
    import torchvision.transforms as T
    rec = []
    orig = T.RandomPerspective.get_params

    def spy(width, height, distortion_scale):
        r = orig(width, height, distortion_scale)
        rec.append(r)
        return r

    T.RandomPerspective.get_params = staticmethod(spy)
    try:
        for ds in (0.2, 0.1):
            outs, pts = [], []
            for i, p in enumerate(pil_list(X["x32"])):
                torch.manual_seed(900 + i)
                rec.clear()
                outs.append(np.array(tc.apply_perspective_warp(p, ds)))
                pts.append(np.array(rec[0], np.int32).reshape(2, 4, 2))  # [start|end][corner][x,y]
            b.case("apply_perspective_warp", "x32", np.stack(outs), args={"distortion_scale": ds},
                   draws={"points": np.stack(pts)})
    finally:
        T.RandomPerspective.get_params = staticmethod(orig)
    b.save()


# ------------------------------------------------------------------------------------------------
# AugMix
# ------------------------------------------------------------------------------------------------
def gen_augmix(X):
    import torch
    b = Book("augmix")
    A = R.augmix()
    x32 = X["x32"]
    t01 = torch.from_numpy(x32).permute(0, 3, 1, 2).float().div(255)           # [0,1]
    tn = torch.stack([A.normalize_tensor(t) for t in t01])                     # normalised (wrap reachable)
    t224 = torch.from_numpy(X["x224"]).permute(0, 3, 1, 2).float().div(255)
    b.put("t01", t01.numpy())
    b.put("tn", tn.numpy())
    b.put("t224", t224.numpy())
    b.case("normalize_tensor", "t01", tn.numpy())
    b.case("denormalize_tensor", "tn", torch.stack([A.denormalize_tensor(t) for t in tn]).numpy())

    def run(key, T, idx, seed, severity, width, depth):
        np.random.seed(seed)
        random.seed(seed)
        out = A.augmix(T[idx], severity=severity, width=width, depth=depth).numpy()
        # replay the draws
        np.random.seed(seed)
        random.seed(seed)
        ws = np.random.dirichlet([1.0] * width)
        m = np.random.beta(1.0, 1.0)
        ops = np.full((width, 3), -1, np.int32)
        signs = np.zeros((width, 3), np.int32)
        for i in range(width):
            d = depth if depth > 0 else np.random.randint(1, 4)
            for j in range(d):
                op = random.choice(A.AUG_OPS)
                ops[i, j] = A.AUG_OPS.index(op)
                if op is A.rotate:
                    signs[i, j] = random.choice([-1, 1])
        b.case("augmix", key, out[None], args={"index": idx, "severity": severity, "width": width, "depth": depth},
               draws={"ws": ws, "m": np.array([m]), "ops": ops, "signs": signs})

    seed = 100
    for idx in range(8):
        for depth in (3, -1):
            run("t01", t01, idx, seed, 3, 3, depth)
            seed += 1
    for idx in range(4):
        run("tn", tn, idx, seed, 3, 3, 3)
        seed += 1
    run("t01", t01, 0, seed, 1, 2, 2)
    run("t01", t01, 5, seed + 1, 5, 3, 3)
    for s in range(3):
        run("t224", t224, 0, seed + 2 + s, 3, 3, 3)

    # the eight ops on their own (u8 -> u8), both rotate signs
    pl = pil_list(x32)
    b.put("x32", x32)
    for name in ("posterize", "shear_x", "shear_y", "translate_x", "translate_y", "equalize", "solarize"):
        for sev in (1, 3, 5):
            b.case("op_" + name, "x32", unpil([getattr(A, name)(p, sev) for p in pl]), args={"severity": sev})
    for sign in (-1, 1):
        class _R:
            @staticmethod
            def choice(_):
                return sign
        real = A.random
        A.random = _R
        try:
            b.case("op_rotate", "x32", unpil([A.rotate(p, 3) for p in pl]), args={"severity": 3, "sign": sign})
        finally:
            A.random = real
    b.put("xedge", X["xedge"])
    b.case("op_equalize", "xedge", unpil([A.equalize(p, 3) for p in pil_list(X["xedge"])]), args={"severity": 3})
    b.save()


# ------------------------------------------------------------------------------------------------
# TTA (tta_initial 16 views, TTA_transforms strategies, datasets)
# ------------------------------------------------------------------------------------------------
def gen_tta(X):
    import torch
    import torchvision.transforms as T
    b = Book("tta")
    ti = R.tta_initial()
    x32 = X["x32"]
    b.put("x32", x32)
    # pil_to_tensor / tensor_to_pil (CIFAR-100 stats)
    tn = torch.stack([ti.pil_to_tensor(p) for p in pil_list(x32)])
    b.put("tn100", tn.numpy())
    b.case("tta_initial.pil_to_tensor", "x32", tn.numpy())
    back = unpil([ti.tensor_to_pil(t) for t in tn])
    b.case("tta_initial.tensor_to_pil", "tn100", back)
    # out-of-range tensor values exercise the clamp
    wild = (tn * 1.7 + 0.3).contiguous()
    b.put("tn100_wild", wild.numpy())
    b.case("tta_initial.tensor_to_pil", "tn100_wild", unpil([ti.tensor_to_pil(t) for t in wild]))

    # the 16-view expansion (tta_initial_claudeFixed.py:218-230): for each aug type, AUG_COUNTS views
    for img in range(4):
        seed = 7000 + img
        np.random.seed(seed)
        pil = ti.tensor_to_pil(tn[img])
        views = []
        for fn, cnt in zip(ti.AUG_TYPES, ti.AUG_COUNTS):
            for _ in range(cnt):
                views.append(ti.pil_to_tensor(fn(pil)).numpy())
        np.random.seed(seed)
        crops = np.array([(np.random.randint(0, 9), np.random.randint(0, 9)) for _ in range(5)], np.int32)
        bright = np.array([np.random.uniform(0.5, 1.5) for _ in range(5)])
        contr = np.array([np.random.uniform(0.8, 1.2) for _ in range(5)])
        b.case("tta_initial.views16", "tn100", np.stack(views)[None], args={"index": img},
               draws={"crop_xy": crops, "brightness": bright, "contrast": contr})

    # TTA_transforms.py strategies: record torchvision's random parameters through get_params spies
    tt = R.tta_transforms()
    b.case("TTA_transforms.base_transform", "x32", np.stack([tt.base_transform(p).numpy() for p in pil_list(x32)]))
    rec = {"crop": [], "jitter": []}
    o_crop, o_jit = T.RandomCrop.get_params, T.ColorJitter.get_params

    def spy_crop(img, output_size):
        r = o_crop(img, output_size)
        rec["crop"].append(r[:2])
        return r

    def spy_jit(brightness, contrast, saturation, hue):
        r = o_jit(brightness, contrast, saturation, hue)
        rec["jitter"].append([v if v is not None else float("nan") for v in r[1:]])
        return r

    T.RandomCrop.get_params = staticmethod(spy_crop)
    T.ColorJitter.get_params = staticmethod(spy_jit)
    try:
        for sname in ("gentle", "multiscale", "color"):
            tl = getattr(tt, "tta_transforms_" + sname)
            outs, crops, jits = [], [], []
            for i, p in enumerate(pil_list(x32)):
                torch.manual_seed(8000 + i)
                rec["crop"].clear()
                rec["jitter"].clear()
                outs.append(np.stack([tr(p).numpy() for tr in tl]))
                crops.append(np.array(rec["crop"], np.int32).reshape(-1, 2))
                jits.append(np.array(rec["jitter"], np.float64).reshape(-1, 4))
            b.case("TTA_transforms." + sname, "x32", np.stack(outs),
                   draws={"crop_ij": np.stack(crops), "jitter": np.stack(jits)})
    finally:
        T.RandomCrop.get_params = staticmethod(o_crop)
        T.ColorJitter.get_params = staticmethod(o_jit)
    assert tt.get_tta_transforms("fog") is tt.tta_transforms_color
    assert tt.get_tta_transforms("pixelate") is tt.tta_transforms_multiscale
    assert tt.get_tta_transforms("gaussian_noise") is tt.tta_transforms_gentle

    # CIFARC_Dataset: severity-major slicing of a (50000,32,32,3) file (TTA_transforms.py:10-36)
    with tempfile.TemporaryDirectory() as d:
        rng = np.random.default_rng(5)
        marks = rng.integers(0, 256, (50000, 4), dtype=np.uint8)
        data = np.zeros((50000, 32, 32, 3), np.uint8)
        data[:, 0, 0, :] = marks[:, :3]
        data[:, 31, 31, 0] = marks[:, 3]
        labels = rng.integers(0, 10, 50000).astype(np.int64)
        np.save(os.path.join(d, "fog.npy"), data)
        np.save(os.path.join(d, "labels.npy"), labels)
        for sev in (1, 3, 5):
            ds = tt.CIFARC_Dataset(d, "fog", sev)
            assert len(ds) == 10000
            idx = [0, 1, 4999, 9999]
            imgs = np.stack([np.array(ds[i][0]) for i in idx])
            labs = np.array([int(ds[i][1]) for i in idx])
            b.case("CIFARC_Dataset", None, imgs, args={"severity": sev, "idx": idx, "seed": 5},
                   draws={"labels": labs})
    b.save()


# ------------------------------------------------------------------------------------------------
# TransformationPool / Individual
# ------------------------------------------------------------------------------------------------
def gen_pool(X):
    import torch
    b = Book("pool")
    P = R.pool()
    TP = P.TransformationPool
    x32 = X["x32"]
    b.put("x32", x32)
    xn = x32[[0, 5]]
    b.put("x32n", xn)
    b.put("xedge", X["xedge"])
    b.put("x224", X["x224"])
    b.put("xodd", X["xodd"])
    for sev in range(1, 6):
        # gaussian: normal(0, std*255, shape) f64
        std = [0.08, 0.12, 0.18, 0.26, 0.38][sev - 1]
        outs, dr = [], []
        for i, p in enumerate(pil_list(xn)):
            np.random.seed(3000 + sev * 10 + i)
            outs.append(np.array(TP.gaussian_noise(p, severity=sev)))
            np.random.seed(3000 + sev * 10 + i)
            dr.append(np.random.normal(0, std * 255, (32, 32, 3)))
        b.case("pool.gaussian_noise", "x32n", np.stack(outs), args={"severity": sev}, draws={"noise": np.stack(dr)})
        outs, dr = [], []
        for i, p in enumerate(pil_list(xn)):
            np.random.seed(3100 + sev * 10 + i)
            outs.append(np.array(TP.impulse_noise(p, severity=sev)))
            np.random.seed(3100 + sev * 10 + i)
            dr.append(np.random.random((32, 32)))
        b.case("pool.impulse_noise", "x32n", np.stack(outs), args={"severity": sev}, draws={"mask": np.stack(dr)})
        lam = [60, 25, 12, 5, 3][sev - 1]
        outs, dr = [], []
        for i, p in enumerate(pil_list(xn)):
            np.random.seed(3200 + sev * 10 + i)
            outs.append(np.array(TP.shot_noise(p, severity=sev)))
            np.random.seed(3200 + sev * 10 + i)
            dr.append(np.random.poisson(np.array(p).astype(np.float32) / 255.0 * lam))
        b.case("pool.shot_noise", "x32n", np.stack(outs), args={"severity": sev},
               draws={"counts": np.stack(dr).astype(np.int64)})
        b.case("pool.defocus_blur", "x32", unpil([TP.defocus_blur(p, severity=sev) for p in pil_list(x32)]),
               args={"severity": sev})
    b.case("pool.defocus_blur", "x224", unpil([TP.defocus_blur(p, severity=2) for p in pil_list(X["x224"])]),
           args={"severity": 2})
    b.case("pool.defocus_blur", "xodd", unpil([TP.defocus_blur(p, severity=1) for p in pil_list(X["xodd"])]),
           args={"severity": 1})
    for name, factors in (("enhance_contrast", (0.0, 0.5, 0.8, 1.0, 1.2, 1.7, 2.0)),
                          ("enhance_brightness", (0.0, 0.5, 0.8, 1.0, 1.2, 1.5, 2.0)),
                          ("enhance_sharpness", (0.0, 0.5, 1.0, 1.7, 3.0)),
                          ("enhance_color", (0.0, 0.5, 1.0, 1.3, 2.0))):
        for f in factors:
            b.case("pool." + name, "x32", unpil([getattr(TP, name)(p, factor=f) for p in pil_list(x32)]),
                   args={"factor": f})
        b.case("pool." + name, "xodd", unpil([getattr(TP, name)(p, factor=1.4) for p in pil_list(X["xodd"])]),
               args={"factor": 1.4})
    for size in (5, 7, 9, 11):
        b.case("pool.motion_blur", "x32", unpil([TP.motion_blur(p, size=size) for p in pil_list(x32)]),
               args={"size": size})
    b.case("pool.motion_blur", "xodd", unpil([TP.motion_blur(p, size=7) for p in pil_list(X["xodd"])]),
           args={"size": 7})
    for key in ("x32", "xedge", "xodd", "x224"):
        b.case("pool.histogram_equalization", key, unpil([TP.histogram_equalization(p) for p in pil_list(X[key])]))

    # tensor_to_pil / pil_to_tensor with CIFAR-10 stats and the Individual chain (base chain, severity 3)
    tn = torch.stack([P.pil_to_tensor(p) for p in pil_list(x32)])
    b.put("tn10", tn.numpy())
    b.case("pool.pil_to_tensor", "x32", tn.numpy())
    b.case("pool.tensor_to_pil", "tn10", unpil([P.tensor_to_pil(t) for t in tn]))
    chain = [("gaussian_noise", {"severity": 3}), ("impulse_noise", {"severity": 3}),
             ("shot_noise", {"severity": 3}), ("defocus_blur", {"severity": 3})]
    chain2 = [("enhance_contrast", {"factor": 1.3}), ("motion_blur", {"size": 5}), ("histogram_equalization", {}),
              ("enhance_sharpness", {"factor": 2.0}), ("enhance_color", {"factor": 0.7}),
              ("enhance_brightness", {"factor": 1.1})]
    for img in range(3):
        seed = 4000 + img
        np.random.seed(seed)
        out = P.Individual(chain).apply_transformations(tn[img]).numpy()
        # replay: draws happen on the intermediate images, so recompute them step by step
        np.random.seed(seed)
        pil = P.tensor_to_pil(tn[img])
        noise = np.random.normal(0, 0.18 * 255, (32, 32, 3))
        pil1 = np.clip(np.array(pil).astype(np.float32) + noise, 0, 255).astype(np.uint8)
        mask = np.random.random((32, 32))
        a = pil1.astype(np.float32)
        a[mask < 0.09 / 2] = 0
        a[mask > 1 - 0.09 / 2] = 255
        pil2 = a.astype(np.uint8)
        counts = np.random.poisson(pil2.astype(np.float32) / 255.0 * 12)
        b.case("pool.individual_base", "tn10", out[None], args={"index": img, "severity": 3},
               draws={"noise": noise, "mask": mask, "counts": counts.astype(np.int64)})
        out2 = P.Individual(chain2).apply_transformations(tn[img]).numpy()
        b.case("pool.individual_enhance", "tn10", out2[None], args={"index": img, "chain": chain2})
    b.save()


# ------------------------------------------------------------------------------------------------
# sign-step family + float corruptions
# ------------------------------------------------------------------------------------------------
def gen_fgsm(X):
    import torch
    b = Book("fgsm")
    F = R.fgsm()
    mean, std = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)
    x = torch.from_numpy(X["x32"]).permute(0, 3, 1, 2).contiguous().float().div(255)
    xn = (x - torch.tensor(mean).view(1, 3, 1, 1)) / torch.tensor(std).view(1, 3, 1, 1)
    gen = torch.Generator().manual_seed(5)
    g = torch.randn(xn.shape, generator=gen)
    g[torch.rand(xn.shape, generator=gen) < 0.01] = 0.0
    g.view(-1)[:8] = torch.tensor([0.0, -0.0, float("nan"), float("inf"), -float("inf"), 1e-45, -1e-45, 3.0])
    xs = xn.clone()
    xs.view(-1)[8:12] = torch.tensor([float("nan"), float("inf"), -float("inf"), 5.0])
    b.put("x", x.numpy())
    b.put("xn", xs.numpy())
    b.put("g", g.numpy())
    eps_list = [4 / 355, 8 / 255, 12 / 255, 0.03, 0.0, 0.1]
    min_vals = torch.tensor([(0 - m) / s for m, s in zip(mean, std)]).view(1, 3, 1, 1)
    max_vals = torch.tensor([(1 - m) / s for m, s in zip(mean, std)]).view(1, 3, 1, 1)
    b.put("min_vals", min_vals.numpy())
    b.put("max_vals", max_vals.numpy())
    for e in eps_list:
        b.case("fgsm_attack", "x", F.fgsm_attack(x, e, g).numpy(), args={"epsilon": e}, draws={"g": "g"})
        b.case("fgsm_attack", "xn", F.fgsm_attack(xs, e, g).numpy(), args={"epsilon": e}, draws={"g": "g"})
        b.case("reverse_fgsm_attack", "xn", F.reverse_fgsm_attack(xs, e, g, min_vals, max_vals).numpy(),
               args={"epsilon": e}, draws={"g": "g"})

    # FGSM(net, x, y, eps) with a stub network whose input gradient is easy to reproduce
    torch.manual_seed(3)
    net = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(3 * 32 * 32, 10))
    y = torch.arange(8) % 10
    for e in (4 / 355, 8 / 255, 12 / 255):
        x_adv, h_adv, y_adv, pert = F.FGSM(net, x, y, e)
        xg = x.clone().requires_grad_(True)
        torch.nn.CrossEntropyLoss()(net(xg), y).backward()
        b.case("FGSM", "x", x_adv.detach().numpy(), args={"eps": e},
               draws={"g": xg.grad.numpy(), "pert": pert.numpy()})

    # pgd_attack: log every (x_i, grad_i) pair through the stub model, then the final image
    class Spy(torch.nn.Module):
        def __init__(self, inner):
            super().__init__()
            self.inner = inner
            self.seen = []

        def forward(self, t):
            self.seen.append(t)
            return self.inner(t)

    spy = Spy(net)
    torch.manual_seed(77)
    eps, alpha, iters = 0.3, 2 / 255.0, 4
    final = F.pgd_attack(spy, xn[:4], y[:4], eps, alpha, iters, min_vals, max_vals)
    torch.manual_seed(77)
    init_noise = torch.empty_like(xn[:4]).uniform_(-eps, eps)
    xs_it = torch.stack([t.detach() for t in spy.seen])
    gs_it = torch.stack([t.grad.detach() for t in spy.seen])
    b.put("xn4", xn[:4].numpy())
    b.case("pgd_attack", "xn4", final.numpy(), args={"epsilon": eps, "alpha": alpha, "num_iter": iters},
           draws={"init_noise": init_noise.numpy(), "x_iter": xs_it.numpy(), "g_iter": gs_it.numpy()})

    # float corruptions (pipenline/angellic.py:34-46, fall_2025/angellic_with_corrupt.py:51-62)
    A = R.angellic()
    torch.manual_seed(21)
    out = A.add_gaussian_noise(x, mean=0.05, std=0.1).numpy()
    torch.manual_seed(21)
    r = torch.randn_like(x).numpy()
    b.case("angellic.add_gaussian_noise", "x", out, args={"mean": 0.05, "std": 0.1}, draws={"randn": r})
    for f in (0.3, -0.2):
        b.case("angellic.add_brightness", "x", A.add_brightness(x, f).numpy(), args={"factor": f})
    for f in (1.5, 0.5):
        b.case("angellic.add_contrast", "x", A.add_contrast(x, f).numpy(), args={"factor": f})
    import torchvision.transforms.functional as TF
    AC = R.slice_module("fall_2025/angellic_with_corrupt.py", {"add_gaussian_noise", "gaussian_blur_img"})
    torch.manual_seed(22)
    out = torch.stack([AC.add_gaussian_noise(t, 0.05) for t in x]).numpy()
    torch.manual_seed(22)
    r = torch.stack([torch.randn_like(t) for t in x]).numpy()
    b.case("angellic_with_corrupt.add_gaussian_noise", "x", out, args={"std": 0.05}, draws={"randn": r})
    b.case("angellic_with_corrupt.gaussian_blur_img", "x", torch.stack([AC.gaussian_blur_img(t, 3) for t in x]).numpy(),
           args={"kernel_size": 3})
    b.save()


def gen_entropy(X):
    """Per-image Shannon entropy gate: the two compute_shannon_entropy variants run on normalised and on
    [0,1] tensors, plus the plain 256-bin counts np.histogram gives for the same values."""
    import torch
    b = Book("entropy")
    E = R.entropy_gate()
    mean, std = E.CIFAR_MEAN, E.CIFAR_STD
    x = torch.from_numpy(X["x32"]).permute(0, 3, 1, 2).contiguous().float().div(255)
    xn = ((x - torch.tensor(mean).view(1, 3, 1, 1)) / torch.tensor(std).view(1, 3, 1, 1)).clone()
    rng = np.random.default_rng(41)
    extra = torch.from_numpy(rng.normal(0.0, 1.5, (6, 3, 32, 32)).astype(np.float32))   # many values clamp to 0 / 1
    extra[0] = 0.25                        # one bin per channel
    extra[1, :, :, :16] = -9.0             # half the image clamps to 0
    extra[2].view(-1)[:512] = torch.arange(512, dtype=torch.float32) / 256.0 - 0.5
    xn = torch.cat([xn, extra])
    unit = torch.cat([x, torch.from_numpy(rng.random((4, 3, 32, 32)).astype(np.float32))])
    unit[-1].view(-1)[:257] = torch.arange(257, dtype=torch.float32) / 256.0     # exact bin edges, 1.0 included
    unit[-2].view(-1)[:6] = torch.tensor([-0.5, 1.5, float("nan"), 1.0, 0.0, -0.0])  # outside [0,1]: not counted
    big = torch.from_numpy(rng.random((1, 3, 224, 224)).astype(np.float32) ** 2)
    b.put("xn", xn.numpy())
    b.put("unit", unit.numpy())
    b.put("big", big.numpy())

    def counts_tent(t):
        img = t.clone()
        for c in range(3):
            img[c] = img[c] * std[c] + mean[c]
        return np.histogram(torch.clamp(img, 0, 1).numpy().flatten(), bins=256, range=(0, 1))[0]

    def counts_unit(t):
        return np.histogram(t.numpy().flatten(), bins=256, range=(0, 1))[0]

    b.case("TENT_Pipelines.compute_shannon_entropy", "xn",
           np.array([float(E.tent_entropy(t)) for t in xn], np.float64),
           draws={"counts": np.stack([counts_tent(t) for t in xn]).astype(np.int64)})
    for key, arr in (("unit", unit), ("big", big)):
        b.case("Initial_Experiments.compute_shannon_entropy", key,
               np.array([float(E.unit_entropy(t)) for t in arr], np.float64),
               draws={"counts": np.stack([counts_unit(t) for t in arr]).astype(np.int64)})
    b.save()


def gen_posthoc(X):
    """Tensor-backend TTA chain: the reference's TestTimeAugmentation()(x) on f32 CHW tensors under
    recorded torch seeds, with the draws it consumed (replayed through torchvision's own get_params), and the mean over
    T+1 'logit' vectors accumulated as post_hoc_aaryaa.py:334-342 does."""
    import torch
    import torchvision.transforms as T
    b = Book("posthoc")
    tta = R.post_hoc().TestTimeAugmentation()
    x = torch.from_numpy(X["x32"]).permute(0, 3, 1, 2).contiguous().float().div(255)
    rng = np.random.default_rng(43)
    xodd = torch.from_numpy(rng.random((2, 3, 37, 53)).astype(np.float32))
    b.put("x", x.numpy())
    b.put("xodd", xodd.numpy())
    for key, arr in (("x", x), ("xodd", xodd)):
        n, _, h, w = arr.shape
        views = 6
        outs, draws = [], {k: [] for k in ("flip", "angle", "scale", "brightness", "persp", "start", "end")}
        for i in range(n):
            for v in range(views):
                seed = 7000 + 100 * i + v
                torch.manual_seed(seed)
                outs.append(tta(arr[i]).numpy())
                torch.manual_seed(seed)                  # replay: the same RNG consumption, draw by draw
                flip = bool(torch.rand(1) < 0.5)
                angle = T.RandomRotation.get_params([-10.0, 10.0])
                _, _, scale, _ = T.RandomAffine.get_params((0.0, 0.0), None, (1.0, 1.1), None, [w, h])
                _, br, _, _, _ = T.ColorJitter.get_params((0.9, 1.1), None, None, None)
                fired = bool(torch.rand(1) < 0.5)
                sp, ep = T.RandomPerspective.get_params(w, h, 0.2) if fired else ([[0, 0]] * 4, [[0, 0]] * 4)
                for k, val in zip(draws, (flip, angle, scale, br, fired, sp, ep)):
                    draws[k].append(val)
        b.case("post_hoc_aaryaa.TestTimeAugmentation", key, np.stack(outs), args={"views": views, "seed0": 7000},
               draws={"flip": np.array(draws["flip"], np.int32), "angle": np.array(draws["angle"], np.float64),
                      "scale": np.array(draws["scale"], np.float64),
                      "brightness": np.array(draws["brightness"], np.float64),
                      "persp": np.array(draws["persp"], np.int32), "start": np.array(draws["start"], np.int64),
                      "end": np.array(draws["end"], np.int64)})
    # mean over T+1 logit vectors (:334-342) and torch.stack(...).mean(0) (TTA_transforms.py:208-209)
    L = [torch.from_numpy(rng.normal(0, 3, (1, 100)).astype(np.float32)) for _ in range(11)]
    l_hat = 0
    for logits in L:
        l_hat = l_hat + logits
    l_hat = l_hat / (10 + 1)
    b.put("logits", torch.stack(L).numpy())
    b.case("post_hoc_aaryaa.mean_logits", "logits", l_hat.numpy())
    probs = [torch.softmax(t, dim=1) for t in L[:4]]
    b.put("probs", torch.stack(probs).numpy())
    b.case("TTA_transforms.mean_probs", "probs", torch.stack(probs).mean(dim=0).numpy())
    b.save()


def gen_adaptive(X):
    """H26: adaptive_transforms of fall_2025/Initial_Experiments.py:88-92 (the live T.Compose) on PIL images under recorded
    torch seeds, with the draws it consumed replayed through torchvision's own get_params (flip rand, RandomAffine.get_params:
    angle, tx, ty, scale; perspective rand, then RandomPerspective.get_params), and the full call site :147-167 (normalised
    tensor -> denormalise, clamp -> ToPILImage -> chain -> ToTensor + Normalize)."""
    import torch
    import torchvision.transforms as T
    from PIL import Image
    b = Book("adaptive")
    A = R.adaptive()
    b.put("x32", X["x32"])
    b.put("xodd", X["xodd"])
    for key in ("x32", "xodd"):
        arr = X[key]
        n, h, w, _ = arr.shape
        outs, D = [], {k: [] for k in ("flip", "angle", "tx", "ty", "scale", "persp", "start", "end")}
        reps = 6
        for i in range(n):
            for v in range(reps):
                seed = 9000 + 100 * i + v
                torch.manual_seed(seed)
                outs.append(np.array(A.adaptive_transforms(Image.fromarray(arr[i]))))
                torch.manual_seed(seed)
                flip = bool(torch.rand(1) < 0.5)
                angle, (tx, ty), scale, shear = T.RandomAffine.get_params([-15.0, 15.0], (0.1, 0.1), (0.9, 1.1), None, [w, h])
                fired = bool(torch.rand(1) < 0.5)
                sp, ep = T.RandomPerspective.get_params(w, h, 0.2) if fired else ([[0, 0]] * 4, [[0, 0]] * 4)
                for k, val in zip(D, (flip, angle, tx, ty, scale, fired, sp, ep)):
                    D[k].append(val)
        b.case("Initial_Experiments.adaptive_transforms", key, np.stack(outs), args={"reps": reps, "seed0": 9000},
               draws={"flip": np.array(D["flip"], np.int32), "angle": np.array(D["angle"], np.float64),
                      "tx": np.array(D["tx"], np.int64), "ty": np.array(D["ty"], np.int64),
                      "scale": np.array(D["scale"], np.float64), "persp": np.array(D["persp"], np.int32),
                      "start": np.array(D["start"], np.int64), "end": np.array(D["end"], np.int64)})
    # the call site: img (normalised) -> denorm, clamp -> ToPILImage -> adaptive_transforms -> transform
    x = torch.from_numpy(X["x32"]).permute(0, 3, 1, 2).contiguous().float().div(255)
    mean, std = torch.tensor(A.CIFAR_MEAN).view(1, 3, 1, 1), torch.tensor(A.CIFAR_STD).view(1, 3, 1, 1)
    xn = (x - mean) / std
    xn = xn + 0.3 * torch.randn(xn.shape, generator=torch.Generator().manual_seed(11))   # some values leave [0,1] after denorm
    b.put("xn", xn.numpy())
    outs = []
    for i in range(len(xn)):
        img = xn[i]
        img_denorm = img.clone()
        for c in range(3):
            img_denorm[c] = img_denorm[c] * A.CIFAR_STD[c] + A.CIFAR_MEAN[c]
        img_denorm = torch.clamp(img_denorm, 0, 1)
        torch.manual_seed(9500 + i)
        outs.append(A.transform(A.adaptive_transforms(T.ToPILImage()(img_denorm))).numpy())
    b.case("Initial_Experiments.adapt_call_site", "xn", np.stack(outs), args={"seed0": 9500})
    b.save()


def gen_signsteps(X):
    """H25: the in-place patch step of fall_2025/angellic_patches_fgsm.py:67-69 and the ODIN perturbation of
    fall_2025/generalizedOdin2.py:47,76, evaluated with the literal torch expressions of those lines on tensors that carry
    +-0, NaN, +-inf and denormals in the gradient."""
    import torch
    b = Book("signsteps")
    gen = torch.Generator().manual_seed(25)
    patch = torch.rand((3, 8, 8), generator=gen)
    imgs = torch.randn((6, 3, 32, 32), generator=gen)
    for t in (patch, imgs):
        t.view(-1)[:3] = torch.tensor([0.0, -0.0, 1.0])
    gp = torch.randn(patch.shape, generator=gen)
    gi = torch.randn(imgs.shape, generator=gen)
    for g in (gp, gi):
        g[torch.rand(g.shape, generator=gen) < 0.02] = 0.0
        g.view(-1)[:8] = torch.tensor([0.0, -0.0, float("nan"), float("inf"), -float("inf"), 1e-45, -1e-45, 3.0])
    b.put("patch", patch.numpy()); b.put("imgs", imgs.numpy())
    for eps in (0.05, 0.01, 0.0):
        p = patch.clone()
        with torch.no_grad():                      # angellic_patches_fgsm.py:67-69
            p -= eps * gp.sign()
            p.clamp_(0, 1)
        b.case("angellic_patches_fgsm.patch_step", "patch", p.numpy(), args={"eps": eps}, draws={"g": gp.numpy()})
    for eps in (0.0014, 0.01, 0.0):
        perturbed = imgs - eps * torch.sign(-gi)   # generalizedOdin2.py:47 / :76
        b.case("generalizedOdin2.perturb", "imgs", perturbed.numpy(), args={"eps": eps}, draws={"g": gi.numpy()})
    b.save()


def gen_fixtures(X):
    """The reference's own CIFAR-10-C samples (CIFAR-10-C/store/*.png, lossless, all severity 1 — SURVEY §2.1) as one
    array per corruption type, ordered by source index (0, 1001, 2002, 3003, 4004)."""
    from PIL import Image
    import re
    out = {}
    for p in sorted(glob.glob(os.path.join(R.REF_ROOT, "CIFAR-10-C", "store", "*.png"))):
        m = re.match(r"(.+)_severity(\d)_idx(\d+)\.png", os.path.basename(p))
        out.setdefault(m.group(1), []).append((int(m.group(3)), np.array(Image.open(p).convert("RGB"))))
    arrays = {k: np.stack([a for _, a in sorted(v, key=lambda t: t[0])]) for k, v in out.items()}
    path = os.path.join(GOLDEN_DIR, "cifar_c_fixtures.npz")
    np.savez_compressed(path, **arrays)
    print(f"{path}: {len(arrays)} types, {os.path.getsize(path) / 1e6:.2f} MB")


def main():
    if not R.available():
        raise SystemExit("the reference is not mounted at " + R.REF_ROOT)
    X = make_inputs()
    which = sys.argv[1:] or ["transformation", "augmix", "tta", "pool", "fgsm", "entropy", "posthoc", "adaptive", "signsteps",
                            "fixtures"]
    for w in which:
        globals()["gen_" + w](X)


if __name__ == "__main__":
    main()
