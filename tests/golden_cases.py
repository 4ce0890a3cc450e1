"""Golden-vector plumbing shared by the CPU (oracle) and GPU (CUDA library) parity tests.

tests/golden/*.npz were produced by oracle/make_golden.py from the LIVE reference functions; each case names the
reference function, its arguments, the input array and the injected RNG draws.  `oracle_eval` replays a case
through oracle/np_ops.py + np_chains.py; tests/test_gpu_golden.py replays the same cases through the CUDA
library's drop-in modules.
"""
from __future__ import annotations

import json
import os

import numpy as np

from oracle import np_chains as CH
from oracle import np_ops as O

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FILES = ["transformation", "augmix", "tta", "pool", "fgsm"]

# tolerance classes: "exact" (bit-identical), "f32" (|a-b| <= 1e-5*max(1,|b|), NaN positions equal), "lsb1" (uint8, <=1)
TOL = {
    "apply_perspective_warp": "lsb1",                    # f32 grid_sample restatement
    "angellic_with_corrupt.gaussian_blur_img": "f32",    # conv2d summation order is torch's
}


class Golden:
    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        hdr = json.loads(bytes(z["manifest"]).decode())
        self.name = name
        self.versions = hdr["versions"]
        self.cases = hdr["cases"]
        self.z = z

    def arr(self, key):
        return self.z[key]

    def draws(self, case):
        return {k: self.z[v] for k, v in case["draws"].items()}


_cache = {}


def golden(name) -> Golden:
    if name not in _cache:
        _cache[name] = Golden(name)
    return _cache[name]


def all_cases():
    out = []
    for f in FILES:
        g = golden(f)
        for i, c in enumerate(g.cases):
            out.append((f, i))
    return out


def case_id(fi):
    f, i = fi
    c = golden(f).cases[i]
    a = ",".join(f"{k}={v}" for k, v in c["args"].items() if k != "chain")
    return f"{f}:{i}:{c['fn']}({a})[{c['in']}]"


def compare(fn, got, want):
    tol = TOL.get(fn, "exact")
    got = np.asarray(got)
    assert got.shape == want.shape, (got.shape, want.shape)
    assert got.dtype == want.dtype, (got.dtype, want.dtype)
    if want.dtype == np.uint8:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        lim = 1 if tol == "lsb1" else 0
        assert d.max(initial=0) <= lim, f"{fn}: max |diff| {d.max()} on {(d > lim).sum()} bytes (limit {lim})"
        return
    if tol == "exact":
        same = (got == want) | (np.isnan(got) & np.isnan(want))
        assert same.all(), f"{fn}: {(~same).sum()} of {same.size} floats differ"
        return
    nan_ok = np.isnan(got) == np.isnan(want)
    assert nan_ok.all()
    m = ~np.isnan(want)
    err = np.abs(got[m] - want[m]) / np.maximum(1.0, np.abs(want[m]))
    assert err.max(initial=0) <= 1e-5, f"{fn}: max rel err {err.max()}"


# ------------------------------------------------------------------------------------------------
# oracle replay
# ------------------------------------------------------------------------------------------------
def oracle_eval(g: Golden, c: dict):
    fn, a = c["fn"], c["args"]
    x = g.arr(c["in"]) if c["in"] else None
    d = g.draws(c)
    # ---- transformation.py / transformations_code ------------------------------------------------
    if fn in ("apply_scale", "apply_random_zoom"):
        return O.apply_scale(x, a["scale_factor"])
    if fn == "apply_camera_distance":
        return O.apply_scale(x, 2.75 / a["distance_factor"])
    if fn in ("apply_rotation", "apply_rotation_3d"):
        return O.pil_rotate_nearest(x, -a["angle"], fill=0)
    if fn == "apply_contrast":
        return O.convert_scale_abs(x, a["contrast_amount"])
    if fn == "apply_shear":
        return O.apply_shear(x, a["shear_factor"])
    if fn == "apply_blur":
        return O.apply_blur(x, a["blur_radius"])
    if fn == "apply_brightness":
        return O.enhance_brightness(x, 1.0 + a["brightness_factor"])
    if fn == "apply_translation":
        return O.translate(x, a["tx"], a["ty"])
    if fn == "apply_xy_translation_3d":
        return O.translate(x, int(a["tx"] * x.shape[2]), int(a["ty"] * x.shape[1]))
    if fn == "apply_gaussian_noise":
        return O.gaussian_noise_f32(x, d["noise"])
    if fn == "vert_flip":
        return O.hflip(x)
    if fn == "rand_crop":
        return O.crop_resize(x, d["xy"][:, 0], d["xy"][:, 1], 24, 32, O.BICUBIC)
    if fn == "apply_perspective_warp":
        return np.concatenate([CH.perspective_warp_u8(x[i:i + 1], d["points"][i, 0], d["points"][i, 1])
                               for i in range(x.shape[0])])
    # ---- AugMix ----------------------------------------------------------------------------------
    if fn == "normalize_tensor":
        return O.normalize(x, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    if fn == "denormalize_tensor":
        return O.denormalize(x, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    if fn == "augmix":
        return CH.augmix(x[a["index"]], a["severity"], d["ws"], float(d["m"][0]), d["ops"], d["signs"])[None]
    if fn.startswith("op_"):
        op = CH.AUG_OP_NAMES.index(fn[3:])
        return CH.augmix_op(x, op, a["severity"], a.get("sign", 1))
    # ---- TTA ---------------------------------------------------------------------------------------
    if fn == "tta_initial.pil_to_tensor":
        return O.to_tensor_normalize(x, CH.CIFAR100_MEAN, CH.CIFAR100_STD)
    if fn == "tta_initial.tensor_to_pil":
        return O.tensor_to_pil(x, CH.CIFAR100_MEAN, CH.CIFAR100_STD)
    if fn == "tta_initial.views16":
        return CH.tta_views16(x[a["index"]], d["crop_xy"], d["brightness"], d["contrast"])[None]
    if fn == "TTA_transforms.base_transform":
        return O.to_tensor_normalize(x, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    if fn.startswith("TTA_transforms."):
        s = fn.split(".")[1]
        return np.stack([CH.tta_strategy(x[i:i + 1], s, d["crop_ij"][i], d["jitter"][i]) for i in range(x.shape[0])])
    if fn == "CIFARC_Dataset":
        return cifarc_synthetic(a["seed"])[0][(a["severity"] - 1) * 10000:a["severity"] * 10000][a["idx"]]
    # ---- TransformationPool -------------------------------------------------------------------------
    if fn == "pool.pil_to_tensor":
        return O.to_tensor_normalize(x, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    if fn == "pool.tensor_to_pil":
        return O.tensor_to_pil(x, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    if fn == "pool.individual_base":
        s = a["severity"]
        chain = [("gaussian_noise", {"severity": s}), ("impulse_noise", {"severity": s}),
                 ("shot_noise", {"severity": s}), ("defocus_blur", {"severity": s})]
        return CH.individual_apply(x[a["index"]], chain, d)[None]
    if fn == "pool.individual_enhance":
        return CH.individual_apply(x[a["index"]], a["chain"], d)[None]
    if fn.startswith("pool."):
        return CH.pool_op(x, fn[5:], a, d)
    # ---- sign-step family -----------------------------------------------------------------------------
    if fn == "fgsm_attack":
        return O.fgsm_step(x, d["g"], a["epsilon"], +1.0, 0.0, 1.0)
    if fn == "reverse_fgsm_attack":
        return O.fgsm_step(x, d["g"], a["epsilon"], -1.0, g.arr("min_vals").reshape(-1), g.arr("max_vals").reshape(-1))
    if fn == "FGSM":
        return O.fgsm_step(x, d["g"], a["eps"], +1.0)
    if fn == "pgd_attack":
        lo, hi = g.arr("min_vals").reshape(-1), g.arr("max_vals").reshape(-1)
        cur = O._clamp_nan((x + d["init_noise"]).astype(np.float32), lo.reshape(1, -1, 1, 1), hi.reshape(1, -1, 1, 1))
        for it in range(a["num_iter"]):
            assert np.array_equal(cur, d["x_iter"][it]), f"pgd iterate {it} differs"
            cur = O.pgd_step(cur, x, d["g_iter"][it], a["alpha"], a["epsilon"], lo, hi)
        return cur
    if fn == "angellic.add_gaussian_noise":
        return CH.float_corrupt(x, "gaussian_noise", a["std"], a["mean"], d["randn"])
    if fn == "angellic.add_brightness":
        return CH.float_corrupt(x, "brightness", a["factor"])
    if fn == "angellic.add_contrast":
        return CH.float_corrupt(x, "contrast", a["factor"])
    if fn == "angellic_with_corrupt.add_gaussian_noise":
        return CH.float_corrupt(x, "add_noise", a["std"], 0.0, d["randn"])
    if fn == "angellic_with_corrupt.gaussian_blur_img":
        return CH.gaussian_blur3_f32(x, a["kernel_size"])
    raise KeyError(fn)


_cifarc = {}


def cifarc_synthetic(seed):
    """The synthetic (50000,32,32,3) file make_golden.gen_tta wrote for the CIFARC_Dataset cases."""
    if seed not in _cifarc:
        rng = np.random.default_rng(seed)
        marks = rng.integers(0, 256, (50000, 4), dtype=np.uint8)
        data = np.zeros((50000, 32, 32, 3), np.uint8)
        data[:, 0, 0, :] = marks[:, :3]
        data[:, 31, 31, 0] = marks[:, 3]
        labels = rng.integers(0, 10, 50000).astype(np.int64)
        _cifarc[seed] = (data, labels)
    return _cifarc[seed]


# ---- tensor-backend TTA chain: shared by the CPU and GPU golden tests --------------------------------
TTA_CHAIN_TOL = 1e-5   # |a-b| <= 1e-5*max(1,|b|) on perspective views (grid_sample summation order);
                       # views where RandomPerspective did not fire are NEAREST gathers + one multiply: bit-exact


def posthoc_chain_cases():
    """(input (n,c,h,w), expected (n*views,c,h,w), views, list of per-output draw dicts) per golden chain case."""
    g = golden("posthoc")
    out = []
    for c in g.cases:
        if c["fn"] != "post_hoc_aaryaa.TestTimeAugmentation":
            continue
        d = g.draws(c)
        draws = []
        for j in range(len(d["flip"])):
            fired = bool(d["persp"][j])
            draws.append({"flip": bool(d["flip"][j]), "angle": float(d["angle"][j]), "scale": float(d["scale"][j]),
                          "brightness": float(d["brightness"][j]),
                          "perspective": (d["start"][j].tolist(), d["end"][j].tolist()) if fired else None})
        out.append((g.arr(c["in"]), g.arr(c["out"]), c["args"]["views"], draws))
    return out


def check_chain_view(got, want, fired):
    if not fired:
        assert np.array_equal(got, want)
    else:
        assert np.all(np.abs(got - want) <= TTA_CHAIN_TOL * np.maximum(1.0, np.abs(want)))
