"""Loader for the *live* reference (TEST INFRASTRUCTURE — never imported by the product).

The reference is a pile of scripts with import-time side effects (matplotlib imports, os.makedirs on the
author's home directory, torch.hub downloads).  This module imports / AST-slices just the hot-path
functions so `oracle/make_golden.py` can run them in THIS container (where /root/reference is mounted) and
freeze golden vectors under tests/golden/.  /root/reference does not exist on the GPU box, so nothing in
tests marked `gpu`, smoke() or bench.py may call into this file.
"""
from __future__ import annotations

import ast
import importlib.machinery
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("IMGTX_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(REF_ROOT)


def _stub_modules():
    for name in ("matplotlib", "matplotlib.pyplot", "tensorflow_datasets"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            sys.modules[name] = m
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]


def load_module(rel_path: str, name: str):
    """Import a reference file whole, with matplotlib stubbed and os.makedirs neutralised."""
    _stub_modules()
    path = os.path.join(REF_ROOT, rel_path)
    loader = importlib.machinery.SourceFileLoader(name, path)
    spec = importlib.util.spec_from_loader(name, loader)
    mod = importlib.util.module_from_spec(spec)
    real_makedirs = os.makedirs
    os.makedirs = lambda *a, **k: None
    try:
        loader.exec_module(mod)
    finally:
        os.makedirs = real_makedirs
    return mod


def slice_module(rel_path: str, names, extra_globals=None):
    """Exec only the top-level defs / assignments whose names are in `names` (AST slice)."""
    path = os.path.join(REF_ROOT, rel_path)
    src = open(path).read()
    tree = ast.parse(src)
    keep = []
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            keep.append(node)
        elif isinstance(node, ast.Assign):
            tg = [t.id for t in node.targets if isinstance(t, ast.Name)]
            if any(t in names for t in tg):
                keep.append(node)
    mod = ast.Module(body=keep, type_ignores=[])
    ns = {}
    import numpy as np
    import torch
    import torchvision.transforms as T
    import torchvision.transforms.functional as TF
    from PIL import Image, ImageEnhance, ImageOps, ImageFilter
    ns.update(dict(np=np, torch=torch, T=T, TF=TF, transforms=T, Image=Image, ImageEnhance=ImageEnhance,
                   ImageOps=ImageOps, ImageFilter=ImageFilter, os=os))
    if extra_globals:
        ns.update(extra_globals)
    exec(compile(mod, path, "exec"), ns)
    return types.SimpleNamespace(**{k: v for k, v in ns.items() if k in names})


_cache = {}


def transformation():
    if "transformation" not in _cache:
        _cache["transformation"] = load_module("transformation.py", "ref_transformation")
    return _cache["transformation"]


def transformations_code():
    if "tc" not in _cache:
        _cache["tc"] = slice_module(
            "fall_2025/transformations_code",
            {"vert_flip", "rand_crop", "apply_random_zoom", "apply_perspective_warp", "apply_scale"},
            extra_globals={"apply_scale": transformation().apply_scale})
    return _cache["tc"]


def augmix():
    if "augmix" not in _cache:
        _cache["augmix"] = load_module("fall_2025/AugMix.py", "ref_augmix")
    return _cache["augmix"]


def tta_transforms():
    if "tta" not in _cache:
        _cache["tta"] = slice_module(
            "fall_2025/TTA_transforms.py",
            {"CIFAR_MEAN", "CIFAR_STD", "base_transform", "tta_transforms_gentle", "tta_transforms_multiscale",
             "tta_transforms_color", "corruption_to_strategy", "get_tta_transforms", "CIFARC_Dataset"},
            extra_globals={"Dataset": __import__("torch").utils.data.Dataset})
    return _cache["tta"]


def tta_initial():
    if "tta_initial" not in _cache:
        _cache["tta_initial"] = slice_module(
            "fall_2025/tta_initial_claudeFixed.py",
            {"CIFAR100_MEAN", "CIFAR100_STD", "denormalize_tensor", "normalize_tensor", "tensor_to_pil",
             "pil_to_tensor", "horiz_flip", "rand_crop", "rand_brightness", "rand_contrast", "AUG_TYPES",
             "AUG_COUNTS"})
    return _cache["tta_initial"]


def pool():
    if "pool" not in _cache:
        _cache["pool"] = load_module("pipenline/cifar_image_transformations.py", "ref_pool")
    return _cache["pool"]


def fgsm():
    if "fgsm" not in _cache:
        ns = types.SimpleNamespace()
        ns.fgsm_attack = slice_module("attacks/fgsm/cifar-10/fgsm_cifar10.py", {"fgsm_attack"}).fgsm_attack
        ns.reverse_fgsm_attack = slice_module("attacks/fgsm/cifar-10/reverse_fgsm_cifar10.py",
                                              {"reverse_fgsm_attack"}).reverse_fgsm_attack
        import torch
        from torch.autograd import Variable
        ns.FGSM = slice_module("attacks/fgsm.py", {"FGSM"},
                               extra_globals={"Variable": Variable}).FGSM
        ns.pgd_attack = slice_module("attacks/fgsm/cifar-10/pgd_cifar10_corrupted.py", {"pgd_attack"},
                                     extra_globals={"criterion": torch.nn.CrossEntropyLoss()}).pgd_attack
        _cache["fgsm"] = ns
    return _cache["fgsm"]


def entropy_gate():
    """compute_shannon_entropy of fall_2025/TENT_Pipelines.py:99-121 (normalised input) and of
    fall_2025/Initial_Experiments.py:95-113 ([0,1] input)."""
    if "entropy_gate" not in _cache:
        import scipy.stats
        ns = types.SimpleNamespace()
        t = slice_module("fall_2025/TENT_Pipelines.py", {"compute_shannon_entropy", "CIFAR_MEAN", "CIFAR_STD"})
        ns.tent_entropy, ns.CIFAR_MEAN, ns.CIFAR_STD = t.compute_shannon_entropy, t.CIFAR_MEAN, t.CIFAR_STD
        ns.unit_entropy = slice_module("fall_2025/Initial_Experiments.py", {"compute_shannon_entropy"},
                                       extra_globals={"entropy": scipy.stats.entropy}).compute_shannon_entropy
        _cache["entropy_gate"] = ns
    return _cache["entropy_gate"]


def adaptive():
    """adaptive_transforms (the T.Compose of fall_2025/Initial_Experiments.py:88-92) and `transform` (:81-84)."""
    if "adaptive" not in _cache:
        import torchvision.transforms as T
        _cache["adaptive"] = slice_module("fall_2025/Initial_Experiments.py",
                                          {"adaptive_transforms", "transform", "CIFAR_MEAN", "CIFAR_STD"}, extra_globals={"T": T})
    return _cache["adaptive"]


def post_hoc():
    """TestTimeAugmentation of fall_2025/post_hoc_aaryaa.py:22-51 (the class only; the file trains models at import)."""
    if "post_hoc" not in _cache:
        _cache["post_hoc"] = slice_module("fall_2025/post_hoc_aaryaa.py", {"TestTimeAugmentation"})
    return _cache["post_hoc"]


def angellic():
    if "angellic" not in _cache:
        _cache["angellic"] = slice_module("pipenline/angellic.py",
                                          {"add_gaussian_noise", "add_brightness", "add_contrast"})
    return _cache["angellic"]
