"""CPU: the oracle restatements of H25 (patch / ODIN sign steps) and H26 (adaptive_transforms) against goldens frozen from the
live reference (oracle/make_golden.py: gen_signsteps, gen_adaptive)."""
import numpy as np

import golden_cases as G
from oracle import np_chains as CH
from oracle import np_ops as O


def _pcoeffs(start, end):
    """torchvision hands Pillow the least-squares solution ROUNDED TO FLOAT32 (_get_perspective_coeffs: .to(torch.float32))."""
    return CH.perspective_coeffs(start, end).astype(np.float32).astype(np.float64)


def adaptive_cases():
    """(image u8 (h,w,3), flip, affine matrix (6,), perspective coeffs or None, expected u8) per golden output."""
    from torchvision.transforms import functional as TF
    g = G.golden("adaptive")
    for c in g.cases:
        if c["fn"] != "Initial_Experiments.adaptive_transforms":
            continue
        x, want, d = g.arr(c["in"]), g.arr(c["out"]), g.draws(c)
        reps = c["args"]["reps"]
        h, w = x.shape[1:3]
        for j in range(len(want)):
            m = TF._get_inverse_affine_matrix([w * 0.5, h * 0.5], float(d["angle"][j]), [float(d["tx"][j]), float(d["ty"][j])],
                                              float(d["scale"][j]), [0.0, 0.0])
            pc = _pcoeffs(d["start"][j].tolist(), d["end"][j].tolist()) if d["persp"][j] else None
            yield x[j // reps], bool(d["flip"][j]), m, pc, want[j]


def test_oracle_adaptive_chain_golden():
    n = fired = flipped = 0
    for u, flip, m, pc, want in adaptive_cases():
        got = CH.adaptive_chain(u, flip, m, pc)
        assert np.array_equal(got, want), (n, flip, pc is not None)
        n += 1; fired += pc is not None; flipped += flip
    assert n == 60 and 10 < fired < 50 and 10 < flipped < 50


def test_oracle_adaptive_call_site_golden():
    """:147-167: denormalise, clamp, ToPILImage (x255, truncate), chain under the recorded seed, ToTensor + Normalize."""
    import torch
    from imagetransformations_b200 import Initial_Experiments as IE   # only its host-side draw replay (no GPU call)
    from torchvision.transforms import functional as TF
    g = G.golden("adaptive")
    c = [c for c in g.cases if c["fn"] == "Initial_Experiments.adapt_call_site"][0]
    xn, want = g.arr(c["in"]), g.arr(c["out"])
    for i in range(len(xn)):
        u = O.tensor_to_pil(xn[i:i + 1], IE.CIFAR_MEAN, IE.CIFAR_STD)[0]
        torch.manual_seed(c["args"]["seed0"] + i)
        d = IE.draw_adaptive_params(32, 32)
        m = TF._get_inverse_affine_matrix([16.0, 16.0], d["angle"], [float(d["translate"][0]), float(d["translate"][1])],
                                          d["scale"], [0.0, 0.0])
        pc = _pcoeffs(d["start"], d["end"]) if d["persp"] else None
        y = CH.adaptive_chain(u, d["flip"], m, pc)
        got = O.to_tensor_normalize(y[None], IE.CIFAR_MEAN, IE.CIFAR_STD)[0]
        assert np.array_equal(got, want[i]), i


def test_oracle_sign_steps_golden():
    g = G.golden("signsteps")
    assert len(g.cases) == 6
    for c in g.cases:
        x, want, grad = g.arr(c["in"]), g.arr(c["out"]), g.draws(c)["g"]
        got = O.patch_step(x, grad, c["args"]["eps"]) if c["fn"].endswith("patch_step") else O.odin_perturb(x, grad, c["args"]["eps"])
        same = (got == want) | (np.isnan(got) & np.isnan(want))
        assert same.all(), (c["fn"], c["args"], int((~same).sum()))
