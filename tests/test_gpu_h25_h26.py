"""GPU: H26 (fused adaptive_transforms chain) and H25 (patch / ODIN sign steps) through the C ABI against goldens frozen from
the live reference: uint8 bit-exact, float32 bit-exact (NaN positions equal)."""
import numpy as np
import pytest
import torch

import golden_cases as G
from test_oracle_h25_h26 import adaptive_cases

pytestmark = pytest.mark.gpu


def test_adaptive_chain_golden_fused(cuda):
    from imagetransformations_b200 import Initial_Experiments as IE
    cases = list(adaptive_cases())
    for shape in {c[0].shape for c in cases}:
        sub = [c for c in cases if c[0].shape == shape]
        h, w = shape[:2]
        recs = np.zeros(len(sub), IE.ADAPTIVE_DTYPE)
        for k, (u, flip, m, pc, want) in enumerate(sub):
            recs[k]["flags"] = (1 if flip else 0) | 2 | (4 if pc is not None else 0)
            recs[k]["a"] = [IE._fix(m[0]), IE._fix(m[1]), IE._fix(m[2] + m[0] * 0.5 + m[1] * 0.5), IE._fix(m[3]), IE._fix(m[4]),
                            IE._fix(m[5] + m[3] * 0.5 + m[4] * 0.5)]
            if pc is not None:
                recs[k]["persp"] = pc
        x = torch.from_numpy(np.stack([c[0] for c in sub])).cuda()
        got = IE.adaptive_transforms_batch(x, recs).cpu().numpy()
        want = np.stack([c[4] for c in sub])
        assert np.array_equal(got, want), (shape, int((got != want).sum()))


def test_adaptive_transforms_pil_dropin_consumes_the_same_draws(cuda):
    """The PIL-signature drop-in under the golden's torch seeds reproduces the reference's outputs byte for byte."""
    from PIL import Image
    from imagetransformations_b200 import Initial_Experiments as IE
    g = G.golden("adaptive")
    c = [c for c in g.cases if c["fn"] == "Initial_Experiments.adaptive_transforms" and c["in"] == "x32"][0]
    x, want, reps = g.arr(c["in"]), g.arr(c["out"]), c["args"]["reps"]
    for i in range(len(x)):
        for v in range(reps):
            torch.manual_seed(c["args"]["seed0"] + 100 * i + v)
            got = np.array(IE.adaptive_transforms(Image.fromarray(x[i])))
            assert np.array_equal(got, want[i * reps + v]), (i, v)


def test_adapt_call_site_golden(cuda):
    """:147-167 for a batch: with the threshold at -inf every image is adapted; the reference seeds per image, so the batch is
    replayed image by image under the same seeds."""
    from imagetransformations_b200 import Initial_Experiments as IE
    g = G.golden("adaptive")
    c = [c for c in g.cases if c["fn"] == "Initial_Experiments.adapt_call_site"][0]
    xn, want = torch.from_numpy(g.arr(c["in"])).cuda(), g.arr(c["out"])
    for i in range(len(xn)):
        torch.manual_seed(c["args"]["seed0"] + i)
        got = IE.adapt_batch(xn[i:i + 1], threshold=-1.0).cpu().numpy()[0]
        assert np.array_equal(got, want[i]), i
    # the gate itself: nothing is touched above the entropy range
    assert torch.equal(IE.adapt_batch(xn, threshold=1e9), xn)


def test_sign_steps_golden(cuda):
    from imagetransformations_b200 import fgsm as F
    g = G.golden("signsteps")
    for c in g.cases:
        x, want, grad = g.arr(c["in"]), g.arr(c["out"]), g.draws(c)["g"]
        xd, gd = torch.from_numpy(x.copy()).cuda(), torch.from_numpy(grad).cuda()
        if c["fn"].endswith("patch_step"):
            ret = F.patch_step_(xd, gd, c["args"]["eps"])
            assert ret.data_ptr() == xd.data_ptr()          # in place, like the reference
            got = xd.cpu().numpy()
        else:
            got = F.odin_perturb(xd, gd, c["args"]["eps"]).cpu().numpy()
        same = (got == want) | (np.isnan(got) & np.isnan(want))
        assert same.all(), (c["fn"], c["args"], int((~same).sum()))


def test_sign_step_post_fused_equals_the_unfused_chain(cuda):
    """BASELINE config 5: sign step + clamp, then tensor_to_pil -> ImageEnhance.Brightness -> pil_to_tensor
    (pipenline/cifar_image_transformations.py:27-34,88-92,141-152), fused into one kernel: bit-identical to the oracle's chain
    of the four golden-pinned stages."""
    from imagetransformations_b200 import batch as B
    from oracle import np_ops as O
    mean, std = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)
    rng = np.random.default_rng(5)
    x8 = rng.integers(0, 256, (40, 32, 32, 3), dtype=np.uint8)
    x = O.to_tensor_normalize(x8, mean, std)
    g = rng.standard_normal(x.shape).astype(np.float32)
    g.reshape(-1)[:6] = [0.0, -0.0, np.nan, np.inf, -np.inf, 1e-45]
    lo = [(0 - m) / s for m, s in zip(mean, std)]
    hi = [(1 - m) / s for m, s in zip(mean, std)]
    eps = [4 / 355, 8 / 255, 12 / 255]
    lut = O.pil_blend_lut(0, 1.3)
    got = B.sign_step_post(torch.from_numpy(x).cuda(), torch.from_numpy(g).cuda(), eps, lut, mean, std, -1.0, lo, hi).cpu().numpy()
    for k, e in enumerate(eps):
        step = O.fgsm_step(x, g, e, -1, lo, hi)
        u = O.tensor_to_pil(step, mean, std)
        want = O.to_tensor_normalize(lut[u], mean, std)
        assert np.array_equal(got[k], want), (k, int((got[k] != want).sum()))
