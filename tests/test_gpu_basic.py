"""GPU parity: pointwise / geometry / tensor / sign-step kernels vs the numpy oracle (bit-exact)."""
import numpy as np
import pytest
import torch

from oracle import np_ops as O

pytestmark = pytest.mark.gpu

SIZES = [(32, 32), (224, 224), (37, 53), (64, 48)]


def _imgs(seed, n, h, w, c=3):
    return np.random.default_rng(seed).integers(0, 256, (n, h, w, c), dtype=np.uint8)


def _B():
    from imagetransformations_b200 import batch as B
    return B


@pytest.mark.parametrize("hw", SIZES)
def test_pointwise_uniform(cuda, hw):
    B = _B()
    x = _imgs(1, 9, *hw)
    xd = torch.from_numpy(x).cuda()
    for f in [0.95, 0.97, 1.0, 1.05, 0.5, 1.5, 0.0, 1.3]:
        assert np.array_equal(B.brightness(xd, f).cpu().numpy(), O.enhance_brightness(x, f)), f
    for a in [0.0, 0.1, 0.30000000000000004, 0.7000000000000001, 1.0, 1.7]:
        assert np.array_equal(B.convert_scale_abs(xd, a).cpu().numpy(), O.convert_scale_abs(x, a)), a
    assert np.array_equal(B.posterize(xd, 3).cpu().numpy(), O.posterize(x, 3))
    assert np.array_equal(B.solarize(xd, 60).cpu().numpy(), O.solarize(x, 60))


def test_pointwise_per_image(cuda):
    B = _B()
    x = _imgs(2, 33, 32, 32)
    xd = torch.from_numpy(x).cuda()
    f = np.random.default_rng(3).uniform(0.5, 1.5, 33)
    assert np.array_equal(B.brightness(xd, f).cpu().numpy(), O.enhance_brightness(x, f))
    a = np.random.default_rng(4).uniform(0.0, 1.2, 33)
    assert np.array_equal(B.convert_scale_abs(xd, a).cpu().numpy(), O.convert_scale_abs(x, a))
    x2 = _imgs(5, 7, 37, 53)
    f2 = np.random.default_rng(6).uniform(0.5, 1.5, 7)
    assert np.array_equal(B.brightness(torch.from_numpy(x2).cuda(), f2).cpu().numpy(), O.enhance_brightness(x2, f2))


def test_pointwise_empty_and_large(cuda):
    B = _B()
    x = torch.empty((0, 32, 32, 3), dtype=torch.uint8, device="cuda")
    assert B.brightness(x, 1.1).shape == (0, 32, 32, 3)
    big = _imgs(8, 5000, 32, 32)
    assert np.array_equal(B.brightness(torch.from_numpy(big).cuda(), 1.04).cpu().numpy(), O.enhance_brightness(big, 1.04))


@pytest.mark.parametrize("hw", SIZES)
def test_rotate(cuda, hw):
    B = _B()
    x = _imgs(10, 5, *hw)
    xd = torch.from_numpy(x).cuda()
    for ang in [-22.5, -2.5, 0.0, 2.5, 12.5, 22.5, 45.0, 90.0, 180.0, 270.0, 3.0, -3.0]:
        got = B.rotate(xd, ang).cpu().numpy() if not (ang in (90.0, 270.0) and hw[0] != hw[1]) else None
        if got is None:
            got = B.rotate(xd, ang).cpu().numpy()
        assert np.array_equal(got, O.pil_rotate_nearest(x, ang)), ang


def test_rotate_per_image_many(cuda):
    B = _B()
    x = _imgs(11, 1000, 32, 32)
    ang = np.random.default_rng(12).choice(np.arange(-22.5, 22.6, 2.5), 1000)
    got = B.rotate(torch.from_numpy(x).cuda(), ang).cpu().numpy()
    assert np.array_equal(got, O.pil_rotate_nearest(x, ang))


@pytest.mark.parametrize("hw", SIZES)
def test_affine_nearest_and_shift(cuda, hw):
    B = _B()
    x = _imgs(13, 4, *hw)
    xd = torch.from_numpy(x).cuda()
    for m in [[1, 0.9, 0, 0, 1, 0], [1, 0, 0, 0.9, 1, 0], [1, 0, 6, 0, 1, 0], [1, 0, 0, 0, 1, 6],
              [1, 0, -6.5, 0, 1, 2.25], [1.1, 0, 3, 0, 0.9, -2]]:
        assert np.array_equal(B.affine_nearest(xd, m).cpu().numpy(), O.affine_nearest(x, m)), m
    for tx, ty in [(-50, 5), (5, -10), (0, 0), (31, 31), (-5, -5), (10, 0)]:
        assert np.array_equal(B.translate(xd, tx, ty).cpu().numpy(), O.translate(x, tx, ty)), (tx, ty)
    assert np.array_equal(B.hflip(xd).cpu().numpy(), O.hflip(x))
    txs = np.array([-5, 0, 7, 40])
    tys = np.array([3, -3, 0, 1])
    assert np.array_equal(B.translate(xd, txs, tys).cpu().numpy(), O.translate(x, txs, tys))


def test_crop_reflect(cuda):
    B = _B()
    x = _imgs(14, 20, 32, 32)
    rng = np.random.default_rng(15)
    i = rng.integers(0, 9, 20)
    j = rng.integers(0, 9, 20)
    got = B.crop_reflect(torch.from_numpy(x).cuda(), i, j, 4, 32).cpu().numpy()
    ref = np.stack([np.pad(x[k], ((4, 4), (4, 4), (0, 0)), mode="reflect")[i[k]:i[k] + 32, j[k]:j[k] + 32] for k in range(20)])
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("hw", SIZES)
def test_tensor_roundtrip(cuda, hw):
    B = _B()
    mean, std = (0.5071, 0.4867, 0.4408), (0.2675, 0.2565, 0.2761)
    x = _imgs(16, 6, *hw)
    xd = torch.from_numpy(x).cuda()
    t = B.to_tensor(xd, mean, std)
    assert np.array_equal(t.cpu().numpy(), O.to_tensor_normalize(x, mean, std))
    assert np.array_equal(B.to_tensor(xd).cpu().numpy(), O.to_tensor(x))
    back = B.to_uint8(t, mean, std, clamp=True).cpu().numpy()
    assert np.array_equal(back, O.tensor_to_pil(t.cpu().numpy(), mean, std))
    # AugMix-style to_pil_image on normalised (out-of-range) floats: modulo-256 wrap
    assert np.array_equal(B.to_uint8(t).cpu().numpy(), O.to_pil_image(t.cpu().numpy()))


def _grad(n, c, h, w, seed=5):
    g = torch.Generator().manual_seed(seed)
    t = torch.randn((n, c, h, w), generator=g)
    flat = t.view(-1)
    flat[::97] = 0.0
    flat[1] = -0.0
    flat[2] = float("nan")
    flat[3] = float("inf")
    flat[4] = float("-inf")
    return t


def test_sign_step(cuda):
    B = _B()
    mean, std = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)
    x8 = _imgs(17, 50, 32, 32)
    x = O.to_tensor_normalize(x8, mean, std)
    g = _grad(50, 3, 32, 32).numpy()
    xd, gd = torch.from_numpy(x).cuda(), torch.from_numpy(g).cuda()
    lo = [(0 - m) / s for m, s in zip(mean, std)]
    hi = [(1 - m) / s for m, s in zip(mean, std)]
    eps_list = [4 / 355, 8 / 255, 12 / 255, 0.03]
    out, pert = B.sign_step(xd, gd, eps_list, +1.0, return_pert=True)
    for k, e in enumerate(eps_list):
        assert np.array_equal(out[k].cpu().numpy(), O.fgsm_step(x, g, e), equal_nan=True)
        assert np.array_equal(pert[k].cpu().numpy(), (np.float32(e) * O.torch_sign(g)).astype(np.float32))
        assert np.array_equal(B.sign_step(xd, gd, e, +1.0, 0.0, 1.0).cpu().numpy(), O.fgsm_step(x, g, e, +1, 0.0, 1.0), equal_nan=True)
        assert np.array_equal(B.sign_step(xd, gd, e, -1.0, lo, hi).cpu().numpy(), O.fgsm_step(x, g, e, -1, lo, hi), equal_nan=True)
    # NaN in x propagates through the clamp
    x2 = x.copy()
    x2[0, 0, 0, 0] = np.nan
    got = B.sign_step(torch.from_numpy(x2).cuda(), gd, 0.03, -1.0, lo, hi).cpu().numpy()
    assert np.array_equal(got, O.fgsm_step(x2, g, 0.03, -1, lo, hi), equal_nan=True)
    # odd plane size -> scalar path
    xs, gs = x[:, :, :5, :7].copy(), g[:, :, :5, :7].copy()
    got = B.sign_step(torch.from_numpy(xs).cuda(), torch.from_numpy(gs).cuda(), 0.03, -1.0, lo, hi).cpu().numpy()
    assert np.array_equal(got, O.fgsm_step(xs, gs, 0.03, -1, lo, hi), equal_nan=True)


def test_pgd_step(cuda):
    B = _B()
    rng = np.random.default_rng(18)
    x0 = rng.random((20, 3, 32, 32), dtype=np.float32)
    x = (x0 + rng.uniform(-0.3, 0.3, x0.shape).astype(np.float32)).astype(np.float32)
    g = _grad(20, 3, 32, 32, seed=9).numpy()
    got = B.pgd_step(torch.from_numpy(x).cuda(), torch.from_numpy(x0).cuda(), torch.from_numpy(g).cuda(),
                     2 / 255.0, 0.3, 0.0, 1.0).cpu().numpy()
    assert np.array_equal(got, O.pgd_step(x, x0, g, 2 / 255.0, 0.3, [0.0], [1.0]), equal_nan=True)
