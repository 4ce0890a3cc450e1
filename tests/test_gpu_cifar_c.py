"""GPU parity of the CIFAR-10-C-style suite: CUDA (injected draws) vs oracle/cifar_c.py, bit-exact; on-chip RNG mode
checked statistically; full-size runs checked through size-independent properties."""
import os

import numpy as np
import pytest
import torch

from oracle import cifar_c as CC

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def imgs(seed=0, n=7):
    z = np.load(os.path.join(GOLDEN, "cifar_c_fixtures.npz"))
    nat = np.concatenate([z["contrast"][:2], z["brightness"][2:3], z["fog"][3:4]])
    rnd = np.random.default_rng(seed).integers(0, 256, (n - 4, 32, 32, 3), dtype=np.uint8)
    return np.concatenate([nat, rnd])


def _C():
    from imagetransformations_b200 import cifar_c
    return cifar_c


@pytest.mark.parametrize("sev", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("name", CC.CORRUPTIONS + ["speckle_noise"])
def test_corruption_matches_oracle(cuda, name, sev):
    C = _C()
    x = imgs(sev)
    rng = np.random.RandomState(1000 + sev)
    if name == "speckle_noise":
        d = {"noise": rng.normal(size=x.shape, scale=CC.PARAMS[name][sev - 1])}
        want = CC.speckle_noise(x, d["noise"])
    else:
        d = CC.draw(name, sev, x, rng)
        want = CC.corrupt(name, sev, x, d)
    got = C.corrupt(torch.from_numpy(x).cuda(), name, sev, draws=d).cpu().numpy()
    diff = np.abs(got.astype(np.int32) - want.astype(np.int32))
    assert diff.max() == 0, f"{name} s{sev}: max diff {diff.max()}, {(diff > 0).mean():.4%} of bytes differ"


def test_group_boundaries_and_batch_independence(cuda):
    """Results must not depend on how images are grouped into CTAs: a 301-image batch equals its images one by one."""
    C = _C()
    x = np.random.default_rng(5).integers(0, 256, (301, 32, 32, 3), dtype=np.uint8)
    xd = torch.from_numpy(x).cuda()
    for name, sev in (("defocus_blur", 5), ("zoom_blur", 2), ("jpeg_compression", 3), ("contrast", 1), ("pixelate", 4),
                      ("brightness", 4)):
        full = C.corrupt(xd, name, sev, mode="exact").cpu().numpy()
        for i in (0, 1, 150, 299, 300):
            one = C.corrupt(xd[i:i + 1], name, sev, mode="exact").cpu().numpy()
            assert np.array_equal(full[i], one[0]), (name, i)
        assert np.array_equal(full[:9], CC.corrupt(name, sev, x[:9]))


def test_empty_batch(cuda):
    C = _C()
    x = torch.empty((0, 32, 32, 3), dtype=torch.uint8, device="cuda")
    for name in CC.CORRUPTIONS:
        assert C.corrupt(x, name, 3).shape == (0, 32, 32, 3)


def test_onchip_rng_statistics(cuda):
    C = _C()
    n = 2000
    x = np.full((n, 32, 32, 3), 128, np.uint8)
    xd = torch.from_numpy(x).cuda()
    g = C.corrupt(xd, "gaussian_noise", 5, seed=7).cpu().numpy().astype(np.float64)
    assert abs(g.mean() - 127.5) < 0.2 and abs(g.std() - 25.5) < 0.5        # sigma = 0.10 * 255; truncation biases by -0.5
    g2 = C.corrupt(xd, "gaussian_noise", 5, seed=7).cpu().numpy()
    assert np.array_equal(g2.astype(np.float64), g)                          # same seed -> same draws
    g3 = C.corrupt(xd, "gaussian_noise", 5, seed=8).cpu().numpy()
    assert (g3 != g2).mean() > 0.9
    s = C.corrupt(xd, "shot_noise", 5, seed=3).cpu().numpy().astype(np.float64)
    lam = 128 / 255 * 50
    assert abs(s.mean() / 255 * 50 - lam) < 0.2 and abs((s / 255 * 50).var() - lam) < 1.5   # Poisson: mean = var
    i = C.corrupt(xd, "impulse_noise", 5, seed=4).cpu().numpy()
    frac = (i != 128).mean()
    assert abs(frac - 0.07) < 0.003 and abs((i == 255).sum() / max((i != 128).sum(), 1) - 0.5) < 0.02
    for name in ("glass_blur", "motion_blur", "snow", "frost", "fog", "elastic_transform"):
        xr = torch.from_numpy(imgs(1, 64)).cuda()
        a = C.corrupt(xr, name, 3, seed=11).cpu().numpy()
        b = C.corrupt(xr, name, 3, seed=11).cpu().numpy()
        c = C.corrupt(xr, name, 3, seed=12).cpu().numpy()
        assert np.array_equal(a, b), name
        assert (a != c).mean() > 0.05, name


def test_onchip_plasma_is_a_plasma(cuda):
    """fog with on-chip plasma: output must be a brightened, lower-contrast version bounded like the injected form."""
    C = _C()
    x = imgs(2, 32)
    out = C.corrupt(torch.from_numpy(x).cuda(), "fog", 5, seed=5).cpu().numpy().astype(np.float64)
    # v' = (v + c0*p) * m/(m+c0) with p in [0,1]: bounded by the p = 0 and p = 1 images
    v = x / 255.0
    m = v.reshape(len(x), -1).max(1).reshape(-1, 1, 1, 1)
    lo = np.floor(np.clip(v * m / (m + 1.5), 0, 1) * 255) - 1
    hi = np.floor(np.clip((v + 1.5) * m / (m + 1.5), 0, 1) * 255) + 1
    assert (out >= lo).all() and (out <= hi).all()
    assert out.std() > 1.0


def test_full_suite_properties_at_scale(cuda):
    """BASELINE config sizes (50000 x 32x32x3), every deterministic type: per-image independence at full size in both modes,
    the exact mode bit-identical to the oracle on 1000 images, the fast (default, benchmarked) mode within 1 LSB of it."""
    C = _C()
    n = 50000
    x = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(3))
    perm = torch.randperm(n, device="cuda", generator=torch.Generator("cuda").manual_seed(4))
    xp = x[perm].contiguous()
    head_x = x[:1000].cpu().numpy()
    for name, sev in (("pixelate", 5), ("jpeg_compression", 5), ("contrast", 5), ("defocus_blur", 1), ("zoom_blur", 5),
                      ("brightness", 3), ("defocus_blur", 5), ("zoom_blur", 2)):
        want = CC.corrupt(name, sev, head_x)
        for mode in ("exact", "fast"):
            a = C.corrupt(x, name, sev, mode=mode)
            b = C.corrupt(xp, name, sev, mode=mode)
            assert torch.equal(a[perm], b), (name, mode)               # per-image independence at full size
            got = a[:1000].cpu().numpy()
            if mode == "exact":
                assert np.array_equal(got, want), name                 # 1000 images bit-identical to the oracle
            else:
                d = np.abs(got.astype(np.int32) - want.astype(np.int32))
                assert d.max() <= 1 and (d > 0).mean() < 0.06, (name, int(d.max()), float((d > 0).mean()))
    # contrast c<1 shrinks the per-image range
    a = C.corrupt(x[:1000], "contrast", 5).float()
    assert bool((a.amax((1, 2)) - a.amin((1, 2)) <= 0.15 * 255 + 1.01).all())   # per channel


def test_write_and_read_back_cifar_c_layout(cuda, tmp_path):
    C = _C()
    from imagetransformations_b200.TENT_transform import CIFARC_Dataset
    x = imgs(3, 10)
    # CIFARC_Dataset slices blocks of 10000: write a 10000-image set made of repeats
    xs = np.tile(x, (1000, 1, 1, 1))
    labels = np.arange(10000) % 10
    C.write_cifar_c(str(tmp_path), xs, labels, corruptions=["contrast", "pixelate"], mode="exact")
    for sev in (1, 5):
        ds = CIFARC_Dataset(str(tmp_path), "contrast", sev)
        assert len(ds) == 10000
        img, lab = ds[7]
        assert int(lab) == 7 and np.array_equal(np.array(img), CC.contrast(x[7:8], CC.PARAMS["contrast"][sev - 1])[0])


def test_onchip_noise_follows_the_oracle_law(cuda):
    """On-chip gaussian / shot / impulse noise: the empirical law of the output byte given the input byte must match the law
    the oracle formula defines (normal / Poisson draws pushed through oracle/cifar_c.py), bin by bin."""
    from scipy import stats
    C = _C()
    n = 1500
    N = n * 3072
    for sev in (1, 3, 5):
        sigma = CC.PARAMS["gaussian_noise"][sev - 1]
        for xv in (0, 128, 250):
            xd = torch.full((n, 32, 32, 3), xv, dtype=torch.uint8, device="cuda")
            o = C.corrupt(xd, "gaussian_noise", sev, seed=sev * 10 + xv).cpu().numpy().ravel()
            emp = np.bincount(o, minlength=256) / N
            edges = (np.arange(257) - xv) / (255.0 * sigma)           # out = b  <=>  255*sigma*z in [b - x, b + 1 - x)
            want = np.diff(stats.norm.cdf(edges))
            want[0] += stats.norm.cdf(edges[0])
            want[255] += 1 - stats.norm.cdf(edges[256])
            tol = 5 * np.sqrt(want * (1 - want) / N) + 2e-6
            assert (np.abs(emp - want) <= tol).all(), ("gaussian", sev, xv, np.abs(emp - want).max())
        c = CC.PARAMS["shot_noise"][sev - 1]
        for xv in (0, 3, 128, 255):
            xd = torch.full((n, 32, 32, 3), xv, dtype=torch.uint8, device="cuda")
            o = C.corrupt(xd, "shot_noise", sev, seed=sev * 10 + xv).cpu().numpy().ravel()
            emp = np.bincount(o, minlength=256) / N
            lam = xv / 255.0 * c
            k = np.arange(0, int(lam + 12 * np.sqrt(lam + 1) + 30))
            pk = stats.poisson.pmf(k, lam) if lam > 0 else (k == 0).astype(float)
            ob = CC.shot_noise(np.zeros((1, 1, 1, len(k)), np.uint8), k.reshape(1, 1, 1, -1), c).ravel().astype(int)
            want = np.zeros(256)
            np.add.at(want, ob, pk)
            tol = 5 * np.sqrt(want * (1 - want) / N) + 2e-6
            assert (np.abs(emp - want) <= tol).all(), ("shot", sev, xv, np.abs(emp - want).max())
        p = CC.PARAMS["impulse_noise"][sev - 1]
        xd = torch.full((n, 32, 32, 3), 77, dtype=torch.uint8, device="cuda")
        o = C.corrupt(xd, "impulse_noise", sev, seed=sev).cpu().numpy().ravel()
        f0, f255 = (o == 0).mean(), (o == 255).mean()
        assert abs(f0 - p / 2) < 5 * np.sqrt(p / 2 / N) and abs(f255 - p / 2) < 5 * np.sqrt(p / 2 / N)
        assert ((o == 0) | (o == 255) | (o == 77)).all()
    # neighbouring elements are independent (no stream reuse inside a word / across words / across images)
    xd = torch.full((64, 32, 32, 3), 128, dtype=torch.uint8, device="cuda")
    g = C.corrupt(xd, "gaussian_noise", 5, seed=99).cpu().numpy().astype(np.float64).reshape(64, -1)
    g = g - g.mean()
    for lag in (1, 2, 3, 4, 768, 1536):
        r = (g[:, :-lag] * g[:, lag:]).mean() / g.var()
        assert abs(r) < 0.01, (lag, r)
    r = (g[:-1] * g[1:]).mean() / g.var()
    assert abs(r) < 0.01


def test_sharded_equals_single_gpu(cuda):
    """shards are independent, so cutting the batch (and its injected draws) at rank boundaries and
    concatenating the results is bit-identical to one launch.  Ranks are emulated on one GPU (ragged world sizes too)."""
    from imagetransformations_b200 import shard as S
    C = _C()
    x = imgs(9, 23)
    for name, sev in (("gaussian_noise", 3), ("glass_blur", 4), ("elastic_transform", 2), ("snow", 5), ("fog", 2),
                      ("motion_blur", 5), ("contrast", 2)):
        d = CC.draw(name, sev, x, np.random.RandomState(7))
        full = C.corrupt(torch.from_numpy(x).cuda(), name, sev, draws=d)
        for ws in (2, 3, 8):
            parts = []
            for r in range(ws):
                xs = S.shard(x, r, ws)
                parts.append(C.corrupt(torch.from_numpy(xs).cuda(), name, sev, draws=S.shard_draws(d, len(x), r, ws)))
            assert torch.equal(torch.cat(parts), full), (name, ws)
            assert torch.equal(S.image_checksums(torch.cat(parts)), S.image_checksums(full))


@pytest.mark.parametrize("sev", [1, 2, 3, 4, 5])
def test_brightness_all_colours(cuda, sev):
    """The fixed-point fast path + literal replay of brightness, over a 1/2 sample of all 2^24 colours (every value of
    every channel, all orderings and ties), bit-exact against the f64 oracle."""
    C = _C()
    idx = np.arange(0, 1 << 24, 2, dtype=np.int64) + (np.arange(1 << 23) >> 11 & 1)   # alternate parity so both halves occur
    x = np.stack([idx & 255, (idx >> 8) & 255, (idx >> 16) & 255], -1).astype(np.uint8).reshape(-1, 32, 32, 3)
    got = C.corrupt(torch.from_numpy(x).cuda(), "brightness", sev, mode="exact").cpu().numpy()
    want = np.concatenate([CC.brightness(x[i:i + 1024], CC.PARAMS["brightness"][sev - 1]) for i in range(0, len(x), 1024)])
    assert np.array_equal(got, want), f"{int((got != want).sum())} bytes differ"


@pytest.mark.parametrize("hw,c", [((32, 32), 0.95), ((32, 32), 0.65), ((64, 48), 0.75), ((16, 40), 0.5), ((48, 64), 0.9),
                                  ((224, 224), 0.65), ((224, 224), 0.95), ((96, 80), 0.75), ((37, 53), 0.85), ((130, 66), 0.6)])
def test_pixelate_gather_form_shapes(cuda, hw, c):
    """BOX pixelate gather forms (tables in shared memory up to 64x64, in device memory above; odd widths too) on non-CIFAR
    shapes against the Pillow-exact oracle."""
    from imagetransformations_b200 import _lib as L
    h, w = hw
    x = np.random.default_rng(h * w).integers(0, 256, (37, h, w, 3), dtype=np.uint8)
    xd = torch.from_numpy(x).cuda()
    out = torch.empty_like(xd)
    sw, sh = int(w * c), int(h * c)
    rc = L.load().imgtx_box_pixelate_u8(xd.data_ptr(), out.data_ptr(), len(x), h, w, 3, sw, sh, 0, None, None)
    assert rc == 0
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), CC.pixelate(x, c))


@pytest.mark.parametrize("name", ["brightness", "fog", "defocus_blur", "glass_blur", "motion_blur", "snow", "elastic_transform"])
def test_fast_path_equals_literal_replay(cuda, name):
    """Every kernel that decides bytes with a fixed-point / float32 evaluation must agree with its own literal replay on the
    same on-chip draws: run the suite pass twice, the second time with every element forced through the replay."""
    C = _C()
    x = torch.from_numpy(np.concatenate([imgs(3, 40), np.zeros((1, 32, 32, 3), np.uint8), np.full((1, 32, 32, 3), 255, np.uint8)])).cuda()
    for sev in (1, 3, 5):
        fast = C.corrupt(x, name, sev, seed=11, mode="exact").cpu().numpy()
        lit = C.corrupt(x, name, sev, seed=11, mode="replay").cpu().numpy()
        assert np.array_equal(fast, lit), f"{name}/{sev}: {int((fast != lit).sum())} bytes differ"


def test_pixelate_imagenette_shape_and_clean_error(cuda):
    """224x224 pixelate has its own gather form; calling the fused shared-memory chain directly on such a shape must fail
    cleanly, not fault."""
    from imagetransformations_b200 import _lib as L, batch as B
    C = _C()
    x = np.random.default_rng(3).integers(0, 256, (3, 224, 224, 3), dtype=np.uint8)
    xd = torch.from_numpy(x).cuda()
    got = C.corrupt(xd, "pixelate", 3, mode="exact").cpu().numpy()
    assert np.array_equal(got, CC.pixelate(x, CC.PARAMS["pixelate"][2]))
    sw = sh = int(224 * 0.85)
    t = [B._resample_table(a, b, L.FILTER_BOX, xd.device) for a, b in ((224, sw), (224, sh), (sw, 224), (sh, 224))]
    out = torch.empty_like(xd)
    rc = L.load().imgtx_resize_chain_u8(xd.data_ptr(), out.data_ptr(), 3, 224, 224, 3, sw, sh, t[0].data_ptr(), t[0].numel(),
                                        t[1].data_ptr(), t[1].numel(), t[2].data_ptr(), t[2].numel(), t[3].data_ptr(),
                                        t[3].numel(), None)
    torch.cuda.synchronize()
    assert rc != 0 and b"large-image" in L.load().imgtx_last_error()
    # the context is still healthy
    assert torch.equal(C.corrupt(xd, "pixelate", 3, mode="exact").cpu(), torch.from_numpy(got))
