"""Law of the ON-CHIP draws of the six random stencil / warp types (glass_blur, motion_blur, snow, frost, fog,
elastic_transform) — the path bench.py times.

No CPU oracle can replay Philox draws bit for bit, so the on-chip mode is held to the LAW the canonical generator defines
(oracle/cifar_c.py:draw: randint(-1, 1) swap deltas, U(-45, 45) / U(-135, -45) angles, N(loc, scale) snow layers, the
diamond-square plasma with its wibble decay, U(-c2, c2) corner jitter, U(-1, 1) displacement fields) with two-sample tests:
the same image is corrupted N times on chip (default = fast mode, and the exact mode) and N times by the oracle fed numpy
draws, and the two output ensembles must agree in

  * the per-element mean and variance maps (Welch z per byte; max |z| and the spread of z bounded),
  * per-image spatial autocovariances at 13 lags (streak direction, plasma roughness, displacement smoothness: a wrong angle
    range, decay or sigma moves these by many standard errors even when every marginal matches),

plus the targeted statistics the law names: the glass displacement histogram, the motion-blur angle distribution (KS against
the oracle's and the U(-45, 45) range), and the uniformity of the frost crop origins.
Tolerances: |z| < 6 for 3072 simultaneous per-byte mean tests (P < 1e-5 under H0; bytes of one image are correlated, so only
the maximum is bounded), |z| < 7 for the per-byte variances (z built with the fourth moment), |z| < 5 for the 13 lag
statistics; chi-square / KS p-values > 1e-4.  The fast mode is allowed its contractual 1 LSB per byte in the mean test.
"""
import os

import numpy as np
import pytest
import torch

from oracle import cifar_c as CC

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
N = 2000
LAGS = [(0, 1), (1, 0), (1, 1), (1, -1), (0, 2), (2, 0), (2, 1), (1, 2), (0, 4), (4, 0), (4, 4), (0, 8), (8, 0)]


def _C():
    from imagetransformations_b200 import cifar_c
    return cifar_c


def natural_image(k=0):
    z = np.load(os.path.join(GOLDEN, "cifar_c_fixtures.npz"))
    return z["contrast"][k]


def lag_stats(y, mean_map):
    """(n, len(LAGS)) per-image autocovariance of the green-weighted grey residual at each lag."""
    g = (y.astype(np.float64) - mean_map).mean(axis=-1)          # (n, 32, 32)
    out = np.empty((len(g), len(LAGS)))
    for k, (dy, dx) in enumerate(LAGS):
        a = g[:, max(0, -dy):32 - max(0, dy), max(0, -dx):32 - max(0, dx)]
        b = g[:, max(0, dy):32 - max(0, -dy), max(0, dx):32 - max(0, -dx)]
        out[:, k] = (a * b).mean(axis=(1, 2))
    return out


def compare_ensembles(a, b, tag, z_max=6.0, lsb=0.0):
    """a, b: (n, 32, 32, 3) uint8 ensembles.  lsb = systematic per-byte offset allowed between the two (1 for the fast mode,
    whose contract is +-1 LSB per byte: on a byte that hardly varies that offset is many standard errors)."""
    a64, b64 = a.astype(np.float64), b.astype(np.float64)
    ma, mb = a64.mean(0), b64.mean(0)
    ca, cb = a64 - ma, b64 - mb
    va, vb = (ca ** 2).mean(0), (cb ** 2).mean(0)
    se = np.sqrt(va / len(a) + vb / len(b))
    live = se > 1e-9
    assert np.abs(ma - mb)[~live].max(initial=0.0) <= 1.0, tag          # bytes that never vary: equal up to the fast-mode LSB
    z = np.maximum(np.abs(ma - mb)[live] - lsb, 0.0) / se[live]
    assert z.max() < z_max, (tag, "mean map", float(z.max()))
    # variances: z-test with the sampling variance of a variance, (m4 - v^2)/n (bytes are far from Gaussian)
    sv = np.sqrt(((ca ** 4).mean(0) - va ** 2) / len(a) + ((cb ** 4).mean(0) - vb ** 2) / len(b))
    big = (va > 4.0) & (vb > 4.0) & (sv > 1e-9)
    if big.any():
        slack = 2.0 * lsb * np.sqrt(np.maximum(va, vb)[big])              # d(var) for a +-lsb perturbation of the values
        zv = np.maximum(np.abs(va - vb)[big] - slack, 0.0) / sv[big]
        assert zv.max() < z_max + 1.0, (tag, "variance map", float(zv.max()))
        assert np.abs(np.median(np.log(va[big] / vb[big]))) < 0.05, (tag, "variance map (median log ratio)")
    pooled = 0.5 * (ma + mb)
    la, lb = lag_stats(a, pooled), lag_stats(b, pooled)
    zl = (la.mean(0) - lb.mean(0)) / np.sqrt(la.var(0, ddof=1) / len(la) + lb.var(0, ddof=1) / len(lb) + 1e-30)
    assert np.abs(zl).max() < 5.0, (tag, "lag autocovariances", dict(zip(map(str, LAGS), np.round(zl, 2))))


def oracle_ensemble(name, sev, img, seed):
    x = np.repeat(img[None], N, 0)
    rng = np.random.RandomState(seed)
    return CC.corrupt(name, sev, x, CC.draw(name, sev, x, rng))


@pytest.mark.parametrize("name,sev", [("glass_blur", 1), ("glass_blur", 2), ("glass_blur", 5), ("motion_blur", 1), ("motion_blur", 5),
                                      ("snow", 2), ("snow", 5), ("fog", 1), ("fog", 3), ("fog", 5), ("elastic_transform", 1),
                                      ("elastic_transform", 2), ("elastic_transform", 4), ("elastic_transform", 5)])
def test_onchip_ensemble_matches_the_oracle_law(cuda, name, sev):
    C = _C()
    img = natural_image(sev % 3)
    want = oracle_ensemble(name, sev, img, 100 + sev)
    xd = torch.from_numpy(np.repeat(img[None], N, 0)).cuda()
    for mode in ("fast", "exact"):
        got = C.corrupt(xd, name, sev, seed=31 + sev, mode=mode).cpu().numpy()
        compare_ensembles(got, want, f"{name}/{sev}/{mode}", lsb=1.0 if mode == "fast" else 0.0)


def test_the_ensemble_test_has_power(cuda):
    """The same comparison must REJECT a wrong law: motion blur with angles U(-20, 20) instead of U(-45, 45), fog with the wrong
    wibble decay, glass with P(delta = -1) = 0.8 instead of 1/2."""
    img = natural_image(1)
    x = np.repeat(img[None], N, 0)
    rng = np.random.RandomState(5)
    good = oracle_ensemble("motion_blur", 5, img, 1)
    bad = CC.corrupt("motion_blur", 5, x, {"angles": rng.uniform(-20, 20, N)})
    with pytest.raises(AssertionError):
        compare_ensembles(bad, good, "motion wrong range")
    good = oracle_ensemble("fog", 3, img, 2)
    bad = CC.fog(x, np.stack([CC.plasma_fractal(32, 2.0, lambda lo, hi, s: rng.uniform(lo, hi, s)) for _ in range(N)]), 3)
    with pytest.raises(AssertionError):
        compare_ensembles(bad, good, "fog wrong decay")
    good = oracle_ensemble("glass_blur", 1, img, 3)
    bad = CC.corrupt("glass_blur", 1, x, {"deltas": -(rng.random_sample((N, 1, 30, 30, 2)) < 0.8).astype(np.int8)})
    with pytest.raises(AssertionError):
        compare_ensembles(bad, good, "glass wrong deltas")


def test_glass_displacement_histogram(cuda):
    """Severity 1 (sigma 0.05: the Gaussians are byte maps) moves pixels without mixing them, so an image that encodes its
    own coordinates reveals where every output pixel came from.  The histogram of displacements under on-chip draws must equal
    the one under randint(-1, 1) deltas (two-sample chi-square)."""
    from scipy import stats
    C = _C()
    yy, xx = np.mgrid[0:32, 0:32]
    img = np.stack([yy * 8 + 4, xx * 8 + 4, np.full_like(yy, 100)], -1).astype(np.uint8)
    n = 1500
    x = np.repeat(img[None], n, 0)
    rng = np.random.RandomState(9)

    def hist(y):
        sy, sx = (y[..., 0].astype(int)) // 8, (y[..., 1].astype(int)) // 8
        dy, dx = np.clip(sy - yy, -4, 4) + 4, np.clip(sx - xx, -4, 4) + 4
        return np.bincount((dy * 9 + dx).ravel(), minlength=81).astype(np.float64)

    for sev in (1,):
        want = hist(CC.corrupt("glass_blur", sev, x, CC.draw("glass_blur", sev, x, rng)))
        for mode in ("fast", "exact"):
            got = hist(C.corrupt(torch.from_numpy(x).cuda(), "glass_blur", sev, seed=77, mode=mode).cpu().numpy())
            keep = (want + got) >= 20
            # pixels of one image are not independent draws: the effective sample is the number of images, so the
            # statistic is computed on per-image-normalised counts
            tbl = np.stack([got[keep], want[keep]]) / 1024.0
            chi2, p, _, _ = stats.chi2_contingency(tbl)
            assert p > 1e-4, (mode, sev, chi2, p)
            assert np.abs(got / got.sum() - want / want.sum()).max() < 0.01


def test_motion_angle_law(cuda):
    """A single bright pixel shows the blur line: its weighted centroid points along (cos a, sin a).  On-chip angles must be
    distributed like the oracle's U(-45, 45) draws (two-sample KS on the estimator) and stay inside the range."""
    from scipy import stats
    C = _C()
    img = np.zeros((32, 32, 3), np.uint8)
    img[16, 24] = 255
    n = 3000
    x = np.repeat(img[None], n, 0)
    yy, xx = np.mgrid[0:32, 0:32]

    def angles(y):
        w = y[..., 0].astype(np.float64)
        cy = (w * (yy - 16)).sum((1, 2)) / w.sum((1, 2))
        cx = (w * (xx - 24)).sum((1, 2)) / w.sum((1, 2))
        return np.degrees(np.arctan2(-cy, -cx))      # out[p] = sum k_i src[p + off_i]: the mass sits at -off_i

    rng = np.random.RandomState(4)
    want = angles(CC.corrupt("motion_blur", 5, x, CC.draw("motion_blur", 5, x, rng)))
    for mode in ("fast", "exact"):
        got = angles(C.corrupt(torch.from_numpy(x).cuda(), "motion_blur", 5, seed=8, mode=mode).cpu().numpy())
        assert got.min() > -50 and got.max() < 50, (mode, got.min(), got.max())
        assert stats.ks_2samp(got, want).pvalue > 1e-4, mode
        assert abs(got.mean()) < 2.0 and abs(got.std() - 90 / np.sqrt(12)) < 1.5, (mode, got.mean(), got.std())


def test_snow_layer_and_angle_law(cuda):
    """Snow on a black image is the blurred flake layer plus its 180-degree rotation (plus a constant): its per-image mass is
    the thresholded N(loc, scale) layer's, and the streaks point along U(-135, -45).  Two-sample KS on the per-image mass and on
    the streak direction estimated from second moments of the autocovariance."""
    from scipy import stats
    C = _C()
    n = 3000
    x = np.zeros((n, 32, 32, 3), np.uint8)
    rng = np.random.RandomState(6)

    def feats(y):
        g = y[..., 0].astype(np.float64)
        g = g - np.median(g, axis=(1, 2), keepdims=True)
        mass = g.sum((1, 2))
        # streak direction from the ratio of vertical to horizontal lag-2 autocovariance
        v = (g[:, 2:, :] * g[:, :-2, :]).mean((1, 2)); h = (g[:, :, 2:] * g[:, :, :-2]).mean((1, 2))
        return mass, (v - h) / (np.abs(v) + np.abs(h) + 1e-9)

    for sev in (3, 5):
        want = feats(CC.corrupt("snow", sev, x, CC.draw("snow", sev, x, rng)))
        for mode in ("fast", "exact"):
            got = feats(C.corrupt(torch.from_numpy(x).cuda(), "snow", sev, seed=17 + sev, mode=mode).cpu().numpy())
            for k, nm in enumerate(("mass", "direction")):
                assert stats.ks_2samp(got[k], want[k]).pvalue > 1e-4, (sev, mode, nm)


def test_frost_crop_origins_are_uniform(cuda):
    """On a black image frost is floor(c1 * crop): every output must be a crop of the texture bank, with the texture index
    and the crop origin uniform over their ranges (chi-square)."""
    from scipy import stats
    C = _C()
    n = 20000
    xd = torch.zeros((n, 32, 32, 3), dtype=torch.uint8, device="cuda")
    bank = C.procedural_frost_bank(xd.device).cpu().numpy()
    c1 = CC.PARAMS["frost"][2][1]
    y = C.corrupt(xd, "frost", 3, seed=123, mode="fast").cpu().numpy()
    exact = C.corrupt(xd, "frost", 3, seed=123, mode="exact").cpu().numpy()
    assert np.abs(y.astype(int) - exact.astype(int)).max() <= 1
    fb = np.floor(c1 * bank.astype(np.float64)).astype(np.uint8)
    T, S = bank.shape[0], bank.shape[1] - 32
    # signature = the 16 bytes at the crop origin; the exact mode's bytes are floor(c1 * crop)
    sig = {}
    for t in range(T):
        for y0 in range(S):
            rows = fb[t, y0]
            for x0 in range(S):
                sig.setdefault(rows[x0:x0 + 6].tobytes(), []).append((t, y0, x0))
    ti, ys, xs = [], [], []
    for im in exact:
        cands = [c for c in sig.get(im[0, :6].tobytes(), []) if np.array_equal(fb[c[0], c[1]:c[1] + 32, c[2]:c[2] + 32], im)]
        assert cands, "output is not a crop of the texture bank"
        if len(cands) == 1:
            ti.append(cands[0][0]); ys.append(cands[0][1]); xs.append(cands[0][2])
    assert len(ti) > 0.9 * n
    for vals, k in ((ti, T), (ys, S), (xs, S)):
        cnt = np.bincount(vals, minlength=k)
        assert stats.chisquare(cnt).pvalue > 1e-4, (k, cnt)
