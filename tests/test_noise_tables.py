"""Host logic of the on-chip noise: the Walker alias tables built by imgtx_cc_noise_table_host must encode exactly (to the
2^-23 threshold rounding and the 1e-10 tail cut) the output-domain laws that the f64 oracle formulas define:
gaussian_noise  out = clip(x + floor(255*sigma*Z))            oracle/cifar_c.py gaussian_noise
shot_noise      out = uint8(clip(Poisson(x/255*c)/c, 0, 1)*255)  oracle/cifar_c.py shot_noise
"""
import numpy as np
import pytest
from scipy import stats

from imagetransformations_b200 import _lib as L
from imagetransformations_b200 import cifar_c as C
from oracle import cifar_c as CC


def decode(tab):
    t = tab.view(np.uint32)
    assert t[0] == 0x31544C41
    n_rows, n_ent, n_out = int(t[2]), int(t[3]), int(t[4])
    rows = t[8:8 + 2 * n_rows].reshape(n_rows, 2)
    ent = t[8 + 2 * n_rows:8 + 2 * n_rows + n_ent]
    out = t[8 + 2 * n_rows + n_ent:].view(np.uint8)[:n_out]
    return rows, ent, out


def row_law(rows, ent, r):
    base, W, lo = int(rows[r, 0]), int(rows[r, 1] >> 16), int(rows[r, 1] & 0xFFFF)
    e = ent[base:base + W]
    thr = (e >> 9).astype(np.float64) / 2 ** 23
    alias = (e & 511).astype(np.int64)
    assert alias.max() < W
    p = np.zeros(W)
    np.add.at(p, np.arange(W), thr / W)
    np.add.at(p, alias, (1 - thr) / W)
    return p, lo


@pytest.mark.parametrize("sigma", CC.PARAMS["gaussian_noise"])
def test_gaussian_table_is_floor_of_scaled_normal(sigma):
    rows, ent, _ = decode(C.noise_alias_table_host(L.CC_GAUSSIAN, sigma))
    p, lo = row_law(rows, ent, 0)
    dlo = lo - 65536 if lo >= 32768 else lo
    d = np.arange(dlo, dlo + len(p))
    s = 255.0 * sigma
    want = stats.norm.cdf((d + 1) / s) - stats.norm.cdf(d / s)
    assert abs(p.sum() - 1) < 1e-9
    assert np.abs(p - want).max() < 2e-7
    assert dlo <= -6.5 * s and dlo + len(p) >= 6.5 * s
    # against the oracle formula itself: x = 100, z on a fine grid of quantiles
    z = stats.norm.ppf((np.arange(200000) + 0.5) / 200000) * sigma
    o = CC.gaussian_noise(np.full((1, 1, 1, 200000), 128, np.uint8), z.reshape(1, 1, 1, -1)).ravel().astype(int) - 128
    emp = np.bincount(o - dlo, minlength=len(p))[:len(p)] / 200000
    inner = np.abs(d) < 120                                   # away from the clip at 0 / 255
    assert np.abs(emp - p)[inner].max() < 2e-5


@pytest.mark.parametrize("c", CC.PARAMS["shot_noise"])
def test_shot_table_is_quantised_poisson(c):
    rows, ent, out = decode(C.noise_alias_table_host(L.CC_SHOT, c))
    assert len(rows) == 256
    for x in (0, 1, 2, 7, 40, 128, 200, 255):
        p, ob = row_law(rows, ent, x)
        ov = out[ob:ob + len(p)].astype(int)
        assert (np.diff(ov) > 0).all()
        lam = x / 255.0 * c
        k = np.arange(0, int(lam + 12 * np.sqrt(lam + 1) + 30))
        pk = stats.poisson.pmf(k, lam) if lam > 0 else (k == 0).astype(float)
        o = CC.shot_noise(np.zeros((1, 1, 1, len(k)), np.uint8), k.reshape(1, 1, 1, -1), c).ravel().astype(int)
        want = np.zeros(256)
        np.add.at(want, o, pk)
        got = np.zeros(256)
        got[ov] = p
        assert abs(got.sum() - 1) < 1e-9
        assert np.abs(got - want).max() < 5e-7, (c, x)


def test_no_table_outside_range():
    assert C.noise_alias_table_host(L.CC_GAUSSIAN, 0.38) is None     # 255*sigma = 97 > 39: general path
    assert C.noise_alias_table_host(L.CC_CONTRAST, 0.5) is None


@pytest.mark.parametrize("c", list(CC.PARAMS["shot_noise"]) + [60, 12, 3])
def test_flat_shot_table_is_the_same_law_to_16_bit_thresholds(c):
    """IMGTX_CC_SHOT_FLAT: 256 rows of K = 2^lg slots, entry = threshold16 << 16 | alias byte << 8 | own byte.  Decoded, every
    row must be the quantised Poisson law of the oracle formula to 2^-17 per outcome (the thresholds carry 16 bits), padded
    slots must never be kept, and the table must be refused when a row needs more than 128 slots (c = 250)."""
    tab = C.noise_alias_table_host(L.CC_SHOT_FLAT, c)
    if c == 250:
        assert tab is None
        return
    t = tab.view(np.uint32)
    lg = int(t[5])
    K = 1 << lg
    assert t[0] == 0x32544C41 and t[2] == 256 and len(t) == 8 + 256 * K and 1 <= lg <= 7
    e = t[8:].reshape(256, K)
    for x in range(256):
        thr = (e[x] >> 16).astype(np.float64) / 65536
        ali, own = (e[x] >> 8) & 255, e[x] & 255
        got = np.zeros(256)
        np.add.at(got, own, thr / K)
        np.add.at(got, ali, (1 - thr) / K)
        lam = x / 255.0 * c
        k = np.arange(0, int(lam + 12 * np.sqrt(lam + 1) + 30))
        pk = stats.poisson.pmf(k, lam) if lam > 0 else (k == 0).astype(float)
        o = CC.shot_noise(np.zeros((1, 1, 1, len(k)), np.uint8), k.reshape(1, 1, 1, -1), c).ravel().astype(int)
        want = np.zeros(256)
        np.add.at(want, o, pk)
        assert abs(got.sum() - 1) < 1e-9
        assert np.abs(got - want).max() < 2.0 ** -17, (c, x)
        assert (got[want == 0] == 0).all(), (c, x)          # nothing outside the support (1e-10 tail cut aside)
