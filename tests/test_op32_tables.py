"""Host side of imgtx_cc_op32_mma_u8 (no GPU): the table builder imagetransformations_b200/cifar_c.py:op32_mma_tables is
decoded back — fragment bytes -> Q16 weights per (tile row, contraction slot), quad table -> source pixels — and evaluated
with numpy exactly as the kernel does ((((bias + S_lo) >> 8) + S_hi) >> 8); the result must be within 1 LSB of
oracle/cifar_c.py:zoom_blur, flat images must come out exact, every weight row must sum to 2^16 (or 2^16 - 1 with the larger
bias when a weight is 1.0), and the quads of a half-step must fall into different shared-memory banks."""
import numpy as np
import pytest

from oracle import cifar_c as CC


def _decode(tab, steps):
    from imagetransformations_b200 import cifar_c as C
    assert tab[0] == C.OP32_MAGIC and tab[1] == steps
    off_meta, off_bias, off_qtab, off_lo, off_hi, words = (int(v) for v in tab[2:8])
    assert words == len(tab) and off_hi == off_lo + steps * 128
    meta = tab[off_meta:off_bias].reshape(64, 4)
    bias = tab[off_bias:off_qtab].view(np.uint32).reshape(64, 8)
    qtab = tab[off_qtab:off_lo].reshape(64, 3, 4, 2)
    lo = tab[off_lo:off_hi].view(np.uint32).reshape(steps, 32, 4)
    hi = tab[off_hi:].view(np.uint32).reshape(64, 3, 32, 4)
    tiles = []
    for t in range(64):
        ns, ab, p0 = int(meta[t, 0]), int(meta[t, 1]), int(meta[t, 2])
        m = np.arange(16)
        pix = p0 + ((m % 8) >> 1) * 32 + ((m % 8) & 1) * 2 + (m >> 3)           # as the kernel's epilogue addresses them
        rb = np.concatenate([bias[t] & 0xFFFF, bias[t] >> 16]).astype(np.int64)
        W = np.zeros((16, ns * 32), np.int64)
        src = np.zeros(ns * 32, np.int64)
        for s in range(ns):
            for slot in range(8):
                boff = int(qtab[t, s, slot & 3, slot >> 2])
                assert boff % 4 == 0
                row, col = divmod(boff // 4, 10)
                assert 0 <= row < 32 and 0 <= col < 8
                src[s * 32 + slot * 4:s * 32 + slot * 4 + 4] = (row * 8 + col) * 4 + np.arange(4)
            for lane in range(32):
                gid, tig = lane >> 2, lane & 3
                for reg in range(4):
                    r, kb = gid + (reg & 1) * 8, (reg >> 1) * 16 + tig * 4
                    for i in range(4):
                        b0 = (int(lo[ab + s, lane, reg]) >> (8 * i)) & 255
                        b1 = (int(hi[t, s, lane, reg]) >> (8 * i)) & 255
                        W[r, s * 32 + kb + i] = b0 | (b1 << 8)
        tiles.append((pix, rb, W, src, qtab[t, :ns]))
    return tiles


def _apply(tiles, x):
    n = x.shape[0]
    xf = x.reshape(n, 1024, 3).astype(np.int64)
    out = np.zeros((n, 1024, 3), np.int64)
    for pix, rb, W, src, _ in tiles:
        B = xf[:, src, :]
        s_lo = np.einsum("mk,nkc->nmc", W & 255, B)
        s_hi = np.einsum("mk,nkc->nmc", W >> 8, B)
        out[:, pix, :] = (((s_lo + rb[None, :, None]) >> 8) + s_hi) >> 8
    assert out.min() >= 0 and out.max() <= 255
    return out.reshape(n, 32, 32, 3).astype(np.uint8)


@pytest.mark.parametrize("sev", [1, 3, 5])
def test_zoom_tables_reproduce_the_oracle_within_one_lsb(sev):
    from imagetransformations_b200 import cifar_c as C
    rp, ci, wt = CC.zoom_operator(32, 32, sev)
    r = C.op32_mma_tables(rp, ci, wt.astype(np.float64))
    assert r is not None
    tab, steps = r
    tiles = _decode(tab, steps)
    assert sorted(int(p) for t in tiles for p in t[0]) == list(range(1024)), "the 64 tiles cover every pixel once"
    for pix, rb, W, src, qt in tiles:
        sums = W.sum(1)
        assert ((sums == 65536) | ((sums == 65535) & (rb == 256))).all()
        # bank of a B-fragment load: 4 * image + word offset (mod 32); the four quads of a half-step should differ mod 4
        bad = sum(len({int(v) // 4 % 4 for v in qt[s, :, h]}) < 4 for s in range(qt.shape[0]) for h in range(2))
        assert bad <= 2, "at most a couple of half-steps with two quads in one bank"
    rng = np.random.default_rng(sev)
    x = rng.integers(0, 256, (12, 32, 32, 3), dtype=np.uint8)
    x[0] = 77
    x[1, :16] = 200
    x[2] = 255
    got = _apply(tiles, x)
    want = CC.corrupt("zoom_blur", sev, x)
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    assert d.max() <= 1 and (d[3:] > 0).mean() < 0.01, (int(d.max()), float((d[3:] > 0).mean()))
    assert (got[0] == 77).all() and (got[2] == 255).all(), "flat images are exact"


def test_tables_refuse_what_the_kernel_cannot_run():
    from imagetransformations_b200 import cifar_c as C
    # a dense row needs 256 quads: more than 3 steps
    indptr = np.concatenate([[0], np.full(1024, 1024)]).astype(np.int64)
    assert C.op32_mma_tables(indptr, np.arange(1024), np.full(1024, 1.0 / 1024)) is None
    # negative weights are outside the unsigned limbs
    ip = np.arange(1025, dtype=np.int64)
    assert C.op32_mma_tables(ip, np.arange(1024), -np.ones(1024)) is None
