"""Fused AugMix (fall_2025/AugMix.py:45-62) against oracle/np_chains.py:augmix on random chains: every op kind in every
position (equalize before / after geometric stages, several equalizes in one chain, chains that end early), image shapes on
the vector path (width and pixel count multiples of 4), on the scalar path (odd sizes) and at 224 x 224 where one image fills
a CTA's shared memory.  f32 results must be bit-identical (tolerance 0, as tests/golden_cases.py uses for the goldens)."""
import numpy as np
import pytest
import torch

from oracle import np_chains as CH

pytestmark = pytest.mark.gpu


def _draws(rs, n, width, depth, force=None):
    ws = rs.dirichlet([1.0] * width, size=n)
    m = rs.beta(1.0, 1.0, size=n)
    ops = np.full((n, width, depth), -1, np.int32)
    signs = np.zeros((n, width, depth), np.int32)
    for i in range(n):
        for a in range(width):
            d = rs.randint(1, depth + 1)
            for b in range(d):
                ops[i, a, b] = rs.randint(0, 8)
                signs[i, a, b] = rs.choice([-1, 1]) if ops[i, a, b] == 0 else 0
    if force is not None:
        for i, chain in enumerate(force):
            ops[i, 0, :] = -1
            ops[i, 0, :len(chain)] = chain
            signs[i, 0] = np.where(ops[i, 0] == 0, 1, 0)
    return ws, m, ops, signs


# rotate 0, posterize 1, shear_x 2, shear_y 3, translate_x 4, translate_y 5, equalize 6, solarize 7 (AugMix.py:30-37)
FORCED = [[6], [6, 6, 6], [0, 6], [6, 0], [0, 6, 4], [4, 5, 6], [1, 6, 7], [7, 0, 6], [6, 2, 6], [3, 3, 3], [1, 7, 1], [5, 6, 0]]


@pytest.mark.parametrize("hw,n", [((32, 32), 48), ((37, 53), 16), ((64, 48), 16), ((30, 36), 12), ((224, 224), 6)])
def test_fused_augmix_equals_the_oracle_on_random_chains(cuda, hw, n):
    from imagetransformations_b200 import AugMix
    rs = np.random.RandomState(hw[0] * 131 + hw[1])
    h, w = hw
    x = rs.uniform(0.0, 1.0, (n, 3, h, w)).astype(np.float32)
    x[0] = np.round(x[0] * 4) / 4            # few distinct levels: equalize with nearly empty histograms
    x[1, :, : h // 2] = 0.5                  # a flat half
    x[2] = x[2] * 2.5 - 0.7                  # out of range floats: to_pil_image wraps modulo 256
    ws, m, ops, signs = _draws(rs, n, 3, 3, FORCED[:n])
    got = AugMix.augmix_batch(torch.from_numpy(x).cuda(), 3, ws, m, ops, signs).cpu().numpy()
    for i in range(n):
        want = CH.augmix(x[i], 3, ws[i], m[i], ops[i], signs[i])
        assert np.array_equal(got[i], want), (hw, i, ops[i].tolist(), float(np.abs(got[i] - want).max()))


def test_fused_augmix_width_and_depth_other_than_three(cuda):
    from imagetransformations_b200 import AugMix
    rs = np.random.RandomState(9)
    for width, depth in [(1, 1), (2, 4), (5, 2)]:
        n = 6
        x = rs.uniform(0.0, 1.0, (n, 3, 32, 32)).astype(np.float32)
        ws, m, ops, signs = _draws(rs, n, width, depth)
        got = AugMix.augmix_batch(torch.from_numpy(x).cuda(), 2, ws, m, ops, signs).cpu().numpy()
        for i in range(n):
            want = CH.augmix(x[i], 2, ws[i], m[i], ops[i], signs[i])
            assert np.array_equal(got[i], want), (width, depth, i, ops[i].tolist())


def test_fused_augmix_when_a_point_op_moves_zero(cuda):
    """solarize with threshold 0 maps 0 -> 255: the fill value of a geometric stage then differs from stage to stage and
    from comp[c][0] (the kernel's general fill path, csrc/augmix.cu zero_is_fixed == false)."""
    import ctypes as C
    from imagetransformations_b200 import AugMix, _lib as L, batch as B
    from oracle import np_ops as O
    f32 = np.float32
    rs = np.random.RandomState(77)
    for (h, w) in [(32, 32), (224, 224), (37, 53)]:
        n = 5
        x = rs.uniform(0.0, 1.0, (n, 3, h, w)).astype(np.float32)
        ws, m, ops, signs = _draws(rs, n, 3, 3, [[7, 0, 6], [0, 7, 4], [7, 6, 2], [5, 7, 7], [0, 6, 7]])
        recs = np.zeros((n, 3, 3), AugMix.AUG_OP_DTYPE)
        for i in range(n):
            for a in range(3):
                for b in range(3):
                    recs[i, a, b] = AugMix.compile_op(int(ops[i, a, b]), 3, int(signs[i, a, b]), w, h)
                    if ops[i, a, b] == 7:
                        recs[i, a, b]["param"] = 0
        wts = np.concatenate([ws.astype(f32), m.astype(f32)[:, None], (1.0 - m).astype(f32)[:, None]], axis=1)
        xd = torch.from_numpy(x).cuda()
        out = torch.empty_like(xd)
        rd = torch.from_numpy(recs.view(np.uint8).reshape(-1).copy()).cuda()
        wd = torch.from_numpy(np.ascontiguousarray(wts)).cuda()
        L.check(L.load().imgtx_augmix_f32(xd.data_ptr(), out.data_ptr(), n, h, w, 3, 3, rd.data_ptr(), wd.data_ptr(), B._stream()))
        got = out.cpu().numpy()
        for i in range(n):
            mix = np.zeros_like(x[i])
            for a in range(3):
                u = O.to_pil_image(x[i][None])
                aug = x[i]
                for b in range(3):
                    op = int(ops[i, a, b])
                    if op < 0:
                        break
                    u = O.solarize(u, 0) if op == 7 else CH.augmix_op(u, op, 3, int(signs[i, a, b]))
                    aug = O.to_tensor(u)[0]
                mix = (mix + (f32(ws[i, a]) * aug).astype(f32)).astype(f32)
            want = ((f32(1 - m[i]) * x[i]).astype(f32) + (f32(m[i]) * mix).astype(f32)).astype(f32)
            assert np.array_equal(got[i], want), ((h, w), i, ops[i].tolist(), float(np.abs(got[i] - want).max()))
