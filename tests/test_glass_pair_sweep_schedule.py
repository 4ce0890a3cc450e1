"""The lane schedule of glass_blur's paired swap sweep (csrc/cifar_c_team2.cuh, OpTGlass::sweep_pair), restated lane for lane
in numpy and checked against the sequential swaps of oracle/cifar_c.py's glass_blur (rows h = 31..2, columns w = 31..2, each
pixel swapped with the one at (h + dy, w + dx), dx, dy in {-1, 0}).

What the device code relies on, and what this test pins without a GPU:
  * 15 lanes per image: lane l takes row l during steps [2l, 2l + 30) and row l + 15 during [2l + 30, 2l + 60), 88 steps in all;
  * within one step the active lanes touch pairwise disjoint map entries (all loads of a step are issued before its stores);
  * a lane that has not started or has finished swaps inside the dummy region only, whatever its registers hold;
  * the draws: column ci of row r uses bit ci of the row's two 32-bit words; X is kept pre-shifted by one (its bit 1 is the
    byte offset of one entry), Y rotated left by six (its bit 6 is the byte offset of one map row);
  * the result equals the sequential sweep, so it equals the map the exact mode applies to the pixels themselves.
The GPU side of the same statement is tests/test_gpu_fast_mode.py::test_glass_paired_sweep_makes_the_maps_of_the_exact_mode."""
import numpy as np
import pytest


def sequential_map(bits_x, bits_y):
    m = np.arange(1024)
    for r in range(30):
        h = 31 - r
        for ci in range(30):
            w = 31 - ci
            dx, dy = -int((int(bits_x[r]) >> ci) & 1), -int((int(bits_y[r]) >> ci) & 1)
            a, b = h * 32 + w, (h + dy) * 32 + (w + dx)
            m[a], m[b] = m[b], m[a]
    return m


def rotl(v, k):
    return ((v << k) | (v >> (32 - k))) & 0xFFFFFFFF


def rotr(v, k):
    return ((v >> k) | (v << (32 - k))) & 0xFFFFFFFF


def paired_sweep_one_image(bits_x, bits_y):
    """16 lanes (15 active) of one half-warp; byte-addressed memory: the map at 0..2047, the dummy region above it."""
    MAP_BYTES, DUMMY_LO, DUMMY_ENTRIES = 2048, 2048 + 16, 128
    mem = np.zeros((DUMMY_LO + 2 * DUMMY_ENTRIES) // 2, dtype=np.int64)
    mem[:1024] = np.arange(1024)
    dm = DUMMY_LO + 2 * 100
    L = 16
    a, X, Y = [dm] * L, [0] * L, [0] * L
    first = [2 * l if l < 15 else 1000 for l in range(L)]
    row1 = [((31 - l) * 32 + 31) * 2 for l in range(L)]
    row2 = [r - 15 * 64 for r in row1]
    x30 = [int(bits_x[l]) & 0x3FFFFFFF if l < 15 else 0 for l in range(L)], [int(bits_x[l + 15]) & 0x3FFFFFFF if l < 15 else 0 for l in range(L)]
    y30 = [int(bits_y[l]) & 0x3FFFFFFF if l < 15 else 0 for l in range(L)], [int(bits_y[l + 15]) & 0x3FFFFFFF if l < 15 else 0 for l in range(L)]
    for t in range(88):
        for l in range(L):
            if t < 30 and t == first[l]:
                a[l], X[l], Y[l] = row1[l], x30[0][l] << 1, rotl(y30[0][l], 6)
            if 30 <= t < 60 and t == first[l] + 30:
                a[l], X[l], Y[l] = row2[l], x30[1][l] << 1, rotl(y30[1][l], 6)
            if t >= 60 and t == first[l] + 60:
                a[l] = dm
        b = [a[l] - (X[l] & 2) - (Y[l] & 64) for l in range(L)]
        active = [l for l in range(L) if a[l] < MAP_BYTES]
        cells = [c for l in active for c in {a[l], b[l]}]
        assert len(cells) == len(set(cells)), f"step {t}: two lanes touch the same entry"
        assert all(0 <= c < MAP_BYTES for c in cells)
        idle = [l for l in range(L) if l not in active]
        assert all(DUMMY_LO <= c < DUMMY_LO + 2 * DUMMY_ENTRIES for l in idle for c in (a[l], b[l])), f"step {t}: idle lane outside the dummy region"
        assert len(active) <= 15
        ta, tb = [mem[a[l] // 2] for l in range(L)], [mem[b[l] // 2] for l in range(L)]
        for l in range(L):
            mem[a[l] // 2] = tb[l]
        for l in range(L):
            mem[b[l] // 2] = ta[l]
        for l in range(L):
            a[l] -= 2
            X[l] >>= 1
            Y[l] = rotr(Y[l], 1)
    return mem[:1024]


@pytest.mark.parametrize("seed", range(6))
def test_paired_sweep_equals_sequential_swaps(seed):
    rng = np.random.default_rng(seed)
    m_seq = m_par = None
    bx = rng.integers(0, 2 ** 32, 30, dtype=np.uint64)
    by = rng.integers(0, 2 ** 32, 30, dtype=np.uint64)
    if seed == 4:
        bx[:] = by[:] = 0xFFFFFFFF       # every swap goes up and left
    if seed == 5:
        bx[:] = by[:] = 0                # every swap is with itself: the identity
    m_seq, m_par = sequential_map(bx, by), paired_sweep_one_image(bx, by)
    assert np.array_equal(m_seq, m_par)
    assert sorted(m_par.tolist()) == list(range(1024))
    if seed == 5:
        assert np.array_equal(m_par, np.arange(1024))


def test_sequential_map_is_the_oracle_permutation():
    """The index-map form of the sweep is the oracle's pixel sweep (oracle/cifar_c.py glass_swaps): applying the map as a
    gather equals swapping the pixels themselves with the same deltas."""
    from oracle import cifar_c as CC
    rng = np.random.default_rng(3)
    x = rng.integers(0, 256, (1, 32, 32, 3), dtype=np.uint8)
    bx = rng.integers(0, 2 ** 30, 30, dtype=np.uint64)
    by = rng.integers(0, 2 ** 30, 30, dtype=np.uint64)
    deltas = np.zeros((1, 1, 30, 30, 2), np.int64)
    for r in range(30):
        for ci in range(30):
            deltas[0, 0, r, ci] = (-int((int(bx[r]) >> ci) & 1), -int((int(by[r]) >> ci) & 1))
    want = CC.glass_swaps(x, deltas)[0]
    got = x[0].reshape(1024, 3)[paired_sweep_one_image(bx, by)].reshape(32, 32, 3)
    assert np.array_equal(got, want)
