"""Fused TransformationPool chain (pipenline/cifar_image_transformations.py:141-152, 203-226): one launch for a
population.  With the reference's draws injected it equals the goldens frozen from the live reference and the oracle chain
bit for bit (f32 tensors: tolerance 0, as tests/golden_cases.py uses for the op-by-op path); with on-chip draws it is
deterministic per seed and follows the law of the reference's numpy draws."""
import numpy as np
import pytest
import torch

import golden_cases as G
from oracle import np_chains as CH
from oracle import np_ops as O

pytestmark = pytest.mark.gpu


def _M():
    from imagetransformations_b200 import cifar_image_transformations as M
    return M


def _ops(M, chains, draw_idx=None):
    max_ops = max(1, max(len(c) for c in chains))
    ops = np.zeros((len(chains), max_ops), M.POOL_OP_DTYPE)
    ops["kind"] = -1
    for i, c in enumerate(chains):
        rec = M.encode_chain(c)
        if draw_idx is not None:
            for k, (name, _) in enumerate(c):
                if name in ("gaussian_noise", "impulse_noise", "shot_noise"):
                    rec[k]["draw"] = draw_idx[i]
        ops[i, :len(c)] = rec
    return ops


def test_fused_chain_equals_the_live_reference_goldens(cuda):
    M = _M()
    g = G.golden("pool")
    n_checked = 0
    for case in g.cases:
        if not case["fn"].startswith("pool.individual"):
            continue
        a = case["args"]
        x = g.arr(case["in"])[a["index"]]
        want = g.arr(case["out"])
        if case["fn"] == "pool.individual_base":
            s = a["severity"]
            chain = [("gaussian_noise", {"severity": s}), ("impulse_noise", {"severity": s}), ("shot_noise", {"severity": s}),
                     ("defocus_blur", {"severity": s})]
            d = g.draws(case)
            draws = {"gauss": d["noise"][None], "impulse": d["mask"][None], "shot": d["counts"][None]}
            ops = _ops(M, [chain], [0])
        else:
            chain = [(nm, pr) for nm, pr in a["chain"]]
            draws, ops = None, _ops(M, [chain])
        got = M.pool_chain_batch(torch.from_numpy(x).cuda(), ops, draws=draws, shared_input=True).cpu().numpy()
        assert np.array_equal(got.reshape(want.shape), want), case["fn"]
        n_checked += 1
    assert n_checked >= 6


def _oracle_population(x, chains, rs, shared):
    """Oracle results and the draws the reference would make, op by op (shot noise draws depend on the image it finds)."""
    outs, gd, md, sd = [], [], [], []
    for i, chain in enumerate(chains):
        tn = x if shared else x[i]
        u = O.tensor_to_pil(tn[None], CH.CIFAR10_MEAN, CH.CIFAR10_STD)
        dg = np.zeros(u.shape[1:], np.float64)
        dm = np.zeros(u.shape[1:3], np.float64)
        ds = np.zeros(u.shape[1:], np.int64)
        for name, params in chain:
            d = {}
            if name == "gaussian_noise":
                dg = rs.normal(0, CH.POOL_GAUSS[params["severity"] - 1] * 255, u.shape[1:])
                d = {"noise": dg}
            elif name == "impulse_noise":
                dm = rs.random_sample(u.shape[1:3])
                d = {"mask": dm}
            elif name == "shot_noise":
                ds = rs.poisson(O.shot_noise_scaled(u, CH.POOL_SHOT[params["severity"] - 1]))[0]
                d = {"counts": ds}
            u = CH.pool_op(u, name, params, d)
        outs.append(O.to_tensor_normalize(u, CH.CIFAR10_MEAN, CH.CIFAR10_STD)[0])
        gd.append(dg), md.append(dm), sd.append(ds)
    return np.stack(outs), {"gauss": np.stack(gd), "impulse": np.stack(md), "shot": np.stack(sd)}


POOL = [("gaussian_noise", "severity"), ("impulse_noise", "severity"), ("shot_noise", "severity"), ("defocus_blur", "severity"),
        ("enhance_contrast", "factor"), ("enhance_brightness", "factor"), ("enhance_sharpness", "factor"),
        ("enhance_color", "factor"), ("motion_blur", "size"), ("histogram_equalization", None)]


def _random_chain(rs):
    """The base chain in a random order (each noise at most once: one draw block per type and individual) + 0-4 others."""
    chain = []
    for j in rs.permutation(4)[: rs.randint(1, 5)]:
        chain.append((POOL[j][0], {"severity": int(rs.randint(1, 6))}))
    for _ in range(rs.randint(0, 5)):
        name, key = POOL[rs.randint(4, 10)]
        p = {} if key is None else ({"size": int(rs.choice([5, 7, 9, 11]))} if key == "size"
                                    else {"factor": float(np.float32(rs.uniform(0.5, 2.5)))})
        chain.insert(rs.randint(0, len(chain) + 1), (name, p))
    return chain


@pytest.mark.parametrize("hw,shared", [((32, 32), True), ((32, 32), False), ((37, 53), False), ((64, 48), True)])
def test_fused_chain_equals_the_oracle_chain(cuda, hw, shared):
    M = _M()
    rs = np.random.RandomState(hw[0] * 7 + shared)
    n = 24
    u8 = rs.randint(0, 256, (1 if shared else n, hw[0], hw[1], 3)).astype(np.uint8)
    x = O.to_tensor_normalize(u8, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    x = x[0] if shared else x
    chains = [_random_chain(rs) for _ in range(n)]
    chains[0] = [("gaussian_noise", {"severity": 3}), ("impulse_noise", {"severity": 3}), ("shot_noise", {"severity": 3}),
                 ("defocus_blur", {"severity": 3})]                      # the GA's base chain (:207-212)
    chains[1] = []                                                       # tensor_to_pil -> pil_to_tensor alone
    want, draws = _oracle_population(x, chains, rs, shared)
    ops = _ops(M, chains, list(range(n)))
    got = M.pool_chain_batch(torch.from_numpy(np.ascontiguousarray(x)).cuda(), ops, draws=draws, shared_input=shared).cpu().numpy()
    for i in range(n):
        assert np.array_equal(got[i], want[i]), (i, chains[i], int((got[i] != want[i]).sum()))


def test_apply_population_on_chip_draws(cuda):
    M = _M()
    import random
    rs = np.random.RandomState(3)
    u8 = rs.randint(0, 256, (1, 32, 32, 3)).astype(np.uint8)
    x = torch.from_numpy(O.to_tensor_normalize(u8, CH.CIFAR10_MEAN, CH.CIFAR10_STD)[0]).cuda()
    base = [("gaussian_noise", {"severity": 3}), ("impulse_noise", {"severity": 3}), ("shot_noise", {"severity": 3}),
            ("defocus_blur", {"severity": 3})]
    pop = [M.Individual(base + [("enhance_contrast", {})] * (i % 3)) for i in range(20)]
    random.seed(5)
    a = M.apply_population(pop, x, seed=1).cpu().numpy()
    random.seed(5)
    b = M.apply_population(pop, x, seed=1).cpu().numpy()
    assert np.array_equal(a, b) and a.shape == (20, 3, 32, 32)
    assert torch.equal(pop[7].transformed_image.cpu(), torch.from_numpy(a[7]))
    random.seed(5)
    c = M.apply_population(pop, x, seed=2).cpu().numpy()
    assert (a != c).mean() > 0.5, "another seed, other noise"
    assert (a[0] != a[3]).mean() > 0.5, "individuals with the same chain draw their own noise"


def test_on_chip_noise_follows_the_reference_law(cuda):
    """Noise part of the base chain, severity 3, on a fixed image: the uint8 result of 600 individuals with on-chip draws
    against 600 oracle individuals with numpy draws — per-byte mean and variance agree (z < 6 over the image)."""
    M = _M()
    rs = np.random.RandomState(11)
    u8 = np.zeros((1, 32, 32, 3), np.uint8)
    u8[0] = np.linspace(0, 255, 32 * 32 * 3).reshape(32, 32, 3).astype(np.uint8)
    x = O.to_tensor_normalize(u8, CH.CIFAR10_MEAN, CH.CIFAR10_STD)[0]
    chain = [("gaussian_noise", {"severity": 1}), ("impulse_noise", {"severity": 3}), ("shot_noise", {"severity": 1})]
    n = 600
    want, _ = _oracle_population(x, [chain] * n, rs, True)
    ops = _ops(M, [chain] * n)
    got = M.pool_chain_batch(torch.from_numpy(x).cuda(), ops, seed=21, shared_input=True).cpu().numpy()
    a, b = want.reshape(n, -1).astype(np.float64), got.reshape(n, -1).astype(np.float64)
    se = np.sqrt(a.var(0, ddof=1) / n + b.var(0, ddof=1) / n) + 1e-12
    z = (a.mean(0) - b.mean(0)) / se
    assert np.abs(z).max() < 6 and abs(z.mean()) < 0.2, (np.abs(z).max(), z.mean())
    ratio = (b.var(0, ddof=1) + 1e-9) / (a.var(0, ddof=1) + 1e-9)
    assert 0.93 < np.median(ratio) < 1.07, np.median(ratio)


def test_fused_chain_refuses_images_that_do_not_fit(cuda):
    M = _M()
    x = torch.zeros((3, 224, 224), dtype=torch.float32, device="cuda")
    ops = _ops(M, [[("defocus_blur", {"severity": 1})]])
    with pytest.raises(Exception, match="does not fit in shared memory"):
        M.pool_chain_batch(x, ops, shared_input=True)
