"""CPU: pin the numpy oracle to the golden vectors frozen from the live reference (oracle/make_golden.py)."""
import numpy as np
import pytest

import golden_cases as G


@pytest.mark.parametrize("fi", G.all_cases(), ids=G.case_id)
def test_oracle_matches_reference_golden(fi):
    f, i = fi
    g = G.golden(f)
    c = g.cases[i]
    got = G.oracle_eval(g, c)
    G.compare(c["fn"], got, g.arr(c["out"]))


def test_golden_versions_recorded():
    for f in G.FILES:
        v = G.golden(f).versions
        assert {"pillow", "opencv", "torch", "torchvision", "numpy"} <= set(v)


ENTROPY_TOL = 1e-13   # f64 entropies: the oracle sums -p ln p with numpy, the reference with scipy.special.entr


def test_oracle_entropy_golden():
    """compute_shannon_entropy of TENT_Pipelines.py / Initial_Experiments.py, frozen from the live
    reference: counts bit-exact, entropy to 1e-13."""
    from oracle import np_ops as O
    g = G.golden("entropy")
    assert len(g.cases) == 3
    for c in g.cases:
        x = g.arr(c["in"])
        tent = c["fn"].startswith("TENT_Pipelines")
        e, h = O.shannon_entropy(x, (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)) if tent else O.shannon_entropy(x)
        assert np.array_equal(h, g.draws(c)["counts"]), c["fn"]
        assert np.max(np.abs(e - g.arr(c["out"]))) <= ENTROPY_TOL, c["fn"]


def _affine_matrix(angle, scale):
    """torchvision.transforms.functional._get_inverse_affine_matrix(center 0, translate 0, shear 0) restated: the
    inverse of rotation-by-angle times scale, in Python doubles."""
    import math
    rot = math.radians(angle)
    a, b, c, d = math.cos(rot), -math.sin(rot), math.sin(rot), math.cos(rot)
    return [d / scale, -b / scale, 0.0, -c / scale, a / scale, 0.0]


def test_oracle_tensor_tta_chain_golden():
    """TestTimeAugmentation()(x) of post_hoc_aaryaa.py frozen under recorded seeds; the oracle's chain with
    the recorded draws reproduces it (bit-exact when perspective did not fire)."""
    from oracle import np_chains as CH
    n_exact = 0
    for x, want, views, draws in G.posthoc_chain_cases():
        for j, d in enumerate(draws):
            pc = CH.perspective_coeffs(*d["perspective"]) if d["perspective"] is not None else None
            got = CH.tensor_tta_chain(x[j // views], d["flip"], _affine_matrix(-d["angle"], 1.0),
                                      _affine_matrix(0.0, d["scale"]), d["brightness"], pc)
            G.check_chain_view(got, want[j], pc is not None)
            n_exact += int(np.array_equal(got, want[j]))
    assert n_exact >= 40   # 60 views in all; about half have no perspective stage and must be identical


def test_oracle_mean_views_golden():
    from oracle import np_chains as CH
    g = G.golden("posthoc")
    for c in g.cases:
        if c["fn"] in ("post_hoc_aaryaa.mean_logits", "TTA_transforms.mean_probs"):
            assert np.array_equal(CH.mean_views(g.arr(c["in"])), g.arr(c["out"])), c["fn"]
