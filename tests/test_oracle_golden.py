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
