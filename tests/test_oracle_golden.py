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
