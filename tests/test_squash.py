"""Squash clustering on the GPU against genesis' SquashClustering::run (fixture from the
compiled reference, tests/golden/make_golden_squash.py): same merge sequence, distances within 1e-10, same
Newick string."""
import os

import numpy as np
import pytest

import gappa_b200
from gappa_b200 import krd

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "squash.npz")


def _plan(z, bins):
    plan = gappa_b200.KrdPlan(bins=bins)
    plan.set_tree(z["parent"], z["child"], z["bl"], int(z["root"]))
    plan.set_masses(z["off"], z["edge"], z["pos"], z["mass"])
    return plan


def test_tree_string_matches_reference_format():
    # host-only: SquashClustering::tree_string (squash_clustering.cpp:251-274) from the reference's own mergers
    z = np.load(GOLD)
    for key in ("b0_p1.0", "b4_p1.0", "b0_p0.5"):
        idx, dst = z["idx_" + key], z["dst_" + key]
        m = np.zeros(len(idx), krd.MERGER_DTYPE)
        m["index_a"], m["index_b"], m["distance_a"], m["distance_b"] = idx[:, 0], idx[:, 1], dst[:, 0], dst[:, 1]
        assert krd.squash_tree_string([str(i) for i in range(len(idx) + 1)], m) == str(z["tree_" + key])


@pytest.mark.gpu
@pytest.mark.parametrize("bins,p", [(0, 1.0), (0, 2.0), (4, 1.0), (0, 0.5)])
def test_squash_matches_reference(bins, p):
    z = np.load(GOLD)
    plan = _plan(z, bins)
    m = plan.squash(p=p)
    key = f"b{bins}_p{p}"
    idx, dst = z["idx_" + key], z["dst_" + key]
    assert np.array_equal(m["index_a"], idx[:, 0]) and np.array_equal(m["index_b"], idx[:, 1])   # same merge sequence
    got = np.stack([m["distance_a"], m["distance_b"]], 1)
    assert np.max(np.abs(got - dst) / np.maximum(dst, 1e-300)) <= 1e-10
    n = len(idx) + 1
    if p >= 1:
        assert krd.squash_tree_string([str(i) for i in range(n)], m) == str(z["tree_" + key])
    # the handle is still usable for the plain matrix afterwards, and squash is repeatable
    d1 = plan.compute(p=p)
    assert d1.shape == (n, n) and np.array_equal(d1, d1.T)
    m2 = plan.squash(p=p)
    assert np.array_equal(m2, m)
    plan.close()


@pytest.mark.gpu
def test_squash_needs_two_samples_and_valid_exponent():
    z = np.load(GOLD)
    plan = _plan(z, 0)
    with pytest.raises(krd.KrdError, match="Invalid exponent value p"):
        plan.squash(p=0.0)
    plan.set_masses(z["off"][:2], z["edge"][: z["off"][1]], z["pos"][: z["off"][1]], z["mass"][: z["off"][1]])
    with pytest.raises(krd.KrdError, match="at least 2 samples"):
        plan.squash()
    plan.close()
