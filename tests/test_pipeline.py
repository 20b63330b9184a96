"""Pipelined ingest (krd_config.pipeline = 1): uploads of sample blocks overlap K0/K1 and the pair tiles that are already
complete.  The result must equal the ordinary path up to the rounding of different partial-sum splits, in every call order
a user can produce (repeat, invalidate, new masses, clustering calls, sharding, errors in a late block)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _data(n, seed=5, edges=301, placements=80):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(edges, seed=seed)
    mass = synth.sample_masses(tree, n, placements, seed=seed)
    return tree, mass


def _plan(tree, mass, bins, **kw):
    import gappa_b200
    plan = gappa_b200.KrdPlan(bins=bins, **kw)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    return plan


def _close(a, b, tol=1e-12):
    off = ~np.eye(len(a), dtype=bool)
    return np.max(np.abs(a[off] - b[off]) / np.maximum(b[off], 1e-300)) <= tol and np.array_equal(np.diag(a), np.diag(b))


@pytest.mark.parametrize("n,bins", [(600, 3), (1000, 4), (512, 1), (777, 7), (1300, 2)])
def test_pipelined_equals_ordinary(n, bins):
    tree, mass = _data(n)
    ref = _plan(tree, mass, bins)
    pip = _plan(tree, mass, bins, pipeline=1)
    for p in (1.0, 0.5, 3.0, 2.0, 1.7):
        want = ref.compute(p=p)
        got = pip.compute(p=p)                    # first call: phased (p = 2 takes the Gram path: all blocks first)
        assert _close(got, want, 1e-12 if p >= 1 else 1e-10), p
        assert np.array_equal(got, got.T)
        again = pip.compute(p=p)                  # F valid, children resident: one launch over all tiles
        assert _close(again, want, 1e-12 if p >= 1 else 1e-10)
        pip.invalidate_F()
        assert np.array_equal(pip.compute(p=p), again)    # rebuild + same schedule: bit-identical
        pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])   # next cohort: phased again
    ref.close(); pip.close()


def test_pipelined_handle_reuse_and_fallbacks():
    import torch
    from gappa_b200 import krd
    tree, mass = _data(700)
    pip = _plan(tree, mass, 2, pipeline=1)
    ref = _plan(tree, mass, 2)
    want = ref.compute(p=1.0)
    assert _close(pip.compute(p=1.0), want)
    # fewer samples than four blocks: the ordinary path inside the same handle
    from gappa_b200 import workloads
    small = workloads.head(mass, 300)
    pip.set_masses(small["off"], small["edge"], small["pos"], small["mass"])
    ref.set_masses(small["off"], small["edge"], small["pos"], small["mass"])
    assert np.array_equal(pip.compute(p=1.0), ref.compute(p=1.0))
    # back to a pipelined cohort, then straight into the clustering calls (they flatten the handle first)
    pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    ref.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    a1 = pip.kmeans(3, [1, 400, 650]); a2 = ref.kmeans(3, [1, 400, 650])
    assert np.array_equal(a1[0], a2[0]) and a1[3] == a2[3] and np.allclose(a1[1], a2[1], rtol=1e-12)
    pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    m1 = pip.squash(p=1.0); m2 = ref.squash(p=1.0)
    assert np.array_equal(m1["index_a"], m2["index_a"]) and np.array_equal(m1["index_b"], m2["index_b"])
    # sharded compute on a pipelined handle (flattened to the ordinary tile order)
    pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    host = krd.unpack_shards_host(700, [pip.compute_shard_host(r, 3, p=1.0) for r in range(3)])
    assert _close(host, want)
    # device output
    pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    d = torch.zeros((700, 700), dtype=torch.float64, device="cuda")
    pip.compute_device(d, p=1.0)
    torch.cuda.synchronize()
    assert _close(d.cpu().numpy(), want)
    # introspection of F before any compute
    pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    F1, l1 = pip.debug_F(); F2, l2 = ref.debug_F()
    assert np.array_equal(F1[:, :700], F2[:, :700]) and np.array_equal(l1, l2)
    pip.close(); ref.close()


def test_error_in_a_late_block_surfaces_at_compute():
    import gappa_b200
    tree, mass = _data(640)
    bad = mass["edge"].copy()
    bad[-1] = 10_000_000                      # last entry of the last sample: block 3
    pip = gappa_b200.KrdPlan(bins=2, pipeline=1)
    pip.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    pip.set_masses(mass["off"], bad, mass["pos"], mass["mass"])        # block 0 is fine
    with pytest.raises(gappa_b200.KrdError, match="outside the tree"):
        pip.compute(p=1.0)
    bad0 = mass["edge"].copy(); bad0[0] = -5
    with pytest.raises(gappa_b200.KrdError, match="outside the tree"):
        pip.set_masses(mass["off"], bad0, mass["pos"], mass["mass"])   # block 0: reported by set_masses itself
    pip.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    ref = _plan(tree, mass, 2)
    assert _close(pip.compute(p=1.0), ref.compute(p=1.0))
    pip.close(); ref.close()


def test_unsorted_late_block_falls_back_to_the_sorting_path():
    """Children after the first are checked without a host round trip (flags read once at the end): entries that are
    not in MassTree order in a late block make the call start over on the ordinary path, which sorts them."""
    import gappa_b200
    tree, mass = _data(700)
    off = mass["off"]
    edge, pos, m = mass["edge"].copy(), mass["pos"].copy(), mass["mass"].copy()
    for i in (650, 699):                       # reverse the entries of two samples of the last block
        a, b = off[i], off[i + 1]
        edge[a:b] = edge[a:b][::-1]; pos[a:b] = pos[a:b][::-1]; m[a:b] = m[a:b][::-1]
    ref = _plan(tree, mass, 2)
    want = ref.compute(p=1.0)
    pip = gappa_b200.KrdPlan(bins=2, pipeline=1)
    pip.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    for p in (1.0, 2.0):
        pip.set_masses(off, edge, pos, m)
        got = pip.compute(p=p)
        assert _close(got, ref.compute(p=p), 1e-11), p
    pip.set_masses(off, mass["edge"], mass["pos"], mass["mass"])     # and a clean cohort afterwards: phased again
    assert _close(pip.compute(p=1.0), want)
    pip.close(); ref.close()


def test_pipelined_blocks_without_masses():
    """Ragged cohorts: a whole late block (and a few scattered samples) without a single mass entry."""
    import gappa_b200
    tree, mass = _data(900)
    off = mass["off"].copy()
    keep = np.ones(len(mass["edge"]), bool)
    for i in list(range(640, 768)) + [3, 130, 899]:          # block 5 entirely, plus singles in the first and last block
        keep[off[i]:off[i + 1]] = False
    counts = np.diff(mass["off"]); counts[[*range(640, 768), 3, 130, 899]] = 0
    off2 = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    sub = dict(off=off2, edge=mass["edge"][keep], pos=mass["pos"][keep], mass=mass["mass"][keep])
    ref = _plan(tree, sub, 3)
    pip = _plan(tree, sub, 3, pipeline=1)
    for p in (1.0, 0.5):
        want, got = ref.compute(p=p), pip.compute(p=p)
        assert _close(got, want, 1e-12 if p >= 1 else 1e-10), p
        assert np.all(got[640:768, 640:768] == 0.0)          # empty samples are at distance 0 from each other
        pip.set_masses(sub["off"], sub["edge"], sub["pos"], sub["mass"])
    ref.close(); pip.close()
