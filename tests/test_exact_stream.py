"""Exact mode (bins = 0 — what `gappa analyze krd` runs, src/commands/analyze/krd.cpp:136-140): the single-write builder
against the legacy scatter + scan builder, and the STREAMED segment axis (krd_config.max_f_mb: F does not have to fit in
HBM) against the resident form and against the port, through the C ABI on a B200."""
import numpy as np
import pytest

from conftest import assert_matches_reference_p_below_one, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _plan(tree, mass, **kw):
    import gappa_b200
    plan = gappa_b200.KrdPlan(bins=0, **kw)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    return plan


@pytest.mark.parametrize("n,E,P", [(70, 150, 300), (200, 97, 50), (130, 5, 40), (33, 20, 0), (257, 61, 120)])
def test_exact_builder_equals_legacy_builder(n, E, P):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(E, seed=31)
    mass = synth.sample_masses(tree, n, P, seed=31)
    a, b = _plan(tree, mass), _plan(tree, mass, builder=1)
    Fa, la = a.debug_F()
    Fb, lb = b.debug_F()
    assert np.array_equal(la, lb)
    assert Fa.shape == Fb.shape and np.max(np.abs(Fa - Fb), initial=0.0) <= 1e-14
    for p in (1.0, 2.0, 3.0):
        assert rel_err(a.compute(p=p), b.compute(p=p)) <= 1e-12, p
    a.close(); b.close()


@pytest.mark.parametrize("n,E,P,mb", [(150, 61, 200, 1), (300, 201, 150, 2), (129, 33, 40, 1), (40, 999, 600, 1)])
def test_streamed_equals_resident_and_port(n, E, P, mb, oracle_mod):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(E, seed=n)
    mass = synth.sample_masses(tree, n, P, seed=n)
    res, st = _plan(tree, mass), _plan(tree, mass, max_f_mb=mb)
    s_res, s_st = res.stats(), st.stats()
    assert s_res["n_segments"] == s_st["n_segments"]
    assert s_st["f_bytes"] <= mb << 20 < s_res["f_bytes"]            # really streamed: several slabs
    sub = list(range(0, n, max(1, n // 20)))[:20]
    o, e, ps, m = [0], [], [], []
    for smp in sub:
        lo, hi = mass["off"][smp], mass["off"][smp + 1]
        e.append(mass["edge"][lo:hi]); ps.append(mass["pos"][lo:hi]); m.append(mass["mass"][lo:hi]); o.append(o[-1] + hi - lo)
    subm = dict(off=np.array(o, np.int64), edge=np.concatenate(e), pos=np.concatenate(ps), mass=np.concatenate(m))
    for p in (1.0, 2.0, 0.5, 3.0, 7.0, 1.7):
        Dr, Ds = res.compute(p=p), st.compute(p=p)
        assert np.array_equal(Ds, Ds.T) and np.all(np.diag(Ds) == 0)
        assert rel_err(Ds, Dr) <= (1e-12 if p >= 1 else 1e-10), p
        assert np.array_equal(st.compute(p=p), Ds)                      # run-to-run bit identity
        if p < 1:     # tiny trees tie: port first, else the binary128 truth within the tie bound
            assert_matches_reference_p_below_one(tree, mass, sub, p, 0, Ds[np.ix_(sub, sub)])
            continue
        want = oracle_mod.port_matrix(tree, subm, p=p)
        assert rel_err(Ds[np.ix_(sub, sub)], want) <= TOL, p
    assert st.stats()["ms_slabs"] > 0
    res.close(); st.close()


def test_streamed_sharded_unsorted_and_signed_lengths(oracle_mod):
    import torch
    from gappa_b200 import synth
    n = 200
    tree = synth.tree_with_edges(77, seed=3)
    mass = synth.sample_masses(tree, n, 90, seed=3)
    pos = mass["pos"].copy()
    pos[11] = tree["bl"][mass["edge"][11]] * 1.3                         # beyond the branch: a negative segment length
    rng = np.random.default_rng(3)
    # shuffle every sample's entries and duplicate one entry per sample (split mass): any order is accepted
    e2, p2, m2, off = [], [], [], [0]
    for i in range(n):
        lo, hi = mass["off"][i], mass["off"][i + 1]
        perm = rng.permutation(hi - lo)
        ee, pp, mm = mass["edge"][lo:hi][perm], pos[lo:hi][perm], mass["mass"][lo:hi][perm].copy()
        mm[0] *= 0.5
        e2.append(np.append(ee, ee[0])); p2.append(np.append(pp, pp[0])); m2.append(np.append(mm, mm[0])); off.append(off[-1] + hi - lo + 1)
    m = dict(off=np.array(off, np.int64), edge=np.concatenate(e2), pos=np.concatenate(p2), mass=np.concatenate(m2))
    st = _plan(tree, m, max_f_mb=1)
    want = {p: oracle_mod.port_matrix(tree, m, p=p) for p in (1.0, 2.0)}
    for p in (1.0, 2.0):
        got = st.compute(p=p)
        fin = np.isfinite(want[p])
        assert np.array_equal(np.isfinite(got), fin)
        assert rel_err(got[fin], want[p][fin]) <= TOL, p
    # sharded: all ranks' shares one after another, unpacked = the single result, bit for bit
    world = 3
    shard = st.shard_doubles(world)
    g = torch.empty(shard * world, dtype=torch.float64, device="cuda")
    for r in range(world):
        st.compute_shard(g[r * shard:(r + 1) * shard], r, world, p=1.0)
    out = torch.full((n, n), -1.0, dtype=torch.float64, device="cuda")
    st.unpack_shards(g, out, world)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), st.compute(p=1.0))
    st.close()


def test_streamed_handle_refuses_what_needs_resident_F_with_sizes():
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(61, seed=4)
    mass = synth.sample_masses(tree, 150, 200, seed=4)
    st = _plan(tree, mass, max_f_mb=1)
    with pytest.raises(gappa_b200.KrdError, match=r"resident \(\d+\.\d+ GB for \d+ segments x 150 samples\)"):
        st.squash(p=1.0)
    with pytest.raises(gappa_b200.KrdError, match="resident"):
        st.debug_F()
    D = st.compute(p=1.0)                                               # the handle still computes
    assert np.all(np.isfinite(D))
    st.close()


def test_survey_probe_shape_1000_samples_30k_exact_positions(oracle_mod):
    """SURVEY's CPU probe shape scaled to a cohort the dense formulation could not hold in round 1: 1,000 samples x 9,999 edges
    x 30k exact positions -> S ~ 3e7 segments, F = 245 GB.  Streamed on one B200; sub-blocks against the port."""
    from gappa_b200 import synth
    n, E, P = 1000, 9_999, 30_000
    tree = synth.tree_with_edges(E, seed=42)
    mass = synth.sample_masses(tree, n, P, seed=42)
    st = _plan(tree, mass)
    s = st.stats()
    assert s["n_segments"] > 2.9e7 and s["f_bytes"] < s["n_segments"] * s["ld_samples"] * 8     # streamed by the auto rule
    D = st.compute(p=1.0)
    st.close()
    assert np.array_equal(D, D.T) and np.all(np.diag(D) == 0) and np.all(np.isfinite(D))
    rng = np.random.default_rng(2)
    for idx in (list(range(8)), list(range(n - 5, n)) + [0], sorted(rng.choice(n, 8, replace=False).tolist())):
        o, e, ps, m = [0], [], [], []
        for smp in idx:
            lo, hi = mass["off"][smp], mass["off"][smp + 1]
            e.append(mass["edge"][lo:hi]); ps.append(mass["pos"][lo:hi]); m.append(mass["mass"][lo:hi]); o.append(o[-1] + hi - lo)
        sub = dict(off=np.array(o, np.int64), edge=np.concatenate(e), pos=np.concatenate(ps), mass=np.concatenate(m))
        want = oracle_mod.port_matrix(tree, sub, p=1.0)
        off = ~np.eye(len(idx), dtype=bool)
        assert np.max(np.abs(D[np.ix_(idx, idx)][off] - want[off]) / want[off]) <= TOL
