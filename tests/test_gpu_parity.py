"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on
the same seeded inputs and against the committed golden fixtures of the compiled reference.
Tolerance: 1e-10 relative (BASELINE.json north_star, FP64)."""
import numpy as np
import pytest

from conftest import load_golden, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def gb():
    import gappa_b200
    return gappa_b200


def dense_F_numpy(tree, mass_canon, seg_off, slot_of):
    """Plain restatement of F for small cases: F[s][i] = subtree mass below + own masses at slots >= t."""
    E = len(tree["bl"])
    n = len(mass_canon["off"]) - 1
    S = seg_off[-1]
    F = np.zeros((S, n))
    kids = [[] for _ in range(E + 1)]
    for e in range(E):
        kids[tree["parent"][e]].append(e)
    import oracle
    order = np.zeros(E, np.int32)
    oracle.port().okrd_postorder(E, np.ascontiguousarray(tree["parent"], np.int32),
                                 np.ascontiguousarray(tree["child"], np.int32), int(tree["root"]), order)
    for i in range(n):
        lo, hi = mass_canon["off"][i], mass_canon["off"][i + 1]
        per_edge = {}
        for m in range(lo, hi):
            per_edge.setdefault(int(mass_canon["edge"][m]), []).append(m)
        top = np.zeros(E)
        for e in order:
            acc = sum(top[c] for c in kids[tree["child"][e]])
            k = seg_off[e + 1] - seg_off[e] - 1
            ms = per_edge.get(int(e), [])
            mp = len(ms) - 1
            for t in range(k, -1, -1):
                while mp >= 0 and slot_of(ms[mp]) >= t:
                    acc += mass_canon["mass"][ms[mp]]
                    mp -= 1
                F[seg_off[e] + t, i] = acc
            top[e] = acc
    return F


def test_golden_test_ab(gb):
    tree, mass, z = load_golden("test_ab.npz")
    for p in (1.0, 2.0, 0.5):
        got = gb.earth_movers_distance(tree, mass, p=p)
        assert rel_err(got, z[f"m_p{p}"]) <= TOL
        assert got[0, 0] == 0.0 and got[0, 1] == got[1, 0]
    assert np.float32(gb.earth_movers_distance(tree, mass, p=1.0)[0, 1]) == np.float32(2.8458333)


@pytest.mark.parametrize("p", [1.0, 2.0, 0.5])
def test_golden_cfg1_exact_mode(gb, p):
    tree, mass, z = load_golden("cfg1.npz")
    plan = gb.KrdPlan(bins=0)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    got = plan.compute(p=p)
    assert rel_err(got, z[f"m_p{p}"]) <= TOL
    assert np.array_equal(got, got.T) and np.all(np.diag(got) == 0)
    # --normalize: krd.cpp:144-150 divides by the diameter
    dia = float(z["diameter"])
    assert rel_err(plan.compute(p=p, divisor=dia), z[f"m_p{p}"] / dia) <= TOL
    st = plan.stats()
    assert st["n_samples"] == 30 and st["n_edges"] == 597 and st["n_segments"] > 597
    plan.close()


def test_golden_cfg1_point_mass(gb):
    tree, mass, z = load_golden("cfg1_pm.npz")
    assert rel_err(gb.earth_movers_distance(tree, mass, p=1.0), z["m_p1.0"]) <= TOL


@pytest.mark.parametrize("bins", [0, 1, 4])
@pytest.mark.parametrize("p", [1.0, 2.0, 0.5, 1.7])
def test_golden_synth_small(gb, bins, p):
    tree, mass, z = load_golden("synth_small.npz")
    assert rel_err(gb.earth_movers_distance(tree, mass, p=p, bins=bins), z[f"m_b{bins}_p{p}"]) <= TOL


@pytest.mark.parametrize("p2_mode", [1, 2])   # 1 = direct on FP64 CUDA cores, 2 = Gram on the FP64 tensor cores (DMMA) + fix-ups
@pytest.mark.parametrize("fixture,bins,key", [("cfg1.npz", 0, "m_p2.0"), ("test_ab.npz", 0, "m_p2.0"),
                                              ("synth_small.npz", 0, "m_b0_p2.0"), ("synth_small.npz", 4, "m_b4_p2.0")])
def test_p2_both_paths_vs_reference(gb, p2_mode, fixture, bins, key):
    tree, mass, z = load_golden(fixture)
    plan = gb.KrdPlan(bins=bins, p2_mode=p2_mode)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    got = plan.compute(p=2.0)
    assert rel_err(got, z[key]) <= TOL
    assert np.array_equal(got, got.T) and np.all(np.diag(got) == 0.0)
    assert np.array_equal(got, plan.compute(p=2.0))       # run-to-run bit identity
    plan.close()


def test_p2_gram_cancellation_is_repaired(gb, oracle_mod):
    """Near-duplicate samples: |a|^2 + |b|^2 - 2ab cancels (distance ~1e-4 of the norms); the flagged cells
    must be recomputed directly so that the Gram path stays inside the 1e-10 bar."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(300, seed=41)
    base = synth.sample_masses(tree, 6, 800, seed=41)
    off, edge, pos, mass = [0], [], [], []
    rng = np.random.default_rng(41)
    for i in range(6):
        lo, hi = base["off"][i], base["off"][i + 1]
        for rep in range(30):                      # 30 copies of each sample with masses jittered by 1e-4
            m = base["mass"][lo:hi] * (1.0 + 1e-4 * rng.standard_normal(hi - lo) * (rep > 0))
            edge.append(base["edge"][lo:hi]); pos.append(base["pos"][lo:hi]); mass.append(m / m.sum())
            off.append(off[-1] + hi - lo)
    masses = dict(off=np.array(off, np.int64), edge=np.concatenate(edge), pos=np.concatenate(pos), mass=np.concatenate(mass))
    want = oracle_mod.port_matrix(tree, masses, p=2.0)
    res = {}
    for mode in (1, 2):
        plan = gb.KrdPlan(p2_mode=mode)
        plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
        res[mode] = plan.compute(p=2.0)
        st = plan.stats()
        plan.close()
        assert rel_err(res[mode], want) <= TOL, mode
    # every within-group pair was flagged and repaired — except in sample 0's own group: sample 0 is the
    # reference column, so those norms are as small as the distances and nothing cancels
    assert st["gram_fixups"] == 5 * 30 * 29 // 2
    assert want[0, 1] < 1e-2 * want[0, 31]              # the test really is in the cancellation regime


@pytest.mark.parametrize("builder", [0, 1])   # 0 = tiled single-write builder, 1 = general scatter + level scan
def test_F_matrix_matches_restatement_binned(gb, oracle_mod, builder):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(61, seed=5)
    mass = synth.sample_masses(tree, 9, 120, seed=5)
    B = 3
    plan = gb.KrdPlan(bins=B, builder=builder)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    F, ln = plan.debug_F()
    E = len(tree["bl"])
    seg_off = np.arange(E + 1) * (B + 1)
    L = oracle_mod.port()
    slot = [L.okrd_bin_index(float(mass["pos"][m]), float(tree["bl"][mass["edge"][m]]), B) for m in range(len(mass["pos"]))]
    want = dense_F_numpy(tree, mass, seg_off, lambda m: slot[m])
    assert F.shape[0] == E * (B + 1)
    assert np.max(np.abs(F[:, :9] - want)) <= (1e-15 if builder else 4e-15)   # tiled: M = D + T rounds differently
    assert np.all(F[:, 9:] == 0.0)
    assert abs(ln.sum() - tree["bl"].sum()) < 1e-12
    # every sample's total mass arrives at the root: sum over root child edges of F[seg0]
    root_edges = [e for e in range(E) if tree["parent"][e] == tree["root"]]
    tot = sum(F[seg_off[e], :9] for e in root_edges)
    assert np.max(np.abs(tot - 1.0)) < 1e-12
    plan.close()


@pytest.mark.parametrize("n,E,P,bins", [(70, 150, 300, 4), (200, 97, 50, 1), (65, 40, 500, 190), (33, 20, 100, 300), (64, 1, 7, 2)])
def test_tiled_builder_equals_general_builder(gb, n, E, P, bins):
    """Binned mode has two F builders; they must agree (different summation trees -> a few ulps)."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(E, seed=23)
    mass = synth.sample_masses(tree, n, P, seed=23)
    res = []
    for builder in (0, 1):
        plan = gb.KrdPlan(bins=bins, builder=builder)
        plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
        F, ln = plan.debug_F()
        m1 = plan.compute(p=1.0)
        plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])   # buffers are reused
        assert np.array_equal(plan.compute(p=1.0), m1)
        res.append((F, ln, m1, plan.compute(p=0.5)))
        plan.close()
    assert np.array_equal(res[0][1], res[1][1])
    assert np.max(np.abs(res[0][0] - res[1][0])) <= 1e-14
    assert rel_err(res[0][2], res[1][2]) <= 1e-12
    # p = 1/2: small trees tie (all of two samples' mass below one root child), and there rounding-level differences
    # between ANY two FP64 evaluations are amplified by sqrt; each builder is held to the port, else to the binary128 truth
    # within the a-priori tie bound (conftest.assert_matches_reference_p_below_one, tests/test_p_below_one.py)
    from conftest import assert_matches_reference_p_below_one
    idx = list(range(0, n, max(1, n // 24)))[:24]
    for k in (0, 1):
        assert_matches_reference_p_below_one(tree, mass, idx, 0.5, bins, res[k][3][np.ix_(idx, idx)])


@pytest.mark.parametrize("p", [3.0, 4.0, 5.0, 6.0, 7.0, 8.0, 11.0, 16.0, 17.0, 2.5, 0.3, 0.75, 1.0000001])
def test_every_exponent_mode_vs_oracle(gb, oracle_mod, p):
    """One exponent per inner-operator variant (P3, P4, INT with every bit pattern class, GENERIC above and below 1) on a
    matrix with diagonal AND off-diagonal tiles (150 samples = 2 tile rows, the second one ragged)."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(61, seed=21)
    mass = synth.sample_masses(tree, 150, 60, seed=21)
    got = gb.earth_movers_distance(tree, mass, p=p, bins=3)
    want = oracle_mod.port_matrix(tree, mass, p=p, bins=3)
    off = ~np.eye(150, dtype=bool)
    assert rel_err(got[off], want[off]) <= TOL
    assert np.array_equal(got, got.T) and np.all(np.diag(got) == 0)


@pytest.mark.parametrize("n,E,P,bins", [(2, 3, 5, 0), (1, 9, 4, 0), (129, 33, 40, 2), (257, 17, 25, 1), (130, 5, 0, 0)])
def test_ragged_shapes_vs_oracle(gb, oracle_mod, n, E, P, bins):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(E, seed=11)
    mass = synth.sample_masses(tree, n, P, seed=11) if P > 0 else dict(
        off=np.zeros(n + 1, np.int64), edge=np.zeros(0, np.int32), pos=np.zeros(0), mass=np.zeros(0))
    for p in (1.0, 3.0):
        got = gb.earth_movers_distance(tree, mass, p=p, bins=bins)
        want = oracle_mod.port_matrix(tree, mass, p=p, bins=bins)
        assert got.shape == (n, n)
        assert np.max(np.abs(got - want)) <= TOL * max(1.0, np.max(want))
        if np.max(want) > 0:
            assert rel_err(got[want > 1e-6], want[want > 1e-6]) <= TOL


def test_unsorted_and_duplicate_input(gb, oracle_mod):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(41, seed=2)
    mass = synth.sample_masses(tree, 20, 200, seed=2)
    rng = np.random.default_rng(1)
    # duplicates: copy some positions inside each sample, then shuffle each sample
    e, x, m = mass["edge"].copy(), mass["pos"].copy(), mass["mass"].copy()
    for i in range(20):
        s = slice(mass["off"][i], mass["off"][i + 1])
        idx = np.arange(s.start, s.stop)
        x[idx[1::7]] = x[idx[0::7][: len(idx[1::7])]]
        e[idx[1::7]] = e[idx[0::7][: len(idx[1::7])]]
        perm = rng.permutation(idx)
        e[s], x[s], m[s] = e[perm], x[perm], m[perm]
    shuf = dict(off=mass["off"], edge=e, pos=x, mass=m)
    for bins in (0, 2):
        got = gb.earth_movers_distance(tree, shuf, p=1.0, bins=bins)
        want = oracle_mod.port_matrix(tree, shuf, p=1.0, bins=bins)
        assert rel_err(got, want) <= TOL


def test_split_invariance_and_determinism(gb):
    from gappa_b200 import synth
    tree = synth.tree_with_edges(301, seed=9)
    mass = synth.sample_masses(tree, 140, 500, seed=9)
    res = []
    for split in (1, 3, 0, 0):
        plan = gb.KrdPlan(bins=4, split=split)
        plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
        res.append(plan.compute(p=1.0))
        plan.close()
    assert rel_err(res[0], res[1]) <= 1e-13 and rel_err(res[0], res[2]) <= 1e-13
    assert np.array_equal(res[2], res[3])  # bit-identical run to run


def test_errors_through_the_abi(gb):
    tree, mass, _ = load_golden("test_ab.npz")
    plan = gb.KrdPlan()
    with pytest.raises(gb.KrdError) as ei:
        plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    assert ei.value.status == gb.krd.KRD_ERR_STATE
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    for bad in (0.0, -1.0, float("nan")):
        with pytest.raises(gb.KrdError, match="Invalid exponent value p") as ei:
            plan.compute(p=bad)
        assert ei.value.status == gb.krd.KRD_ERR_INVALID_EXPONENT
    bad_edge = mass["edge"].copy(); bad_edge[0] = 99
    with pytest.raises(gb.KrdError):
        plan.set_masses(mass["off"], bad_edge, mass["pos"], mass["mass"])
    with pytest.raises(gb.KrdError) as ei:   # two parents for one node
        plan.set_tree(np.array([0, 0, 1], np.int32), np.array([1, 2, 2], np.int32), np.ones(3), 0)
    assert ei.value.status == gb.krd.KRD_ERR_INCOMPATIBLE_TREES
    plan.close()


def test_full_size_properties_cfg2_shape(gb):
    """BASELINE cfg2 shape (N=1000, E=9999, bins=4) with fewer placements: properties that need no
    oracle — symmetry, zero diagonal, triangle inequality (p=1 KR is a metric), scaling linearity in
    branch length, and an oracle spot check on a 24-sample subset."""
    import oracle
    from gappa_b200 import synth
    tree = synth.tree_with_edges(9999, seed=42)
    mass = synth.sample_masses(tree, 1000, 2000, seed=42)
    plan = gb.KrdPlan(bins=4)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    D = plan.compute(p=1.0)
    assert np.array_equal(D, D.T) and np.all(np.diag(D) == 0) and np.all(D[np.triu_indices(1000, 1)] > 0)
    rng = np.random.default_rng(0)
    i, j, k = rng.integers(0, 1000, (3, 20000))
    assert np.all(D[i, j] <= D[i, k] + D[k, j] + 1e-12)
    sub = dict(off=mass["off"][:25] - mass["off"][0], edge=mass["edge"][: mass["off"][24]],
               pos=mass["pos"][: mass["off"][24]], mass=mass["mass"][: mass["off"][24]])
    want = oracle.port_matrix(tree, sub, 1.0, bins=4)
    assert rel_err(D[:24, :24], want) <= TOL
    # scaling all branch lengths and positions by 2 doubles every distance (p = 1)
    plan2 = gb.KrdPlan(bins=4)
    plan2.set_tree(tree["parent"], tree["child"], tree["bl"] * 2.0, tree["root"])
    plan2.set_masses(mass["off"], mass["edge"], mass["pos"] * 2.0, mass["mass"])
    assert rel_err(plan2.compute(p=1.0), 2.0 * D) <= 1e-12
    plan.close(); plan2.close()


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_compute_equals_single(gb, world):
    """All ranks' shares computed one after another on this GPU (no inter-rank waiting), concatenated
    like an all-gather, unpacked: must equal the single-GPU matrix bit for bit."""
    import torch
    from gappa_b200 import synth
    tree = synth.tree_with_edges(201, seed=4)
    mass = synth.sample_masses(tree, 300, 300, seed=4)
    plan = gb.KrdPlan(bins=2)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    single = torch.zeros((300, 300), dtype=torch.float64, device="cuda")
    plan.compute_device(single, p=1.0)
    shard = plan.shard_doubles(world)
    gathered = torch.empty(shard * world, dtype=torch.float64, device="cuda")
    for r in range(world):
        plan.compute_shard(gathered[r * shard:(r + 1) * shard], r, world, p=1.0)
    out = torch.full((300, 300), -1.0, dtype=torch.float64, device="cuda")
    plan.unpack_shards(gathered, out, world)
    torch.cuda.synchronize()
    assert torch.equal(out, single)
    # host twin (krd_compute_shard_host + krd_unpack_shards_host, what `gappa-b200 --gpus G` runs): same bits
    from gappa_b200 import krd
    for p in (1.0, 2.0):
        host = krd.unpack_shards_host(300, [plan.compute_shard_host(r, world, p=p) for r in range(world)])
        assert np.array_equal(host, plan.compute(p=p))
    plan.close()


def test_exponents_beyond_the_table_operator_use_library_pow(gb, oracle_mod):
    """The reference accepts ANY p > 0 (emd.cpp:186-190).  The table-driven general-p operator covers p <= 30000; beyond
    that the pair kernel runs a library-pow body (slow, correct).  Large legal exponents stay inside the tolerance."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(99, seed=3)
    masses = synth.sample_masses(tree, 12, 50, seed=3)
    plan = gb.KrdPlan(bins=0)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
    for p in (250.5, 37.25, 1e5, 3.5e6):
        got = plan.compute(p=p)
        want = oracle_mod.port_matrix(tree, masses, p=p, bins=0)
        assert np.all(np.isfinite(got))
        assert rel_err(got, want) <= 1e-10, p
    plan.close()


def test_underflowing_terms_are_flushed_to_zero(gb, oracle_mod):
    """Near-duplicate samples and a large general exponent: every |d|^p underflows, the reference returns 0 (or a
    denormal-sized value); the table operator must not return its clamped minimum 2^-1021 per term."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(51, seed=13)
    base = synth.sample_masses(tree, 3, 120, seed=13)
    off, edge, pos, mass = [0], [], [], []
    rng = np.random.default_rng(13)
    for i in range(3):
        lo, hi = base["off"][i], base["off"][i + 1]
        for rep in range(4):                       # copies jittered by 1e-6: |F_i - F_j| ~ 1e-6, ^250 underflows
            m = base["mass"][lo:hi] * (1.0 + 1e-6 * rng.standard_normal(hi - lo) * (rep > 0))
            edge.append(base["edge"][lo:hi]); pos.append(base["pos"][lo:hi]); mass.append(m / m.sum())
            off.append(off[-1] + hi - lo)
    masses = dict(off=np.array(off, np.int64), edge=np.concatenate(edge), pos=np.concatenate(pos), mass=np.concatenate(mass))
    for p in (250.5, 120.25):
        got = gb.earth_movers_distance(tree, masses, p=p)
        want = oracle_mod.port_matrix(tree, masses, p=p)
        within = np.zeros((12, 12), bool)
        for g in range(3):
            within[4 * g:4 * g + 4, 4 * g:4 * g + 4] = True
        np.fill_diagonal(within, False)
        assert np.all(want[within] == 0.0)                       # every term underflows in the reference's std::pow
        assert np.all(got[within] == 0.0), (p, got[within].max())
        assert rel_err(got[~within], want[~within]) <= 1e-10, p


@pytest.mark.parametrize("scale", [1.0, 1e-150, 1e120])
@pytest.mark.parametrize("p", [0.3, 1.7])
def test_general_exponent_identical_samples_and_extreme_magnitudes(gb, oracle_mod, p, scale):
    """The pair kernel's general-p operator (kr_pow_acc16) takes |d| apart as integers: identical samples (d = 0 on every
    segment) must give exactly 0 for p below and above 1, and unnormalised masses far from 1 (exponent fields near both
    ends of the table arithmetic) must stay within the bar."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(47, seed=5)
    base = synth.sample_masses(tree, 70, 50, seed=5)
    n0 = 70
    off = np.concatenate([base["off"], base["off"][-1] + base["off"][1:11]])            # samples 70..79 = copies of 0..9
    m10 = int(base["off"][10])
    masses = dict(off=off, edge=np.concatenate([base["edge"], base["edge"][:m10]]),
                  pos=np.concatenate([base["pos"], base["pos"][:m10]]),
                  mass=np.concatenate([base["mass"], base["mass"][:m10]]) * scale)
    got = gb.earth_movers_distance(tree, masses, p=p, bins=2)
    want = oracle_mod.port_matrix(tree, masses, p=p, bins=2)
    for k in range(10):
        assert got[k, n0 + k] == 0.0 and got[n0 + k, k] == 0.0 and want[k, n0 + k] == 0.0
    off_diag = ~np.eye(80, dtype=bool)
    assert np.all(np.isfinite(got))
    assert rel_err(got[off_diag], want[off_diag]) <= TOL, (p, scale)


@pytest.mark.parametrize("p2_mode", [1, 2])
def test_positions_outside_the_branch_keep_signed_lengths(gb, oracle_mod, p2_mode):
    """Exact mode: a jplace distal/proximal length outside the branch gives pos < 0 or pos > bl and the reference keeps the
    signed segment length (emd.cpp:156,166).  The direct p = 2 path must not scale by sqrt(len) then (NaN for every pair)."""
    from gappa_b200 import synth
    tree = synth.tree_with_edges(41, seed=8)
    mass = synth.sample_masses(tree, 20, 40, seed=8)
    pos = mass["pos"].copy()
    bl = tree["bl"][mass["edge"]]
    pos[5] = bl[5] * 1.25            # beyond the branch
    pos[77] = bl[77] * 1.5
    m2 = dict(off=mass["off"], edge=mass["edge"], pos=pos, mass=mass["mass"])
    m2 = oracle_mod.canonicalize(m2)
    want1 = oracle_mod.port_matrix(tree, m2, p=1.0)
    want2 = oracle_mod.port_matrix(tree, m2, p=2.0)
    plan = gb.KrdPlan(p2_mode=p2_mode)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(m2["off"], m2["edge"], m2["pos"], m2["mass"])
    got1, got2 = plan.compute(p=1.0), plan.compute(p=2.0)
    plan.close()
    assert np.all(np.isfinite(got1)) and rel_err(got1, want1) <= TOL
    fin = np.isfinite(want2)                      # (a negative total would be NaN in the reference as well)
    assert np.array_equal(np.isfinite(got2), fin)
    assert rel_err(got2[fin], want2[fin]) <= TOL


@pytest.mark.parametrize("entry", ["device", "shard", "host"])
def test_gram_fixup_overflow_is_handled_by_every_entry_point(gb, oracle_mod, entry):
    """More near-duplicate pairs than the Gram path's fix-up list holds (2^20): the launch is repeated on the direct path
    inside the library for EVERY entry point, not only krd_compute (round-1 advice)."""
    import torch
    from gappa_b200 import synth
    n = 1500                                              # 1500 * 1499 / 2 = 1,124,250 > 2^20 cancelling pairs
    tree = synth.tree_with_edges(31, seed=5)
    base = synth.sample_masses(tree, 2, 60, seed=5)
    rng = np.random.default_rng(5)
    lo, hi = base["off"][1], base["off"][2]
    off, edge, pos, mass = [0, int(base["off"][1])], [base["edge"][:lo]], [base["pos"][:lo]], [base["mass"][:lo]]
    for rep in range(n - 1):                              # sample 0 is the Gram reference column; the other n-1 are near-copies of sample 1
        m = base["mass"][lo:hi] * (1.0 + 1e-5 * rng.standard_normal(hi - lo))
        edge.append(base["edge"][lo:hi]); pos.append(base["pos"][lo:hi]); mass.append(m / m.sum())
        off.append(off[-1] + hi - lo)
    masses = dict(off=np.array(off, np.int64), edge=np.concatenate(edge), pos=np.concatenate(pos), mass=np.concatenate(mass))
    direct = gb.KrdPlan(p2_mode=1)
    direct.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    direct.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
    want = direct.compute(p=2.0)
    direct.close()
    sub = list(range(0, n, 97))
    idx = np.concatenate([masses["off"][s] + np.arange(masses["off"][s + 1] - masses["off"][s]) for s in sub])
    sub_m = dict(off=np.concatenate([[0], np.cumsum([masses["off"][s + 1] - masses["off"][s] for s in sub])]).astype(np.int64),
                 edge=masses["edge"][idx], pos=masses["pos"][idx], mass=masses["mass"][idx])
    assert rel_err(want[np.ix_(sub, sub)], oracle_mod.port_matrix(tree, sub_m, p=2.0)) <= TOL
    plan = gb.KrdPlan(p2_mode=2)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
    if entry == "host":
        got = plan.compute(p=2.0)
    elif entry == "device":
        d = torch.zeros((n, n), dtype=torch.float64, device="cuda")
        plan.compute_device(d, p=2.0)
        torch.cuda.synchronize()
        got = d.cpu().numpy()
    else:
        world = 1                                         # (one rank's share must hold > 2^20 cancelling pairs)
        shard = plan.shard_doubles(world)
        g = torch.empty(shard * world, dtype=torch.float64, device="cuda")
        for r in range(world):
            plan.compute_shard(g[r * shard:(r + 1) * shard], r, world, p=2.0)
        d = torch.zeros((n, n), dtype=torch.float64, device="cuda")
        plan.unpack_shards(g, d, world)
        torch.cuda.synchronize()
        got = d.cpu().numpy()
    assert plan.stats()["gram_fixups"] == -1              # the list did overflow ...
    plan.close()
    assert rel_err(got, want) <= TOL                      # ... and the result is the direct path's
