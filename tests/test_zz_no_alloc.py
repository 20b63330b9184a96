"""krd_matrix_from_F_noalloc (include/krd_b200.h): the allocation-free pair reduction added
(no counterpart in the reference).  Called through the C ABI on DEVICE buffers, twice in a row with DIFFERENT
cohorts (the shared static scratch keeps the first call's contents: the second result must not see them), and checked
against the existing entry point krd_compute on the same samples within the suite's 1e-10 tolerance.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _cohort(n, edges, placements, bins, seed):
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(edges, seed=seed)
    mass = synth.sample_masses(tree, n, placements, seed=seed)
    plan = gappa_b200.KrdPlan(bins=bins)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    F, ln = plan.debug_F()                      # n_segments x ld, n_segments (host)
    want = {p: plan.compute(p=p) for p in (1.0, 2.0, 0.5, 1.7)}
    st = plan.stats()
    plan.close()
    return F, ln, st, want


def test_no_alloc_entry_point_matches_krd_compute_on_two_calls_in_a_row():
    import torch
    from gappa_b200 import krd
    cohorts = [_cohort(120, 201, 150, 2, seed=3), _cohort(57, 61, 90, 0, seed=4)]      # binned, then exact and smaller
    # every device buffer exists BEFORE the calls under test (the entry point itself allocates nothing)
    bufs = []
    for F, ln, st, _ in cohorts:
        n = int(st["n_samples"])
        bufs.append((torch.from_numpy(F).cuda(), torch.from_numpy(ln).cuda(), torch.full((n, n), -1.0, dtype=torch.float64, device="cuda")))
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    # (the first call makes the driver load the module that holds the static scratch array: its 8 MiB appear then, once)
    krd.matrix_from_F_noalloc(bufs[1][0], bufs[1][1], cohorts[1][0].shape[0], int(cohorts[1][2]["n_samples"]), cohorts[1][0].shape[1],
                              bufs[1][2], p=1.0, stream=stream)
    torch.cuda.synchronize()
    for p in (1.0, 2.0, 0.5, 1.7):
        free_before = torch.cuda.mem_get_info()[0]
        for (F, ln, st, want), (dF, dl, dout) in zip(cohorts, bufs):              # two calls in a row on one stream
            krd.matrix_from_F_noalloc(dF, dl, F.shape[0], int(st["n_samples"]), F.shape[1], dout, p=p, stream=stream)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before                        # nothing was allocated on the device
        for (F, ln, st, want), (dF, dl, dout) in zip(cohorts, bufs):
            got = dout.cpu().numpy()
            off = ~np.eye(got.shape[0], dtype=bool)
            assert np.array_equal(got, got.T) and np.all(np.diag(got) == 0.0)
            assert np.max(np.abs(got[off] - want[p][off]) / want[p][off]) <= TOL, p
    # divisor (--normalize) and a repeat of the first cohort after the second: same bits as its first result
    F, ln, st, want = cohorts[0]
    dF, dl, dout = bufs[0]
    krd.matrix_from_F_noalloc(dF, dl, F.shape[0], int(st["n_samples"]), F.shape[1], dout, p=1.0, divisor=2.5, stream=stream)
    torch.cuda.synchronize()
    got = dout.cpu().numpy()
    off = ~np.eye(got.shape[0], dtype=bool)
    assert np.max(np.abs(got[off] * 2.5 - want[1.0][off]) / want[1.0][off]) <= TOL


def test_no_alloc_entry_point_rejects_what_does_not_fit():
    import torch
    from gappa_b200 import krd
    n = 129
    dF = torch.zeros((4, 256), dtype=torch.float64, device="cuda"); dl = torch.ones(4, dtype=torch.float64, device="cuda")
    dout = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    with pytest.raises(krd.KrdError, match="at most 128 samples"):
        krd.matrix_from_F_noalloc(dF, dl, 4, n, 256, dout)
    with pytest.raises(krd.KrdError, match="Has to be > 0.0"):
        krd.matrix_from_F_noalloc(dF, dl, 4, 100, 256, dout, p=0.0)
    with pytest.raises(krd.KrdError, match="bad sizes"):
        krd.matrix_from_F_noalloc(dF, dl, 4, 100, 64, dout)
