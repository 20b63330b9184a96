"""Needs >= 2 B200s in one box (gpurun --gpus 2; skipped on the 1-GPU test box): the product's NCCL path —
krd_compute_allgather_host from one host thread per GPU with communicators from krd_nccl_init_all, which is what
`gappa-b200 analyze krd --gpus G --exchange nccl` runs — must reproduce the single-GPU matrix (to the rounding of a different partial-sum split)."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _need_gpus(n):
    import torch
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs (gpurun --gpus {n})")


@pytest.mark.parametrize("bins,p", [(2, 1.0), (0, 1.0), (4, 2.0), (1, 0.5)])
def test_nccl_allgather_in_one_process_equals_single_gpu(bins, p):
    _need_gpus(2)
    import torch
    import gappa_b200
    from gappa_b200 import synth, krd
    G = min(torch.cuda.device_count(), 4)
    n = 333
    tree = synth.tree_with_edges(201, seed=4)
    mass = synth.sample_masses(tree, n, 300, seed=4)
    plans = []
    for g in range(G):
        pl = gappa_b200.KrdPlan(device=g, bins=bins)
        pl.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        pl.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
        plans.append(pl)
    single = plans[0].compute(p=p)
    comms = krd.nccl_init_all(list(range(G)))
    out = np.full((n, n), -1.0)
    errs = []

    def work(g):
        try:
            r0, r1 = g * n // G, (g + 1) * n // G
            plans[g].compute_allgather_host(g, G, comms[g], r0, r1, p=p, out_rows=out[r0:r1])
        except Exception as ex:          # noqa: BLE001
            errs.append(ex)
    th = [threading.Thread(target=work, args=(g,)) for g in range(G)]
    [t.start() for t in th]; [t.join() for t in th]
    krd.nccl_destroy(comms)
    for pl in plans:
        pl.close()
    assert not errs, errs
    # (a rank's tile share is cut between its CTAs differently than the whole list: partial sums round differently)
    assert np.max(np.abs(out - single) / np.maximum(single, 1e-300)) <= 1e-12
    assert np.array_equal(out, out.T) and np.all(np.diag(out) == 0)


def test_peer_scatter_across_two_gpus_equals_single_gpu():
    """krd_scatter_shard_rows over NVLink peer mappings, one process driving two GPUs (krd_p2p_enable_peer): every GPU
    computes its tile share and stores tile + mirror straight into the row block of the GPU that owns those rows."""
    _need_gpus(2)
    import torch
    import gappa_b200
    from gappa_b200 import synth, krd
    G, n = 2, 700
    tree = synth.tree_with_edges(201, seed=4)
    mass = synth.sample_masses(tree, n, 200, seed=4)
    krd.p2p_enable_peer(0, 1); krd.p2p_enable_peer(1, 0)
    bounds = np.array([g * n // G for g in range(G + 1)], np.int32)
    plans, blocks, packed = [], [], []
    for g in range(G):
        pl = gappa_b200.KrdPlan(device=g, bins=2)
        pl.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        pl.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
        plans.append(pl)
        blocks.append(torch.full((int(bounds[g + 1] - bounds[g]), n), -3.0, dtype=torch.float64, device=f"cuda:{g}"))
        packed.append(torch.empty(pl.shard_doubles(G), dtype=torch.float64, device=f"cuda:{g}"))
    single = plans[0].compute(p=1.0)
    ptrs = [b.data_ptr() for b in blocks]
    for g in range(G):
        plans[g].compute_shard(packed[g], g, G, p=1.0)
        plans[g].scatter_shard_rows(g, G, packed[g], bounds, ptrs)
    for g in range(G):
        torch.cuda.synchronize(g)
    got = np.concatenate([b.cpu().numpy() for b in blocks])
    for pl in plans:
        pl.close()
    assert np.max(np.abs(got - single) / np.maximum(single, 1e-300)) <= 1e-12
    assert np.array_equal(got, got.T) and np.all(np.diag(got) == 0)
