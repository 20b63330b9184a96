"""Compact binned ingest (krd_set_masses_binned: cell = edge * bins + bin, 12 B per entry) and the row-range unpack of the
sharded result (krd_unpack_shards_rows), through the C ABI on a B200.  The compact form must give the SAME bits as
krd_set_masses fed with the binified mass trees, and the same matrix as the reference path (port oracle)."""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _binified(tree, n, placements, bins, seed):
    from gappa_b200 import synth
    raw = synth.sample_masses(tree, n, placements, seed=seed)
    return raw, synth.binify_host(tree, raw, bins)


@pytest.mark.parametrize("n,E,P,bins,pipeline", [(150, 61, 80, 3, 0), (70, 150, 300, 4, 0), (33, 20, 100, 40, 0),
                                                 (700, 301, 200, 4, 1), (1300, 97, 60, 1, 1)])
def test_compact_form_equals_full_form(n, E, P, bins, pipeline):
    import gappa_b200
    import oracle
    from gappa_b200 import synth
    tree = synth.tree_with_edges(E, seed=n)
    raw, b = _binified(tree, n, P, bins, seed=n)
    cell = gappa_b200.krd.binned_cells(tree, b, bins)
    for i in range(0, n, 17):        # the binified trees are in MassTree order: cells strictly ascending within a sample
        assert np.all(np.diff(cell[b["off"][i]:b["off"][i + 1]].astype(np.int64)) > 0)
    full = gappa_b200.KrdPlan(bins=bins)
    full.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    full.set_masses(b["off"], b["edge"], b["pos"], b["mass"])
    comp = gappa_b200.KrdPlan(bins=bins, pipeline=pipeline)
    comp.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    comp.set_masses_binned(b["off"], cell, b["mass"])
    for p in (1.0, 0.5, 2.0):
        Df, Dc = full.compute(p=p), comp.compute(p=p)
        if pipeline:                   # phase-major tile order: different partial-sum splits (DESIGN.md "Pipelined ingest")
            assert rel_err(Dc, Df) <= (1e-12 if p >= 1 else 1e-10)
        else:
            assert np.array_equal(Dc, Df), p
    Ff, _ = full.debug_F()
    Fc, _ = comp.debug_F()
    assert np.array_equal(Ff, Fc)
    # and against the reference path on the RAW placements (binning included)
    idx = list(range(min(n, 40)))
    sub = dict(off=raw["off"][: len(idx) + 1], edge=raw["edge"][: raw["off"][len(idx)]], pos=raw["pos"][: raw["off"][len(idx)]],
               mass=raw["mass"][: raw["off"][len(idx)]])
    want = oracle.port_matrix(tree, sub, p=1.0, bins=bins)
    assert rel_err(comp.compute(p=1.0)[np.ix_(idx, idx)], want) <= 1e-10
    full.close(); comp.close()


@pytest.mark.parametrize("n,E,P,bins,pipeline", [(150, 61, 80, 3, 0), (700, 301, 200, 4, 1), (1300, 97, 60, 1, 1)])
def test_16_bit_cells_equal_32_bit_cells(n, E, P, bins, pipeline):
    """krd_set_masses_binned16 (10 B per entry) is the same input in a narrower transport form: identical bits out, host and
    device variants, ordinary and pipelined handles; a tree with more than 65,536 cells is refused with a sized message."""
    import torch
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(E, seed=n)
    raw, b = _binified(tree, n, P, bins, seed=n)
    cell = gappa_b200.krd.binned_cells(tree, b, bins)
    assert cell.max() < 65536
    res = {}
    for name, c in (("u32", cell), ("u16", cell.astype(np.uint16))):
        plan = gappa_b200.KrdPlan(bins=bins, pipeline=pipeline)
        plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        plan.set_masses_binned(b["off"], c, b["mass"])
        res[name] = [plan.compute(p=p) for p in (1.0, 0.5, 2.0)] + [plan.debug_F()[0]]
        plan.close()
    for a, c in zip(res["u32"], res["u16"]):
        assert np.array_equal(a, c)
    # device arrays (2-byte tensor -> krd_set_masses_binned16_device)
    d_cell = torch.from_numpy(cell.astype(np.uint16).view(np.int16)).cuda()
    d_mass = torch.from_numpy(b["mass"]).cuda()
    plan = gappa_b200.KrdPlan(bins=bins)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses_binned_device(b["off"], d_cell, d_mass)
    got = plan.compute(p=1.0)
    if pipeline:       # (this handle is not pipelined: default tile order, different partial-sum splits than res[...])
        assert rel_err(got, res["u32"][0]) <= 1e-12
    else:
        assert np.array_equal(got, res["u32"][0])
    plan.close()


def test_16_bit_cells_are_refused_for_large_trees():
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(20001, seed=1)
    plan = gappa_b200.KrdPlan(bins=4)                       # 80,004 cells > 65,536
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    with pytest.raises(gappa_b200.KrdError, match="krd_set_masses_binned16 needs n_edges \\* bins <= 65536"):
        plan.set_masses_binned(np.array([0, 1], np.int64), np.array([5], np.uint16), np.array([1.0]))
    plan.set_masses_binned(np.array([0, 1, 2], np.int64), np.array([5, 70000], np.uint32), np.array([1.0, 1.0]))   # 32-bit cells still work
    assert plan.compute(p=1.0).shape == (2, 2)
    plan.close()


def test_compact_form_duplicates_are_summed_and_errors_are_reported():
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(21, seed=2)
    raw, b = _binified(tree, 9, 50, 2, seed=2)
    cell = gappa_b200.krd.binned_cells(tree, b, 2)
    # every entry split into two halves with the same cell: same tree
    off2 = b["off"] * 2
    cell2 = np.repeat(cell, 2); mass2 = np.repeat(b["mass"] * 0.5, 2)
    a = gappa_b200.KrdPlan(bins=2); a.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    a.set_masses_binned(b["off"], cell, b["mass"]); Da = a.compute()
    a.set_masses_binned(off2, cell2, mass2); Db = a.compute()
    assert rel_err(Db, Da) <= 1e-13
    bad = cell.copy(); bad[[3, 4]] = bad[[4, 3]]                 # out of order inside sample 0
    with pytest.raises(gappa_b200.KrdError, match="sorted"):
        a.set_masses_binned(b["off"], bad, b["mass"])
    bad = cell.copy(); bad[7] = 21 * 2 + 5                        # edge 23 of a 21-edge tree
    with pytest.raises(gappa_b200.KrdError, match="outside the tree"):
        a.set_masses_binned(b["off"], bad, b["mass"])
    a.close()
    e = gappa_b200.KrdPlan(bins=0); e.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    with pytest.raises(gappa_b200.KrdError, match="bins"):
        e.set_masses_binned(b["off"], cell, b["mass"])
    e.close()


@pytest.mark.parametrize("world", [2, 5])
def test_row_range_unpack_equals_full_unpack(world):
    import torch
    import gappa_b200
    from gappa_b200 import synth
    n = 333
    tree = synth.tree_with_edges(77, seed=6)
    raw, b = _binified(tree, n, 90, 2, seed=6)
    plan = gappa_b200.KrdPlan(bins=2)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses_binned(b["off"], gappa_b200.krd.binned_cells(tree, b, 2), b["mass"])
    shard = plan.shard_doubles(world)
    g = torch.empty(shard * world, dtype=torch.float64, device="cuda")
    for r in range(world):
        plan.compute_shard(g[r * shard:(r + 1) * shard], r, world)
    full = torch.full((n, n), -1.0, dtype=torch.float64, device="cuda")
    plan.unpack_shards(g, full, world)
    rows = torch.full((n, n), -2.0, dtype=torch.float64, device="cuda")
    for r in range(world):                                    # every rank unpacks its own row block
        r0, r1 = r * n // world, (r + 1) * n // world
        plan.unpack_shards_rows(g, rows[r0:r1], world, r0, r1)
    torch.cuda.synchronize()
    assert torch.equal(rows, full)
    assert np.array_equal(full.cpu().numpy(), plan.compute())
    plan.close()


@pytest.mark.parametrize("world,n", [(3, 333), (8, 1000)])
def test_peer_scatter_equals_full_unpack(world, n):
    """krd_scatter_shard_rows with every "rank" on this one GPU (the row blocks are then ordinary local buffers): the blocks,
    stacked, are the unpacked matrix — tile, mirror and zero diagonal, row bounds that do not fall on tile borders."""
    import torch
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(77, seed=6)
    raw, b = _binified(tree, n, 60, 2, seed=6)
    plan = gappa_b200.KrdPlan(bins=2)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses_binned(b["off"], gappa_b200.krd.binned_cells(tree, b, 2), b["mass"])
    shard = plan.shard_doubles(world)
    bounds = np.array([r * n // world for r in range(world + 1)], np.int32)
    blocks = [torch.full((int(bounds[r + 1] - bounds[r]), n), -3.0, dtype=torch.float64, device="cuda") for r in range(world)]
    packed = torch.empty(shard, dtype=torch.float64, device="cuda")
    for r in range(world):
        plan.compute_shard(packed, r, world)
        plan.scatter_shard_rows(r, world, packed, bounds, [blk.data_ptr() for blk in blocks])
    torch.cuda.synchronize()
    got = torch.cat(blocks).cpu().numpy()
    assert np.array_equal(got, plan.compute())
    plan.close()


@pytest.mark.parametrize("world,k,n", [(1, 1, 200), (3, 2, 333), (2, 4, 1000)])
def test_copy_engine_tiles_to_host_equals_compute(world, k, n):
    """krd_shard_to_host_async: every (virtual) rank's packed tiles go straight into one pinned n x n host matrix — tile, mirror
    (transposed on the device), symmetric diagonal tiles, ragged edges — through 2-D copies on a second stream."""
    import torch
    import gappa_b200
    from gappa_b200 import synth
    tree = synth.tree_with_edges(77, seed=6)
    raw, b = _binified(tree, n, 60, 2, seed=6)
    plan = gappa_b200.KrdPlan(bins=2)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses_binned(b["off"], gappa_b200.krd.binned_cells(tree, b, 2), b["mass"])
    W = world * k
    shard = plan.shard_doubles(W)
    host = torch.full((n, n), -7.0, dtype=torch.float64).pin_memory()
    side = torch.cuda.Stream()
    bufs = [(torch.empty(shard, dtype=torch.float64, device="cuda"), torch.empty(shard, dtype=torch.float64, device="cuda")) for _ in range(W)]
    for r in range(W):
        plan.compute_shard(bufs[r][0], r, W)
        plan.shard_to_host_async(r, W, bufs[r][0], bufs[r][1], host, side.cuda_stream)
    torch.cuda.synchronize()
    want = plan.compute()
    assert np.array_equal(host.numpy(), want)
    # the same through arbitrary tile ranges (an uneven cut of the list)
    host.fill_(-7.0)
    nt = len(gappa_b200.krd.tile_list(n)[0])
    cuts = [0, max(1, (nt * 7) // 8), nt] if nt > 1 else [0, nt]
    for a, c in zip(cuts[:-1], cuts[1:]):
        pk = torch.empty((c - a) * 128 * 128, dtype=torch.float64, device="cuda"); sc = torch.empty_like(pk)
        plan.compute_tile_range(pk, a, c)
        plan.tile_range_to_host_async(a, c, pk, sc, host, side.cuda_stream)
        torch.cuda.synchronize()
    got = host.numpy()
    assert np.max(np.abs(got - want) / np.maximum(want, 1e-300)) <= 1e-12 and np.array_equal(got, got.T)
    plan.close()


def test_late_block_tile_order_and_two_part_scatter():
    """krd_set_late_blocks: the reordered tile list is still the whole upper triangle (same matrix from every entry point);
    after the first part alone, the rectangle outside the late blocks is already final."""
    import torch
    import gappa_b200
    from gappa_b200 import synth, krd
    n, world = 1000, 3
    tree = synth.tree_with_edges(77, seed=6)
    raw, b = _binified(tree, n, 60, 2, seed=6)
    plan = gappa_b200.KrdPlan(bins=2)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses_binned(b["off"], krd.binned_cells(tree, b, 2), b["mass"])
    want = plan.compute()
    T = (n + 127) // 128
    split = plan.set_late_blocks(2)
    assert split == (T - 2) * (T - 1) // 2 and plan.stats()["n_tiles"] == T * (T + 1) // 2
    got = plan.compute()
    assert np.max(np.abs(got - want) / np.maximum(want, 1e-300)) <= 1e-12 and np.array_equal(got, got.T)
    cA = (T - 2) * 128
    bounds = np.array([r * n // world for r in range(world + 1)], np.int32)
    blocks = [torch.full((int(bounds[r + 1] - bounds[r]), n), -3.0, dtype=torch.float64, device="cuda") for r in range(world)]
    ptrs = [blk.data_ptr() for blk in blocks]
    nt = T * (T + 1) // 2
    for r in range(world):                                   # part A of every "rank"
        a0, a1 = krd.rank_tiles(split, r, world)
        pk = torch.empty(max(1, a1 - a0) * 128 * 128, dtype=torch.float64, device="cuda")
        plan.compute_tile_range(pk, a0, a1)
        plan.scatter_tile_range_rows(a0, a1, pk, bounds, ptrs)
    torch.cuda.synchronize()
    early = torch.cat(blocks).cpu().numpy()
    assert np.max(np.abs(early[:cA, :cA] - want[:cA, :cA]) / np.maximum(want[:cA, :cA], 1e-300)) <= 1e-12
    assert np.all(early[cA:, :] == -3.0) and np.all(early[:, cA:] == -3.0)      # nothing of the late strip yet
    for r in range(world):                                   # part B
        b0, b1 = (split + q for q in krd.rank_tiles(nt - split, r, world))
        pk = torch.empty(max(1, b1 - b0) * 128 * 128, dtype=torch.float64, device="cuda")
        plan.compute_tile_range(pk, b0, b1)
        plan.scatter_tile_range_rows(b0, b1, pk, bounds, ptrs)
    torch.cuda.synchronize()
    full = torch.cat(blocks).cpu().numpy()
    assert np.max(np.abs(full - want) / np.maximum(want, 1e-300)) <= 1e-12 and np.array_equal(full, full.T)
    # rectangle copy helper + back to the default order
    host = torch.zeros((n, n), dtype=torch.float64).pin_memory()
    krd.copy_rect_to_host_async(host.data_ptr() + 5 * 8, n, blocks[0].data_ptr() + 5 * 8, n, 100, 7, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(host.numpy()[:7, 5:105], full[:7, 5:105]) and host.numpy()[7:].sum() == 0
    assert plan.set_late_blocks(0) == nt
    assert np.array_equal(plan.compute(), want)
    plan.close()
