"""CPU test (gloo, world_size 2 and 3) of the multi-GPU sharding contract: contiguous shares of the
tile list (host-only C-ABI helpers), equal-size packed tiles, all-gather, unpack.
The per-rank tile contents come from the CPU oracle standing in for the pair kernel; what is under
test is the partition / pack / gather / unpack logic that bench.py runs over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT

T = 128


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def pack_rank(D, n, rank, world, k):
    """What krd_compute_shard leaves in d_packed: this rank's tiles, (i<j) ? v : 0, zero padded."""
    ti, tj = k.tile_list(n)
    t0, t1 = k.rank_tiles(len(ti), rank, world)
    per = (len(ti) + world - 1) // world
    Dp = np.zeros((((n + T - 1) // T) * T,) * 2)
    Dp[:n, :n] = D
    out = np.zeros((per, T, T))
    for s, t in enumerate(range(t0, t1)):
        blk = Dp[ti[t] * T:(ti[t] + 1) * T, tj[t] * T:(tj[t] + 1) * T].copy()
        if ti[t] == tj[t]:
            blk = np.triu(blk, 1)
        out[s] = blk
    return out.reshape(-1)


def unpack(g, n, k):
    """k_unpack: gathered tiles (rank-major == tile order) -> symmetric n x n."""
    ti, tj = k.tile_list(n)
    g = g.reshape(-1, T, T)
    out = np.zeros((((n + T - 1) // T) * T,) * 2)
    for t in range(len(ti)):
        blk = g[t]
        i0, j0 = ti[t] * T, tj[t] * T
        if ti[t] == tj[t]:
            blk = np.triu(blk, 1)
            out[i0:i0 + T, j0:j0 + T] += blk + blk.T
        else:
            out[i0:i0 + T, j0:j0 + T] = blk
            out[j0:j0 + T, i0:i0 + T] = blk.T
    return out[:n, :n]


def _worker(rank, world, port, n, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    import gappa_b200.krd as k
    from gappa_b200 import synth
    tree = synth.tree_with_edges(31, seed=1)
    mass = synth.sample_masses(tree, n, 20, seed=1)
    D = oracle.port_matrix(tree, mass, 1.0)            # every rank holds the replicated inputs
    mine = torch.from_numpy(pack_rank(D, n, rank, world, k))
    gathered = torch.empty(mine.numel() * world, dtype=torch.float64)
    dist.all_gather_into_tensor(gathered, mine)
    full = unpack(gathered.numpy(), n, k)
    ok = bool(np.array_equal(full, D))
    t = torch.tensor([int(ok)]); dist.all_reduce(t)
    if rank == 0:
        q.put(int(t.item()))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 300), (3, 513), (2, 100)])
def test_shard_gather_unpack_gloo(world, n):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get() == world


def test_rank_shares_cover_every_tile_once():
    import gappa_b200.krd as k
    for n in (1, 128, 129, 1000, 2500):
        ti, tj = k.tile_list(n)
        assert len(ti) == k.lib().krd_tile_count(n)
        assert len(set(zip(ti.tolist(), tj.tolist()))) == len(ti) and np.all(ti <= tj)
        for world in (1, 2, 3, 8):
            spans = [k.rank_tiles(len(ti), r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == len(ti)
            assert all(spans[r][1] == spans[r + 1][0] for r in range(world - 1))
            per = (len(ti) + world - 1) // world
            assert all(b - a <= per for a, b in spans)


@pytest.mark.parametrize("n,late", [(1000, 1), (10000, 5), (20000, 10), (130, 3), (129, 1), (128, 4), (5, 0)])
def test_two_part_tile_order_host_logic(n, late):
    """krd_set_late_blocks' order (host-only mirror krd_tile_list_late): a permutation of the default list; part A = exactly the
    tiles inside the leading T - L sample blocks, in the default order of that smaller matrix; part B = the tiles touching a
    late block, column block by column block.  With every rank taking its share of A and then of B (bench.py --exchange
    p2p_rows), each tile is computed once, and after part A every cell of the leading square has an owner that wrote it."""
    import gappa_b200.krd as k
    T = (n + 127) // 128
    L = max(0, min(late, T - 1))
    ti, tj, split = k.tile_list_late(n, late)
    di, dj = k.tile_list(n)
    assert sorted(zip(ti.tolist(), tj.tolist())) == sorted(zip(di.tolist(), dj.tolist()))
    Ta = T - L
    assert split == Ta * (Ta + 1) // 2
    ai, aj = k.tile_list(Ta * 128)
    assert np.array_equal(ti[:split], ai) and np.array_equal(tj[:split], aj)
    assert np.all(tj[:split] < Ta) and np.all(tj[split:] >= Ta)
    assert np.all(np.diff(tj[split:]) >= 0)                       # column block by column block
    for world in (1, 2, 8):
        done = np.zeros((T, T), np.int32)
        lead = np.zeros((T, T), bool)
        for r in range(world):
            a0, a1 = k.rank_tiles(split, r, world)
            b0, b1 = k.rank_tiles(len(ti) - split, r, world)
            for t in list(range(a0, a1)) + list(range(split + b0, split + b1)):
                done[ti[t], tj[t]] += 1
            for t in range(a0, a1):
                lead[ti[t], tj[t]] = lead[tj[t], ti[t]] = True
        assert np.array_equal(done, np.triu(np.ones((T, T), np.int32)))
        assert lead[:Ta, :Ta].all()


@pytest.mark.parametrize("world,n", [(1, 5), (2, 300), (3, 513), (8, 1000)])
def test_unpack_shards_host_matches_device_contract(world, n):
    # krd_unpack_shards_host is the host twin of k_unpack (used by `gappa-b200 --gpus G`): no GPU needed
    import gappa_b200.krd as k
    rng = np.random.default_rng(n)
    D = rng.random((n, n)); D = np.triu(D, 1); D = D + D.T
    packed = [pack_rank(D, n, r, world, k) for r in range(world)]
    assert all(len(p) == len(packed[0]) for p in packed)
    assert np.array_equal(k.unpack_shards_host(n, packed), D)
