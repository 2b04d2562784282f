"""world_size-2 (and 3) test of the camera-sharded visibility exchange on CPU
(gloo): shard_csr -> per-rank records -> gather_records must reproduce the
single-process per-pair table in full pair order on every rank.  The compute
callback is replaced by a deterministic function of (block, camera) -- the
CUDA visibility pass itself is covered by the `gpu` tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import ptgs_b200  # noqa: F401
from ptgs_b200 import multigpu as M


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_csr(seed, L=13, C=57):
    rng = np.random.default_rng(seed)
    lists = [np.sort(rng.choice(C, size=rng.integers(0, 12), replace=False)) for _ in range(L)]
    lists[3] = np.zeros(0, np.int64)                      # an empty block
    off = np.zeros(L + 1, np.int64)
    off[1:] = np.cumsum([len(l) for l in lists])
    idx = np.concatenate(lists).astype(np.int32)
    return torch.from_numpy(off), torch.from_numpy(idx)


def _fake_records(off, idx):
    """Deterministic per-pair 'visibility result' from (block, camera)."""
    L = off.shape[0] - 1
    blk = torch.repeat_interleave(torch.arange(L), off[1:] - off[:-1])
    cam = idx.to(torch.int64)
    n_hit = (blk * 1000 + cam).to(torch.int32)
    ratio = (cam.to(torch.float64) + 0.5) / 64.0 + blk.to(torch.float64) * 1e-9
    area = ratio * 1920000.0
    keep = (ratio >= 0.45).to(torch.uint8)
    flags = ((cam % 5 == 0).to(torch.uint8) * 16) | (blk % 2).to(torch.uint8)
    return n_hit, ratio, area, keep, flags


def _block_cost(L, seed):
    return torch.from_numpy(np.random.default_rng(100 + seed).integers(1, 5000, L).astype(np.int64))


def _worker(rank, world, port, seed, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        off, idx = _make_csr(seed)
        my_off, my_idx, pos = M.shard_csr(off, idx, rank, world)
        # shard invariants
        assert bool(((my_idx % world) == rank).all())
        assert int(my_off[-1]) == my_idx.shape[0] == pos.shape[0]
        assert torch.equal(idx[pos], my_idx)
        nh, rt, ar, kp, fl = _fake_records(my_off, my_idx)
        full = M.gather_records(M.pack_records(nh, rt, ar, kp, fl), idx, world, pos=pos)
        full2 = M.gather_records(M.pack_records(nh, rt, ar, kp, fl), idx, world, rank=rank)
        assert torch.equal(full, full2)
        got = M.unpack_records(full)
        want = _fake_records(off, idx)
        for g, w in zip(got, want):
            assert torch.equal(g, w), "rank %d: gathered table differs" % rank
        # contiguous equal-cost ranges (the default sharding of the visibility pass)
        r_off, r_idx, r_pos = M.shard_csr_ranges(off, idx, rank, world, _block_cost(off.shape[0] - 1, seed))
        assert int(r_off[-1]) == r_idx.shape[0] == r_pos.shape[0] and torch.equal(idx[r_pos], r_idx)
        if r_pos.numel():
            assert torch.equal(r_pos, torch.arange(int(r_pos[0]), int(r_pos[-1]) + 1))
        nh, rt, ar, kp, fl = _fake_records(r_off, r_idx)
        full3 = M.gather_records(M.pack_records(nh, rt, ar, kp, fl), idx, world, pos=r_pos)
        for g, w in zip(M.unpack_records(full3), want):
            assert torch.equal(g, w), "rank %d: gathered table differs (ranges)" % rank
        # the same ranges through the plan + ONE padded all-gather of the contiguous slices (the product path)
        cost = _block_cost(off.shape[0] - 1, seed)
        plan = M.RangePlan(off, idx, off.numpy(), cost.numpy(), rank, world)
        assert torch.equal(plan.my_off, r_off) and torch.equal(plan.my_idx, r_idx)
        assert (plan.lo, plan.hi) == M.range_bounds(off.numpy(), cost.numpy(), rank, world)
        assert plan.slot == int(np.diff(plan.bounds).max()) and plan.bounds[0] == 0 and plan.bounds[-1] == idx.shape[0]
        got4 = M.exchange_ranges(_fake_records(plan.my_off, plan.my_idx), plan, M.pack_slot_torch, M.unpack_slot_torch)
        for g, w in zip(got4, want):
            assert torch.equal(g, w), "rank %d: gathered table differs (slot exchange)" % rank
        if plan.hi > plan.lo:   # the block span covers exactly the blocks with a pair in the range
            blk = torch.repeat_interleave(torch.arange(off.shape[0] - 1), off[1:] - off[:-1])[plan.lo:plan.hi]
            assert int(blk.min()) == plan.bl0 and int(blk.max()) == plan.bl1 - 1
        # several interleaved ranges per rank (the default of the visibility step: evens out the pairs that
        # stay undecided): rank's pair list = its ranges in pair order, same exchange
        for chunks in (2, 5):
            cp = M.RangePlan(off, idx, off.numpy(), cost.numpy(), rank, world, chunks=chunks)
            assert int(cp.my_off[-1]) == cp.my_idx.shape[0] == cp.n_local
            pos = torch.cat([torch.arange(int(cp.bounds[c]), int(cp.bounds[c + 1])) for c in cp.mine])
            assert torch.equal(idx[pos], cp.my_idx)
            cnt = torch.from_numpy(cp.local_counts())
            assert torch.equal(cp.my_off[1:] - cp.my_off[:-1], cnt)
            got5 = M.exchange_ranges(_fake_records(cp.my_off, cp.my_idx), cp, M.pack_slot_torch, M.unpack_slot_torch)
            for g, w in zip(got5, want):
                assert torch.equal(g, w), "rank %d: gathered table differs (%d chunks)" % (rank, chunks)
        # coverage sweep exchange (c5): contiguous camera ranges, counts gathered, seen flags merged
        C = 57
        first, cnt = M.camera_range(C, rank, world)
        counts = (torch.arange(first, first + cnt) * 7 + 3).to(torch.int32)
        allc = M.gather_camera_counts(counts, C, world)
        assert torch.equal(allc, (torch.arange(C) * 7 + 3).to(torch.int32))
        seen = ((torch.arange(1000) % world) == rank).to(torch.uint8)
        assert bool(M.merge_seen(seen, world).all())
        # point-sharded coverage exchange: counts summed, seen slices gathered
        npts = 1003
        p0, p1 = M.point_range(npts, rank, world)
        assert p0 % 4 == 0
        part = torch.zeros(C, dtype=torch.int32)
        part[rank::world] = (torch.arange(C)[rank::world] + 1).to(torch.int32)
        assert torch.equal(M.sum_camera_counts(part, world), (torch.arange(C) + 1).to(torch.int32))
        flags = (torch.arange(p0, p1) % 3 == 0).to(torch.uint8)
        assert torch.equal(M.gather_seen(flags, npts, world), (torch.arange(npts) % 3 == 0).to(torch.uint8))
        torch.save([g.clone() for g in got], os.path.join(out_dir, "r%d.pt" % rank))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,seed", [(2, 1), (2, 2), (3, 3)])
def test_sharded_records_gloo(world, seed, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, seed, str(tmp_path)), nprocs=world, join=True)
    tabs = [torch.load(os.path.join(tmp_path, "r%d.pt" % r)) for r in range(world)]
    for t in tabs[1:]:
        for a, b in zip(tabs[0], t):
            assert torch.equal(a, b)


def test_all_range_bounds_equals_per_rank_loop():
    rng = np.random.default_rng(0)
    for trial in range(200):
        L, world = int(rng.integers(1, 40)), int(rng.integers(1, 9))
        cnt = rng.integers(0, 9, L)
        cost = rng.integers(0, 5000, L).astype(np.int64)
        if trial % 5 == 0:
            cnt[rng.integers(0, L)] = 0
        if trial % 7 == 0:
            cost[rng.integers(0, L)] = 0
        co = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
        b = M.all_range_bounds(co, cost, world)
        assert b[0] == 0 and b[-1] == co[-1] and bool((np.diff(b) >= 0).all())
        for r in range(world):
            assert (int(b[r]), int(b[r + 1])) == M.range_bounds(co, cost, r, world)


def test_camera_range_partitions_cameras():
    for C in (0, 1, 7, 16000):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                f, c = M.camera_range(C, r, world)
                got.extend(range(f, f + c))
            assert got == list(range(C))


def test_point_range_partitions_points():
    for n in (0, 3, 1003, 100_000_000):
        for world in (1, 2, 3, 8):
            edges = [M.point_range(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] or edges[i + 1][0] == n for i in range(world - 1))
            assert all(a % 4 == 0 or a == n for a, _ in edges)


def test_shard_csr_single_process_identity():
    off, idx = _make_csr(9)
    o, i, pos = M.shard_csr(off, idx, 0, 1)
    assert torch.equal(o, off) and torch.equal(i, idx) and torch.equal(pos, torch.arange(idx.shape[0]))
    # the shards of a world partition the pair list
    for world in (2, 4, 8):
        seen = torch.zeros(idx.shape[0], dtype=torch.int64)
        for r in range(world):
            o, i, pos = M.shard_csr(off, idx, r, world)
            seen[pos] += 1
            cnt = o[1:] - o[:-1]
            blk = torch.repeat_interleave(torch.arange(off.shape[0] - 1), off[1:] - off[:-1])
            assert torch.equal(cnt, torch.bincount(blk[pos], minlength=off.shape[0] - 1))
        assert bool((seen == 1).all())
        # ranges: a partition of the pair list into contiguous pieces of near-equal cost
        cost = _block_cost(off.shape[0] - 1, 9)
        seen = torch.zeros(idx.shape[0], dtype=torch.int64)
        loads = []
        blk = torch.repeat_interleave(torch.arange(off.shape[0] - 1), off[1:] - off[:-1])
        for r in range(world):
            o, i, pos = M.shard_csr_ranges(off, idx, r, world, cost)
            seen[pos] += 1
            loads.append(int(cost[blk[pos]].sum()))
            assert torch.equal(o[1:] - o[:-1], torch.bincount(blk[pos], minlength=off.shape[0] - 1))
        assert bool((seen == 1).all())
        assert max(loads) - min(loads) <= 2 * int(cost.max())
