"""GPU tests at BASELINE.json sizes and multi-rank equivalence.

* config 1 (1 M points / 300 cameras / threshold 500 000, the reference's own
  CPU-runnable case): leaves, expanded blocks, camera assignment and every
  pair's hit count against the oracle (bit-exact); keep decisions against the
  oracle's exact hull ratio for a sample of pairs and through the bound
  invariants (ratio_lo <= exact <= ratio_hi) for all of them.
* size-independent properties at a larger size: determinism (two runs are
  identical), default (bound-decided) mode == exact mode, and camera sharding
  for world sizes 2 / 3 / 8 (ranks run one after the other on this GPU)
  == the unsharded result."""
import numpy as np
import pytest
import torch

import ptgs_oracle as O
from helpers import scene_cameras
import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, multigpu as M, workload as W

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _lists(off, idx):
    off = off.cpu().numpy()
    idx = idx.cpu().numpy()
    return [idx[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def test_config1_against_oracle():
    w = W.make("c1", 1.0)
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    centers = D.camera_centers(w.R, w.T)
    res = D.partition_device(torch.from_numpy(w.xyz).to(DEV), recs, centers, w.threshold)
    cams = O.CameraSet(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    want = O.run_partition(w.xyz, cams, w.threshold, centers=centers, visibility=False)
    assert np.array_equal(res.keep_mask.cpu().numpy().astype(bool), want.keep_mask)
    assert res.leaf_ids == [l.partition_id for l in want.leaves] and len(res.leaf_ids) >= 2
    assert [tuple(b) for b in res.leaf_bounds.tolist()] == [l.bounds for l in want.leaves]
    for g, wv in zip(_lists(res.pt_offsets, res.pt_index), want.block_index):
        assert np.array_equal(g, wv)
    cam_lists = _lists(res.cam_offsets, res.cam_index)
    for g, wv in zip(cam_lists, want.cam_index):
        assert np.array_equal(g, wv)
    # voxel down-sampling of every block, bit for bit (set comparison through first-occurrence order)
    ds_off = res.ds_offsets.cpu().numpy()
    ds = np.stack([res.ds_x.cpu().numpy(), res.ds_y.cpu().numpy(), res.ds_z.cpu().numpy()], 1)
    n_hit = res.vis.n_hit.cpu().numpy()
    keep = res.vis.keep.cpu().numpy().astype(bool)
    flags = res.vis.flags.cpu().numpy()
    rlo = res.vis.ratio_lo.cpu().numpy()
    rhi = res.vis.ratio_hi.cpu().numpy()
    p = 0
    rng = np.random.default_rng(0)
    checked = 0
    for b, cl in enumerate(cam_lists):
        d = ds[ds_off[b]:ds_off[b + 1]]
        wd = O.voxel_downsample(w.xyz[want.block_index[b]])
        assert len(d) == len(wd)
        du, wu = d.view(np.uint32), wd.view(np.uint32)
        du = du[np.lexsort((du[:, 2], du[:, 1], du[:, 0]))]
        wu = wu[np.lexsort((wu[:, 2], wu[:, 1], wu[:, 0]))]
        assert np.array_equal(du, wu)
        for c in cl:
            hit, u, v = O.project_port(d, recs[c])
            assert n_hit[p] == int(hit.sum()), (b, c)
            assert rlo[p] <= rhi[p] + 1e-12
            if n_hit[p] > O.MIN_HITS and not (flags[p] & 3) and rng.random() < 0.04:
                rt = O.run_graham_scan(np.stack([u[hit], v[hit]], 1).astype(np.float64),
                                       float(recs[c][16]), float(recs[c][17]))
                if not O.first_ray_collinear(np.stack([u[hit], v[hit]], 1).astype(np.float64)):
                    assert rlo[p] <= rt * (1 + 1e-12) + 1e-15 and rhi[p] >= rt * (1 - 1e-12), (b, c)
                    assert keep[p] == (rt >= O.VISIBLE_RATE), (b, c, rt)
                    checked += 1
            if n_hit[p] <= O.MIN_HITS:
                assert not keep[p]
            p += 1
    assert p == len(n_hit) and checked >= 3
    assert 0 < keep.sum() < len(keep)


def _scene_state(scale=0.2):
    w = W.make("c3", scale)                      # many leaves (deep recursion config)
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    centers = D.camera_centers(w.R, w.T)
    res = D.partition_device(torch.from_numpy(w.xyz).to(DEV), recs, centers, w.threshold, man_trans=w.man_trans)
    return w, recs, res


def test_determinism_and_modes():
    w, recs, res = _scene_state()
    assert len(res.leaf_ids) >= 16
    cams_d = torch.from_numpy(recs).to(DEV)
    args = (res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets, res.cam_index)
    a = D.visibility(*args)
    b = D.visibility(*args)
    for f in ("n_hit", "keep", "flags", "ratio"):
        assert torch.equal(getattr(a, f), getattr(b, f)), f
    e = D.visibility(*args, exact_ratio=True)
    n = D.visibility(*args, interior=False)          # without the interior-region fast path
    assert torch.equal(a.n_hit, e.n_hit) and torch.equal(a.keep, e.keep)
    assert torch.equal(a.n_hit, n.n_hit) and torch.equal(a.keep, n.keep) and torch.equal(a.ratio, n.ratio)
    assert bool((a.ratio_lo <= e.ratio * (1 + 1e-12) + 1e-15).all())
    assert bool((a.ratio_hi >= e.ratio * (1 - 1e-12)).all())
    assert 0 < int(a.keep.sum()) < a.keep.numel()


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("mode", ["ranges", "modulo"])
def test_camera_sharding_equals_single_gpu(world, mode):
    """Rank r evaluates its share of the pairs (contiguous equal-cost ranges, or cameras
    c % world == r); scattering every rank's records to their global pair positions
    reproduces the unsharded table."""
    w, recs, res = _scene_state()
    cams_d = torch.from_numpy(recs).to(DEV)
    full = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets, res.cam_index)
    P = res.cam_index.shape[0]
    table = torch.zeros((P, M.REC_COLS), dtype=torch.float64, device=DEV)
    seen = torch.zeros(P, dtype=torch.int32, device=DEV)
    for r in range(world):
        if mode == "ranges":
            off, idx, pos = M.shard_csr_ranges(res.cam_offsets, res.cam_index, r, world,
                                               res.ds_offsets[1:] - res.ds_offsets[:-1])
        else:
            off, idx, pos = M.shard_csr(res.cam_offsets, res.cam_index, r, world)
        v = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, off, idx)
        table[pos] = M.pack_records(v.n_hit, v.ratio, v.area, v.keep, v.flags)
        seen[pos] += 1
    assert bool((seen == 1).all())
    n_hit, ratio, area, keep, flags = M.unpack_records(table)
    assert torch.equal(n_hit, full.n_hit) and torch.equal(keep, full.keep) and torch.equal(flags, full.flags)
    assert torch.equal(ratio, full.ratio)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_host_entry_rank_local_blocks(world):
    """visibility_select_host with world > 1: every rank uploads and sweeps only the blocks of its
    pair range; the union of the ranks' records equals the single-process result (the collective
    is replaced by a recording stand-in, ranks run one after the other on this GPU)."""
    w, recs, res = _scene_state()
    x, y, z = D.aos_to_soa(torch.from_numpy(w.xyz).to(DEV))
    idx = res.pt_index.to(torch.int64)
    blocks = torch.stack([x[idx], y[idx], z[idx]], 1).contiguous().cpu()
    blk_off = res.pt_offsets.cpu().numpy()
    cam_off, cam_idx = res.cam_offsets.cpu().numpy(), res.cam_index.cpu().numpy()
    recs_t = torch.from_numpy(recs)
    one, _ = D.visibility_select_host(blocks, blk_off, recs_t, cam_off, cam_idx, torch.device(DEV))
    P = len(cam_idx)
    table = torch.zeros((P, M.REC_COLS), dtype=torch.float64, device=DEV)
    seen = torch.zeros(P, dtype=torch.int32, device=DEV)

    def exchange(local, pos):
        table[pos] = local
        seen[pos] += 1
        return table.clone()
    out = None
    for r in range(world):
        out, _ = D.visibility_select_host(blocks, blk_off, recs_t, cam_off, cam_idx, torch.device(DEV),
                                          rank=r, world=world, exchange=exchange)
    assert bool((seen == 1).all())
    # after the last rank the table is complete: that call's outputs are the full result
    for a, b in zip(one, out):
        assert np.array_equal(a, b)


def test_host_entry_offsets_cover_camera_less_blocks():
    """Blocks without cameras at either end of the block list are skipped by the host entry
    (nothing to upload), but kept_offsets still has one entry per block + 1 and lines up with the
    global block index, for world 1 and world > 1 (device.py visibility_select_host)."""
    w, recs, res = _scene_state()
    x, y, z = D.aos_to_soa(torch.from_numpy(w.xyz).to(DEV))
    idx = res.pt_index.to(torch.int64)
    blocks = torch.stack([x[idx], y[idx], z[idx]], 1).contiguous().cpu()
    blk_off = res.pt_offsets.cpu().numpy()
    cam_off, cam_idx = res.cam_offsets.cpu().numpy().copy(), res.cam_index.cpu().numpy()
    L = len(cam_off) - 1
    cnt = np.diff(cam_off)
    keep_pairs = np.ones(len(cam_idx), bool)
    for b in (0, 1, L - 1):                                   # two leading and one trailing empty list
        keep_pairs[cam_off[b]:cam_off[b + 1]] = False
        cnt[b] = 0
    cam_idx = np.ascontiguousarray(cam_idx[keep_pairs])
    cam_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    recs_t = torch.from_numpy(recs)
    one, _ = D.visibility_select_host(blocks, blk_off, recs_t, cam_off, cam_idx, torch.device(DEV))
    k_off, k_idx, n_hit, ratio, keep, flags = one
    assert len(k_off) == L + 1 and k_off[0] == 0 and k_off[1] == 0 and k_off[2] == 0
    assert k_off[L] == k_off[L - 1] == len(k_idx) == int(keep.sum())
    # per block: the kept list is the block's camera list filtered by keep
    for b in range(L):
        sl = slice(cam_off[b], cam_off[b + 1])
        assert np.array_equal(k_idx[k_off[b]:k_off[b + 1]], cam_idx[sl][keep[sl].astype(bool)]), b
    P = len(cam_idx)
    table = torch.zeros((P, M.REC_COLS), dtype=torch.float64, device=DEV)

    def exchange(local, pos):
        table[pos] = local
        return table.clone()
    out = None
    for r in range(3):
        out, _ = D.visibility_select_host(blocks, blk_off, recs_t, cam_off, cam_idx, torch.device(DEV),
                                          rank=r, world=3, exchange=exchange)
    assert len(out[0]) == L + 1
    for a, b in zip(one, out):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("world,chunks", [(2, 1), (3, 1), (8, 1), (2, 4), (3, 3), (8, 4)])
def test_slot_exchange_kernels_equal_single_gpu(world, chunks):
    """The product exchange of the range-sharded pass -- ptgs_pack_records per rank, the ranks'
    padded slots laid out as the all-gather delivers them, ptgs_unpack_records -- reproduces the
    unsharded result vectors (ranks run one after the other on this GPU; the collective itself is
    covered by the gloo test and the multi-GPU bench)."""
    w, recs, res = _scene_state()
    cams_d = torch.from_numpy(recs).to(DEV)
    full = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets, res.cam_index)
    slots, plan = [], None
    for r in range(world):
        plan = D.range_plan(res.ds_offsets, res.cam_offsets, res.cam_index, r, world, chunks=chunks)
        assert int(plan.my_off[-1]) == plan.my_idx.shape[0] == plan.n_local == int(plan.local_counts().sum())
        v = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, plan.my_off, plan.my_idx)
        n = plan.n_local
        slots.append(D.pack_records_dev((v.n_hit[:n], v.ratio[:n], v.area[:n], v.keep[:n], v.flags[:n]), plan.slot))
    assert plan.bounds[0] == 0 and plan.bounds[-1] == res.cam_index.shape[0]
    n_hit, ratio, area, keep, flags = D.unpack_records_dev(torch.cat(slots), plan)
    assert torch.equal(n_hit, full.n_hit) and torch.equal(keep, full.keep) and torch.equal(flags, full.flags)
    assert torch.equal(ratio, full.ratio) and torch.equal(area, full.area)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("f64", [False, True])
def test_slab_sharded_radius_filter_equals_single_gpu(world, f64):
    """Multi-GPU A1: every rank filters its X slab (+ one-cell halo); the union of the ranks' partial
    masks (what the MAX all-reduce produces) is the unsharded mask, bit for bit.  Ranks run one
    after the other on this GPU."""
    w = W.make("c3", 0.05)
    x, y, z = D.aos_to_soa(torch.from_numpy(w.xyz).to(DEV))
    if f64:
        man = np.eye(4)
        a = np.deg2rad(33.0)
        man[:2, :2] = [[np.cos(a), -np.sin(a)], [np.sin(a), np.cos(a)]]
        x, y, z = D.manhattan_apply(x, y, z, man)
    full, n_full = D.radius_outlier_mask(x, y, z)
    bounds = D.slab_bounds(x, world)
    assert bounds[0] == -np.inf and bounds[-1] == np.inf and bool((np.diff(bounds) >= 0).all())
    acc = torch.zeros_like(full)
    owned = torch.zeros(full.shape[0], dtype=torch.int32, device=DEV)
    for r in range(world):
        part, _ = D.radius_outlier_mask_sharded(x, y, z, r, world, bounds=bounds, reduce=False)
        acc = torch.maximum(acc, part)
        xs = x.to(torch.float64)
        owned += ((xs >= bounds[r]) & (xs < bounds[r + 1])).to(torch.int32)
        assert int(part[~((xs >= bounds[r]) & (xs < bounds[r + 1]))].sum()) == 0    # zero outside the own slab
    assert bool((owned == 1).all())
    assert torch.equal(acc, full) and int(acc.sum()) == int(n_full.item())
    assert 0 < int(full.sum()) < full.numel()


@pytest.mark.parametrize("world", [2, 5])
def test_slab_sharded_coverage_equals_single_gpu(world):
    """Config-c5 sharding: every rank sweeps the X slab between two quantiles of x; per-camera counts
    add up to the unsharded sweep's, per-point seen flags scatter back through the slab's index list
    (the collectives are replaced by sums here; ranks run one after the other on this GPU)."""
    w = W.make("c3", 0.03)
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    recs[::2, 18] = 2.0
    cams_d = torch.from_numpy(recs).to(DEV)
    x, y, z = D.aos_to_soa(torch.from_numpy(w.xyz).to(DEV))
    want, wseen = D.coverage_sweep(x, y, z, cams_d, want_seen=True)
    bounds = D.slab_bounds(x, world)
    total = torch.zeros_like(want)
    seen = torch.zeros_like(wseen)
    n_sel = 0
    for r in range(world):
        sub, sidx, m_dev = D.slab_select(x, y, z, bounds[r], bounds[r + 1])
        m = int(m_dev.item())
        n_sel += m
        idx = sidx[:m].to(torch.int64)
        assert bool((idx[1:] > idx[:-1]).all())                       # index order kept
        assert torch.equal(sub[0, :m], x[idx]) and torch.equal(sub[2, :m], z[idx])
        nh, sl = D.coverage_sweep(sub[0, :m], sub[1, :m], sub[2, :m], cams_d, want_seen=True)
        total += nh
        seen[idx] = torch.maximum(seen[idx], sl)
    assert n_sel == x.shape[0]
    assert torch.equal(total, want) and torch.equal(seen, wseen) and int(want.sum()) > 1000
