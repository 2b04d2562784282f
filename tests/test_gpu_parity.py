"""GPU parity tests: every stage of libptgs (called through the C ABI via
ptgs_b200.device) against the CPU oracle on the same seeded inputs.
Bit-exact for masks, split lines, index sets, voxel means, hit counts and keep
decisions; hull areas to 1e-12 relative (float64 summation order only)."""
import numpy as np
import pytest
import torch

import ptgs_oracle as O
from helpers import scene_cameras, small_scene
from ptgs_b200 import device as D

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _dev_xyz(sc):
    return torch.from_numpy(sc.xyz).to(DEV)


def _csr_lists(off, idx):
    off = off.cpu().numpy()
    idx = idx.cpu().numpy()
    return [idx[off[i]:off[i + 1]] for i in range(len(off) - 1)]


@pytest.mark.parametrize("n,seed,zfar", [(0, 1, None), (1, 2, None), (37, 3, None), (5000, 4, None),
                                         (60000, 5, None), (5000, 6, 1.2e8), (5000, 7, 3.0e9)])
def test_radius_mask(n, seed, zfar):
    """zfar: one point far above the cloud -- 1.2e8 needs coarse (4-cell) height buckets in the
    32-bit sort keys, 3e9 does not fit them at all (64-bit keys with the exact height order)."""
    rng = np.random.default_rng(seed)
    pts = (rng.uniform(0, max(1.0, (n / 48.0) ** 0.5), (n, 3)) * [1, 1, 0.2]).astype(np.float32)
    if n > 100:
        pts[:50] += 500.0   # isolated points
        pts[50:60] = pts[60]  # exact duplicates
        pts[100:160, 2] += rng.uniform(0, 40, 60).astype(np.float32)   # tall columns
    if zfar is not None:
        pts[200, 2] = zfar
    x, y, z = D.aos_to_soa(torch.from_numpy(pts).to(DEV)) if n else (torch.empty(0, device=DEV),) * 3
    keep, n_keep = D.radius_outlier_mask(x, y, z)
    want = O.radius_outlier_mask(pts.astype(np.float64))
    assert np.array_equal(keep.cpu().numpy().astype(bool), want)
    assert int(n_keep.item()) == int(want.sum())


def test_radius_mask_boundary_distance():
    # neighbours at exactly d^2 == r^2 are NOT counted (strict <)
    pts = np.zeros((7, 3), dtype=np.float32)
    pts[1:6, 0] = [1.0, -1.0, 0.5, -0.5, 0.25]
    pts[6] = [0.0, 0.75, 0.0]
    x, y, z = D.aos_to_soa(torch.from_numpy(pts).to(DEV))
    keep, _ = D.radius_outlier_mask(x, y, z)
    assert np.array_equal(keep.cpu().numpy().astype(bool), O.radius_outlier_mask_bruteforce(pts))


@pytest.mark.parametrize("f64", [False, True])
def test_split_expand_assign(f64):
    sc = small_scene(n=80000, c=40, thr=6000, seed=21)
    cams = scene_cameras(sc)
    x, y, z = D.aos_to_soa(_dev_xyz(sc))
    pts = sc.xyz
    if f64:
        man = np.eye(4)
        a = np.deg2rad(17.0)
        man[:2, :2] = [[np.cos(a), -np.sin(a)], [np.sin(a), np.cos(a)]]
        x, y, z = D.manhattan_apply(x, y, z, man)
        pts = O.manhattan_points(sc.xyz, man)
        got = np.stack([x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy()], 1)
        assert np.array_equal(got, pts), "manhattan rotation must be bit-identical to NumPy"
        # the fused layout + rotation pass (what partition_device runs), incl. a ragged tail
        for n in (len(sc.xyz), 4099, 3, 0):
            fx_, fy_, fz_ = D.aos_to_soa_manhattan(_dev_xyz(sc)[:n].contiguous(), man)
            got = np.stack([fx_.cpu().numpy(), fy_.cpu().numpy(), fz_.cpu().numpy()], 1)
            assert np.array_equal(got, pts[:n]), n
    keep, _ = D.radius_outlier_mask(x, y, z)
    want_keep = O.radius_outlier_mask(pts.astype(np.float64))
    assert np.array_equal(keep.cpu().numpy().astype(bool), want_keep)
    fx, fy, nf = D.compact_xy(x, y, keep)
    root = D.bounds_xy(fx, fy, nf)
    filt = pts.astype(np.float64)[want_keep]
    assert tuple(root.cpu().tolist()) == O.bounds_xy(filt)
    lb, lp, ld, lc = D.balanced_split(fx, fy, nf, root, sc.threshold)
    leaves = O.quad_tree_partition(filt, O.bounds_xy(filt), sc.threshold)
    assert len(leaves) == len(lb) and len(leaves) > 4
    for i, lf in enumerate(leaves):
        assert tuple(lb[i].tolist()) == lf.bounds
        assert D.leaf_id(lp[i], ld[i]) == lf.partition_id
        assert int(lc[i]) == lf.point_num
    boxes = D.expanded_boxes(lb)
    assert [tuple(b) for b in boxes.tolist()] == [O.expanded_box(l.bounds) for l in leaves]
    off, idx = D.expand_select(x, y, boxes)
    got = _csr_lists(off, idx)
    for b, box in enumerate(boxes.tolist()):
        assert np.array_equal(got[b], np.nonzero(O.expand_mask(pts, tuple(box)))[0])
    centers = D.camera_centers(cams.R, cams.T)
    assert np.array_equal(centers, O.camera_centers(cams.R, cams.T))
    coff, cidx = D.assign_cameras(centers, boxes, DEV)
    want = O.assign_cameras([tuple(b) for b in boxes.tolist()], centers)
    for g, w in zip(_csr_lists(coff, cidx), want):
        assert np.array_equal(g, w)


def test_split_many_ties():
    # heavy exact ties on both axes: points on split lines belong to both children
    rng = np.random.default_rng(5)
    n = 40000
    pts = np.zeros((n, 3), dtype=np.float32)
    pts[:, 0] = np.round(rng.uniform(0, 20, n) * 4) / 4
    pts[:, 1] = np.round(rng.uniform(0, 15, n) * 4) / 4
    pts[::7, 0] = -0.0
    x, y, z = D.aos_to_soa(torch.from_numpy(pts).to(DEV))
    nf = torch.full((1,), n, dtype=torch.int64, device=DEV)
    root = D.bounds_xy(x, y, nf)
    p64 = pts.astype(np.float64)
    assert tuple(root.cpu().tolist()) == O.bounds_xy(p64)
    lb, lp, ld, lc = D.balanced_split(x, y, nf, root, 300)
    leaves = O.quad_tree_partition(p64, O.bounds_xy(p64), 300)
    assert len(leaves) == len(lb)
    for i, lf in enumerate(leaves):
        assert tuple(lb[i].tolist()) == lf.bounds
        assert D.leaf_id(lp[i], ld[i]) == lf.partition_id
        assert int(lc[i]) == lf.point_num


def test_split_single_leaf_and_depth_cap():
    rng = np.random.default_rng(6)
    pts = rng.uniform(0, 10, (3000, 3)).astype(np.float32)
    x, y, z = D.aos_to_soa(torch.from_numpy(pts).to(DEV))
    nf = torch.full((1,), len(pts), dtype=torch.int64, device=DEV)
    root = D.bounds_xy(x, y, nf)
    lb, lp, ld, lc = D.balanced_split(x, y, nf, root, 100000)
    assert len(lb) == 1 and D.leaf_id(lp[0], ld[0]) == "0" and int(lc[0]) == 3000
    # threshold 1 -> recursion stops at depth 10 (density_partition.py:32, :163)
    lb, lp, ld, lc = D.balanced_split(x, y, nf, root, 1)
    leaves = O.quad_tree_partition(pts.astype(np.float64), O.bounds_xy(pts.astype(np.float64)), 1)
    assert len(lb) == len(leaves)
    assert [D.leaf_id(p, d) for p, d in zip(lp, ld)] == [l.partition_id for l in leaves]
    assert [int(c) for c in lc] == [l.point_num for l in leaves]
    assert all(tuple(lb[i].tolist()) == leaves[i].bounds for i in range(len(leaves)))


def test_voxel_downsample():
    sc = small_scene(n=50000, c=8, thr=9000, seed=31)
    # densify so that voxels hold several points
    sc.xyz[:, 2] = np.round(sc.xyz[:, 2] * 2) / 2
    sc.xyz[:, :2] = np.round(sc.xyz[:, :2] * 8) / 8
    x, y, z = D.aos_to_soa(_dev_xyz(sc))
    boxes = np.array([[-30, -30, 5, 5], [0, 0, 40, 40], [100, 100, 101, 101], [-5, -5, 5, 5.0]])
    off, idx = D.expand_select(x, y, boxes)
    ds_off, dx, dy, dz, dfirst = D.voxel_downsample(x, y, z, off, idx)
    ds_off = ds_off.cpu().numpy()
    got = np.stack([dx.cpu().numpy(), dy.cpu().numpy(), dz.cpu().numpy()], 1)
    first = dfirst.cpu().numpy()
    lists = _csr_lists(off, idx)
    for b in range(len(boxes)):
        want, wfirst = O.voxel_downsample(sc.xyz[lists[b]], return_first=True)
        g = got[ds_off[b]:ds_off[b + 1]]
        f = first[ds_off[b]:ds_off[b + 1]]
        order = np.argsort(f, kind="stable")
        assert len(g) == len(want)
        assert np.array_equal(f[order], wfirst)
        assert np.array_equal(g[order].view(np.uint32), want.view(np.uint32))
    assert any(len(l) > 0 and ds_off[b + 1] - ds_off[b] < len(l) for b, l in enumerate(lists))
    # rows as the only source (what the host-buffer entry uploads): blocks stored back to back
    rows = _dev_xyz(sc)[idx.to(torch.int64)].contiguous()
    got_rows = D.voxel_downsample(None, None, None, off, None, n_entries=idx.shape[0], xyz_aos=rows)
    for a, b in zip(got_rows, (torch.from_numpy(ds_off).to(DEV), dx, dy, dz, dfirst)):
        assert torch.equal(a, b)


def _vis_compare(vis, ds_lists, cam_lists, recs, thr=O.VISIBLE_RATE):
    """Per pair against the oracle.  Exact pairs: ratio to 1e-12; pairs decided
    from the area bounds (flags bit4): keep identical and lo <= exact <= hi."""
    n_hit = vis.n_hit.cpu().numpy()
    ratio = vis.ratio.cpu().numpy()
    keep = vis.keep.cpu().numpy().astype(bool)
    flags = vis.flags.cpu().numpy()
    rlo = vis.ratio_lo.cpu().numpy()
    rhi = vis.ratio_hi.cpu().numpy()
    p = 0
    n_kept = 0
    for b, cl in enumerate(cam_lists):
        for c in cl:
            nh, rt, kp = O.camera_visibility(ds_lists[b], recs[c], threshold=thr)
            assert n_hit[p] == nh, (b, c)
            if flags[p] & 2:          # first-ray collinear: atan2-tie dependent, reported apart
                p += 1
                continue
            if nh > O.MIN_HITS:
                hit, u, v = O.project_port(ds_lists[b], recs[c])
                collinear = O.first_ray_collinear(np.stack([u[hit], v[hit]], 1).astype(np.float64))
                if not collinear:
                    assert rlo[p] <= rt * (1 + 1e-12) + 1e-15, (b, c, rlo[p], rt)
                    assert rhi[p] >= rt * (1 - 1e-12), (b, c, rhi[p], rt)
            if flags[p] & 16:
                assert keep[p] == kp, (b, c, ratio[p], rt)
            else:
                assert abs(ratio[p] - rt) <= 1e-12 * max(1.0, abs(rt)), (b, c, ratio[p], rt)
                if not (flags[p] & 1):
                    assert keep[p] == kp, (b, c, ratio[p], rt)
            n_kept += int(kp)
            p += 1
    assert p == len(n_hit)
    return n_kept


@pytest.mark.parametrize("seed,obl", [(41, 0.0), (42, 0.5)])
def test_visibility_blocks(seed, obl):
    sc = small_scene(n=120000, c=60, thr=20000, seed=seed, oblique_fraction=obl)
    cams = scene_cameras(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    assert np.array_equal(recs[3], O.camera_record(cams.R[3], cams.T[3], cams.FoVx[3], cams.FoVy[3],
                                                   sc.width, sc.height))
    centers = D.camera_centers(cams.R, cams.T)
    res = D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold, exact_ratio=True)
    assert not bool((res.vis.flags & 16).any())
    ds_off = res.ds_offsets.cpu().numpy()
    ds = np.stack([res.ds_x.cpu().numpy(), res.ds_y.cpu().numpy(), res.ds_z.cpu().numpy()], 1)
    ds_lists = [ds[ds_off[b]:ds_off[b + 1]] for b in range(len(ds_off) - 1)]
    cam_lists = _csr_lists(res.cam_offsets, res.cam_index)
    assert sum(len(c) for c in cam_lists) > 30
    n_kept = _vis_compare(res.vis, ds_lists, cam_lists, recs)
    kept = _csr_lists(res.kept_offsets, res.kept_index)
    keep = res.vis.keep.cpu().numpy().astype(bool)
    p = 0
    for b, cl in enumerate(cam_lists):
        assert np.array_equal(kept[b], cl[keep[p:p + len(cl)]])
        p += len(cl)
    assert (res.vis.n_hit > 3).sum().item() > 10
    # a threshold that splits the cameras into kept and dropped
    cams_d = torch.from_numpy(recs).to(DEV)
    vis2 = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets,
                        res.cam_index, threshold=0.93, cam_split=3)
    k2 = _vis_compare(vis2, ds_lists, cam_lists, recs, thr=0.93)
    assert 0 < k2 < len(vis2.keep)
    assert torch.equal(vis2.n_hit, res.vis.n_hit)
    # default mode: pairs decided by the area bounds skip the exact hull; the
    # keep decisions are identical to the exact pass for every threshold
    for thr in (0.2, O.VISIBLE_RATE, 0.7, 0.93):
        ve = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets,
                          res.cam_index, threshold=thr, exact_ratio=True)
        vd = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets,
                          res.cam_index, threshold=thr)
        assert torch.equal(ve.keep, vd.keep), thr
        assert vd.n_exact <= ve.n_exact
        ex = (vd.flags & 16) == 0
        assert torch.equal(ve.ratio[ex], vd.ratio[ex])
        assert bool((vd.ratio_lo <= ve.ratio * (1 + 1e-12) + 1e-15).all())
        assert bool((vd.ratio_hi >= ve.ratio * (1 - 1e-12)).all())
    _vis_compare(vd, ds_lists, cam_lists, recs, thr=0.93)


def _unit_camera(W=64.0, H=64.0, front_axis=1):
    """u = x/z, v = y/z exactly for z = 1: adversarial image-plane inputs."""
    rec = np.zeros(20, dtype=np.float32)
    rec[[0, 4, 8]] = 1.0
    rec[12] = rec[13] = 1.0
    rec[16], rec[17], rec[18] = W, H, front_axis
    return rec


def _run_single(points, recs):
    pts = np.asarray(points, dtype=np.float32)
    x, y, z = D.aos_to_soa(torch.from_numpy(pts).to(DEV))
    off = torch.tensor([0, len(pts)], dtype=torch.int64, device=DEV)
    cams = torch.from_numpy(np.stack(recs)).to(DEV)
    coff = torch.tensor([0, len(recs)], dtype=torch.int64, device=DEV)
    cidx = torch.arange(len(recs), dtype=torch.int32, device=DEV)
    return D.visibility(x, y, z, off, cams, coff, cidx, cand_capacity=8, exact_ratio=True, cand_pool=300), pts


def test_visibility_adversarial_hulls():
    rng = np.random.default_rng(9)
    cases = []
    # pivot row shared by several hits (farthest one is skipped by the reference)
    a = np.array([[2, 1, 1], [60, 1, 1], [30, 1, 1], [61, 30, 1], [3, 50, 1], [40, 63, 1], [20, 20, 1]], float)
    cases.append(a)
    # duplicates of the pivot and of other vertices
    cases.append(np.vstack([a, a[[0, 0, 3, 4]]]))
    # everything collinear -> area 0
    cases.append(np.array([[i, i + 1, 1] for i in range(1, 40)], float))
    # integer grid (massive collinearity, pivot-row ties)
    g = np.stack(np.meshgrid(np.arange(1, 30), np.arange(1, 25)), -1).reshape(-1, 2)
    cases.append(np.column_stack([g, np.ones(len(g))]).astype(float))
    # points on a circle (every point is a hull vertex -> candidate overflow path)
    t = rng.uniform(0, 2 * np.pi, 700)
    cases.append(np.column_stack([32 + 30 * np.cos(t), 32 + 30 * np.sin(t), np.ones(700)]))
    # <= 3 hits
    cases.append(np.array([[1, 1, 1], [5, 9, 1], [9, 2, 1]], float))
    # y <= 0 (not in front for the reference's Y test), z < 0 mirrored points
    cases.append(np.array([[5, -1, 1], [5, 0, 1], [-10, 3, -1], [-20, 8, -1], [-30, 4, -1], [-12, 30, -1],
                           [-40, 40, -1], [7, 7, 0]], float))
    # random cloud with a few thousand hits
    cases.append(np.column_stack([rng.uniform(-5, 70, 6000), rng.uniform(-5, 70, 6000), np.ones(6000)]))
    rec = _unit_camera()
    for k, pts in enumerate(cases):
        vis, p32 = _run_single(pts, [rec, _unit_camera(40.0, 48.0), _unit_camera(64.0, 64.0, 2)])
        for j, r in enumerate([rec, _unit_camera(40.0, 48.0), _unit_camera(64.0, 64.0, 2)]):
            nh, rt, kp = O.camera_visibility(p32, r)
            assert int(vis.n_hit[j]) == nh, (k, j)
            fl = int(vis.flags[j])
            assert not (fl & 12), (k, j, fl)
            hit, u, v = O.project_port(p32, r)
            uv = np.stack([u[hit], v[hit]], 1).astype(np.float64)
            if nh > 3 and O.first_ray_collinear(uv):
                if fl & 2:
                    continue
                # unflagged is only legitimate when every hit is collinear
                # (degenerate hull: area 0 whatever the tie order)
                assert rt == 0.0 and float(vis.ratio[j]) == 0.0, (k, j)
                continue
            assert not (fl & 2), (k, j)
            assert abs(float(vis.ratio[j]) - rt) <= 1e-12 * max(1.0, rt), (k, j, float(vis.ratio[j]), rt)
            if not (fl & 1):
                assert bool(vis.keep[j]) == kp, (k, j)


def test_candidate_pass_variants_agree():
    """The warp-per-run candidate pass (default) and the tile-staged sweep (PTGS_SWEEP2=tiles) emit
    the same candidate sets: identical counts per pair and identical hull results."""
    import os
    sc = small_scene(n=140000, c=70, thr=20000, seed=43, oblique_fraction=0.4)
    cams = scene_cameras(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    centers = D.camera_centers(cams.R, cams.T)
    res = D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold, do_visibility=False)
    ds_off, dx, dy, dz, _ = D.voxel_downsample(*D.aos_to_soa(_dev_xyz(sc)), res.pt_offsets, res.pt_index)
    cams_d = torch.from_numpy(recs).to(DEV)
    out = {}
    for mode in ("runs", "tiles"):
        for exact in (True, False):
            if mode == "tiles":
                os.environ["PTGS_SWEEP2"] = "tiles"
            try:
                out[mode, exact] = D.visibility(dx, dy, dz, ds_off, cams_d, res.cam_offsets, res.cam_index,
                                                exact_ratio=exact)
            finally:
                os.environ.pop("PTGS_SWEEP2", None)
    # the sample pass as a tile sweep instead of one warp per pair: a different (equally valid) interior
    # polygon at most, identical results
    os.environ["PTGS_SAMPLE_PASS"] = "sweep"
    try:
        alt = D.visibility(dx, dy, dz, ds_off, cams_d, res.cam_offsets, res.cam_index, exact_ratio=True)
    finally:
        os.environ.pop("PTGS_SAMPLE_PASS", None)
    for f in ("n_hit", "area", "ratio", "keep", "flags"):
        assert torch.equal(getattr(alt, f), getattr(out["runs", True], f)), f
    for exact in (True, False):
        a, b = out["runs", exact], out["tiles", exact]
        if exact:
            assert int(a.cand_count.sum()) > 1000 and a.n_exact == len(a.keep)
        for f in ("n_hit", "area", "ratio", "keep", "flags", "cand_count"):
            assert torch.equal(getattr(a, f), getattr(b, f)), (f, exact)


def test_visibility_hull_prefilter_large_lists():
    """Pairs with long candidate lists go through the multi-CTA pre-filter before the hull: the
    answers equal those of the single-CTA path bit for bit, and the oracle's."""
    rng = np.random.default_rng(23)
    n_disc, n_ring = 250000, 150000
    r = np.concatenate([29.0 * np.sqrt(rng.uniform(0, 1, n_disc)), rng.uniform(29.0, 30.0, n_ring)])
    t = rng.uniform(0, 2 * np.pi, len(r))
    pts = np.column_stack([32 + r * np.cos(t), 32 + r * np.sin(t), np.ones(len(r))]).astype(np.float32)
    recs = [_unit_camera(), _unit_camera(56.0, 60.0), _unit_camera(64.0, 64.0, 2)]
    x, y, z = D.aos_to_soa(torch.from_numpy(pts).to(DEV))
    off = torch.tensor([0, len(pts)], dtype=torch.int64, device=DEV)
    cams = torch.from_numpy(np.stack(recs)).to(DEV)
    coff = torch.tensor([0, len(recs)], dtype=torch.int64, device=DEV)
    cidx = torch.arange(len(recs), dtype=torch.int32, device=DEV)
    out = {}
    for pf in (True, False):
        out[pf] = D.visibility(x, y, z, off, cams, coff, cidx, exact_ratio=True, hull_prefilter=pf)
    a, b = out[True], out[False]
    assert int(a.cand_count.max()) > 3 * 6144, a.cand_count      # several chunks per pair
    assert not bool((a.flags & 12).any())
    for f in ("n_hit", "area", "ratio", "keep", "flags", "cand_count"):
        assert torch.equal(getattr(a, f), getattr(b, f)), f
    for j, rec in enumerate(recs):
        nh, rt, kp = O.camera_visibility(pts, rec)
        assert int(a.n_hit[j]) == nh
        assert abs(float(a.ratio[j]) - rt) <= 1e-12 * max(1.0, rt), (j, float(a.ratio[j]), rt)


def test_coverage_sweep():
    sc = small_scene(n=70001, c=37, thr=50000, seed=51)
    cams = scene_cameras(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    recs[::3, 18] = 2.0   # mix of Y-front (reference) and Z-front cameras
    x, y, z = D.aos_to_soa(_dev_xyz(sc))
    cams_d = torch.from_numpy(recs).to(DEV)
    n_hit, seen = D.coverage_sweep(x, y, z, cams_d, want_seen=True)
    want, wseen = O.coverage_counts(sc.xyz, recs, want_seen=True)
    assert np.array_equal(n_hit.cpu().numpy(), want)
    assert np.array_equal(seen.cpu().numpy().astype(bool), wseen)
    part = D.coverage_sweep(x, y, z, cams_d, cam_first=5, cam_count=11)
    assert np.array_equal(part.cpu().numpy(), want[5:16])
    assert want.sum() > 1000
    # file order (no spatial sort) and a reused ordering give the same answers
    n2, s2 = D.coverage_sweep(x, y, z, cams_d, want_seen=True, spatial_order_points=False)
    assert torch.equal(n2, n_hit) and torch.equal(s2, seen)
    od = D.spatial_order(x, y, z)
    n3, s3 = D.coverage_sweep(x, y, z, cams_d, want_seen=True, ordered=od)
    assert torch.equal(n3, n_hit) and torch.equal(s3, seen)
    od2 = D.spatial_order(x, y, z, xyz_aos=_dev_xyz(sc))       # gather from the (N,3) rows: same order
    assert all(torch.equal(a, b) for a, b in zip(od, od2))


def test_partition_results_to_host_early_copy():
    """partition_device(host_results=True) starts the index list's device -> host copy under A6-A9;
    the host result set equals the one copied afterwards."""
    sc = small_scene(n=90000, c=40, thr=12000, seed=71)
    cams = scene_cameras(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    centers = D.camera_centers(cams.R, cams.T)
    a = D.partition_results_to_host(D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold, host_results=True))
    b = D.partition_results_to_host(D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold))
    assert set(a) == set(b) and len(a["pt_index"]) > len(sc.xyz) // 2
    for k in a:
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k


def test_spatial_order_is_a_permutation_and_local():
    rng = np.random.default_rng(7)
    n = 300001
    pts = rng.uniform(-50, 50, size=(n, 3)).astype(np.float32)
    pts[:, 2] *= 0.1
    pts[5] = [np.nan, 1.0, 2.0]            # non-finite coordinates must not break the grid
    pts[6] = [np.inf, -np.inf, 0.0]
    x, y, z = (torch.from_numpy(np.ascontiguousarray(pts[:, a])).to(DEV) for a in range(3))
    sx, sy, sz, perm = D.spatial_order(x, y, z)
    p = perm.cpu().numpy().view(np.uint32).astype(np.int64)
    assert np.array_equal(np.sort(p), np.arange(n))
    got = np.stack([sx.cpu().numpy(), sy.cpu().numpy(), sz.cpu().numpy()], 1)
    assert np.array_equal(got.view(np.uint32), pts[p].view(np.uint32))
    # locality: 256-point runs of the sorted cloud are far smaller than the cloud
    fin = np.isfinite(got).all(1)
    runs = got[: (n // 256) * 256].reshape(-1, 256, 3)
    ok = fin[: (n // 256) * 256].reshape(-1, 256).all(1)
    ext = (runs.max(1) - runs.min(1))[ok]
    assert np.median(ext[:, 0]) < 10.0 and np.median(ext[:, 1]) < 10.0
    # empty input
    e = torch.empty(0, dtype=torch.float32, device=DEV)
    assert D.spatial_order(e, e, e)[3].numel() == 0


def test_full_partition_vs_oracle():
    sc = small_scene(n=150000, c=80, thr=25000, seed=61)
    cams = scene_cameras(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    centers = D.camera_centers(cams.R, cams.T)
    res = D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold, exact_ratio=True)
    want = O.run_partition(sc.xyz, cams, sc.threshold, centers=centers)
    assert np.array_equal(res.keep_mask.cpu().numpy().astype(bool), want.keep_mask)
    assert res.root_bounds == want.root_bounds
    assert res.leaf_ids == [l.partition_id for l in want.leaves]
    assert [tuple(b) for b in res.leaf_bounds.tolist()] == [l.bounds for l in want.leaves]
    assert [int(c) for c in res.leaf_count] == [l.point_num for l in want.leaves]
    for g, w in zip(_csr_lists(res.pt_offsets, res.pt_index), want.block_index):
        assert np.array_equal(g, w)
    for g, w in zip(_csr_lists(res.cam_offsets, res.cam_index), want.cam_index):
        assert np.array_equal(g, w)
    n_hit = res.vis.n_hit.cpu().numpy()
    assert np.array_equal(n_hit, np.concatenate(want.n_hit))
    flags = res.vis.flags.cpu().numpy()
    ok = (flags & 3) == 0
    assert np.array_equal(res.vis.keep.cpu().numpy().astype(bool)[ok], np.concatenate(want.keep)[ok])
    np.testing.assert_allclose(res.vis.ratio.cpu().numpy()[ok], np.concatenate(want.ratio)[ok], rtol=1e-12, atol=1e-15)
    kept = _csr_lists(res.kept_offsets, res.kept_index)
    if ok.all():
        for g, w in zip(kept, want.kept_cameras):
            assert set(g.tolist()) == set(w.tolist())
        # the default (bounds-decided) mode selects the same cameras
        res2 = D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold)
        for g, w in zip(_csr_lists(res2.kept_offsets, res2.kept_index), want.kept_cameras):
            assert set(g.tolist()) == set(w.tolist())


def test_voxel_downsample_float64():
    """vox_*_kernel<double>: the instantiation every Manhattan configuration (bench c2) runs.
    Reference: densify:572-605 on a float64 cloud (shen_data_read.py:126-130)."""
    sc = small_scene(n=50000, c=8, thr=9000, seed=32)
    sc.xyz[:, 2] = np.round(sc.xyz[:, 2] * 2) / 2
    sc.xyz[:, :2] = np.round(sc.xyz[:, :2] * 8) / 8
    man = np.eye(4)
    a = np.deg2rad(21.0)
    man[:2, :2] = [[np.cos(a), -np.sin(a)], [np.sin(a), np.cos(a)]]
    man[:3, 3] = [3.0, -1.5, 0.25]
    x, y, z = D.manhattan_apply(*D.aos_to_soa(_dev_xyz(sc)), man)
    pts = O.manhattan_points(sc.xyz, man)
    assert x.dtype == torch.float64
    assert np.array_equal(np.stack([x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy()], 1), pts)
    boxes = np.array([[-30, -30, 5, 5], [0, 0, 40, 40], [100, 100, 101, 101], [-5, -5, 5, 5.0]])
    off, idx = D.expand_select(x, y, boxes)
    ds_off, dx, dy, dz, dfirst = D.voxel_downsample(x, y, z, off, idx)
    assert dx.dtype == torch.float32
    ds_off = ds_off.cpu().numpy()
    got = np.stack([dx.cpu().numpy(), dy.cpu().numpy(), dz.cpu().numpy()], 1)
    first = dfirst.cpu().numpy()
    lists = _csr_lists(off, idx)
    for b in range(len(boxes)):
        want, wfirst = O.voxel_downsample(pts[lists[b]], return_first=True)
        g = got[ds_off[b]:ds_off[b + 1]]
        f = first[ds_off[b]:ds_off[b + 1]]
        order = np.argsort(f, kind="stable")
        assert len(g) == len(want)
        assert np.array_equal(f[order], wfirst)
        assert np.array_equal(g[order].view(np.uint32), want.view(np.uint32))
    assert any(len(l) > 0 and ds_off[b + 1] - ds_off[b] < len(l) for b, l in enumerate(lists))


def _assert_partition_equals_oracle(res, want, exact=True):
    assert np.array_equal(res.keep_mask.cpu().numpy().astype(bool), want.keep_mask)
    assert res.root_bounds == want.root_bounds
    assert res.leaf_ids == [l.partition_id for l in want.leaves]
    assert [tuple(b) for b in res.leaf_bounds.tolist()] == [l.bounds for l in want.leaves]
    assert [int(c) for c in res.leaf_count] == [l.point_num for l in want.leaves]
    assert [tuple(b) for b in res.boxes.tolist()] == list(want.boxes)
    for g, w in zip(_csr_lists(res.pt_offsets, res.pt_index), want.block_index):
        assert np.array_equal(g, w)
    for g, w in zip(_csr_lists(res.cam_offsets, res.cam_index), want.cam_index):
        assert np.array_equal(g, w)
    # A6: float32 bit patterns per block in first-occurrence order
    ds_off = res.ds_offsets.cpu().numpy()
    ds = np.stack([res.ds_x.cpu().numpy(), res.ds_y.cpu().numpy(), res.ds_z.cpu().numpy()], 1)
    first = res.ds_first.cpu().numpy()
    for b, wd in enumerate(want.ds_points):
        order = np.argsort(first[ds_off[b]:ds_off[b + 1]], kind="stable")
        assert np.array_equal(ds[ds_off[b]:ds_off[b + 1]][order].view(np.uint32), wd.view(np.uint32)), b
    assert np.array_equal(res.vis.n_hit.cpu().numpy(), np.concatenate(want.n_hit))
    flags = res.vis.flags.cpu().numpy()
    ok = (flags & 3) == 0
    assert np.array_equal(res.vis.keep.cpu().numpy().astype(bool)[ok], np.concatenate(want.keep)[ok])
    if exact:
        np.testing.assert_allclose(res.vis.ratio.cpu().numpy()[ok], np.concatenate(want.ratio)[ok],
                                   rtol=1e-12, atol=1e-15)
    if ok.all():
        for g, w in zip(_csr_lists(res.kept_offsets, res.kept_index), want.kept_cameras):
            assert set(g.tolist()) == set(w.tolist())
    return int(ok.sum())


@pytest.mark.parametrize("semantic", ["consistent", "literal"])
def test_full_partition_manhattan_vs_oracle(semantic):
    """run_DataPartition under a Manhattan transform, A1-A9 end to end on a float64 cloud
    (what bench config c2 times), against the oracle:
      consistent  cloud and cameras rotated before A1 (the declared c2 semantic);
      literal     the shen path as written -- A1-A3 on the un-rotated PLY (utils-shen:62),
                  A4 / A6-A9 on the rotated self.pcd (:86, :91), cameras as read."""
    sc = small_scene(n=160000, c=80, thr=25000, seed=62, manhattan_deg=21.0, height_scale=6.0)
    man = sc.man_trans
    cams = scene_cameras(sc)
    rot = O.manhattan_points(sc.xyz, man)
    if semantic == "consistent":
        Rm, Tm = O.manhattan_cameras(cams.R, cams.T, man)
        cams = O.CameraSet(Rm, Tm, cams.FoVx, cams.FoVy, cams.width, cams.height)
        split_np, split_dev = rot, None
    else:
        split_np, split_dev = sc.xyz, _dev_xyz(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    centers = D.camera_centers(cams.R, cams.T)
    want = O.run_partition(rot, cams, sc.threshold, centers=centers, split_points=split_np)
    assert len(want.leaves) >= 4 and sum(len(c) for c in want.cam_index) > 40
    res = D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold, man_trans=man, split_xyz=split_dev,
                             exact_ratio=True)
    assert res.ds_x.dtype == torch.float32 and res.leaf_bounds.dtype == np.float64
    n_ok = _assert_partition_equals_oracle(res, want)
    assert n_ok > 40
    kept = np.concatenate(want.keep)
    assert 0 < kept.sum() < len(kept)
    # default (bound-decided) mode: same cameras
    res2 = D.partition_device(_dev_xyz(sc), recs, centers, sc.threshold, man_trans=man, split_xyz=split_dev)
    _assert_partition_equals_oracle(res2, want, exact=False)


@pytest.mark.parametrize("f64", [False, True])
def test_expand_grid_path_equals_lockstep(f64):
    """ptgs_expand_*_grid (cell-grid candidates) == ptgs_expand_* (every point x every rectangle),
    and both == the NumPy closed-box mask, for many overlapping rectangles."""
    rng = np.random.default_rng(23)
    n, L = 200003, 150
    pts = rng.uniform(-100, 100, (n, 2))
    pts[::97] = np.round(pts[::97])                      # points on rectangle edges
    dt = np.float64 if f64 else np.float32
    pts = pts.astype(dt)
    pts[11] = [np.nan, 0.0]
    pts[12] = [1e9, -1e9]                                # far outside every rectangle
    lo = np.round(rng.uniform(-100, 80, (L, 2)))
    hi = lo + np.round(rng.uniform(1, 40, (L, 2)))
    boxes = np.stack([lo[:, 0], lo[:, 1], hi[:, 0], hi[:, 1]], 1).astype(np.float64)
    x = torch.from_numpy(np.ascontiguousarray(pts[:, 0])).to(DEV)
    y = torch.from_numpy(np.ascontiguousarray(pts[:, 1])).to(DEV)
    o1, i1 = D.expand_select(x, y, boxes, use_grid=False)
    o2, i2 = D.expand_select(x, y, boxes, use_grid=True)
    assert D._cell_grid(np.stack([boxes[:, 0], boxes[:, 2], boxes[:, 1], boxes[:, 3]], 1)) is not None
    assert torch.equal(o1, o2) and torch.equal(i1, i2)
    off, idx = o2.cpu().numpy(), i2.cpu().numpy()
    for b in range(L):
        m = (pts[:, 0] >= dt(boxes[b, 0])) & (pts[:, 0] <= dt(boxes[b, 2])) & \
            (pts[:, 1] >= dt(boxes[b, 1])) & (pts[:, 1] <= dt(boxes[b, 3]))
        assert np.array_equal(idx[off[b]:off[b + 1]], np.nonzero(m)[0]), b
    assert off[-1] > n // 2


def test_coverage_sweep_many_cameras_supertile_lists():
    """More than 320 cameras: ptgs_coverage_sweep builds per super-tile camera lists first.
    Counts, seen flags and a camera sub-range against the oracle, in file and Morton order."""
    sc = small_scene(n=90001, c=420, thr=50000, seed=53, oblique_fraction=0.3)
    cams = scene_cameras(sc)
    recs = D.camera_records(cams.R, cams.T, cams.FoVx, cams.FoVy, cams.width, cams.height)
    recs[::2, 18] = 2.0
    x, y, z = D.aos_to_soa(_dev_xyz(sc))
    cams_d = torch.from_numpy(recs).to(DEV)
    want, wseen = O.coverage_counts(sc.xyz, recs, want_seen=True)
    for order in (True, False):
        n_hit, seen = D.coverage_sweep(x, y, z, cams_d, want_seen=True, spatial_order_points=order)
        assert np.array_equal(n_hit.cpu().numpy(), want)
        assert np.array_equal(seen.cpu().numpy().astype(bool), wseen)
    part = D.coverage_sweep(x, y, z, cams_d, cam_first=17, cam_count=380)
    assert np.array_equal(part.cpu().numpy(), want[17:397])
    assert want.sum() > 10000 and (want == 0).sum() < len(want)
