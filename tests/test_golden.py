"""Golden-vector tests.  tests/golden/*.npz hold inputs and the outputs of the
REFERENCE's own code (made by oracle/make_golden.py in the build container,
where /root/reference exists).  They travel to the GPU box, so

  * CPU (`-m "not gpu"`): the oracle restatement is pinned against them;
  * GPU (`-m gpu`): the CUDA path (through the C ABI) is compared with the
    reference's outputs directly, not only with the oracle.

Bit-exact: ids, split rectangles, counts, expanded index sets, camera
assignment, voxel means (fp32 bit patterns), hit counts, kept camera sets.
Projected (u, v): <= 1e-5 relative (north star).  Hull ratio: 1e-12 relative
(fp64 summation order)."""
import os

import numpy as np
import pytest

import ptgs_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SPLITS = ["uniform", "ties", "single_leaf", "depth_cap"]


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def _lists(off, idx):
    return [idx[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def _camset(g):
    return O.CameraSet(g["R"], g["T"], g["FoVx"], g["FoVy"], g["width"], g["height"])


def _records(g):
    return np.stack([O.camera_record(g["R"][i], g["T"][i], g["FoVx"][i], g["FoVy"][i],
                                     int(g["width"][i]), int(g["height"][i])) for i in range(len(g["R"]))])


# ----------------------------------------------------------------- oracle (CPU)
@pytest.mark.parametrize("name", SPLITS)
def test_oracle_split(name):
    g = _load("split_" + name)
    pts = g["points"].astype(np.float64)
    assert O.bounds_xy(pts) == tuple(g["root"].tolist())
    leaves = O.quad_tree_partition(pts, tuple(g["root"].tolist()), int(g["threshold"]))
    assert [l.partition_id for l in leaves] == g["ids"].tolist()
    assert np.array_equal(np.array([l.bounds for l in leaves]).reshape(-1, 4), g["bounds"])
    assert [l.point_num for l in leaves] == g["point_num"].tolist()


def test_oracle_expand_assign():
    g = _load("expand_assign")
    pts = g["points"]
    leaves = O.quad_tree_partition(pts.astype(np.float64), tuple(g["root"].tolist()), int(g["threshold"]))
    assert [l.partition_id for l in leaves] == g["ids"].tolist()
    assert np.array_equal(np.array([l.bounds for l in leaves]), g["origin_bounds"])
    boxes = [O.expanded_box(l.bounds) for l in leaves]
    assert np.array_equal(np.array(boxes), g["extend_bounds"])
    want = _lists(g["pt_offsets"], g["pt_index"])
    for box, w, n in zip(boxes, want, g["point_num"]):
        m = O.expand_mask(pts, box)
        assert np.array_equal(np.nonzero(m)[0], w) and int(m.sum()) == int(n)
    centers = O.camera_centers(g["R"], g["T"])
    assert np.array_equal(centers, g["centers"])
    for got, w in zip(O.assign_cameras(boxes, centers), _lists(g["cam_offsets"], g["cam_index"])):
        assert np.array_equal(got, w)


def test_oracle_voxel():
    g = _load("voxel")
    got = O.voxel_downsample(g["points"], float(g["voxel"]))
    assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), g["out"].view(np.uint32))


def test_oracle_projection():
    g = _load("projection")
    recs = _records(g)
    off = g["uv_offsets"]
    exact = 0
    for i in range(len(recs)):
        want = g["uv"][off[i]:off[i + 1]]
        ref_style = O.project_points_to_camera_ref(g["points"], g["R"][i], g["T"][i], g["FoVx"][i], g["FoVy"][i],
                                                   int(g["width"][i]), int(g["height"][i]))
        assert np.array_equal(ref_style, want)
        hit, u, v = O.project_port(g["points"], recs[i])
        assert abs(int(hit.sum()) - len(want)) <= 2
        if int(hit.sum()) == len(want):
            uv = np.stack([u[hit], v[hit]], 1)
            assert np.allclose(uv, want, rtol=1e-5, atol=1e-5 * float(g["width"][i]))
            exact += int(np.array_equal(uv, want))
    assert off[-1] > 500 and exact >= len(recs) - 1


def test_oracle_graham():
    g = _load("graham")
    off = g["pts_offsets"]
    for i, want in enumerate(g["rate"]):
        pts = g["pts"][off[i]:off[i + 1]].astype(np.float64)
        got = O.run_graham_scan(pts, float(g["width"]), float(g["height"]))
        assert got == pytest.approx(float(want), rel=1e-13, abs=1e-15)


def test_oracle_selection():
    g = _load("selection")
    res = O.run_partition(g["points"], _camset(g), int(g["threshold"]), radius_filter=False)
    assert [l.partition_id for l in res.leaves] == g["ids"].tolist()
    assert [len(b) for b in res.block_index] == g["point_num"].tolist()
    for got, w in zip(res.cam_index, _lists(g["cam_offsets"], g["cam_index"])):
        assert np.array_equal(got, w)
    assert np.array_equal(np.concatenate(res.n_hit), g["n_hit"])
    np.testing.assert_allclose(np.concatenate(res.ratio), g["rate"], rtol=1e-12, atol=1e-15)
    for got, w in zip(res.kept_cameras, _lists(g["kept_offsets"], g["kept_index"])):
        assert sorted(got.tolist()) == w.tolist()
    kept = int(g["kept_offsets"][-1])
    assert 0 < kept < int(g["cam_offsets"][-1])


def _man_cams(g, sem):
    """Camera arrays of the fixture's semantic: consistent = W2C @ inv(man_trans), literal = as read."""
    if sem == "consistent":
        return O.CameraSet(g["R_rot"], g["T_rot"], g["FoVx"], g["FoVy"], g["width"], g["height"])
    return _camset(g)


def test_oracle_manhattan_rotation():
    """oracle.manhattan_points / manhattan_cameras == the reference's fetchPly output
    (shen_data_read.py:126-130, float64) and dataset_readers.py:96-104."""
    g = _load("manhattan")
    rot = O.manhattan_points(g["points"], g["man_trans"])
    assert rot.dtype == np.float64 and np.array_equal(rot, g["rotated"])
    Rm, Tm = O.manhattan_cameras(g["R"], g["T"], g["man_trans"])
    assert np.array_equal(Rm, g["R_rot"]) and np.array_equal(Tm, g["T_rot"])


@pytest.mark.parametrize("sem", ["consistent", "literal"])
def test_oracle_manhattan_selection(sem):
    """A3-A9 on a float64 (Manhattan-rotated) cloud against the reference's own run, under both
    semantics of SURVEY.md A13 (literal: split on the un-rotated PLY, utils-shen:62 vs :86)."""
    g = _load("manhattan")
    k = lambda name: g[sem + "_" + name]
    rot = g["rotated"]
    split = rot if sem == "consistent" else g["points"]
    res = O.run_partition(rot, _man_cams(g, sem), int(g["threshold"]), radius_filter=False, split_points=split)
    assert res.root_bounds == tuple(k("root").tolist())
    assert [l.partition_id for l in res.leaves] == k("ids").tolist()
    assert np.array_equal(np.array([l.bounds for l in res.leaves]), k("leaf_bounds"))
    assert [l.point_num for l in res.leaves] == k("leaf_num").tolist()
    assert np.array_equal(np.array(res.boxes), k("extend_bounds"))
    for got, w in zip(res.block_index, _lists(k("pt_offsets"), k("pt_index"))):
        assert np.array_equal(got, w)
    for got, w in zip(res.cam_index, _lists(k("cam_offsets"), k("cam_index"))):
        assert np.array_equal(got, w)
    # A6 on float64 input: float32 bit patterns, first-occurrence order (densify:572-605)
    for got, w in zip(res.ds_points, _lists(k("ds_offsets"), k("ds_points"))):
        assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), w.view(np.uint32))
    assert np.array_equal(np.concatenate(res.n_hit), k("n_hit"))
    np.testing.assert_allclose(np.concatenate(res.ratio), k("rate"), rtol=1e-12, atol=1e-15)
    for got, w in zip(res.kept_cameras, _lists(k("kept_offsets"), k("kept_index"))):
        assert sorted(got.tolist()) == w.tolist()
    assert 0 < int(k("kept_offsets")[-1]) < int(k("cam_offsets")[-1])


def test_golden_output_layout():
    """The reference's on-disk layout for the selection fixture (A10)."""
    g = _load("selection")
    files = set(g["files"].tolist())
    for pid in g["ids"].tolist():
        assert "%s/partition_%s.pkl" % (pid, pid) in files
        for f in ("images.bin", "points3D.bin", "cameras.bin"):
            assert "%s/partition_%s/sparse/0/%s" % (pid, pid, f) in files


# ------------------------------------------------------------------- CUDA (GPU)
DEV = "cuda:0"


def _soa(pts):
    import torch
    from ptgs_b200 import device as D
    return D.aos_to_soa(torch.from_numpy(np.ascontiguousarray(pts, dtype=np.float32)).to(DEV))


@pytest.mark.gpu
@pytest.mark.parametrize("name", SPLITS)
def test_cuda_split(name):
    import torch
    from ptgs_b200 import device as D
    g = _load("split_" + name)
    x, y, z = _soa(g["points"])
    nf = torch.full((1,), len(g["points"]), dtype=torch.int64, device=DEV)
    root = D.bounds_xy(x, y, nf)
    assert np.array_equal(root.cpu().numpy(), g["root"])
    lb, lp, ld, lc = D.balanced_split(x, y, nf, root, int(g["threshold"]))
    assert [D.leaf_id(p, d) for p, d in zip(lp, ld)] == g["ids"].tolist()
    assert np.array_equal(np.asarray(lb, np.float64).reshape(-1, 4), g["bounds"])
    assert [int(c) for c in lc] == g["point_num"].tolist()


@pytest.mark.gpu
def test_cuda_expand_assign():
    from ptgs_b200 import device as D
    g = _load("expand_assign")
    x, y, z = _soa(g["points"])
    boxes = D.expanded_boxes(g["origin_bounds"])
    assert np.array_equal(boxes, g["extend_bounds"])
    off, idx = D.expand_select(x, y, boxes)
    assert np.array_equal(off.cpu().numpy(), g["pt_offsets"])
    assert np.array_equal(idx.cpu().numpy().astype(np.int64), g["pt_index"])
    centers = D.camera_centers(g["R"], g["T"])
    assert np.array_equal(centers, g["centers"])
    coff, cidx = D.assign_cameras(centers, boxes, DEV)
    assert np.array_equal(coff.cpu().numpy(), g["cam_offsets"])
    assert np.array_equal(cidx.cpu().numpy().astype(np.int64), g["cam_index"])


@pytest.mark.gpu
def test_cuda_voxel():
    import torch
    from ptgs_b200 import device as D
    g = _load("voxel")
    x, y, z = _soa(g["points"])
    n = len(g["points"])
    off = torch.tensor([0, n], dtype=torch.int64, device=DEV)
    idx = torch.arange(n, dtype=torch.int32, device=DEV)
    ds_off, dx, dy, dz, first = D.voxel_downsample(x, y, z, off, idx, voxel=float(g["voxel"]))
    got = np.stack([dx.cpu().numpy(), dy.cpu().numpy(), dz.cpu().numpy()], 1)
    order = np.argsort(first.cpu().numpy(), kind="stable")       # first-occurrence order (densify:576-581)
    assert np.array_equal(got[order].view(np.uint32), g["out"].view(np.uint32))


@pytest.mark.gpu
def test_cuda_projection_hits():
    """Hit counts of the CUDA sweep against the reference's (u, v) lists."""
    import torch
    from ptgs_b200 import device as D
    g = _load("projection")
    recs = D.camera_records(g["R"], g["T"], g["FoVx"], g["FoVy"], g["width"], g["height"])
    x, y, z = _soa(g["points"])
    n_hit = D.coverage_sweep(x, y, z, torch.from_numpy(recs).to(DEV)).cpu().numpy()
    want = np.diff(g["uv_offsets"])
    assert np.abs(n_hit - want).max() <= 2          # a projection within 1e-5 of an image edge may flip
    assert (n_hit == want).sum() >= len(want) - 1


@pytest.mark.gpu
def test_cuda_selection():
    import torch
    from ptgs_b200 import device as D
    g = _load("selection")
    recs = D.camera_records(g["R"], g["T"], g["FoVx"], g["FoVy"], g["width"], g["height"])
    centers = D.camera_centers(g["R"], g["T"])
    xyz = torch.from_numpy(g["points"]).to(DEV)
    for exact in (True, False):
        res = D.partition_device(xyz, recs, centers, int(g["threshold"]), radius_filter=False, exact_ratio=exact)
        assert res.leaf_ids == g["ids"].tolist()
        assert np.array_equal(np.diff(res.pt_offsets.cpu().numpy()), g["point_num"])
        assert np.array_equal(res.cam_offsets.cpu().numpy(), g["cam_offsets"])
        assert np.array_equal(res.cam_index.cpu().numpy().astype(np.int64), g["cam_index"])
        assert np.array_equal(res.vis.n_hit.cpu().numpy().astype(np.int64), g["n_hit"])
        flags = res.vis.flags.cpu().numpy()
        assert not (flags & 3).any()
        if exact:
            np.testing.assert_allclose(res.vis.ratio.cpu().numpy(), g["rate"], rtol=1e-12, atol=1e-15)
        koff = res.kept_offsets.cpu().numpy()
        kidx = res.kept_index.cpu().numpy().astype(np.int64)
        for got, w in zip(_lists(koff, kidx), _lists(g["kept_offsets"], g["kept_index"])):
            assert sorted(got.tolist()) == w.tolist()


@pytest.mark.gpu
@pytest.mark.parametrize("sem", ["consistent", "literal"])
def test_cuda_manhattan(sem):
    """partition_device(man_trans=...) -- float64 cloud from A13 on: Manhattan kernel, f64 split /
    expansion / voxel kernels, visibility -- against the reference's own run (tests/golden/manhattan.npz).
    literal = split_xyz=<un-rotated cloud> and un-rotated cameras (the shen path as written)."""
    import torch
    from ptgs_b200 import device as D
    g = _load("manhattan")
    k = lambda name: g[sem + "_" + name]
    cs = _man_cams(g, sem)
    recs = D.camera_records(cs.R, cs.T, cs.FoVx, cs.FoVy, cs.width, cs.height)
    centers = D.camera_centers(cs.R, cs.T)
    xyz = torch.from_numpy(g["points"]).to(DEV)
    x, y, z = D.manhattan_apply(*D.aos_to_soa(xyz), g["man_trans"])
    got_rot = np.stack([x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy()], 1)
    assert np.array_equal(got_rot, g["rotated"])
    for exact in (True, False):
        res = D.partition_device(xyz, recs, centers, int(g["threshold"]), man_trans=g["man_trans"],
                                 radius_filter=False, exact_ratio=exact,
                                 split_xyz=xyz if sem == "literal" else None)
        assert res.root_bounds == tuple(k("root").tolist())
        assert res.leaf_ids == k("ids").tolist()
        assert np.array_equal(res.leaf_bounds, k("leaf_bounds"))
        assert [int(c) for c in res.leaf_count] == k("leaf_num").tolist()
        assert np.array_equal(res.boxes, k("extend_bounds"))
        assert np.array_equal(res.pt_offsets.cpu().numpy(), k("pt_offsets"))
        assert np.array_equal(res.pt_index.cpu().numpy().astype(np.int64), k("pt_index"))
        assert np.array_equal(res.cam_offsets.cpu().numpy(), k("cam_offsets"))
        assert np.array_equal(res.cam_index.cpu().numpy().astype(np.int64), k("cam_index"))
        # A6 <double>: float32 bit patterns in the reference's first-occurrence order
        ds_off = res.ds_offsets.cpu().numpy()
        assert np.array_equal(ds_off, k("ds_offsets"))
        ds = np.stack([res.ds_x.cpu().numpy(), res.ds_y.cpu().numpy(), res.ds_z.cpu().numpy()], 1)
        first = res.ds_first.cpu().numpy()
        for b in range(len(ds_off) - 1):
            order = np.argsort(first[ds_off[b]:ds_off[b + 1]], kind="stable")
            assert np.array_equal(ds[ds_off[b]:ds_off[b + 1]][order].view(np.uint32),
                                  k("ds_points")[ds_off[b]:ds_off[b + 1]].view(np.uint32)), b
        assert np.array_equal(res.vis.n_hit.cpu().numpy().astype(np.int64), k("n_hit"))
        flags = res.vis.flags.cpu().numpy()
        assert not (flags & 3).any()
        if exact:
            np.testing.assert_allclose(res.vis.ratio.cpu().numpy(), k("rate"), rtol=1e-12, atol=1e-15)
        koff = res.kept_offsets.cpu().numpy()
        kidx = res.kept_index.cpu().numpy().astype(np.int64)
        for got, w in zip(_lists(koff, kidx), _lists(k("kept_offsets"), k("kept_index"))):
            assert sorted(got.tolist()) == w.tolist()
