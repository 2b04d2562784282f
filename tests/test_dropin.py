"""The drop-in layer (ptgs_b200.partition / compat/): the reference's function
names and types, driven the way shen_partition_utils.run_DataPartition drives
them (utils-shen:57-93), checked against the golden fixture produced by the
reference's own code and against the oracle."""
import os
import pickle
import sys

import numpy as np
import pytest

import ptgs_oracle as O
from helpers import scene_cameras, small_scene
import ptgs_b200  # noqa: F401
from ptgs_b200 import colmap_io, partition as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def _lists(off, idx):
    return [idx[off[i]:off[i + 1]] for i in range(len(off) - 1)]


def _cams_from(g):
    return [P.SimpleCamera(colmap_id=i + 1, R=g["R"][i], T=g["T"][i], FoVx=float(g["FoVx"][i]),
                           FoVy=float(g["FoVy"][i]), image_name="img_%06d" % i, uid=i,
                           width=int(g["width"][i]), height=int(g["height"][i])) for i in range(len(g["R"]))]


# ------------------------------------------------------------------------- CPU
def test_rect_semantics():
    r = P.Rect(0, 1, 2, 3)
    assert r.bounds == (0.0, 1.0, 2.0, 3.0) and all(isinstance(v, float) for v in r.bounds)
    assert r.contains(P.Point(1, 2)) and not r.contains(P.Point(0, 2)) and not r.contains(P.Point(1, 3))
    assert r.area == 4.0


def test_types_match_reference_fields():
    assert P.Partition._fields == ("partition_id", "camera", "point_cloud", "origin_box", "point_num",
                                   "extend_rate", "extend_box")
    assert P.BasicPointCloud._fields == ("points", "colors", "normals")
    assert P.CameraPose._fields == ("camera", "pose")
    assert P.Partition._field_defaults == {"extend_rate": 0, "extend_box": None}


def test_read_scene_roundtrip(tmp_path):
    sc = small_scene(n=3000, c=7, thr=1000, seed=3)
    colmap_io.write_synthetic_colmap(sc, str(tmp_path))
    si = P.read_scene(str(tmp_path))
    cams = scene_cameras(sc)
    assert [c.image_name for c in si.train_cameras] == sorted("img_%06d" % i for i in range(7))
    assert np.array_equal(np.stack([c.R for c in si.train_cameras]), cams.R)
    assert np.array_equal(np.stack([c.T for c in si.train_cameras]), cams.T)
    assert si.train_cameras[0].FovX == cams.FoVx[0]
    assert si.point_cloud.points.dtype == np.float32 and np.array_equal(si.point_cloud.points, sc.xyz)
    assert np.array_equal(si.point_cloud.colors, sc.rgb / 255.0)
    sims = P.cameraList_from_camInfos_partition(si.train_cameras)
    assert [s.uid for s in sims] == list(range(7)) and sims[0].colmap_id == 1
    assert np.array_equal(np.stack([s.camera_center.numpy() for s in sims]), O.camera_centers(cams.R, cams.T))
    assert os.path.exists(tmp_path / "sparse" / "0" / "points3D.ply")


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError):
        P.quad_tree_partition(np.zeros((10, 3)), P.Rect(0, 0, 1, 1), 5)


@pytest.fixture
def compat_modules():
    """compat/ first on sys.path for one test only (the reference harness imports the
    reference's modules of the same names)."""
    path = os.path.join(ROOT, "compat")
    names = ("shen_partation_densify", "shen_partition_utils")
    saved = {n: sys.modules.pop(n) for n in names if n in sys.modules}
    types_ = (P.BasicPointCloud, P.Partition, P.CameraPose)
    mods = [t.__module__ for t in types_]          # the compat module renames them for pickle
    sys.path.insert(0, path)
    try:
        import shen_partation_densify as d
        import shen_partition_utils as u
        yield d, u
    finally:
        sys.path.remove(path)
        for n in names:
            sys.modules.pop(n, None)
        sys.modules.update(saved)
        for t, m in zip(types_, mods):
            t.__module__ = m


def test_compat_module_names(compat_modules):
    d, u = compat_modules
    assert d.Partition.__module__ == "shen_partation_densify"
    for name in ("quad_tree_partition", "expand_partitions", "assign_cameras_to_partitions",
                 "visibility_based_camera_selection", "downsample_point_cloud", "save_partition_as_pkl"):
        assert callable(getattr(d, name))
    assert u.ProgressiveDataPartitioning is P.ProgressiveDataPartitioning
    p = d.Partition("01", [], None, d.box(0, 0, 1, 1), 3)
    q = pickle.loads(pickle.dumps(p))
    assert q.partition_id == "01" and tuple(q.origin_box.bounds) == (0.0, 0.0, 1.0, 1.0)
    assert b"shen_partation_densify" in pickle.dumps(p)


# ------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_function_level_pipeline_vs_reference_golden(tmp_path):
    """run_DataPartition's call sequence (utils-shen:80-91) through the drop-in
    functions, against the reference's own outputs for the same input."""
    g = np.load(os.path.join(GOLDEN, "selection.npz"))
    pts = g["points"]
    cams = _cams_from(g)
    vis_dir = tmp_path / "model" / "split_result" / "visible"
    os.makedirs(vis_dir)
    os.makedirs(tmp_path / "images")
    os.makedirs(tmp_path / "sparse" / "0")
    colmap_io.write_cameras_binary({1: colmap_io.ColmapCamera(1, "PINHOLE", 1600, 1200,
                                                              np.array([1200.0, 1200.0, 800.0, 600.0]))},
                                   str(tmp_path / "sparse" / "0" / "cameras.bin"))
    p64 = pts.astype(np.float64)
    parts = P.quad_tree_partition(p64, P.box(*O.bounds_xy(p64)), int(g["threshold"]))
    assert [p.partition_id for p in parts] == g["ids"].tolist()
    pcd = P.BasicPointCloud(points=pts, colors=g["rgb"] / 255.0, normals=np.zeros_like(pts))
    parts = P.expand_partitions(parts, pcd)
    assert [p.point_num for p in parts] == g["point_num"].tolist()
    assert all(p.extend_rate == 80 for p in parts)
    parts = P.assign_cameras_to_partitions(parts, cams)
    for p, want in zip(parts, _lists(g["cam_offsets"], g["cam_index"])):
        assert [cp.camera.uid for cp in p.camera] == want.tolist()
    out = P.visibility_based_camera_selection(parts, str(vis_dir), pcd)
    for p, want in zip(out, _lists(g["kept_offsets"], g["kept_index"])):
        assert sorted(cp.camera.uid for cp in p.camera) == want.tolist()
    # the on-disk layout is the reference's
    files = sorted(os.path.relpath(os.path.join(dp, f), vis_dir) for dp, _, fs in os.walk(vis_dir) for f in fs)
    assert files == g["files"].tolist()
    # files are readable and carry the block's data
    pid = out[0].partition_id
    sp = vis_dir / pid / ("partition_%s" % pid) / "sparse" / "0"
    xyz, rgb, err = colmap_io.read_points3D_binary(str(sp / "points3D.bin"))
    assert np.array_equal(xyz, out[0].point_cloud.points.astype(np.float64))
    assert np.array_equal(rgb, colmap_io.colors_to_u8(out[0].point_cloud.colors)) and not err.any()
    imgs = colmap_io.read_images_binary(str(sp / "images.bin"))
    assert sorted(imgs.ids.tolist()) == sorted(cp.camera.uid for cp in out[0].camera)
    assert all(n.endswith(".jpg") for n in imgs.names)
    with open(vis_dir / pid / ("partition_%s.pkl" % pid), "rb") as f:
        q = P.load_compat(f)          # the file carries the reference's class paths (test_io_pickle.py)
    assert q.partition_id == pid and q.point_num == out[0].point_num
    assert np.array_equal(q.point_cloud.points, out[0].point_cloud.points)


@pytest.mark.gpu
def test_downsample_point_cloud_vs_reference_golden():
    g = np.load(os.path.join(GOLDEN, "voxel.npz"))
    pc = P.BasicPointCloud(g["points"], np.zeros_like(g["points"]), np.zeros_like(g["points"]))
    got = P.downsample_point_cloud(pc, float(g["voxel"])).points
    assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), g["out"].view(np.uint32))


@pytest.mark.gpu
def test_downsample_point_cloud_attribute_means():
    """Colours / normals of the down-sampled cloud are the per-voxel means (normals re-normalised),
    densify:593-603, restated literally here on a cloud with many multi-member voxels."""
    rng = np.random.default_rng(12)
    n = 4000
    pts = (np.round(rng.uniform(0, 3, (n, 3)) * 4) / 4).astype(np.float32)
    cols = rng.uniform(0, 1, (n, 3))
    nrm = rng.normal(size=(n, 3)).astype(np.float32)
    nrm[::50] = 0.0
    got = P.downsample_point_cloud(P.BasicPointCloud(pts, cols, nrm), 0.1)
    vid = np.floor((pts - pts.min(0)) / 0.1).astype(np.int32)
    groups = {}
    for i, v in enumerate(vid):
        groups.setdefault((v[0], v[1], v[2]), []).append(i)
    wp, wc, wn = [], [], []
    for idx in groups.values():
        wp.append(np.mean(pts[idx], axis=0))
        wc.append(np.mean(cols[idx], axis=0))
        m = np.mean(nrm[idx], axis=0)
        ln = np.linalg.norm(m)
        wn.append(m / ln if ln > 1e-12 else m)
    assert len(groups) < n // 2
    assert np.array_equal(got.points.view(np.uint32), np.array(wp, dtype=np.float32).view(np.uint32))
    assert got.colors.dtype == np.float32 and np.array_equal(got.colors, np.array(wc, dtype=np.float32))
    np.testing.assert_allclose(got.normals, np.array(wn, dtype=np.float32), rtol=2e-7, atol=1e-12)


@pytest.mark.gpu
def test_cli_entry_end_to_end(tmp_path):
    """python shen_partition_utils.py <sfm>: COLMAP files in, split_result/visible out,
    same leaves / blocks / kept cameras as the oracle (radius filter included)."""
    sc = small_scene(n=60000, c=40, thr=9000, seed=71)
    colmap_io.write_synthetic_colmap(sc, str(tmp_path))
    pdp = P.main([str(tmp_path), str(sc.threshold)])
    want = O.run_partition(sc.xyz, scene_cameras(sc), sc.threshold)
    assert [p.partition_id for p in pdp.partitions] == [l.partition_id for l in want.leaves]
    assert [tuple(p.origin_box.bounds) for p in pdp.partitions] == [l.bounds for l in want.leaves]
    assert [tuple(p.extend_box.bounds) for p in pdp.partitions] == want.boxes
    flags = pdp.device_result.vis.flags.cpu().numpy()
    for p, idx, kept in zip(pdp.partitions, want.block_index, want.kept_cameras):
        assert p.point_num == len(idx) and np.array_equal(p.point_cloud.points, sc.xyz[idx])
        if not (flags & 3).any():
            assert sorted(cp.camera.uid for cp in p.camera) == sorted(kept.tolist())
        d = tmp_path / "model" / "split_result" / "visible" / p.partition_id
        assert os.path.exists(d / ("partition_%s.pkl" % p.partition_id))
        assert os.path.exists(d / ("partition_%s" % p.partition_id) / "sparse" / "0" / "points3D.bin")
        assert os.path.exists(d / ("partition_%s" % p.partition_id) / "sparse" / "0" / "cameras.bin")
    assert os.path.exists(tmp_path / "test" / "sparse" / "0" / "images.bin")
    assert sum(len(p.camera) for p in pdp.partitions) > 0
