"""The camera-grid ("Vast") path, SURVEY.md row A14: ptgs_b200.vast (CUDA through the C
ABI) against the oracle restatement (oracle/ptgs_oracle_vast.py, itself pinned against the
reference's own code) and against tests/golden/vast.npz written by the reference code."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

import ptgs_oracle_vast as OV
from helpers import VastCamera, small_scene, vast_inputs
import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, vast as V

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEV = "cuda:0"


def _cameras(vc):
    return [VastCamera(vc.R[c], vc.T[c], float(vc.FoVx[c]), float(vc.FoVy[c]), int(vc.width[c]),
                       int(vc.height[c]), vc.names[c], vc.center[c]) for c in range(len(vc.names))]


def _scene_info(pts, colors):
    return SimpleNamespace(point_cloud=V.BasicPointCloud(points=pts, colors=colors, normals=np.zeros_like(pts)))


def test_vast_no_cuda_raises():
    if torch.cuda.is_available():
        pytest.skip("needs a machine without CUDA")
    sc = small_scene(n=2000, c=12, thr=6000, seed=3)
    pts, colors, vc = vast_inputs(sc)
    with pytest.raises(RuntimeError):
        V.ProgressiveDataPartitioning(_scene_info(pts, colors), _cameras(vc), "/tmp/unused", 2, 2, write_outputs=False)


def _check(dp, pts, colors, vc, parts, sel):
    name_to_idx = {nm: i for i, nm in enumerate(vc.names)}
    n_added = 0
    assert len(dp.partition_scene) == len(parts)
    for got, p, (cam_list, uniq, first, added) in zip(dp.partition_scene, parts, sel):
        assert got.partition_id == p.partition_id
        assert [float(v) for v in got.ori_camera_bbox] == [float(v) for v in p.ori_camera_bbox]
        assert [float(v) for v in got.extend_camera_bbox] == [float(v) for v in p.extend_camera_bbox]
        assert [float(v) for v in got.ori_point_bbox] == [float(v) for v in p.ori_point_bbox]
        assert [float(v) for v in got.extend_point_bbox] == [float(v) for v in p.extend_point_bbox]
        assert [name_to_idx[cp.camera.image_name] for cp in got.cameras] == list(cam_list)
        assert got.point_cloud.points.dtype == pts.dtype
        assert np.array_equal(got.point_cloud.points, uniq)
        assert np.array_equal(got.point_cloud.colors, colors[first])
        n_added += len(added)
    return n_added


@pytest.mark.gpu
@pytest.mark.parametrize("seed,m,n,rate,f64", [(71, 2, 2, 0.05, False), (72, 3, 2, 0.1, False), (74, 2, 4, 0.25, True)])
def test_vast_against_oracle(tmp_path, seed, m, n, rate, f64):
    sc = small_scene(n=60000, c=67, thr=6000, seed=seed, oblique_fraction=0.5, height_scale=3.0)
    pts, colors, vc = vast_inputs(sc)
    if f64:                                   # a Manhattan-transformed cloud is float64
        pts = pts.astype(np.float64) * 1.0000001
    pts[100] = pts[7]                         # duplicate rows with different colours
    pts[2000] = pts[7]
    parts, sel = OV.run_vast(pts, vc, m, n, 0.2, rate)
    dp = V.ProgressiveDataPartitioning(_scene_info(pts, colors), _cameras(vc), str(tmp_path), m, n, 0.2, rate)
    assert _check(dp, pts, colors, vc, parts, sel) > 0
    # airspace rates of the appended cameras
    want = {}
    for p, (_, _, _, added) in zip(parts, sel):
        for (_, c, r) in added:
            want[(p.partition_id, vc.names[c])] = r
    assert len(dp.appended) == len(want)
    for (pi_, pj, nm, r) in dp.appended:
        assert abs(r - want[(pi_, nm)]) <= 1e-9 * max(1.0, r)
    # output files of data_partition.py
    vis = os.path.join(str(tmp_path), "partition_point_cloud", "visible")
    for p in parts:
        assert os.path.exists(os.path.join(vis, "%s_visible.ply" % p.partition_id))
        assert os.path.exists(os.path.join(vis, "%s_camera_centers.ply" % p.partition_id))
    assert os.path.exists(os.path.join(str(tmp_path), "partition_data.pkl"))
    # second construction loads the pickle (data_partition.py:91-99)
    again = V.ProgressiveDataPartitioning(_scene_info(pts, colors), _cameras(vc), str(tmp_path), m, n, 0.2, rate)
    assert [p.partition_id for p in again.partition_scene] == [p.partition_id for p in parts]


@pytest.mark.gpu
def test_vast_airspace_kernel_vs_oracle():
    sc = small_scene(n=2000, c=64, thr=6000, seed=73, oblique_fraction=0.5)
    pts, colors, vc = vast_inputs(sc)
    rng = np.random.default_rng(5)
    B = 40
    lo = rng.uniform(-60, 20, (B, 3))
    hi = lo + rng.uniform(0.0, 80, (B, 3))
    hi[3] = lo[3]                                           # degenerate box: a single point
    hi[4, 1] = lo[4, 1]                                     # flat box
    boxes = np.stack([lo[:, 0], hi[:, 0], lo[:, 1], hi[:, 1], lo[:, 2], hi[:, 2]], 1).astype(np.float32).astype(np.float64)
    recs = D.vast_camera_records(vc.R, vc.T, vc.FoVx, vc.FoVy, vc.width, vc.height)
    cams_d = torch.from_numpy(recs).to(DEV)
    C = len(vc.names)
    pb = np.repeat(np.arange(B), C).astype(np.int32)
    pc = np.tile(np.arange(C), B).astype(np.int32)
    got = D.vast_airspace(cams_d, boxes, pb, pc)
    nz = 0
    for k in range(len(pb)):
        want = OV.airspace_rate(vc, int(pc[k]), OV.corner_points(boxes[pb[k]]))
        assert abs(got[k] - want) <= 1e-9 * max(1.0, want), (k, got[k], want)
        nz += want > 0
    assert nz > 100


@pytest.mark.gpu
def test_vast_golden_from_reference(tmp_path):
    g = np.load(os.path.join(GOLDEN, "vast.npz"), allow_pickle=False)
    sc = small_scene(n=int(g["n"]), c=int(g["c"]), thr=6000, seed=int(g["seed"]), oblique_fraction=0.5)
    pts, colors, vc = vast_inputs(sc)
    dp = V.ProgressiveDataPartitioning(_scene_info(pts, colors), _cameras(vc), str(tmp_path), int(g["m"]), int(g["n_region"]),
                                       float(g["extend_rate"]), float(g["visible_rate"]), write_outputs=False)
    ids = [str(s) for s in g["ids"]]
    assert [p.partition_id for p in dp.partition_scene] == ids
    for b, p in enumerate(dp.partition_scene):
        assert np.array_equal(np.asarray(p.ori_camera_bbox, np.float64), g["ori_camera_bbox"][b])
        assert np.array_equal(np.asarray(p.extend_camera_bbox, np.float64), g["extend_camera_bbox"][b])
        assert np.array_equal(np.asarray(p.ori_point_bbox, np.float64), g["ori_point_bbox"][b])
        assert np.array_equal(np.asarray(p.extend_point_bbox, np.float64), g["extend_point_bbox"][b])
        cams = np.array([vc.names.index(cp.camera.image_name) for cp in p.cameras])
        assert np.array_equal(cams, g["cams_%d" % b])
        assert np.array_equal(p.point_cloud.points, g["points_%d" % b])
        assert np.array_equal(p.point_cloud.colors, g["colors_%d" % b])


def test_vast_golden_oracle():
    """CPU: the oracle restatement against the fixture written by the reference's own code."""
    g = np.load(os.path.join(GOLDEN, "vast.npz"), allow_pickle=False)
    sc = small_scene(n=int(g["n"]), c=int(g["c"]), thr=6000, seed=int(g["seed"]), oblique_fraction=0.5)
    pts, colors, vc = vast_inputs(sc)
    parts, sel = OV.run_vast(pts, vc, int(g["m"]), int(g["n_region"]), float(g["extend_rate"]), float(g["visible_rate"]))
    assert [p.partition_id for p in parts] == [str(s) for s in g["ids"]]
    tot = 0
    for b, (p, (cam_list, uniq, first, added)) in enumerate(zip(parts, sel)):
        assert np.array_equal(np.asarray(p.extend_point_bbox, np.float64), g["extend_point_bbox"][b])
        assert np.array_equal(cam_list, g["cams_%d" % b])
        assert np.array_equal(uniq, g["points_%d" % b])
        assert np.array_equal(colors[first], g["colors_%d" % b])
        tot += len(added)
    assert tot > 0
