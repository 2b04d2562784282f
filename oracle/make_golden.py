#!/usr/bin/env python
"""Generates tests/golden/*.npz by executing the REFERENCE's own code
(/root/reference, unchanged, through oracle/ref_harness.py) on small seeded
inputs.  Test infrastructure only; build-container only (the GPU box has no
/root/reference -- it uses the committed fixtures through tests/test_golden.py).

    python oracle/make_golden.py            # rewrites tests/golden/

Every fixture stores the INPUT arrays next to the reference's OUTPUT, so the
consumers (oracle parity tests on CPU, CUDA parity tests on the GPU box) never
depend on this repo's scene generator staying bit-stable.

Reference entry points executed (file:line under /root/reference):
  quad_tree_partition            scene/ptgs/shen_partation_densify.py:37
    -> density_partition         scene/ptgs/density_partition.py:148
  expand_partitions              scene/ptgs/shen_partation_densify.py:182
  assign_cameras_to_partitions   scene/ptgs/shen_partation_densify.py:362
  SimpleCamera.camera_center     scene/cameras.py:56-58
  downsample_point_cloud         scene/ptgs/shen_partation_densify.py:541
  project_points_to_camera       scene/ptgs/shen_partation_densify.py:410
  run_graham_scan                scene/ptgs/shen_partation_densify.py:481
  visibility_based_camera_selection  scene/ptgs/shen_partation_densify.py:911
  fetchPly (Manhattan rotation)  scene/ptgs/shen_data_read.py:120-133
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import ref_harness as H          # noqa: E402
import ptgs_oracle as O          # noqa: E402  (only for input-side helpers: bounds, qvec2rotmat)
from helpers import scene_cameras, small_scene   # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def _rect(b):
    minx, miny, maxx, maxy = b
    return H._Polygon([(minx, miny), (maxx, miny), (maxx, maxy), (minx, maxy)])


def _ref_cameras(ref, sc):
    cams = scene_cameras(sc)
    out = [H.make_simple_camera(ref, i, i + 1, cams.R[i], cams.T[i], float(cams.FoVx[i]), float(cams.FoVy[i]),
                                "img_%06d" % i, sc.width, sc.height) for i in range(len(cams.R))]
    return cams, out


def _cam_arrays(cams):
    return dict(R=cams.R, T=cams.T, FoVx=cams.FoVx, FoVy=cams.FoVy,
                width=np.asarray(cams.width), height=np.asarray(cams.height))


def _ragged(lists, dtype):
    off = np.zeros(len(lists) + 1, np.int64)
    off[1:] = np.cumsum([len(l) for l in lists])
    flat = np.concatenate([np.asarray(l, dtype=dtype).reshape(-1) for l in lists]) if lists else np.zeros(0, dtype)
    return off, flat.astype(dtype)


def _ids(parts):
    return np.array([p.partition_id for p in parts], dtype="U16")


def split_cases(ref):
    """A3: ids, closed rectangles and point counts of the median split."""
    cases = {}
    for name, seed, n, thr, ties in (("uniform", 1, 9000, 400, False), ("ties", 2, 9000, 400, True),
                                     ("single_leaf", 3, 2000, 100000, False), ("depth_cap", 4, 3000, 1, False)):
        rng = np.random.default_rng(seed)
        pts = rng.uniform(0, 40, (n, 3))
        if ties:
            pts[:, :2] = np.round(pts[:, :2] * 2) / 2
        pts = pts.astype(np.float32).astype(np.float64)
        b = O.bounds_xy(pts)
        parts = ref.densify.quad_tree_partition(pts, _rect(b), thr)
        cases[name] = dict(points=pts.astype(np.float32), root=np.asarray(b, np.float64), threshold=np.int64(thr),
                           ids=_ids(parts), bounds=np.array([p.origin_box.bounds for p in parts], np.float64),
                           point_num=np.array([p.point_num for p in parts], np.int64))
    return cases


def expand_assign_case(ref):
    """A4 + A5 on a synthetic scene (unfiltered cloud, fp32 compares)."""
    sc = small_scene(n=16000, c=40, thr=1600, seed=12)
    cams, rcams = _ref_cameras(ref, sc)
    p64 = sc.xyz.astype(np.float64)
    b = O.bounds_xy(p64)
    parts = ref.densify.quad_tree_partition(p64, _rect(b), sc.threshold)
    pcd = ref.BasicPointCloud(points=sc.xyz, colors=sc.rgb / 255.0, normals=np.zeros_like(sc.xyz))
    exp = ref.densify.expand_partitions(parts, pcd)
    asg = ref.densify.assign_cameras_to_partitions(exp, rcams)
    # index sets of the expanded blocks: recover indices by matching rows in order
    idx_lists = []
    for e in exp:
        minx, miny, maxx, maxy = e.extend_box.bounds
        m = (sc.xyz[:, 0] >= minx) & (sc.xyz[:, 0] <= maxx) & (sc.xyz[:, 1] >= miny) & (sc.xyz[:, 1] <= maxy)
        assert np.array_equal(sc.xyz[m], e.point_cloud.points)     # the reference's own mask (densify:201-205)
        idx_lists.append(np.nonzero(m)[0])
    p_off, p_idx = _ragged(idx_lists, np.int64)
    c_off, c_idx = _ragged([[cp.camera.uid for cp in a.camera] for a in asg], np.int64)
    d = dict(points=sc.xyz, threshold=np.int64(sc.threshold), root=np.asarray(b, np.float64), ids=_ids(exp),
             origin_bounds=np.array([p.origin_box.bounds for p in exp], np.float64),
             extend_bounds=np.array([p.extend_box.bounds for p in exp], np.float64),
             point_num=np.array([p.point_num for p in exp], np.int64), pt_offsets=p_off, pt_index=p_idx,
             centers=np.stack([c.camera_center.numpy() for c in rcams]), cam_offsets=c_off, cam_index=c_idx)
    d.update(_cam_arrays(cams))
    return d


def voxel_case(ref):
    """A6: voxel means (fp32 bit patterns) in first-occurrence order."""
    rng = np.random.default_rng(5)
    pts = rng.uniform(-3, 9, (6000, 3)).astype(np.float32)
    pts[:, :2] = np.round(pts[:, :2] * 8) / 8
    pc = ref.BasicPointCloud(points=pts, colors=np.zeros_like(pts), normals=np.zeros_like(pts))
    out = ref.densify.downsample_point_cloud(pc, 0.1).points
    return dict(points=pts, voxel=np.float64(0.1), out=np.asarray(out, np.float32))


def projection_case(ref):
    """A7: in-image (u, v) lists per camera, torch CPU fp32."""
    sc = small_scene(n=8000, c=8, thr=8000, seed=14, oblique_fraction=0.5)
    cams, rcams = _ref_cameras(ref, sc)
    uv = [np.asarray(ref.densify.project_points_to_camera(sc.xyz, cam, device="cpu"), np.float32).reshape(-1, 2)
          for cam in rcams]
    off, flat = _ragged([u.reshape(-1) for u in uv], np.float32)
    d = dict(points=sc.xyz, uv_offsets=off // 2, uv=flat.reshape(-1, 2))
    d.update(_cam_arrays(cams))
    return d


def graham_cases(ref):
    """A8: hull-area ratio, incl. collinear and degenerate inputs."""
    rng = np.random.default_rng(7)
    cases = [rng.uniform(0, 100, (500, 2)),
             np.stack(np.meshgrid(np.arange(1, 12), np.arange(1, 9)), -1).reshape(-1, 2).astype(float),
             np.array([[2, 1], [60, 1], [30, 1], [61, 30], [3, 50], [40, 63], [20, 20]], float),
             np.array([[i, i + 1] for i in range(1, 30)], float),
             np.array([[1, 1], [5, 9]], float),
             np.array([[5, 5], [5, 5], [5, 5], [5, 5]], float),
             np.array([[1, 1], [9, 1], [9, 9], [1, 9], [5, 1], [9, 5], [5, 9], [1, 5], [5, 5]], float)]
    rates = []
    for pts in cases:
        pts = pts.astype(np.float32).astype(np.float64)
        r = ref.densify.run_graham_scan(pts.tolist(), 100, 80)
        rates.append(float(r["intersection_rate"] if isinstance(r, dict) else r))
    off, flat = _ragged([c.astype(np.float32).reshape(-1) for c in cases], np.float32)
    return dict(pts_offsets=off // 2, pts=flat.reshape(-1, 2), width=np.float64(100), height=np.float64(80),
                rate=np.array(rates, np.float64))


def selection_case(ref):
    """A3-A9 end to end through visibility_based_camera_selection (the
    reference's thread pool and output writers run too).  The reference's
    device='cuda' default (densify:410) is redirected to 'cpu'."""
    orig = ref.densify.project_points_to_camera
    ref.densify.project_points_to_camera = lambda points, camera, device="cpu": orig(points, camera, device="cpu")
    try:
        sc = small_scene(n=16000, c=24, thr=4500, seed=16, height_scale=3.0, oblique_fraction=0.5)
        cams, rcams = _ref_cameras(ref, sc)
        p64 = sc.xyz.astype(np.float64)
        b = O.bounds_xy(p64)
        parts = ref.densify.quad_tree_partition(p64, _rect(b), sc.threshold)
        pcd = ref.BasicPointCloud(points=sc.xyz, colors=sc.rgb / 255.0, normals=np.zeros_like(sc.xyz))
        parts = ref.densify.expand_partitions(parts, pcd)
        parts = ref.densify.assign_cameras_to_partitions(parts, rcams)
        assigned = [[cp.camera.uid for cp in p.camera] for p in parts]
        with tempfile.TemporaryDirectory() as tmp:
            vis_dir = os.path.join(tmp, "model", "split_result", "visible")
            os.makedirs(vis_dir)
            os.makedirs(os.path.join(tmp, "images"))
            from ptgs_b200 import colmap_io
            colmap_io.write_synthetic_colmap(sc, tmp)
            out = ref.densify.visibility_based_camera_selection(parts, vis_dir, pcd, max_workers=4)
            # per-pair hit count and ratio, recomputed with the reference's functions one pair at a time
            n_hit, rate = [], []
            for p in parts:
                ds = ref.densify.downsample_point_cloud(p.point_cloud, 0.1).points
                for cp in p.camera:
                    uv = ref.densify.project_points_to_camera(ds, cp.camera)
                    k = len(uv)
                    n_hit.append(k if not isinstance(uv, tuple) else 0)
                    if k <= 3 or isinstance(uv, tuple):
                        rate.append(0.0)
                    else:
                        r = ref.densify.run_graham_scan(uv, cp.camera.image_width, cp.camera.image_height)
                        rate.append(float(r["intersection_rate"]))
            files = sorted(os.path.relpath(os.path.join(dp, f), vis_dir)
                           for dp, _, fs in os.walk(vis_dir) for f in fs)
    finally:
        ref.densify.project_points_to_camera = orig
    a_off, a_idx = _ragged(assigned, np.int64)
    k_off, k_idx = _ragged([sorted(cp.camera.uid for cp in p.camera) for p in out], np.int64)
    d = dict(points=sc.xyz, rgb=sc.rgb, threshold=np.int64(sc.threshold), ids=_ids(out),
             cam_offsets=a_off, cam_index=a_idx, kept_offsets=k_off, kept_index=k_idx,
             n_hit=np.array(n_hit, np.int64), rate=np.array(rate, np.float64),
             point_num=np.array([p.point_num for p in out], np.int64), files=np.array(files, dtype="U128"))
    d.update(_cam_arrays(cams))
    return d


class _FakePly:
    """Stand-in for plyfile.PlyData in shen_data_read.fetchPly (the module is absent here): .read()
    returns the vertex columns of an in-memory cloud, so the reference's own Manhattan formula
    (shen_data_read.py:126-130) runs unchanged."""
    cloud = None

    @classmethod
    def read(cls, path):
        xyz, rgb = cls.cloud
        z = np.zeros(len(xyz), np.float32)
        return {"vertex": {"x": xyz[:, 0], "y": xyz[:, 1], "z": xyz[:, 2], "red": rgb[:, 0], "green": rgb[:, 1],
                           "blue": rgb[:, 2], "nx": z, "ny": z, "nz": z}}


def ref_fetch_ply(ref, xyz, rgb, man_trans):
    """fetchPly (scene/ptgs/shen_data_read.py:120-133) on an in-memory cloud."""
    import contextlib
    import io
    mod = ref.shen_data_read
    orig = mod.PlyData
    _FakePly.cloud = (xyz, rgb)
    mod.PlyData = _FakePly
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            return mod.fetchPly("in-memory", man_trans=man_trans)
    finally:
        mod.PlyData = orig


def _selection_run(ref, split_p64, pcd, rcams, sc):
    """run_DataPartition's steps 2-5 (utils-shen:65-91) with the reference's functions: bounds and
    split on `split_p64` (what open3d read from the PLY), expansion / assignment / selection on `pcd`
    (self.pcd).  No radius filter (open3d is absent).  Returns plain arrays."""
    b = O.bounds_xy(split_p64)
    parts = ref.densify.quad_tree_partition(split_p64, _rect(b), THR_MAN)
    leaf_bounds = np.array([p.origin_box.bounds for p in parts], np.float64)
    leaf_num = np.array([p.point_num for p in parts], np.int64)
    parts = ref.densify.expand_partitions(parts, pcd)
    pts = pcd.points
    idx_lists = []
    for e in parts:
        minx, miny, maxx, maxy = e.extend_box.bounds
        m = (pts[:, 0] >= minx) & (pts[:, 0] <= maxx) & (pts[:, 1] >= miny) & (pts[:, 1] <= maxy)
        assert np.array_equal(pts[m], e.point_cloud.points)
        idx_lists.append(np.nonzero(m)[0])
    parts = ref.densify.assign_cameras_to_partitions(parts, rcams)
    assigned = [[cp.camera.uid for cp in p.camera] for p in parts]
    n_hit, rate, ds_all = [], [], []
    for p in parts:
        ds = ref.densify.downsample_point_cloud(p.point_cloud, 0.1).points
        ds_all.append(ds)
        for cp in p.camera:
            uv = ref.densify.project_points_to_camera(ds, cp.camera)
            k = len(uv)
            n_hit.append(k if not isinstance(uv, tuple) else 0)
            if k <= 3 or isinstance(uv, tuple):
                rate.append(0.0)
            else:
                rate.append(float(ref.densify.run_graham_scan(uv, cp.camera.image_width,
                                                              cp.camera.image_height)["intersection_rate"]))
    with tempfile.TemporaryDirectory() as tmp:
        vis_dir = os.path.join(tmp, "model", "split_result", "visible")
        os.makedirs(vis_dir)
        os.makedirs(os.path.join(tmp, "images"))
        from ptgs_b200 import colmap_io
        colmap_io.write_synthetic_colmap(sc, tmp)
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            out = ref.densify.visibility_based_camera_selection(parts, vis_dir, pcd, max_workers=4)
    p_off, p_idx = _ragged(idx_lists, np.int64)
    a_off, a_idx = _ragged(assigned, np.int64)
    k_off, k_idx = _ragged([sorted(cp.camera.uid for cp in p.camera) for p in out], np.int64)
    d_off, d_pts = _ragged([d.reshape(-1) for d in ds_all], np.float32)
    return dict(ids=_ids(out), root=np.asarray(b, np.float64), leaf_bounds=leaf_bounds, leaf_num=leaf_num,
                extend_bounds=np.array([p.extend_box.bounds for p in out], np.float64),
                point_num=np.array([p.point_num for p in out], np.int64), pt_offsets=p_off, pt_index=p_idx,
                cam_offsets=a_off, cam_index=a_idx, kept_offsets=k_off, kept_index=k_idx,
                ds_offsets=d_off // 3, ds_points=d_pts.reshape(-1, 3),
                n_hit=np.array(n_hit, np.int64), rate=np.array(rate, np.float64))


THR_MAN = 3000


def manhattan_case(ref):
    """A13: the cloud rotated by the reference's fetchPly (float64 from there on), then A3-A9 by the
    reference's functions under both semantics of SURVEY.md row A13:
      consistent_*  cameras W2C @ inv(man_trans) (scene/dataset_readers.py:96-104), split on the rotated cloud;
      literal_*     the shen path as written: cameras un-rotated (shen_data_read.py:62-70 is commented
                    out), split on the UN-rotated PLY that open3d re-reads (utils-shen:62), expansion
                    and visibility on the rotated self.pcd (utils-shen:86, :91).
    Also pins A6 on a float64 cloud (densify:572-605 with f64 arithmetic, f32 cast at :605)."""
    from ptgs_b200 import scene_synth
    orig = ref.densify.project_points_to_camera
    ref.densify.project_points_to_camera = lambda points, camera, device="cpu": orig(points, camera, device="cpu")
    try:
        sc = scene_synth.make_scene(14000, 20, THR_MAN, seed=18, height_scale=10.0, oblique_fraction=0.3,
                                    manhattan_deg=21.0)
        man = sc.man_trans
        pcd = ref_fetch_ply(ref, sc.xyz, sc.rgb, man)
        assert pcd.points.dtype == np.float64
        cams = scene_cameras(sc)
        Rm, Tm = O.manhattan_cameras(cams.R, cams.T, man)
        mk = lambda R, T: [H.make_simple_camera(ref, i, i + 1, R[i], T[i], float(cams.FoVx[i]), float(cams.FoVy[i]),
                                                "img_%06d" % i, sc.width, sc.height) for i in range(len(R))]
        d = dict(points=sc.xyz, rgb=sc.rgb, man_trans=man, threshold=np.int64(THR_MAN), rotated=pcd.points,
                 R_rot=Rm, T_rot=Tm)
        d.update(_cam_arrays(cams))
        for k, v in _selection_run(ref, pcd.points.astype(np.float64), pcd, mk(Rm, Tm), sc).items():
            d["consistent_" + k] = v
        for k, v in _selection_run(ref, sc.xyz.astype(np.float64), pcd, mk(cams.R, cams.T), sc).items():
            d["literal_" + k] = v
    finally:
        ref.densify.project_points_to_camera = orig
    return d


def vast_case(ref):
    """A14: the reference's camera-grid partitioner (scene/ptgs/data_partition.py), run unchanged."""
    from types import SimpleNamespace
    from helpers import VastCamera, vast_inputs
    n, c, seed, m, n_region, er, vr = 20000, 37, 75, 2, 2, 0.2, 0.25
    sc = small_scene(n=n, c=c, thr=6000, seed=seed, oblique_fraction=0.5)
    pts, colors, vc = vast_inputs(sc)
    dp = ref.vast_data_partition
    cams = [VastCamera(vc.R[k], vc.T[k], float(vc.FoVx[k]), float(vc.FoVy[k]), int(vc.width[k]),
                       int(vc.height[k]), vc.names[k], vc.center[k]) for k in range(c)]
    pcd = dp.BasicPointCloud(points=pts, colors=colors, normals=np.zeros_like(pts))
    with tempfile.TemporaryDirectory() as tmp:
        out = dp.ProgressiveDataPartitioning(SimpleNamespace(point_cloud=pcd), cams, tmp, m, n_region, er, vr)
    res = out.partition_scene
    d = dict(n=np.int64(n), c=np.int64(c), seed=np.int64(seed), m=np.int64(m), n_region=np.int64(n_region),
             extend_rate=np.float64(er), visible_rate=np.float64(vr),
             ids=np.array([p.partition_id for p in res], dtype="U16"),
             ori_camera_bbox=np.array([p.ori_camera_bbox for p in res], np.float64),
             extend_camera_bbox=np.array([p.extend_camera_bbox for p in res], np.float64),
             ori_point_bbox=np.array([p.ori_point_bbox for p in res], np.float64),
             extend_point_bbox=np.array([p.extend_point_bbox for p in res], np.float64))
    for b, p in enumerate(res):
        d["cams_%d" % b] = np.array([vc.names.index(cp.camera.image_name) for cp in p.cameras], np.int64)
        d["points_%d" % b] = p.point_cloud.points
        d["colors_%d" % b] = p.point_cloud.colors
    return d


def main():
    if not H.available():
        raise SystemExit("make_golden.py needs /root/reference (build container only)")
    ref = H.load()
    os.makedirs(GOLDEN, exist_ok=True)
    for name, d in split_cases(ref).items():
        np.savez_compressed(os.path.join(GOLDEN, "split_%s.npz" % name), **d)
    np.savez_compressed(os.path.join(GOLDEN, "expand_assign.npz"), **expand_assign_case(ref))
    np.savez_compressed(os.path.join(GOLDEN, "voxel.npz"), **voxel_case(ref))
    np.savez_compressed(os.path.join(GOLDEN, "projection.npz"), **projection_case(ref))
    np.savez_compressed(os.path.join(GOLDEN, "graham.npz"), **graham_cases(ref))
    np.savez_compressed(os.path.join(GOLDEN, "selection.npz"), **selection_case(ref))
    np.savez_compressed(os.path.join(GOLDEN, "vast.npz"), **vast_case(ref))
    np.savez_compressed(os.path.join(GOLDEN, "manhattan.npz"), **manhattan_case(ref))
    for f in sorted(os.listdir(GOLDEN)):
        print("%-28s %8d B" % (f, os.path.getsize(os.path.join(GOLDEN, f))))


if __name__ == "__main__":
    main()
