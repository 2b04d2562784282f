#!/usr/bin/env python
"""Build-container only: times the REFERENCE's own code (/root/reference, unchanged, through
oracle/ref_harness.py) on BASELINE.json's synthetic configurations -- the CPU baseline SURVEY.md
section 8(d) asks for beside the GPU number.  Measurement infrastructure, never imported by the
product path; writes one JSON line per configuration.

    python tools/time_reference.py c1            # full reference end to end (A3-A10)
    python tools/time_reference.py c2 c3 c4      # A3 / A4 in full; A6-A9 on the declared sub-sample
                                                 # (first 4 leaves x first 16 cameras), extrapolated

What runs (file:line under /root/reference/scene/ptgs):
  quad_tree_partition (shen_partation_densify.py:37 -> density_partition.py:148), expand_partitions
  (:182), assign_cameras_to_partitions (:362), visibility_based_camera_selection (:911, incl. the
  48-thread pool, downsample_point_cloud :541, project_points_to_camera :410 with device='cpu' --
  this container has no GPU --, run_graham_scan :481, and the pkl / COLMAP writers).
A1 (open3d remove_radius_outlier) is absent from the reference tree and this image: the filtered
cloud handed to the split is the oracle's restated mask (untimed, declared).
"""
from __future__ import annotations

import contextlib
import io
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np
import torch

import ref_harness as H
import ptgs_oracle as O
import ptgs_b200  # noqa: F401
from ptgs_b200 import colmap_io, scene_synth, workload as W

SUB_LEAVES, SUB_CAMERAS = 4, 16


def _rect(b):
    minx, miny, maxx, maxy = b
    return H._Polygon([(minx, miny), (maxx, miny), (maxx, maxy), (minx, maxy)])


def run(name: str, full: bool):
    ref = H.load()
    torch.set_num_threads(os.cpu_count() or 1)
    orig = ref.densify.project_points_to_camera
    ref.densify.project_points_to_camera = lambda points, camera, device="cpu": orig(points, camera, device="cpu")
    quiet = contextlib.redirect_stdout(io.StringIO())
    line = {"config": name, "host_cores": os.cpu_count(), "torch_threads": torch.get_num_threads(),
            "reference": "unmodified /root/reference code under oracle/ref_harness.py import stubs, device='cpu'"}
    t0 = time.perf_counter()
    sc = scene_synth.make_config(name)
    w = W.from_scene(sc, name)
    line["scene_generation_s"] = round(time.perf_counter() - t0, 2)
    pts = w.xyz if w.man_trans is None else O.manhattan_points(w.xyz, w.man_trans)
    t0 = time.perf_counter()
    keep = O.radius_outlier_mask(pts.astype(np.float64))
    line["A1_oracle_mask_s_untimed_for_reference"] = round(time.perf_counter() - t0, 2)
    filt = pts.astype(np.float64)[keep]
    rcams = [H.make_simple_camera(ref, i, i + 1, w.R[i], w.T[i], float(w.FoVx[i]), float(w.FoVy[i]),
                                  "img_%06d" % i, int(w.width[i]), int(w.height[i])) for i in range(w.n_cameras)]
    stages = {}
    with quiet:
        t0 = time.perf_counter()
        parts = ref.densify.quad_tree_partition(filt, _rect(O.bounds_xy(filt)), w.threshold)
        stages["A3_quad_tree_partition"] = time.perf_counter() - t0
    pcd = ref.BasicPointCloud(points=pts, colors=sc.rgb / 255.0, normals=np.zeros_like(pts))
    with quiet:
        t0 = time.perf_counter()
        parts = ref.densify.expand_partitions(parts, pcd)
        stages["A4_expand_partitions"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        parts = ref.densify.assign_cameras_to_partitions(parts, rcams)
        stages["A5_assign_cameras"] = time.perf_counter() - t0
    L = len(parts)
    N_b = [len(p.point_cloud.points) for p in parts]
    C_b = [len(p.camera) for p in parts]
    line.update({"points": int(len(pts)), "filtered_points": int(keep.sum()), "leaves": L,
                 "sum_expanded_points": int(sum(N_b)), "pairs": int(sum(C_b))})
    if full:
        with tempfile.TemporaryDirectory() as tmp:
            vis_dir = os.path.join(tmp, "model", "split_result", "visible")
            os.makedirs(vis_dir)
            colmap_io.write_synthetic_colmap(sc, tmp)
            with quiet:
                t0 = time.perf_counter()
                out = ref.densify.visibility_based_camera_selection(parts, vis_dir, pcd, max_workers=48)
                stages["A6_A10_visibility_based_camera_selection"] = time.perf_counter() - t0
        M_b = [len(O.voxel_downsample(p.point_cloud.points)) for p in parts]
        tests = int(sum(m * c for m, c in zip(M_b, C_b)))
        kept = int(sum(len(p.camera) for p in out))
        line.update({"tests": tests, "kept_cameras": kept,
                     "visibility_tests_per_s": tests / stages["A6_A10_visibility_based_camera_selection"],
                     "note": "A6-A10 time includes the reference's per-point output writers (save_partition.py)"})
    else:
        # declared sub-sample: A6 on the first SUB_LEAVES leaves, A7-A8 on their first SUB_CAMERAS cameras
        t_ds = t_proj = t_hull = 0.0
        n_ds_pts = n_proj_tests = n_pairs = 0
        for p in parts[:SUB_LEAVES]:
            with quiet:
                t0 = time.perf_counter()
                ds = ref.densify.downsample_point_cloud(p.point_cloud, 0.1).points
                t_ds += time.perf_counter() - t0
            n_ds_pts += len(p.point_cloud.points)
            for cp in p.camera[:SUB_CAMERAS]:
                with quiet:
                    t0 = time.perf_counter()
                    uv = ref.densify.project_points_to_camera(ds, cp.camera)
                    t_proj += time.perf_counter() - t0
                    t0 = time.perf_counter()
                    if not isinstance(uv, tuple) and len(uv) > 3:
                        ref.densify.run_graham_scan(uv, cp.camera.image_width, cp.camera.image_height)
                    t_hull += time.perf_counter() - t0
                n_proj_tests += len(ds)
                n_pairs += 1
        stages.update({"A6_downsample_sub": t_ds, "A7_project_sub": t_proj, "A8_graham_sub": t_hull})
        tests_total = float(sum(n * c for n, c in zip(N_b, C_b)))         # ~ M_b * C_b (M_b / N_b > 0.99)
        ext = t_ds / max(n_ds_pts, 1) * sum(N_b) + (t_proj + t_hull) / max(n_proj_tests, 1) * tests_total
        line.update({"sub_sample": "A6 on the first %d leaves (%d points), A7-A8 on their first %d cameras (%d pairs)"
                                   % (SUB_LEAVES, n_ds_pts, SUB_CAMERAS, n_pairs),
                     "tests_approx": tests_total,
                     "A6_A9_extrapolated_s_single_thread": ext,
                     "visibility_tests_per_s_extrapolated": tests_total / ext if ext > 0 else None,
                     "note": "linear extrapolation of the sub-sample, single thread (the reference's 48-thread pool "
                             "runs under the GIL: its Python loops do not scale)"})
    line["stage_s"] = {k: round(v, 3) for k, v in stages.items()}
    ref.densify.project_points_to_camera = orig
    return line


def main():
    if not H.available():
        raise SystemExit("time_reference.py needs /root/reference (build container only)")
    names = sys.argv[1:] or ["c1"]
    for n in names:
        print(json.dumps(run(n, full=(n == "c1"))), flush=True)


if __name__ == "__main__":
    main()
