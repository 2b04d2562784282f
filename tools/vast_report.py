"""GPU box: time the camera-grid ("Vast") path of ptgs_b200.vast on a synthetic scene.
usage: python tools/vast_report.py [n_points] [n_cameras] [m] [n]"""
import json
import os
import sys
import tempfile
import time
from types import SimpleNamespace

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import scene_synth, vast as V, workload as W

n_pts = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
n_cam = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
m = int(sys.argv[3]) if len(sys.argv) > 3 else 2
n = int(sys.argv[4]) if len(sys.argv) > 4 else 4


class Cam:
    def __init__(self, R, T, fx, fy, w, h, name, center):
        self.R, self.T, self.FoVx, self.FoVy = R, T, fx, fy
        self.image_width, self.image_height, self.image_name = w, h, name
        self.camera_center = torch.tensor(center, dtype=torch.float32)


t0 = time.perf_counter()
sc = scene_synth.make_scene(n_pts, n_cam, 500000, 1010, oblique_fraction=0.5, height_scale=3.0)
P = np.array([[1.0, 0, 0], [0, 0, 1.0], [0, 1.0, 0]])        # ground plane XY -> XZ
pts = np.ascontiguousarray(sc.xyz[:, [0, 2, 1]])
cams = []
fovx, fovy = W.focal2fov(sc.fx, sc.width), W.focal2fov(sc.fy, sc.height)
for c in range(n_cam):
    R_wc = W.qvec2rotmat(sc.qvec[c])
    T = np.asarray(sc.tvec[c], np.float64)
    centre = (P @ (-R_wc.T @ T)).astype(np.float32)
    cams.append(Cam((R_wc @ P).T, T, fovx, fovy, sc.width, sc.height, "img_%06d" % c, centre))
colors = sc.rgb.astype(np.float64) / 255.0
info = SimpleNamespace(point_cloud=V.BasicPointCloud(points=pts, colors=colors, normals=np.zeros_like(pts)))
gen_s = time.perf_counter() - t0
best = None
for rep in range(3):
    with tempfile.TemporaryDirectory() as tmp:
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dp = V.ProgressiveDataPartitioning(info, cams, tmp, m, n, 0.2, 0.25, write_outputs=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    best = dt if best is None else min(best, dt)
print(json.dumps({"points": n_pts, "cameras": n_cam, "grid": [m, n], "scene_generation_s": round(gen_s, 1),
                  "wall_s_best_of_3": round(best, 3), "appended_cameras": len(dp.appended),
                  "block_cameras": [len(p.cameras) for p in dp.partition_scene],
                  "block_points": [int(len(p.point_cloud.points)) for p in dp.partition_scene]}))
