"""GPU box: run BASELINE.json's configurations through the device pipeline and print one
JSON line per config (scene statistics SURVEY.md 8(d) asks for + timings).
usage: python tools/configs_report.py [c1 c2 c3 c4 c5] [--c5-cameras N]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, workload as W

names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c1", "c2", "c3"]
c5_cams = None
for i, a in enumerate(sys.argv):
    if a == "--c5-cameras":
        c5_cams = int(sys.argv[i + 1])
dev = torch.device("cuda", 0)


def timed(fn, reps=3):
    best = None
    out = None
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return out, best * 1e3


for name in names:
    t0 = time.perf_counter()
    w = W.make(name, 1.0)
    gen_s = time.perf_counter() - t0
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    xyz = torch.from_numpy(w.xyz).to(dev)
    line = {"config": W.describe(name, 1.0, w), "scene_generation_s": round(gen_s, 1)}
    if name == "c5":
        # all points x all cameras coverage sweep (no blocks, no down-sampling)
        x, y, z = D.aos_to_soa(xyz)
        cams_d = torch.from_numpy(recs).to(dev)
        C = len(recs) if c5_cams is None else min(c5_cams, len(recs))
        od, oms = timed(lambda: D.spatial_order(x, y, z), reps=2)
        n_hit, ms = timed(lambda: D.coverage_sweep(x, y, z, cams_d, cam_first=0, cam_count=C, ordered=od), reps=2)
        tests = float(len(w.xyz)) * C
        line.update({"cameras_swept": C, "tests": tests, "spatial_order_ms": round(oms, 3), "sweep_ms": round(ms, 3),
                     "tests_per_s": tests / ((ms + oms) * 1e-3),
                     "hits": int(n_hit.sum().item()), "cameras_with_hits": int((n_hit > 0).sum().item())})
        if "--c5-file-order" in sys.argv:
            n2, ms2 = timed(lambda: D.coverage_sweep(x, y, z, cams_d, cam_first=0, cam_count=C,
                                                     spatial_order_points=False), reps=1)
            line.update({"file_order_sweep_ms": round(ms2, 3), "file_order_equal": bool(torch.equal(n2, n_hit))})
    else:
        centers = D.camera_centers(w.R, w.T)
        res, ms = timed(lambda: D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans,
                                                   time_stages=True))
        M_b = np.diff(res.ds_offsets.cpu().numpy())
        C_b = np.diff(res.cam_offsets.cpu().numpy())
        N_b = np.diff(res.pt_offsets.cpu().numpy())
        pairs = int(C_b.sum())
        line.update({
            "points": int(len(w.xyz)), "filtered_points": int(res.n_filtered), "leaves": len(res.leaf_ids),
            "max_depth": max(len(i) for i in res.leaf_ids), "sum_expanded_points": int(N_b.sum()),
            "sum_downsampled_points": int(M_b.sum()), "pairs": pairs, "tests": int((M_b * C_b).sum()),
            "kept_cameras": int(res.vis.keep.sum().item()), "kept_fraction": float(res.vis.keep.float().mean().item()),
            "flagged_near_or_collinear": int(((res.vis.flags & 3) != 0).sum().item()),
            "exact_hull_pairs": int(res.vis.n_exact), "cameras_unassigned": int(len(recs) - len(torch.unique(res.cam_index))),
            "partition_wall_ms": round(ms, 3), "stage_ms": {k: round(v, 3) for k, v in res.timings_ms.items()}})
        cams_d = torch.from_numpy(recs).to(dev)
        _, vms = timed(lambda: D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets,
                                            res.cam_index))
        line["visibility_ms"] = round(vms, 3)
        line["visibility_tests_per_s"] = line["tests"] / (vms * 1e-3)
        del res
    print(json.dumps(line), flush=True)
    del xyz
    torch.cuda.empty_cache()
