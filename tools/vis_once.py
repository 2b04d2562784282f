"""GPU box: one partition of a BASELINE.json configuration, then `reps` visibility passes (A7-A9) on
the resident down-sampled blocks -- the short command the ncu captures wrap.
usage: python tools/vis_once.py [c4] [reps]      prints the tests / pairs of one pass as JSON"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, workload as W

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda", 0)
w = W.make(name, 1.0)
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
res = D.partition_device(torch.from_numpy(w.xyz).to(dev), recs, centers, w.threshold, man_trans=w.man_trans,
                         time_stages=True)
cams_d = torch.from_numpy(recs).to(dev)
M_b = np.diff(res.ds_offsets.cpu().numpy())
C_b = np.diff(res.cam_offsets.cpu().numpy())
for _ in range(reps):
    v = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets, res.cam_index)
torch.cuda.synchronize()
print(json.dumps({"workload": name, "tests": int((M_b * C_b).sum()), "pairs": int(C_b.sum()),
                  "downsampled_points": int(M_b.sum()), "kept": int(v.keep.sum().item()),
                  "partition_stage_ms": {k: round(t, 3) for k, t in res.timings_ms.items()}}))
