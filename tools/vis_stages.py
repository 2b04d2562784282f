"""GPU box: stage times of the visibility pass (A7-A9) of a configuration, L2 flushed between passes.
usage: python tools/vis_stages.py [c2] [reps] [noprefilter] [sample_stride]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, workload as W

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
prefilter = not (len(sys.argv) > 3 and sys.argv[3] == "noprefilter")
stride = int(sys.argv[4]) if len(sys.argv) > 4 else D.SAMPLE_STRIDE
dev = torch.device("cuda", 0)
w = W.make(name, 1.0)
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
res = D.partition_device(torch.from_numpy(w.xyz).to(dev), recs, centers, w.threshold, man_trans=w.man_trans)
cams_d = torch.from_numpy(recs).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(reps):
    flush.fill_(i)
    tm = D.StageTimer(True)
    tm.mark("start")
    v = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets, res.cam_index, timer=tm,
                     hull_prefilter=prefilter, sample_stride=stride)
    tm.mark("end")
    print(json.dumps({"workload": name, "rep": i, "hull_prefilter": prefilter, "sample_stride": stride,
                      "kept": int(v.keep.sum().item()), "ratio_sum": float(v.ratio.sum().item()), "rounds": v.rounds, "n_exact": v.n_exact,
                      "cand_total": int(v.cand_count.sum().item()), "cand_max": int(v.cand_count.max().item()),
                      "stage_ms": {k: round(t, 3) for k, t in tm.result().items()}}))
