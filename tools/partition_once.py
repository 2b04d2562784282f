"""One warm partition_device run of a BASELINE config inside a cudaProfiler range
(ncu --profile-from-start off ...).  usage: python tools/partition_once.py [c2]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, workload as W

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
w = W.make(name, 1.0)
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
xyz = torch.from_numpy(w.xyz).to("cuda:0")
for i in range(3):
    if i == 2:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    res = D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans, time_stages=(i == 1))
    torch.cuda.synchronize()
    if i == 1:
        print({k: round(v, 3) for k, v in res.timings_ms.items()})
torch.cuda.profiler.stop()
