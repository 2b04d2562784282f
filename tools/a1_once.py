"""GPU box: A1 (radius outlier filter) alone on a configuration's cloud, L2 flushed between passes.
usage: python tools/a1_once.py [c4] [reps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, workload as W

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
w = W.make(name, 1.0)
xyz = torch.from_numpy(w.xyz).to(dev)
x, y, z = D.aos_to_soa(xyz)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for i in range(reps + 1):
    flush.fill_(i)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    keep, n_keep = D.radius_outlier_mask(x, y, z, xyz_aos=xyz)
    b.record()
    torch.cuda.synchronize()
    if i:
        ts.append(a.elapsed_time(b))
print(json.dumps({"workload": name, "points": int(xyz.shape[0]), "kept": int(n_keep.item()),
                  "mask_checksum": int(keep.to(torch.int64).mul(torch.arange(keep.shape[0], device=dev) % 1000003).sum().item()),
                  "ms_best": round(min(ts), 3), "ms_mean": round(sum(ts) / len(ts), 3)}))
