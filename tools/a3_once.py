"""GPU box: A3 (balanced median split) alone on a configuration's filtered cloud.
usage: python tools/a3_once.py [c4] [reps]"""
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
if w.man_trans is not None:
    x, y, z = D.aos_to_soa_manhattan(xyz, w.man_trans)
else:
    x, y, z = D.aos_to_soa(xyz)
keep, _ = D.radius_outlier_mask(x, y, z)
fx, fy, nf = D.compact_xy(x, y, keep)
root = D.bounds_xy(fx, fy, nf)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for i in range(reps + 1):
    flush.fill_(i)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    lb, lp, ld, lc = D.balanced_split(fx, fy, nf, root, w.threshold)
    b.record()
    torch.cuda.synchronize()
    if i:
        ts.append(a.elapsed_time(b))
import numpy as np
print(json.dumps({"workload": name, "dtype": str(fx.dtype), "filtered": int(nf.item()), "leaves": len(lb),
                  "count_sum": int(np.asarray(lc).sum()), "bounds_checksum": float(np.asarray(lb).sum()),
                  "ms_best": round(min(ts), 3), "ms_mean": round(sum(ts) / len(ts), 3)}))
