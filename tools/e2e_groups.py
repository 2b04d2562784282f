"""GPU box: where the end-to-end (host buffers in, host results out) visibility call spends its time.
Times (a) the bare pinned host -> device copy of the expanded blocks, (b) visibility_select_host for
several upload-group counts.  usage: python tools/e2e_groups.py [c4] [groups ...]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import device as D, workload as W

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
dev = torch.device("cuda", 0)
w = W.make(name, 1.0)
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
xyz = torch.from_numpy(w.xyz).to(dev)
res = D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans)
x, y, z = D.aos_to_soa(xyz)
if w.man_trans is not None:
    x, y, z = D.manhattan_apply(x, y, z, w.man_trans)
idx = res.pt_index.to(torch.int64)
host = torch.empty((idx.shape[0], 3), dtype=x.dtype, pin_memory=True)
for s in range(0, idx.shape[0], 8 << 20):
    i = idx[s:s + (8 << 20)]
    host[s:s + (8 << 20)].copy_(torch.stack([x[i], y[i], z[i]], dim=1))
torch.cuda.synchronize()
blk_off = res.pt_offsets.cpu().numpy()
recs_pinned = torch.from_numpy(recs).pin_memory()
cam_off = res.cam_offsets.cpu().numpy()
cam_idx = res.cam_index.cpu().numpy()
del x, y, z, idx

nbytes = host.numel() * host.element_size()
dst = torch.empty_like(host, device=dev)
for _ in range(2):
    dst.copy_(host, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    dst.copy_(host, non_blocking=True)
torch.cuda.synchronize()
h2d_ms = (time.perf_counter() - t0) / 3 * 1e3
print(json.dumps({"workload": name, "h2d_bytes": nbytes, "bare_h2d_ms": round(h2d_ms, 3),
                  "h2d_GBps": round(nbytes / h2d_ms / 1e6, 2)}))
del dst
configs = [(4, 2), ([0.35, 0.3, 0.2, 0.1, 0.05], 2), ([0.4, 0.3, 0.2, 0.1], 2), ([0.3, 0.3, 0.2, 0.12, 0.08], 2),
           ([0.45, 0.3, 0.15, 0.1], 2), ([0.3, 0.25, 0.2, 0.15, 0.1], 2), ([0.35, 0.3, 0.2, 0.1, 0.05], 3),
           ([0.5, 0.3, 0.2], 2), ([0.3, 0.25, 0.2, 0.15, 0.1], 3)]
for g, streams in configs:
    one = lambda: D.visibility_select_host(host, blk_off, recs_pinned, cam_off, cam_idx, dev, groups=g, streams=streams)
    for _ in range(2):
        ref = one()
    ts = []
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        outp = one()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    print(json.dumps({"groups": g, "streams": streams, "ms_best": round(min(ts), 3),
                      "ms_mean": round(float(np.mean(ts)), 3), "kept": int(outp[0][4].sum())}), flush=True)
tr = []
D.visibility_select_host(host, blk_off, recs_pinned, cam_off, cam_idx, dev, groups=[0.35, 0.3, 0.2, 0.1, 0.05], streams=2, trace=tr)
print(json.dumps({"trace_decreasing_streams2": tr}))
