"""GPU box: A6 (voxel down-sampling of every block) alone, index-list form as partition_device calls it
and rows form as the host-buffer entry calls it.
usage: python tools/a6_once.py [c4] [reps]"""
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
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
xyz = torch.from_numpy(w.xyz).to(dev)
res = D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans)
x, y, z = D.aos_to_soa(xyz)
rows_src = xyz
if w.man_trans is not None:
    x, y, z = D.manhattan_apply(x, y, z, w.man_trans)
    rows_src = torch.stack([x, y, z], dim=1).contiguous()
rows = rows_src[res.pt_index.to(torch.int64)].contiguous()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn):
    ts = []
    for i in range(reps):
        flush.fill_(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return round(min(ts), 3), round(sum(ts) / len(ts), 3), int(out[0][-1].item())


for mode in ("pairs",):
    r1 = timed(lambda: D.voxel_downsample(x, y, z, res.pt_offsets, res.pt_index, xyz_aos=rows_src))
    r2 = timed(lambda: D.voxel_downsample(None, None, None, res.pt_offsets, None, n_entries=rows.shape[0],
                                          xyz_aos=rows))
    print(json.dumps({"workload": name, "sort": mode, "entries": int(rows.shape[0]),
                      "index_list_ms_best_mean_voxels": r1, "rows_ms_best_mean_voxels": r2}))
