"""GPU box: e2e time of visibility_select_host for several group counts."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ptgs_b200
from ptgs_b200 import device as D, workload as W
w = W.make("c2", 1.0)
dev = torch.device("cuda", 0)
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
xyz = torch.from_numpy(w.xyz).to(dev)
res = D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans)
x, y, z = D.aos_to_soa(xyz)
x, y, z = D.manhattan_apply(x, y, z, w.man_trans)
idx = res.pt_index.to(torch.int64)
pts = torch.stack([x[idx], y[idx], z[idx]], 1).contiguous()
host = torch.empty(pts.shape, dtype=pts.dtype, pin_memory=True); host.copy_(pts)
blk_off = res.pt_offsets.cpu().numpy(); cam_off = res.cam_offsets.cpu().numpy(); cam_idx = res.cam_index.cpu().numpy()
recs_p = torch.from_numpy(recs).pin_memory()
torch.cuda.synchronize()
t0 = time.perf_counter(); d = host.to(dev, non_blocking=True); torch.cuda.synchronize(); print("H2D alone %.3f ms (%d MB)" % ((time.perf_counter() - t0) * 1e3, host.numel() * host.element_size() >> 20))
for G in (1, 2, 3, 4, 8):
    for rep in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        out, m = D.visibility_select_host(host, blk_off, recs_p, cam_off, cam_idx, dev, groups=G)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("groups %d: %.3f ms  kept %d" % (G, dt * 1e3, out[0][-1]))
