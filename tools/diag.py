"""Diagnostics on the GPU box: per-stage host/device timing and candidate statistics."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ptgs_b200
from ptgs_b200 import device as D, workload as W

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hs = float(sys.argv[3]) if len(sys.argv) > 3 else None
dev = torch.device("cuda", 0)
w = W.make(name, scale, **({} if hs is None else dict(height_scale=hs)))
recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
centers = D.camera_centers(w.R, w.T)
xyz = torch.from_numpy(w.xyz).to(dev)


def T(label, fn, rep=3):
    out = None
    for i in range(rep):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print("%-28s run %d: %9.3f ms" % (label, i, dt * 1e3), flush=True)
    return out

res = T("partition_device", lambda: D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans))
x, y, z = D.aos_to_soa(xyz)
if w.man_trans is not None:
    x, y, z = T("manhattan", lambda: D.manhattan_apply(x, y, z, w.man_trans))
keep, nk = T("radius", lambda: D.radius_outlier_mask(x, y, z))
fx, fy, nf = T("compact", lambda: D.compact_xy(x, y, keep))
root = T("bounds", lambda: D.bounds_xy(fx, fy, nf))
lb, lp, ld, lc = T("split", lambda: D.balanced_split(fx, fy, nf, root, w.threshold))
boxes = D.expanded_boxes(lb)
off, idx = T("expand", lambda: D.expand_select(x, y, boxes))
coff, cidx = T("assign", lambda: D.assign_cameras(centers, boxes, dev))
ds = T("voxel", lambda: D.voxel_downsample(x, y, z, off, idx))
cams_d = torch.from_numpy(recs).to(dev)
vis = T("visibility", lambda: D.visibility(ds[1], ds[2], ds[3], ds[0], cams_d, coff, cidx))
T("select", lambda: D.select_cameras(vis.keep, coff, cidx))
nh = vis.n_hit.cpu().numpy(); cc = vis.cand_count.cpu().numpy(); rt = vis.ratio.cpu().numpy()
print("leaves", len(lb), "pairs", len(nh), "rounds", vis.rounds)
print("n_hit   pct[0,10,50,90,99,100]:", np.percentile(nh, [0, 10, 50, 90, 99, 100]))
print("cand    pct[0,10,50,90,99,100]:", np.percentile(cc, [0, 10, 50, 90, 99, 100]))
print("ratio   pct[0,10,50,90,99,100]:", np.percentile(rt, [0, 10, 50, 90, 99, 100]))
print("kept fraction", float(vis.keep.float().mean()), "flags", np.bincount(vis.flags.cpu().numpy(), minlength=16)[:16])
M = np.diff(ds[0].cpu().numpy()); Cb = np.diff(coff.cpu().numpy())
print("M_b", M.min(), M.max(), "C_b", Cb.min(), Cb.max(), "tests", int((M * Cb).sum()))
print("hits", int(nh.sum()), "hit rate", float(nh.sum()) / float((M * Cb).sum()))
ex = (vis.flags.cpu().numpy() & 16) == 0
print("rounds", vis.rounds, "n_exact", vis.n_exact, "sum cand (exact pairs)", int(cc[ex].sum()), "max", int(cc.max()),
      "sum min(nh,max(2048,nh/4)) over exact", int(np.minimum(nh[ex], np.maximum(2048, nh[ex] // 4)).sum()))
gap = (vis.ratio_hi - vis.ratio_lo).cpu().numpy()
print("bound gap pct[50,90,99,100]", np.percentile(gap, [50, 90, 99, 100]))
import ctypes
from ptgs_b200 import _lib
lib = _lib.load(); st4 = (ctypes.c_uint64 * 8)()
lib.ptgs_debug_sweep_stats(None, 1)
tm = D.StageTimer(True); tm.mark("start")
D.visibility(ds[1], ds[2], ds[3], ds[0], cams_d, coff, cidx, timer=tm)
print({k: round(v, 3) for k, v in tm.result().items()})
lib.ptgs_debug_sweep_stats(st4, 0)
print("sweep stats [presented, after CTA cull, swept s1, swept s2, warp all-in, tile all-in, slow passes, slow groups]", list(st4))
i = int(np.argmax(cc))
print("worst pair", i, "n_hit", nh[i], "cand", cc[i], "rlo", float(vis.ratio_lo[i]), "rhi", float(vis.ratio_hi[i]),
      "ratio", rt[i], "flags", int(vis.flags[i]))
for stride in (16, 32, 64, 128, 256):
    best = None
    for rep in range(3):
        tm = D.StageTimer(True); tm.mark("start")
        D.visibility(ds[1], ds[2], ds[3], ds[0], cams_d, coff, cidx, timer=tm, sample_stride=stride)
        r = tm.result()
        t = r["vis_interior_rects"] + r["vis_sweep1_count"]
        best = t if best is None else min(best, t)
    print("sample stride %4d: interior + sweep1 = %.3f ms (last: %.3f + %.3f)" % (stride, best, r["vis_interior_rects"], r["vis_sweep1_count"]))
