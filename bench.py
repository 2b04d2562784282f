#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native point-count partitioner.

Metric (BASELINE.json): point-camera visibility tests/s (one test = one
(voxel-down-sampled point, assigned camera) evaluation of
scene/ptgs/shen_partation_densify.py:429-471), plus the end-to-end partition
wall time of the same scene (`partition` object of the JSON line).

A "step" is one visibility pass (A7-A9: sweep 1, inner polygons, sweep 2, exact
hulls, camera selection, and for N > 1 the all-gather of the per-pair records)
over every (block, camera) pair of the workload, with the down-sampled blocks
resident in HBM.  `e2e` is the same metric through the host-buffer entry
(`visibility_select_host`: H2D of the expanded blocks' points and the camera
records, A6 voxel down-sampling, A7-A9, D2H of the kept-camera lists).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2] [--scale s]
  python bench.py --impl reference ...     (CPU arm: the oracle port, rank 0 only)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "point_camera_visibility_tests_per_s"
UNIT = "tests/s"
L2_FLUSH_BYTES = 256 << 20        # > 126 MB L2
SLOTS_PER_TEST = 34               # FP32 issue slots per test


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=20.0)
    ap.add_argument("--sample-file", default=None, help="internal: cpu_baseline leg input")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, local, world


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(v, world, dev):
    if world <= 1:
        return v
    import torch.distributed as dist
    t = torch.tensor([v], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sweep_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of one sweep_kernel<1> launch on
    the bench workload, from the committed ncu capture (profiles/r01_sweep_traffic.json)."""
    p = os.path.join(ROOT, "profiles", "r01_sweep_traffic.json")
    try:
        with open(p) as f:
            return float(json.load(f)["dram_bytes_per_launch"])
    except Exception:
        return None


def run_ours(args):
    import ptgs_b200  # noqa: F401
    from ptgs_b200 import _lib, device as D, workload as W
    rank, local, world = dist_setup(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    lib = _lib.load(build_if_missing=False)
    hbm_peak, sm_max, peak_kind = peaks()

    w = W.make(args.workload, args.scale)
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    centers = D.camera_centers(w.R, w.T)
    xyz_pinned = torch.from_numpy(w.xyz).pin_memory()
    xyz = xyz_pinned.to(dev)

    # ---- full partition once (also the warm-up of every kernel) -------------
    res = D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans)
    torch.cuda.synchronize()
    L = len(res.leaf_ids)
    ds_off_h = res.ds_offsets.cpu().numpy()
    cam_off_h = res.cam_offsets.cpu().numpy()
    M_b = np.diff(ds_off_h)
    C_b = np.diff(cam_off_h)
    tests = int((M_b * C_b).sum())
    pairs = int(C_b.sum())
    cams_d = torch.from_numpy(recs).to(dev)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)

    def vis_step(timer=None):
        v = D.visibility_sharded(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d,
                                 res.cam_offsets, res.cam_index, rank, world, timer=timer)
        k_off, k_idx = D.select_cameras(v.keep, res.cam_offsets, res.cam_index)
        return v, k_off, k_idx

    for _ in range(max(args.warmup, 3)):
        flush.fill_(1)
        vis_step()
    barrier(world)

    # ---- timed region: K steps, each bracketed by events, L2 flushed between --
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    lib.ptgs_reset_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    timers = []
    barrier(world)
    wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.fill_(i & 0xff)
        tm = D.StageTimer(True)
        ev[i][0].record()
        tm.mark("start")
        v, k_off, k_idx = vis_step(tm)
        tm.mark("select")
        ev[i][1].record()
        timers.append(tm)
    barrier(world)
    wall = time.perf_counter() - wall0
    launches = int(lib.ptgs_launch_count())
    clocks = sampler.stop() if rank == 0 else None
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = max_over_ranks(float(sum(step_ms)), world, dev)
    ms_per_step = total_ms / args.steps
    value = tests / (ms_per_step * 1e-3)
    stage = {}
    for tm in timers:
        for k, t in tm.result().items():
            stage[k] = stage.get(k, 0.0) + t / args.steps
    n_kept = int(k_off[-1].item())

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": W.describe(args.workload, args.scale, w),
            "leaves": L, "pairs": pairs, "tests_per_step": tests,
            "downsampled_points": int(M_b.sum()), "kept_cameras": n_kept,
            "kept_fraction": n_kept / max(pairs, 1),
            "parallelism": "(block, camera) pairs sharded in %d contiguous equal-cost ranges, per-pair records all-gathered" % world if world > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MB write)",
        },
        "gpu_launches": launches, "clocks": clocks, "wall_s_timed_region": wall,
        "stage_ms": {k: round(t, 4) for k, t in stage.items()},
    }
    # ---- roofline of the dominant kernel (sweep 1: sweep_kernel<1>) -----------
    k_ms = stage.get("vis_sweep1_count", float("nan"))
    shard_tests = tests / world
    alg_bytes = float(M_b.sum()) * 12 + float(C_b.sum()) / world * 80 + pairs / world * 16
    lane_peak = 148 * 128 * sm_max * 1e6
    out["roofline"] = {
        "kernel": "sweep_kernel<1> (ptgs_visibility_count)", "bound": "hbm",
        "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
        "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak, "traffic": sweep_traffic_bytes(),
        "peak_kind": peak_kind,
        "algorithmic_bytes": alg_bytes, "kernel_ms": k_ms,
        "note": "the sweep re-uses every staged point tile for all cameras of the block, so its HBM traffic "
                "is 12 B/point by construction and the binding roof is FP32 issue (see fp32; DESIGN.md 4.1); "
                "`traffic` = dram read+write bytes of one launch from the ncu --set full capture in profiles/",
        "fp32": {"slots_per_test": SLOTS_PER_TEST,
                 "bruteforce_roof_tests_per_s": lane_peak / SLOTS_PER_TEST,
                 "achieved_tests_per_s": shard_tests / (k_ms * 1e-3),
                 "ratio": shard_tests / (k_ms * 1e-3) / (lane_peak / SLOTS_PER_TEST),
                 "peak_kind": "148 SMs x 128 lanes x sm_max_mhz / 34 issue slots per test",
                 "note": "roof of a kernel that evaluates EVERY (point, camera) test at full FP32 issue rate; the "
                         "sweep decides most tests per box (tile / warp run) without touching the points, so the "
                         "ratio can exceed 1 -- it is an algorithmic speed, not an issue-slot utilisation "
                         "(that is in profiles/: smsp__issue_active, inst_executed)"},
    }
    ctx = dict(w=w, res=res, recs=recs, centers=centers, xyz_pinned=xyz_pinned, xyz=xyz, dev=dev,
               tests=tests, world=world, rank=rank, flush=flush, lib=lib, D=D)
    measure_e2e(args, out, ctx)
    measure_partition(args, out, ctx)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        measure_cpu_baseline(args, out, ctx)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)


def _host_blocks(ctx):
    """Expanded blocks' points back to back on the host (pinned), in the frame
    and dtype the reference's blocks carry (float64 under a Manhattan
    transform, shen_data_read.py:126-130).  Untimed preparation."""
    D, res, w, dev = ctx["D"], ctx["res"], ctx["w"], ctx["dev"]
    x, y, z = D.aos_to_soa(ctx["xyz"])
    if w.man_trans is not None:
        x, y, z = D.manhattan_apply(x, y, z, w.man_trans)
    idx = res.pt_index.to(torch.int64)
    pts = torch.stack([x[idx], y[idx], z[idx]], dim=1).contiguous()
    host = torch.empty(pts.shape, dtype=pts.dtype, pin_memory=True)
    host.copy_(pts)
    torch.cuda.synchronize()
    return host, res.pt_offsets.cpu().numpy()


def measure_e2e(args, out, ctx):
    D, res, dev, world, rank = ctx["D"], ctx["res"], ctx["dev"], ctx["world"], ctx["rank"]
    blocks, blk_off = _host_blocks(ctx)
    ctx["blocks"], ctx["blk_off"] = blocks, blk_off
    recs_pinned = torch.from_numpy(ctx["recs"]).pin_memory()
    cam_off = res.cam_offsets.cpu().numpy()
    cam_idx = res.cam_index.cpu().numpy()
    steps = max(2, min(args.steps, 10))

    def one():
        return D.visibility_select_host(blocks, blk_off, recs_pinned, cam_off, cam_idx, dev,
                                        rank=rank, world=world)
    for _ in range(2):
        outp, m = one()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for i in range(steps):
        ctx["flush"].fill_(i & 0xff)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        outp, m = one()              # ends with D2H copies (synchronising)
        torch.cuda.synchronize()
        tot += time.perf_counter() - t0
    barrier(world)
    tot = max_over_ranks(tot, world, dev)
    if world > 1:
        # every rank uploads only the blocks its pair range touches: job total = sum over ranks
        from ptgs_b200 import multigpu as M
        import torch.distributed as dist
        lo, hi = M.range_bounds(cam_off, np.diff(blk_off), rank, world)
        b0 = int(np.searchsorted(cam_off, lo, side="right")) - 1 if hi > lo else 0
        b1 = int(np.searchsorted(cam_off, hi, side="left")) if hi > lo else 0
        mine = int(blk_off[b1] - blk_off[b0]) * blocks.shape[1] * blocks.element_size() + \
            recs_pinned.numel() * 4 + cam_off.nbytes + cam_idx.nbytes + (b1 - b0 + 1) * 16
        t = torch.tensor([mine], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        h2d = int(t.item())
    else:
        h2d = blocks.numel() * blocks.element_size() + recs_pinned.numel() * 4 + blk_off.nbytes + \
            cam_off.nbytes + cam_idx.nbytes
    d2h = sum(a.nbytes for a in outp)
    out["e2e"] = {"value": ctx["tests"] / (tot / steps), "unit": UNIT, "ms_per_step": tot / steps * 1e3,
                  "steps": steps, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                  "api": "ptgs_b200.device.visibility_select_host (host blocks -> A6 -> A7-A9 -> host kept lists)",
                  "timing": "host wall clock around the call, synchronised both sides, max over ranks"}


def measure_partition(args, out, ctx):
    """End-to-end partition wall time (second half of BASELINE.json's metric):
    device-resident cloud + host camera records -> every per-leaf / per-pair
    result (A1-A9).  Rank r evaluates its camera shard; A1-A6 are replicated."""
    D, w, dev, world, rank = ctx["D"], ctx["w"], ctx["dev"], ctx["world"], ctx["rank"]
    runs = []
    stages = None
    for i in range(4):
        ctx["flush"].fill_(i)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = D.partition_device(ctx["xyz"], ctx["recs"], ctx["centers"], w.threshold, man_trans=w.man_trans,
                               time_stages=True, cam_shard=(rank, world) if world > 1 else None)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if i > 0:
            runs.append(dt)
            stages = r.timings_ms
    best = max_over_ranks(min(runs), world, dev)
    # HBM fractions of the streaming stages (algorithmic bytes of DESIGN.md section 4 / stage time)
    res = ctx["res"]
    es = 8 if w.man_trans is not None else 4            # coordinate bytes (float64 under a Manhattan transform)
    N = float(w.xyz.shape[0])
    Nf = float(res.keep_mask.sum().item())
    Nb = float(res.pt_index.shape[0])
    levels = max(1, max(len(i) for i in res.leaf_ids))
    passes = 6 if es == 8 else 3
    hbm_peak = peaks()[0]
    alg = {"layout": N * (12 + 3 * es) if es == 8 else N * 24,
           "A1_radius": N * (3 * es + 8 + 4) * 2,
           "A2_bounds": Nf * 2 * es,
           "A3_split": levels * Nf * (es * passes + 4 * es),
           "A4_expand": N * 2 * es * 2 + Nb * 4,
           "A6_voxel": Nb * (3 * es + 8 + 4) * 2}
    stage_roof = {}
    for k, b in alg.items():
        t = (stages or {}).get(k)
        if t:
            stage_roof[k] = {"algorithmic_bytes": b, "GBps": b / (t * 1e-3) / 1e9,
                             "hbm_frac": b / (t * 1e-3) / 1e9 / hbm_peak}
    out["partition_stage_rooflines"] = stage_roof
    out["partition"] = {"wall_ms_best": best * 1e3, "wall_ms_mean": float(np.mean(runs)) * 1e3, "runs": len(runs),
                        "stage_ms": {k: round(v, 3) for k, v in (stages or {}).items()},
                        "scope": "A1-A9 from the device-resident cloud to device-resident results (+ host scalars)"}


def measure_cpu_baseline(args, out, ctx):
    """Oracle port of A6-A9 on a bounded sample, in a subprocess (fork pool
    without a CUDA context)."""
    res, blocks, blk_off = ctx["res"], ctx["blocks"], ctx["blk_off"]
    cam_off = res.cam_offsets.cpu().numpy()
    cam_idx = res.cam_index.cpu().numpy()
    b = int(np.argmax(np.diff(cam_off) > 0))
    path = "/tmp/ptgs_cpu_sample_%d.npz" % os.getpid()
    np.savez(path, points=blocks[blk_off[b]:blk_off[b + 1]].numpy(),
             recs=ctx["recs"][cam_idx[cam_off[b]:cam_off[b + 1]]])
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--sample-file", path,
                            "--steps", "2", "--warmup", "0", "--cpu-seconds", str(args.cpu_seconds)],
                           capture_output=True, text=True, timeout=600)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")]
        out["cpu_baseline"] = json.loads(line[-1])["cpu_baseline"] if line else {"error": r.stderr[-400:]}
    except Exception as e:  # the baseline is informative; never fail the bench on it
        out["cpu_baseline"] = {"error": repr(e)}
    finally:
        if os.path.exists(path):
            os.remove(path)


def _reference_prep(args):
    """CPU-only preparation of the reference arm: scene -> (first non-empty
    block's expanded points, its cameras' records) with the oracle (A1-A5).
    Cached in /tmp because the exact radius filter costs about a minute."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ptgs_oracle as O
    import ptgs_b200  # noqa: F401
    from ptgs_b200 import workload as W
    cache = "/tmp/ptgs_refprep_%s_%g.npz" % (args.workload, args.scale)
    if os.path.exists(cache):
        z = np.load(cache, allow_pickle=False)
        return z["points"], z["recs"], str(z["desc"]), int(z["cores"])
    w = W.make(args.workload, args.scale)
    pts = w.xyz if w.man_trans is None else O.manhattan_points(w.xyz, w.man_trans)
    keep = O.radius_outlier_mask(pts.astype(np.float64))
    filt = pts.astype(np.float64)[keep]
    leaves = O.quad_tree_partition(filt, O.bounds_xy(filt), w.threshold)
    centers = O.camera_centers(w.R, w.T)
    boxes = [O.expanded_box(l.bounds) for l in leaves]
    cams = O.assign_cameras(boxes, centers)
    b = next(i for i, c in enumerate(cams) if len(c) > 0)
    block = pts[O.expand_mask(pts, boxes[b])]
    recs = np.stack([O.camera_record(w.R[c], w.T[c], w.FoVx[c], w.FoVy[c], int(w.width[c]), int(w.height[c]))
                     for c in cams[b]])
    desc = W.describe(args.workload, args.scale, w)
    np.savez(cache, points=block, recs=recs, desc=desc, cores=os.cpu_count() or 1)
    return block, recs, desc, os.cpu_count() or 1


def run_reference(args):
    """CPU arm: the oracle port of the visibility stage (A6-A9) on the host
    cores, on a bounded sample of the same workload.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_arm
    if args.sample_file:
        z = np.load(args.sample_file)
        block, recs, desc = z["points"], z["recs"], "sample file"
        budget = args.cpu_seconds
    else:
        block, recs, desc, _ = _reference_prep(args)
        budget = 90.0
    r = cpu_arm.run(block, recs, args.steps, args.warmup, budget)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "sample": r["sample"]},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
