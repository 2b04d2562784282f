#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native point-count partitioner.

Metric (BASELINE.json): point-camera visibility tests/s (one test = one
(voxel-down-sampled point, assigned camera) evaluation of
scene/ptgs/shen_partation_densify.py:429-471), plus the end-to-end partition
wall time of the same scene (`partition` object of the JSON line).

Default workload: **c4** (50 M points / 8 000 cameras / threshold 500 000), the
largest block-partition configuration of BASELINE.json -- it fits one B200.
A "step" is one visibility pass (A7-A9: sample pass, sweep 1, area bounds,
sweep 2, exact hulls, camera selection, and for N > 1 the exchange of the
per-pair records) over every (block, camera) pair of the workload, with the
down-sampled blocks resident in HBM.  `e2e` is the same metric through the
host-buffer entry (`visibility_select_host`: H2D of the expanded blocks' points
and the camera records, A6 voxel down-sampling, A7-A9, D2H of the kept-camera
lists).  The same line carries `sweep_c5` (config c5: 100 M points x 16 000
cameras, Morton order + coverage sweep, point-sharded at N > 1), `partition`
(A1-A9 wall time, to host-resident results), `parity` (the CPU arm's per-pair
results compared with the GPU's for the same pairs) and `cpu_baseline`.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4] [--scale s]
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
    ap.add_argument("--workload", default="c4", choices=["c1", "c2", "c3", "c4"])
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c5", action="store_true", help="skip the sweep_c5 object")
    ap.add_argument("--c5-scale", type=float, default=1.0)
    ap.add_argument("--secondary", default="c2", help="second workload reported at N=1 ('' to skip)")
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
                 "--format=csv,noheader,nounits", "-lms", "20"],
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


def committed_profile(name):
    """A JSON summary of a committed ncu capture under profiles/ (None when absent)."""
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f)
    except Exception:
        return None


def load_workload(name, scale, rank, world, dev):
    """Rank 0 generates the synthetic scene (6 s at c4, 13 s at c5, several GB of host memory) and
    the other ranks receive the cloud over NCCL and the (small) camera arrays as an object, instead
    of every rank running the generator.  Returns (HostWorkload [xyz None on ranks > 0], xyz device)."""
    import ptgs_b200  # noqa: F401
    from ptgs_b200 import workload as W
    if world <= 1:
        w = W.make(name, scale)
        return w, torch.from_numpy(w.xyz).pin_memory().to(dev)
    import copy
    import torch.distributed as dist
    box = [None]
    w = None
    if rank == 0:
        w = W.make(name, scale)
        light = copy.copy(w)
        light.xyz = None
        box = [(light, len(w.xyz))]
    dist.broadcast_object_list(box, src=0)
    light, n = box[0]
    if rank == 0:
        xyz = torch.from_numpy(w.xyz).pin_memory().to(dev)
    else:
        w = light
        xyz = torch.empty((n, 3), dtype=torch.float32, device=dev)
    dist.broadcast(xyz, src=0)
    return w, xyz


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

    w, xyz = load_workload(args.workload, args.scale, rank, world, dev)
    n_points = int(xyz.shape[0])
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    centers = D.camera_centers(w.R, w.T)

    # clocks / throttle reasons are sampled (20 ms period) from here -- the warm-up partition and
    # visibility passes -- to the end of the timed region: the timed steps alone last a few tens of
    # milliseconds, less than nvidia-smi needs to start
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
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
    # static part of the sharded pass (every rank's pair range; L-sized host arithmetic), built once
    # per partition result
    plan = D.range_plan(res.ds_offsets, res.cam_offsets, res.cam_index, rank, world,
                        cam_off_h=cam_off_h, block_cost_h=M_b) if world > 1 else None

    def vis_step(timer=None):
        v = D.visibility_sharded(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d,
                                 res.cam_offsets, res.cam_index, rank, world, timer=timer, plan=plan)
        k_off, k_idx = D.select_cameras(v.keep, res.cam_offsets, res.cam_index, sync=False)
        return v, k_off, k_idx

    for _ in range(max(args.warmup, 3)):
        flush.fill_(1)
        vis_step()
    barrier(world)

    # ---- timed region: K steps, each bracketed by events, L2 flushed between --
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
    flags_h = v.flags.cpu().numpy()
    stage_ranks = None
    if world > 1:
        # per-rank stage times (which rank the others wait for at the exchange)
        import torch.distributed as dist
        keys = sorted(stage)
        mine_t = torch.tensor([stage[k] for k in keys], dtype=torch.float64, device=dev)
        allt = [torch.empty_like(mine_t) for _ in range(world)]
        dist.all_gather(allt, mine_t)
        tab = torch.stack(allt).cpu().numpy()
        stage_ranks = {k: [round(float(x), 4) for x in tab[:, i]] for i, k in enumerate(keys)}

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": W.describe(args.workload, args.scale, w, n_points),
            "leaves": L, "pairs": pairs, "tests_per_step": tests,
            "downsampled_points": int(M_b.sum()), "kept_cameras": n_kept,
            "kept_fraction": n_kept / max(pairs, 1),
            "parallelism": ("(block, camera) pairs sharded in %d x %d interleaved contiguous equal-cost ranges, "
                            "per-pair records exchanged with one padded all-gather" % (world, plan.chunks))
            if world > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MB write)",
        },
        "gpu_launches": launches, "clocks": clocks, "wall_s_timed_region": wall,
        "stage_ms": {k: round(t, 4) for k, t in stage.items()},
    }
    if stage_ranks is not None:
        out["stage_ms_per_rank"] = stage_ranks
    # ---- roofline of the dominant kernel (sweep 1: sweep_kernel<1>), this rank's launch ---------
    k_ms = stage.get("vis_sweep1_count", float("nan"))
    if world > 1:
        mine = plan.local_counts()                       # this rank's pairs per block
        my_pts = float(M_b[mine > 0].sum())
        my_pairs = float(mine.sum())
        my_tests = float((mine * M_b).sum())
    else:
        my_pts, my_pairs, my_tests = float(M_b.sum()), float(pairs), float(tests)
    alg_bytes = my_pts * 12 + my_pairs * 80 + my_pairs * 16
    lane_peak = 148 * 128 * sm_max * 1e6
    ncu = committed_profile("r02_sweep1_ncu.json") or {}
    out["roofline"] = {
        "kernel": "sweep_kernel<1> (ptgs_visibility_count), rank 0's launch", "bound": "hbm",
        "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
        "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak, "traffic": ncu.get("dram_bytes_per_launch"),
        "peak_kind": peak_kind,
        "algorithmic_bytes": alg_bytes, "kernel_ms": k_ms,
        "note": "the sweep re-uses every staged point tile for all cameras of the block, so its HBM traffic "
                "is 12 B/point by construction and the binding roof is FP32 issue (see fp32; DESIGN.md 4.1); "
                "`traffic` = dram read+write bytes of one launch from the ncu --set full capture in profiles/",
        "fp32": {"issue_active_pct": ncu.get("issue_active_pct"), "fma_pipe_pct": ncu.get("fma_pipe_pct"),
                 "alu_pipe_pct": ncu.get("alu_pipe_pct"),
                 "lane_ops_per_test": ncu.get("lane_ops_per_test"),
                 "source": ncu.get("source", "no committed capture"),
                 "slots_per_test_bruteforce": SLOTS_PER_TEST,
                 "bruteforce_roof_tests_per_s": lane_peak / SLOTS_PER_TEST,
                 "achieved_tests_per_s": my_tests / (k_ms * 1e-3),
                 "note": "issue_active / fma_pipe / lane_ops_per_test come from the committed ncu capture of this "
                         "kernel; bruteforce_roof = 148 SMs x 128 lanes x sm_max / 34 slots is the rate of a kernel "
                         "that evaluates EVERY test -- the sweep decides most tests per box, so achieved can exceed "
                         "it (algorithmic speed, not utilisation)"},
    }
    fp = measure_fp32_peak(D, dev)
    if fp:
        out["roofline"]["fp32"]["measured_fma_lane_ops_per_s"] = fp
    ctx = dict(w=w, res=res, recs=recs, centers=centers, xyz=xyz, dev=dev, tests=tests, world=world,
               rank=rank, flush=flush, lib=lib, D=D, vis=v, cam_off_h=cam_off_h, flags_h=flags_h, args=args)
    measure_e2e(args, out, ctx)
    measure_partition(args, out, ctx)
    out["parity"] = {"near_threshold_pairs_all": int(((flags_h & 1) != 0).sum()),
                     "first_ray_collinear_pairs_all": int(((flags_h & 2) != 0).sum()),
                     "pairs_all": int(len(flags_h))}
    c5 = None
    if not args.no_c5:
        c5 = prepare_c5(args, ctx)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        measure_cpu_baseline(args, out, ctx, c5)
    # release the c4 state before the 100 M-point sweep
    for k in ("res", "vis", "xyz", "blocks"):
        ctx.pop(k, None)
    del res, v, xyz
    torch.cuda.empty_cache()
    if c5 is not None:
        measure_c5(args, out, ctx, c5)
    if rank == 0 and world == 1 and args.secondary and args.secondary != args.workload:
        measure_secondary(args, out, ctx)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)


def measure_fp32_peak(D, dev):
    """FP32 FMA issue peak measured live (ptgs_debug_fma_probe: dependent-chain-free FMA loop on
    every SM) -> lane-ops/s.  BASELINE.md section 3 asks for the measured figure."""
    try:
        return D.fma_probe(dev)
    except Exception:
        return None


def _host_blocks(ctx):
    """Expanded blocks' points back to back on the host (pinned), in the frame
    and dtype the reference's blocks carry (float64 under a Manhattan
    transform, shen_data_read.py:126-130).  Untimed preparation."""
    D, res, w, dev = ctx["D"], ctx["res"], ctx["w"], ctx["dev"]
    x, y, z = D.aos_to_soa(ctx["xyz"])
    if w.man_trans is not None:
        x, y, z = D.manhattan_apply(x, y, z, w.man_trans)
    idx = res.pt_index.to(torch.int64)
    host = torch.empty((idx.shape[0], 3), dtype=x.dtype, pin_memory=True)
    step = 8 << 20                                   # bounded device staging
    for s in range(0, idx.shape[0], step):
        i = idx[s:s + step]
        host[s:s + step].copy_(torch.stack([x[i], y[i], z[i]], dim=1))
    torch.cuda.synchronize()
    return host, res.pt_offsets.cpu().numpy()


def measure_e2e(args, out, ctx):
    D, res, dev, world, rank = ctx["D"], ctx["res"], ctx["dev"], ctx["world"], ctx["rank"]
    blocks, blk_off = _host_blocks(ctx)
    ctx["blocks"], ctx["blk_off"] = blocks, blk_off
    recs_pinned = torch.from_numpy(ctx["recs"]).pin_memory()
    cam_off = res.cam_offsets.cpu().numpy()
    cam_idx = res.cam_index.cpu().numpy()
    steps = max(2, min(args.steps, 6))
    # upload groups per rank: each group costs ~0.9 ms of fixed work (A6's layout read-back, sample pass, hull
    # tail), so a rank that uploads little uses fewer (about one group per 96 MB of its own blocks, 1..4)
    local_bytes = blocks.numel() * blocks.element_size() / max(world, 1)
    groups = int(min(4, max(1, round(local_bytes / (96 << 20)))))

    def one():
        return D.visibility_select_host(blocks, blk_off, recs_pinned, cam_off, cam_idx, dev,
                                        rank=rank, world=world, groups=groups)
    for _ in range(2):
        outp, m = one()
    barrier(world)
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
    # the host entry must agree with the device-resident pass (same pairs, same decisions)
    same = bool(np.array_equal(outp[2], ctx["vis"].n_hit.cpu().numpy()) and
                np.array_equal(outp[4], ctx["vis"].keep.cpu().numpy()))
    out["e2e"] = {"value": ctx["tests"] / (tot / steps), "unit": UNIT, "ms_per_step": tot / steps * 1e3,
                  "steps": steps, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                  "upload_groups": groups, "equals_device_pass": same,
                  "api": "ptgs_b200.device.visibility_select_host (host blocks -> A6 -> A7-A9 -> host kept lists)",
                  "timing": "host wall clock around the call, synchronised both sides, max over ranks"}


def measure_partition(args, out, ctx):
    """End-to-end partition wall time (second half of BASELINE.json's metric):
    device-resident cloud + host camera records -> every per-leaf / per-pair result (A1-A9) ON THE
    HOST (leaf rectangles / ids / counts, expanded index CSR, per-pair n_hit / ratio / keep, kept
    CSR).  With N > 1 the visibility pass (A6-A9) is range-sharded and the records exchanged;
    A1-A5 are replicated."""
    D, w, dev, world, rank = ctx["D"], ctx["w"], ctx["dev"], ctx["world"], ctx["rank"]
    runs, runs_dev = [], []
    stages = None
    d2h = 0
    for i in range(4):
        ctx["flush"].fill_(i)
        barrier(world)
        t0 = time.perf_counter()
        r = D.partition_device(ctx["xyz"], ctx["recs"], ctx["centers"], w.threshold, man_trans=w.man_trans,
                               time_stages=True, shard=(rank, world) if world > 1 else None,
                               host_results=(rank == 0))
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        # every rank holds the full result; one host copy of it is the job's output (rank 0)
        host = D.partition_results_to_host(r) if rank == 0 else {}
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        if i > 0:
            runs.append(t2 - t0)
            runs_dev.append(t1 - t0)
            stages = r.timings_ms
        if rank == 0:
            d2h = sum(a.nbytes for a in host.values() if hasattr(a, "nbytes"))
        del r, host
    best = max_over_ranks(min(runs), world, dev)
    best_dev = max_over_ranks(min(runs_dev), world, dev)
    # HBM fractions of the streaming stages (algorithmic bytes of DESIGN.md section 4 / stage time)
    res = ctx["res"]
    es = 8 if w.man_trans is not None else 4            # coordinate bytes (float64 under a Manhattan transform)
    N = float(ctx["xyz"].shape[0])
    Nf = float(res.n_filtered)
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
    if world == 1:
        out["partition_stage_rooflines"] = stage_roof
    out["partition"] = {"wall_ms_best": best * 1e3, "wall_ms_mean": float(np.mean(runs)) * 1e3, "runs": len(runs),
                        "device_resident_ms_best": best_dev * 1e3, "d2h_bytes": int(d2h),
                        "stage_ms": {k: round(v, 3) for k, v in (stages or {}).items()},
                        "scope": "A1-A9 from the device-resident cloud to HOST-resident results (leaf tables, "
                                 "expanded index CSR, per-pair n_hit / ratio / keep, kept CSR); "
                                 "device_resident_ms_best stops before the device->host copies"}


def prepare_c5(args, ctx):
    """Config c5 inputs: the 100 M-point cloud and 16 000 camera records (Z-front test of the Vast
    path, data_partition.py:412)."""
    D, dev, rank, world = ctx["D"], ctx["dev"], ctx["rank"], ctx["world"]
    from ptgs_b200 import workload as W
    t0 = time.perf_counter()
    w5, xyz5 = load_workload("c5", args.c5_scale, rank, world, dev)
    recs5 = D.camera_records(w5.R, w5.T, w5.FoVx, w5.FoVy, w5.width, w5.height)
    recs5[:, 18] = 2.0
    x, y, z = D.aos_to_soa(xyz5)
    n = int(xyz5.shape[0])
    desc = W.describe("c5", args.c5_scale, w5, n)
    # parity sample: a 1 M-point slice x 16 random cameras, counted by the CPU oracle in the
    # cpu_baseline leg and by the GPU sweep here
    sample = None
    if rank == 0:
        rng = np.random.default_rng(5)
        s0 = int(rng.integers(0, max(n - 1_000_000, 1)))
        s1 = min(n, s0 + 1_000_000)
        # cameras near the slice would need the geometry; random cameras mostly see nothing of a
        # random file-order slice -- which is exactly the all-pairs workload (file order = random space)
        cams = np.sort(rng.choice(len(recs5), size=min(16, len(recs5)), replace=False))
        sample = dict(points=xyz5[s0:s1].cpu().numpy(), recs=recs5[cams], s0=s0, s1=s1, cams=cams)
    return dict(x=x, y=y, z=z, xyz=xyz5, recs=recs5, n=n, desc=desc, sample=sample, gen_s=time.perf_counter() - t0)


def measure_c5(args, out, ctx, c5):
    """sweep_c5: every point x every camera.  One step = Morton order of the (rank's slice of the)
    cloud + coverage sweep; N > 1 shards the POINTS (counts summed with one all-reduce)."""
    D, dev, rank, world = ctx["D"], ctx["dev"], ctx["rank"], ctx["world"]
    x, y, z, n = c5["x"], c5["y"], c5["z"], c5["n"]
    cams_d = torch.from_numpy(c5["recs"]).to(dev)
    C = int(cams_d.shape[0])
    from ptgs_b200 import multigpu as M
    p0, p1 = M.point_range(n, rank, world)
    xs, ys, zs = x[p0:p1], y[p0:p1], z[p0:p1]
    rows = c5["xyz"][p0:p1]                      # the same slice as (n,3) rows: cheaper gather source

    def step(timer=None):
        mark = timer.mark if timer is not None else (lambda name: None)
        if world > 1:
            # X-slab sharding: every rank selects, orders and sweeps a spatially compact n / N points
            nh = D.coverage_sweep_sharded(x, y, z, cams_d, rank, world, shard="slabs")
            mark("c5_sharded_step")
            return nh
        od = D.spatial_order(xs, ys, zs, xyz_aos=rows)
        mark("c5_order")
        nh = D.coverage_sweep(xs, ys, zs, cams_d, ordered=od)
        mark("c5_sweep")
        return nh
    step()
    barrier(world)
    reps = 3
    tot, stage = 0.0, {}
    for i in range(reps):
        ctx["flush"].fill_(i)
        tm = D.StageTimer(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(world)
        e0.record()
        tm.mark("start")
        n_hit = step(tm)
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
        for k, t in tm.result().items():
            stage[k] = stage.get(k, 0.0) + t / reps
    ms = max_over_ranks(tot / reps, world, dev)
    tests = float(n) * C
    hbm_peak = peaks()[0]
    nh64 = n_hit.to(torch.int64)
    obj = {"workload": c5["desc"], "ms_per_step": ms, "tests_per_step": tests, "tests_per_s": tests / (ms * 1e-3),
           "order_ms": round(stage.get("c5_order", 0.0), 3), "sweep_ms": round(stage.get("c5_sweep", 0.0), 3),
           "hits": int(nh64.sum().item()),
           "hits_checksum": int((nh64 * torch.arange(1, C + 1, device=dev)).sum().item()),
           "cameras_with_hits": int((n_hit > 0).sum().item()),
           "sharding": "X slabs (every rank selects the points between two quantiles of x, orders and sweeps them "
                       "against all cameras; int32 sum all-reduce)" if world > 1 else "single GPU",
           "scene_generation_s": round(c5["gen_s"], 1),
           "hbm": {"algorithmic_bytes_survey_Bc64": float(np.ceil(C / 64.0) * n * 12 + C * 84),
                   "frac_of_peak_at_that_figure": (np.ceil(C / 64.0) * n * 12 + C * 84) / world / (ms * 1e-3) / 1e9 / hbm_peak,
                   "note": "SURVEY.md 8(d)'s byte figure (camera batches of 64 re-reading the cloud); the sweep reads "
                           "the cloud once (n x 12 B) after the Morton sort and culls per tile, so this fraction is "
                           "an algorithmic speed relative to that schedule, not a DRAM utilisation"}}
    # parity: the GPU's counts of the sample slice x cameras (compared with the oracle's in `parity`)
    if rank == 0 and c5.get("sample") is not None and c5["sample"].get("want") is not None:
        s = c5["sample"]
        px, py, pz = D.aos_to_soa(torch.from_numpy(s["points"]).to(dev))
        got = D.coverage_sweep(px, py, pz, torch.from_numpy(s["recs"]).to(dev)).cpu().numpy()
        want = np.asarray(s["want"], dtype=np.int64)
        out["parity"]["c5_slice"] = {"points": int(len(s["points"])), "cameras": int(len(s["recs"])),
                                     "count_mismatch": int((got.astype(np.int64) != want).sum()),
                                     "hits": int(want.sum())}
    out["sweep_c5"] = obj


def measure_secondary(args, out, ctx):
    """A second BASELINE.json configuration at N = 1 (default c2: Manhattan rotation, float64 cloud):
    one partition + the visibility step, so the line also carries the round-1 headline workload."""
    D, dev = ctx["D"], ctx["dev"]
    from ptgs_b200 import workload as W
    name = args.secondary
    w = W.make(name, 1.0)
    recs = D.camera_records(w.R, w.T, w.FoVx, w.FoVy, w.width, w.height)
    centers = D.camera_centers(w.R, w.T)
    xyz = torch.from_numpy(w.xyz).to(dev)
    best = None
    for i in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = D.partition_device(xyz, recs, centers, w.threshold, man_trans=w.man_trans, time_stages=(i == 2))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None or (i > 0 and dt < best) else best
    M_b = np.diff(res.ds_offsets.cpu().numpy())
    C_b = np.diff(res.cam_offsets.cpu().numpy())
    tests = int((M_b * C_b).sum())
    cams_d = torch.from_numpy(recs).to(dev)
    ms = []
    for i in range(8):          # 3 untimed passes: the allocator cache was emptied before the c5 sweep
        ctx["flush"].fill_(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        v = D.visibility(res.ds_x, res.ds_y, res.ds_z, res.ds_offsets, cams_d, res.cam_offsets, res.cam_index)
        D.select_cameras(v.keep, res.cam_offsets, res.cam_index, sync=False)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            ms.append(e0.elapsed_time(e1))
    out["secondary"] = {"workload": W.describe(name, 1.0, w), "leaves": len(res.leaf_ids), "pairs": int(C_b.sum()),
                        "tests_per_step": tests, "visibility_ms_per_step": float(np.mean(ms)),
                        "tests_per_s": tests / (float(np.mean(ms)) * 1e-3), "partition_device_ms_best": best * 1e3,
                        "partition_stage_ms": {k: round(t, 3) for k, t in res.timings_ms.items()},
                        "kept_cameras": int(v.keep.sum().item())}


def measure_cpu_baseline(args, out, ctx, c5):
    """Oracle port of A6-A9 on a bounded sample, in a subprocess (fork pool without a CUDA
    context).  The same leg is the CHECKER of the at-size parity report: the per-camera
    (n_hit, ratio, keep) it computes for its block x cameras are compared with the GPU's results
    for the same pairs, and the c5 slice is counted by the oracle's coverage_counts."""
    res, blocks, blk_off = ctx["res"], ctx["blocks"], ctx["blk_off"]
    cam_off = ctx["cam_off_h"]
    cam_idx = res.cam_index.cpu().numpy()
    b = int(np.argmax(np.diff(cam_off) > 0))
    path = "/tmp/ptgs_cpu_sample_%d.npz" % os.getpid()
    arrays = dict(points=blocks[blk_off[b]:blk_off[b + 1]].numpy(),
                  recs=ctx["recs"][cam_idx[cam_off[b]:cam_off[b + 1]]])
    if c5 is not None and c5.get("sample") is not None:
        arrays["c5_points"], arrays["c5_recs"] = c5["sample"]["points"], c5["sample"]["recs"]
    np.savez(path, **arrays)
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--sample-file", path,
                            "--steps", "2", "--warmup", "0", "--cpu-seconds", str(args.cpu_seconds)],
                           capture_output=True, text=True, timeout=900)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")]
        if not line:
            out["cpu_baseline"] = {"error": r.stderr[-400:]}
            return
        d = json.loads(line[-1])
        out["cpu_baseline"] = d["cpu_baseline"]
        chk = d.get("check") or {}
        if "n_hit" in chk:
            k = len(chk["n_hit"])
            p0 = int(cam_off[b])
            g_hit = ctx["vis"].n_hit[p0:p0 + k].cpu().numpy()
            g_keep = ctx["vis"].keep[p0:p0 + k].cpu().numpy().astype(bool)
            fl = ctx["flags_h"][p0:p0 + k]
            ok = (fl & 3) == 0                      # near-threshold / first-ray-collinear pairs are reported apart
            c_hit, c_keep = np.asarray(chk["n_hit"]), np.asarray(chk["keep"], dtype=bool)
            c_ratio = np.asarray(chk["ratio"])
            out["parity"].update({
                "pairs_checked": int(k), "n_hit_mismatch": int((g_hit != c_hit).sum()),
                "keep_mismatch": int((g_keep[ok] != c_keep[ok]).sum()),
                "near_threshold": int((np.abs(c_ratio - 0.45) <= 1e-6).sum()),
                "edge_flips": int(chk.get("edge_sensitive", 0)),
                "checker": "oracle port (cpu_baseline leg) on block %d x its first %d cameras; edge_flips = hits of "
                           "those pairs whose projection lies within 1e-5 (relative to W / H) of an image edge, i.e. "
                           "that a different BLAS rounding could flip; near_threshold = pairs with |ratio - 0.45| <= 1e-6"
                           % (b, k)})
        if "c5_counts" in chk and c5 is not None:
            c5["sample"]["want"] = chk["c5_counts"]
        out["cpu_baseline"]["reference_cuda"] = measure_reference_cuda(ctx, b, cam_idx[cam_off[b]:cam_off[b + 1]])
        out["cpu_baseline"]["reference_code"] = reference_code_timing(args.workload)
    except Exception as e:  # the baseline is informative; never fail the bench on it
        out["cpu_baseline"] = {"error": repr(e)}
    finally:
        if os.path.exists(path):
            os.remove(path)


def reference_code_timing(workload):
    """The REFERENCE's own code timed in the build container (tools/time_reference.py, committed under
    profiles/): the full run at c1, the declared sub-sample extrapolation at c2-c4.  Not measured on
    this box (/root/reference does not travel); quoted beside the port for scale."""
    for name in ("r02_reference_cpu_c1.jsonl", "r02_reference_cpu_c2_c4.jsonl"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                for line in f:
                    d = json.loads(line)
                    if d.get("config") == workload:
                        v = d.get("visibility_tests_per_s") or d.get("visibility_tests_per_s_extrapolated")
                        return {"tests_per_s": v, "host_cores": d.get("host_cores"),
                                "how": "full reference run (A6-A10)" if "visibility_tests_per_s" in d else
                                       "sub-sample (4 leaves x 16 cameras) extrapolated linearly, single thread",
                                "A3_s": d["stage_s"].get("A3_quad_tree_partition"),
                                "A4_s": d["stage_s"].get("A4_expand_partitions"),
                                "where": "build container, profiles/" + name}
        except Exception:
            pass
    return None


def measure_reference_cuda(ctx, b, cams, max_cams=16):
    """The reference's OWN GPU path for A7 (project_points_to_camera(device='cuda'), densify:410,
    :431-475), restated call for call by the oracle: per camera one torch.tensor(points, device)
    upload of the block's down-sampled points, two K = 4 matmuls, the divide / mask, and the
    .tolist() of the in-image points; A8 (the Python Graham scan) is NOT included, so this is an
    upper bound of the reference's GPU-assisted rate.  Same block as the CPU arm, its first cameras."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import math
        import ptgs_oracle as O
        res, recs = ctx["res"], ctx["recs"]
        ds_off = res.ds_offsets.cpu().numpy()
        lo, hi = int(ds_off[b]), int(ds_off[b + 1])
        pts = torch.stack([res.ds_x[lo:hi], res.ds_y[lo:hi], res.ds_z[lo:hi]], 1).cpu().numpy()
        cams = list(cams[:max_cams])

        def one(c):
            r = recs[c]
            W_, H_ = int(r[16]), int(r[17])
            fovx, fovy = 2 * math.atan(W_ / (2 * float(r[12]))), 2 * math.atan(H_ / (2 * float(r[13])))
            return O.project_points_to_camera_ref(pts, np.asarray(r[0:9], np.float64).reshape(3, 3),
                                                  np.asarray(r[9:12], np.float64), fovx, fovy, W_, H_,
                                                  device="cuda").tolist()
        one(cams[0])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for c in cams:
            one(c)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        return {"value": len(pts) * len(cams) / dt, "unit": UNIT, "kind": "port of the reference's device='cuda' "
                "projection (A7 only, no Graham scan)", "seconds_per_camera": dt / len(cams),
                "sample": "block %d (%d voxels) x %d cameras" % (b, len(pts), len(cams))}
    except Exception as e:   # informative
        return {"error": repr(e)}


def _reference_prep(args):
    """CPU-only preparation of the reference arm: a bounded sample of the workload -- one
    leaf-sized block and its cameras.  The exact leaf rectangles need the radius filter and the
    median split over the whole cloud (minutes of CPU at c4), so the sample block is DECLARED as:
    the scene rectangle cut into the expected leaf grid (gx x gy cells, gx * gy = expected leaves),
    the cell nearest the scene centre expanded by 5 % per side (densify:222-232), all cloud points
    inside it, and the cameras whose centre lies strictly inside it (densify:388)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ptgs_oracle as O
    import ptgs_b200  # noqa: F401
    from ptgs_b200 import workload as W
    cache = "/tmp/ptgs_refprep2_%s_%g.npz" % (args.workload, args.scale)
    if os.path.exists(cache):
        z = np.load(cache, allow_pickle=False)
        return z["points"], z["recs"], str(z["desc"])
    w = W.make(args.workload, args.scale)
    pts = w.xyz if w.man_trans is None else O.manhattan_points(w.xyz, w.man_trans)
    Lx, Ly = float(w.meta["Lx"]), float(w.meta["Ly"])
    leaves = max(1, int(2 ** np.ceil(np.log2(max(1.0, 0.88 * len(pts) / (1.2 * w.threshold))))))
    gx = int(2 ** np.ceil(np.log2(leaves) / 2))
    gy = max(1, leaves // gx)
    cw, ch = Lx / gx, Ly / gy
    x0, y0 = -Lx / 2 + (gx // 2) * cw, -Ly / 2 + (gy // 2) * ch
    box = (x0 - 0.05 * cw, y0 - 0.05 * ch, x0 + 1.05 * cw, y0 + 1.05 * ch)
    block = pts[O.expand_mask(pts, box)]
    centers = O.camera_centers(w.R, w.T)
    cams = O.assign_cameras([box], centers)[0]
    recs = np.stack([O.camera_record(w.R[c], w.T[c], w.FoVx[c], w.FoVy[c], int(w.width[c]), int(w.height[c]))
                     for c in cams])
    desc = W.describe(args.workload, args.scale, w)
    np.savez(cache, points=block, recs=recs, desc=desc)
    return block, recs, desc


def run_reference(args):
    """CPU arm: the oracle port of the visibility stage (A6-A9) on the host
    cores, on a bounded sample of the same workload.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_arm
    check = None
    if args.sample_file:
        z = np.load(args.sample_file)
        block, recs, desc = z["points"], z["recs"], "sample file"
        budget = args.cpu_seconds
        r = cpu_arm.run(block, recs, args.steps, args.warmup, budget, want_results=True)
        check = r.pop("results")
        if "c5_points" in z.files:
            import ptgs_oracle as O
            check["c5_counts"] = [int(v) for v in O.coverage_counts(z["c5_points"], z["c5_recs"])]
    else:
        block, recs, desc = _reference_prep(args)
        r = cpu_arm.run(block, recs, args.steps, args.warmup, 90.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "sample": r["sample"]},
        "cpu_baseline": dict({k: r[k] for k in ("value", "unit", "cores", "kind", "sample") if k in r},
                             reference_code=reference_code_timing(args.workload)),
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if check is not None:
        line["check"] = check
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
