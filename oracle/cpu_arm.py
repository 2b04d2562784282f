"""TEST / BASELINE INFRASTRUCTURE ONLY -- the CPU arm timed by bench.py
(`cpu_baseline` leg and `--impl reference`): the oracle port of the
reference's visibility stage (A6 voxel down-sampling, A7 projection, A8 Graham
scan, A9 decision; scene/ptgs/shen_partation_densify.py:911-954) run on the
host cores.  Never imported by the product path.

The reference's own code cannot travel to the GPU box (/root/reference does not
exist there), so the timed CPU implementation is the NumPy restatement in
ptgs_oracle.py (`kind: "port"`).  It is considerably FASTER than the
reference's literal code (vectorised A6 instead of a per-point Python dict,
no per-camera torch.tensor()/.tolist() round trip), so speed-ups quoted against
it are conservative.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import time

import numpy as np

import ptgs_oracle as O

_G = {}


def _one_camera(c):
    return O.camera_visibility(_G["ds"], _G["recs"][c])


def _edge_sensitive(c):
    """Hits of camera c whose projection lies within 1e-5 (relative to the image size) of an image
    edge: the hits a different SGEMM rounding could flip (north star: reported separately)."""
    rec = _G["recs"][c]
    hit, u, v = O.project_port(_G["ds"], rec)
    W, H = float(rec[16]), float(rec[17])
    u, v = u[hit].astype(np.float64), v[hit].astype(np.float64)
    e = (np.minimum(u, W - u) <= 1e-5 * W) | (np.minimum(v, H - v) <= 1e-5 * H)
    return int(e.sum())


def visibility_step(block_points, recs, workers, want_results=False):
    """One block: A6 once, then every camera of `recs` (A7-A9) over a fork
    pool of `workers` processes (the reference uses a 48-thread pool under the
    GIL, densify:932).  Returns (tests, n_kept, ds_count[, per-camera results])."""
    ds = O.voxel_downsample(block_points)
    _G["ds"], _G["recs"] = ds, recs
    n = len(recs)
    if workers > 1 and n > 1:
        with mp.get_context("fork").Pool(min(workers, n)) as pool:
            out = pool.map(_one_camera, range(n), chunksize=max(1, n // (4 * workers)))
    else:
        out = [_one_camera(c) for c in range(n)]
    if want_results:
        return len(ds) * n, sum(1 for r in out if r[2]), len(ds), out
    return len(ds) * n, sum(1 for r in out if r[2]), len(ds)


def run(block_points, recs, steps, warmup, budget_s, workers=None, want_results=False):
    """Times `steps` visibility steps on a bounded sample: the given block and
    as many of its cameras as fit `budget_s` seconds in total (calibrated on
    the first cameras).  Returns a dict for the JSON line."""
    workers = workers or os.cpu_count() or 1
    recs = np.asarray(recs)
    # calibration: A6 + `workers` cameras
    t0 = time.perf_counter()
    ds = O.voxel_downsample(block_points)
    t_a6 = time.perf_counter() - t0
    _G["ds"], _G["recs"] = ds, recs
    ncal = min(len(recs), 2)
    t0 = time.perf_counter()
    for c in range(ncal):
        _one_camera(c)
    t_cam = (time.perf_counter() - t0) / max(ncal, 1)
    per_step = budget_s / max(steps + warmup, 1)
    n_cam = int(max(1, min(len(recs), (per_step - t_a6) / max(t_cam, 1e-6) * min(workers, 8))))
    use = recs[:n_cam]
    for _ in range(warmup):
        visibility_step(block_points, use, workers)
    times, tests, results = [], 0, None
    for _ in range(steps):
        t0 = time.perf_counter()
        tests, kept, m, results = visibility_step(block_points, use, workers, want_results=True)
        times.append(time.perf_counter() - t0)
    t = float(np.mean(times)) if times else float("nan")
    out = {
        "value": tests / t if times else None, "unit": "tests/s", "cores": int(min(workers, max(n_cam, 1))),
        "kind": "port", "seconds_per_step": t, "tests_per_step": int(tests),
        "sample": "one block (%d expanded points -> %d voxels) x its first %d cameras, A6-A9 per step" % (
            len(block_points), len(ds), n_cam),
    }
    if want_results and results is not None:
        _G["ds"], _G["recs"] = ds, use
        out["results"] = {"n_hit": [int(r[0]) for r in results], "ratio": [float(r[1]) for r in results],
                          "keep": [bool(r[2]) for r in results],
                          "edge_sensitive": int(sum(_edge_sensitive(c) for c in range(len(use))))}
    return out
