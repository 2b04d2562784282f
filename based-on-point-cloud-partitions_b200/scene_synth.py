"""Seeded synthetic COLMAP-format scenes.

Workload generator shared by the tests, bench.py and the golden-fixture
script.  It is NOT part of the reference: the reference ships no data.  The
scene is returned as plain NumPy arrays in exactly the precision the COLMAP
files would carry (xyz is float32-valued because shen_data_read.storePly,
scene/ptgs/shen_data_read.py:135-148, writes float32 and fetchPly reads it
back as float32).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

RHO = 48.0  # points per unit^2


@dataclass
class SynthScene:
    xyz: np.ndarray            # (N,3) float32
    rgb: np.ndarray            # (N,3) uint8
    qvec: np.ndarray           # (C,4) float64, COLMAP [w,x,y,z], world->camera
    tvec: np.ndarray           # (C,3) float64
    names: list                # image names "img_%06d.jpg"
    width: int = 1600
    height: int = 1200
    fx: float = 1200.0
    fy: float = 1200.0
    cx: float = 800.0
    cy: float = 600.0
    threshold: int = 500000
    man_trans: np.ndarray | None = None
    meta: dict = field(default_factory=dict)


CONFIGS = {
    # name: (N, C, threshold, seed, oblique_fraction, manhattan_deg_z)
    "c1": (1_000_000, 300, 500_000, 1001, 0.0, None),
    "c2": (10_000_000, 2_000, 500_000, 1002, 0.0, 21.0),
    "c3": (10_000_000, 2_000, 100_000, 1003, 0.5, None),
    "c4": (50_000_000, 8_000, 500_000, 1004, 0.0, None),
    "c5": (100_000_000, 16_000, 500_000, 1005, 0.0, None),
}


def rotmat2qvec(R: np.ndarray) -> np.ndarray:
    """Unit quaternion [w,x,y,z] of a rotation matrix (Shepperd's method)."""
    m = R
    tr = m[0, 0] + m[1, 1] + m[2, 2]
    if tr > 0:
        s = math.sqrt(tr + 1.0) * 2
        q = [0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s]
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        s = math.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
        q = [(m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s]
    elif m[1, 1] > m[2, 2]:
        s = math.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
        q = [(m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s]
    else:
        s = math.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
        q = [(m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s]
    q = np.asarray(q, dtype=np.float64)
    if q[0] < 0:
        q = -q
    return q / np.linalg.norm(q)


def _rot(axis: int, a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    if axis == 0:
        return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)
    if axis == 1:
        return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def manhattan_z(deg: float) -> np.ndarray:
    """utils/manhattan_utils.py:3-29 evaluated for pos "0 0 0", rot "0 0 deg"."""
    rz = _rot(2, np.deg2rad(-deg))
    m = np.zeros((4, 4))
    m[:3, :3] = rz.transpose()
    m[3, 3] = 1.0
    return m


def make_scene(n_points: int, n_cameras: int, threshold: int, seed: int,
               oblique_fraction: float = 0.0, manhattan_deg: float | None = None,
               height_scale: float = 1.0, ties: bool = True) -> SynthScene:
    rng = np.random.Generator(np.random.PCG64(seed))
    N, C = int(n_points), int(n_cameras)
    Lx = math.sqrt(N / RHO * 4.0 / 3.0)
    Ly = 0.75 * Lx
    n_out = int(round(0.005 * N))
    n_str = int(round(0.295 * N))
    n_gnd = N - n_out - n_str

    xyz = np.empty((N, 3), dtype=np.float64)
    g = xyz[:n_gnd]
    g[:, 0] = rng.uniform(-Lx / 2, Lx / 2, n_gnd)
    g[:, 1] = rng.uniform(-Ly / 2, Ly / 2, n_gnd)
    g[:, 2] = rng.uniform(0.0, 4.0, n_gnd)
    K = max(8, N // 250_000)
    centres = np.stack([rng.uniform(-Lx / 2, Lx / 2, K), rng.uniform(-Ly / 2, Ly / 2, K)], axis=1)
    which = rng.integers(0, K, n_str)
    s = xyz[n_gnd:n_gnd + n_str]
    s[:, 0] = np.clip(centres[which, 0] + rng.normal(0.0, 15.0, n_str), -Lx / 2, Lx / 2)
    s[:, 1] = np.clip(centres[which, 1] + rng.normal(0.0, 15.0, n_str), -Ly / 2, Ly / 2)
    s[:, 2] = rng.uniform(0.0, 30.0, n_str)
    o = xyz[n_gnd + n_str:]
    o[:, 0] = rng.uniform(-1.5 * Lx, 1.5 * Lx, n_out)
    o[:, 1] = rng.uniform(-1.5 * Ly, 1.5 * Ly, n_out)
    o[:, 2] = rng.uniform(-50.0, 100.0, n_out)
    perm = rng.permutation(N)
    xyz = xyz[perm]
    if ties:
        snap = rng.random(N) < 0.01
        xyz[snap, 0] = np.round(xyz[snap, 0] * 16.0) / 16.0
        snap = rng.random(N) < 0.01
        xyz[snap, 1] = np.round(xyz[snap, 1] * 16.0) / 16.0
    man = None
    if manhattan_deg is not None:
        # pre-rotate the scene rectangle by -deg about Z so that the Manhattan
        # transform (+deg) re-aligns it with the axes.
        pre = _rot(2, np.deg2rad(-manhattan_deg))
        xyz = xyz @ pre.T
        man = manhattan_z(manhattan_deg)
    rgb = rng.integers(0, 256, (N, 3), dtype=np.uint8)
    xyz32 = xyz.astype(np.float32)

    # cameras: jittered grid over the rectangle
    gx = int(math.ceil(math.sqrt(C * 4.0 / 3.0)))
    gy = int(math.ceil(math.sqrt(C * 3.0 / 4.0)))
    while gx * gy < C:
        gx += 1
    cellx, celly = Lx / gx, Ly / gy
    ii, jj = np.meshgrid(np.arange(gx), np.arange(gy), indexing="ij")
    cx_ = (-Lx / 2 + (ii.ravel()[:C] + 0.5) * cellx) + rng.uniform(-0.3, 0.3, C) * cellx
    cy_ = (-Ly / 2 + (jj.ravel()[:C] + 0.5) * celly) + rng.uniform(-0.3, 0.3, C) * celly
    h0 = 0.225 * math.sqrt(threshold / RHO) * height_scale
    cz_ = h0 * rng.uniform(0.8, 1.25, C)
    yaw = rng.uniform(0.0, 2 * math.pi, C)
    pitch = rng.normal(0.0, np.deg2rad(10.0), C)
    roll = rng.normal(0.0, np.deg2rad(10.0), C)
    obl = rng.random(C) < oblique_fraction
    pitch = np.where(obl, rng.normal(np.deg2rad(45.0), np.deg2rad(10.0), C), pitch)
    nadir = np.array([[1, 0, 0], [0, -1, 0], [0, 0, -1]], dtype=np.float64)  # cam->world
    qvec = np.empty((C, 4))
    tvec = np.empty((C, 3))
    centre = np.stack([cx_, cy_, cz_], axis=1)
    if manhattan_deg is not None:
        centre = centre @ _rot(2, np.deg2rad(-manhattan_deg)).T
    for c in range(C):
        R_cw = _rot(2, yaw[c]) @ _rot(0, pitch[c]) @ _rot(1, roll[c]) @ nadir
        if manhattan_deg is not None:
            R_cw = _rot(2, np.deg2rad(-manhattan_deg)) @ R_cw
        R_wc = R_cw.T
        qvec[c] = rotmat2qvec(R_wc)
        tvec[c] = -R_wc @ centre[c]
    names = ["img_%06d.jpg" % i for i in range(C)]
    return SynthScene(xyz=xyz32, rgb=rgb, qvec=qvec, tvec=tvec, names=names,
                      threshold=int(threshold), man_trans=man,
                      meta=dict(n_points=N, n_cameras=C, seed=seed, Lx=Lx, Ly=Ly,
                                cam_height=h0, oblique_fraction=oblique_fraction,
                                manhattan_deg=manhattan_deg))


# Camera height multiplier applied to SURVEY.md 8(d)'s h = 0.225*sqrt(thr/rho).
# With the formula's own height (x1) 98 % of the assigned cameras pass the 0.45
# hull-area test at c2; SURVEY asks for a kept fraction of 20-80 % and allows
# only `h` to change: x3 gives about 60 % (BASELINE.md records the measured
# fraction per configuration).
HEIGHT_SCALE = 3.0


def make_config(name: str, scale: float = 1.0, **kw) -> SynthScene:
    """One of the BASELINE.json configurations (c1..c5); `scale` < 1 shrinks
    both the point and the camera count (same density, same camera height)."""
    N, C, thr, seed, obl, man = CONFIGS[name]
    N = max(1000, int(N * scale))
    C = max(4, int(C * scale))
    kw.setdefault("height_scale", HEIGHT_SCALE)
    sc = make_scene(N, C, thr if scale == 1.0 else kw.pop("threshold", thr), seed,
                    oblique_fraction=obl, manhattan_deg=man, **kw)
    sc.meta["config"] = name
    sc.meta["scale"] = scale
    return sc
