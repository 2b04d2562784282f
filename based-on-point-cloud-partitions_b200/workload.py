"""Workload preparation shared by bench.py, __graft_entry__.smoke() and the
tests: a synthetic scene (scene_synth) turned into the host arrays the device
pipeline takes, with the camera conventions of the reference's reader
(scene/ptgs/shen_data_read.py:33-99: R = qvec2rotmat(q) is world->camera,
FoV = focal2fov(f, pixels))."""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

from . import scene_synth


def qvec2rotmat(q):
    """scene/colmap_loader.py:43-53."""
    return np.array([
        [1 - 2 * q[2] ** 2 - 2 * q[3] ** 2, 2 * q[1] * q[2] - 2 * q[0] * q[3], 2 * q[3] * q[1] + 2 * q[0] * q[2]],
        [2 * q[1] * q[2] + 2 * q[0] * q[3], 1 - 2 * q[1] ** 2 - 2 * q[3] ** 2, 2 * q[2] * q[3] - 2 * q[0] * q[1]],
        [2 * q[3] * q[1] - 2 * q[0] * q[2], 2 * q[2] * q[3] + 2 * q[0] * q[1], 1 - 2 * q[1] ** 2 - 2 * q[2] ** 2]])


def focal2fov(focal, pixels):
    """utils/graphics_utils.py:76."""
    return 2 * math.atan(pixels / (2 * focal))


def manhattan_cameras(R, T, man_trans):
    """scene/dataset_readers.py:96-104: W2C @ inv(man_trans) per camera
    (the 'consistent Manhattan' semantic of SURVEY.md row A13)."""
    inv = np.linalg.inv(man_trans)
    Ro = np.empty_like(R)
    To = np.empty_like(T)
    for i in range(len(R)):
        W2C = np.zeros((4, 4))
        W2C[:3, :3] = R[i]
        W2C[:3, -1] = T[i]
        W2C[3, 3] = 1.0
        W2nC = W2C @ inv
        Ro[i] = W2nC[:3, :3]
        To[i] = W2nC[:3, -1]
    return Ro, To


@dataclass
class HostWorkload:
    name: str
    xyz: np.ndarray            # (N,3) float32 scene cloud (pre-Manhattan)
    man_trans: np.ndarray      # 4x4 or None
    R: np.ndarray              # (C,3,3) float64 world->camera (in the Manhattan frame when set)
    T: np.ndarray              # (C,3)
    FoVx: np.ndarray
    FoVy: np.ndarray
    width: np.ndarray
    height: np.ndarray
    threshold: int
    meta: dict

    @property
    def n_points(self):
        return len(self.xyz)

    @property
    def n_cameras(self):
        return len(self.R)


def from_scene(sc: scene_synth.SynthScene, name="custom") -> HostWorkload:
    C = len(sc.qvec)
    R = np.stack([qvec2rotmat(q) for q in sc.qvec]) if C else np.zeros((0, 3, 3))
    T = np.asarray(sc.tvec, dtype=np.float64).reshape(C, 3)
    if sc.man_trans is not None:
        R, T = manhattan_cameras(R, T, sc.man_trans)
    return HostWorkload(
        name=name, xyz=sc.xyz, man_trans=sc.man_trans, R=R, T=T,
        FoVx=np.full(C, focal2fov(sc.fx, sc.width)), FoVy=np.full(C, focal2fov(sc.fy, sc.height)),
        width=np.full(C, sc.width), height=np.full(C, sc.height), threshold=sc.threshold,
        meta=dict(sc.meta))


def make(name: str, scale: float = 1.0, **kw) -> HostWorkload:
    """One of BASELINE.json's configurations (c1..c5), optionally scaled."""
    return from_scene(scene_synth.make_config(name, scale, **kw), name)


def describe(name: str, scale: float, w: HostWorkload, n_points: int | None = None) -> str:
    N, C, thr, seed, obl, man = scene_synth.CONFIGS[name]
    s = "%s: synthetic COLMAP scene %d points / %d cameras, threshold_value=%d, seed %d" % (
        name, w.n_points if n_points is None else n_points, w.n_cameras, w.threshold, seed)
    if man is not None:
        s += ", Manhattan rotation %g deg about Z (consistent: points and cameras)" % man
    if obl:
        s += ", %d%% oblique cameras" % int(obl * 100)
    if scale != 1.0:
        s += ", scale %g" % scale
    return s
