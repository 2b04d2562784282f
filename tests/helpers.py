"""Shared test helpers: synthetic scene -> camera arrays in the reference's
field conventions."""
import numpy as np

import ptgs_b200  # noqa: F401
from ptgs_b200 import scene_synth
import ptgs_oracle as O


def scene_cameras(sc) -> O.CameraSet:
    """CameraInfo fields as shen_data_read.readColmapCamerasPartition builds
    them (scene/ptgs/shen_data_read.py:57-85): R = qvec2rotmat(q) (W2C), T,
    FoV = focal2fov(f, pixels)."""
    C = len(sc.qvec)
    R = np.stack([O.qvec2rotmat(q) for q in sc.qvec])
    T = np.asarray(sc.tvec, dtype=np.float64)
    fovx = np.full(C, O.focal2fov(sc.fx, sc.width))
    fovy = np.full(C, O.focal2fov(sc.fy, sc.height))
    return O.CameraSet(R, T, fovx, fovy, np.full(C, sc.width), np.full(C, sc.height))


def small_scene(n=60000, c=24, thr=12000, seed=11, **kw):
    return scene_synth.make_scene(n, c, thr, seed, **kw)
