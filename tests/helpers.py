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


class VastCamera:
    """The attributes scene/ptgs/data_partition.py reads from a camera (scene/cameras.py:18-71)."""

    def __init__(self, R, T, FoVx, FoVy, width, height, name, center):
        import torch
        self.R, self.T, self.FoVx, self.FoVy = R, T, FoVx, FoVy
        self.image_width, self.image_height, self.image_name = width, height, name
        self.camera_center = torch.tensor(center, dtype=torch.float32)


def vast_inputs(sc):
    """The synthetic scene in the Vast path's conventions: ground plane XZ (y and z swapped),
    camera.R = transposed W2C rotation (scene/dataset_readers.py:90), camera_center = the
    true centre as float32.  Returns (points (N,3) f32, colors (N,3) f64, VastCameras)."""
    import ptgs_oracle_vast as OV
    cams = scene_cameras(sc)
    P = np.array([[1.0, 0, 0], [0, 0, 1.0], [0, 1.0, 0]])
    pts = np.ascontiguousarray(sc.xyz[:, [0, 2, 1]])
    C = len(cams.R)
    R_store = np.empty((C, 3, 3))
    center = np.empty((C, 3), np.float32)
    for c in range(C):
        R_wc = cams.R[c] @ P                      # W2C of the swapped world
        R_store[c] = R_wc.T
        center[c] = (P @ (-cams.R[c].T @ cams.T[c])).astype(np.float32)
    names = ["img_%06d" % i for i in range(C)]
    vc = OV.VastCameras(center, R_store, cams.T.copy(), cams.FoVx.copy(), cams.FoVy.copy(),
                        np.asarray(cams.width), np.asarray(cams.height), names)
    return pts, sc.rgb.astype(np.float64) / 255.0, vc
