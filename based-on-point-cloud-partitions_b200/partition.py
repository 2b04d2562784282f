"""Host-side mirror of the reference's partition interface -- the drop-in layer.

Same function names, argument meaning, return types and error behaviour as
module `shen_partation_densify` (scene/ptgs/shen_partation_densify.py) and class
`ProgressiveDataPartitioning` (scene/ptgs/shen_partition_utils.py:13-101) of
92401/Based-on-point-cloud-partitions; every function uploads its host inputs,
runs the libptgs CUDA kernels through the C ABI (ptgs_b200.device) and returns
host objects of the reference's NamedTuple types.  No CPU fallback: without a
CUDA device / libptgs.so these functions raise.

    quad_tree_partition               densify:37-49  -> density_partition.py:148-166
    expand_partitions                 densify:182-252
    assign_cameras_to_partitions      densify:362-408
    downsample_point_cloud            densify:541-614
    visibility_based_camera_selection densify:911-954 (+ save_partition_as_pkl :617-639)
    ProgressiveDataPartitioning       shen_partition_utils.py:13-101
    main                              shen_partition_utils.py:102-126 (the script body)

`compat/shen_partation_densify.py` and `compat/shen_partition_utils.py` re-export
this module under the reference's module names.
"""
from __future__ import annotations

import os
import pickle
import shutil
import sys
from typing import List, NamedTuple, Optional

import numpy as np
import torch

from . import colmap_io, device as D

try:  # real shapely rectangles when the target machine has shapely
    from shapely.geometry import Polygon as _ShapelyPolygon, Point as _ShapelyPoint, box as _shapely_box
    HAVE_SHAPELY = True
except Exception:  # pragma: no cover - shapely is absent in the build image
    HAVE_SHAPELY = False


class Point:
    """shapely.geometry.Point stand-in (x, y only)."""
    def __init__(self, *a):
        if len(a) == 1:
            a = tuple(a[0])
        self.x, self.y = float(a[0]), float(a[1])


class Rect:
    """Axis-aligned rectangle with the part of shapely.geometry.Polygon's
    interface the reference and its downstream tools use: `.bounds` ->
    (minx, miny, maxx, maxy) Python floats, `.contains(Point)` = strict
    interior (densify:388), `.exterior.xy`, `.area`."""
    def __init__(self, minx, miny, maxx, maxy):
        self._b = (float(minx), float(miny), float(maxx), float(maxy))

    @property
    def bounds(self):
        return self._b

    def contains(self, pt):
        minx, miny, maxx, maxy = self._b
        return minx < pt.x < maxx and miny < pt.y < maxy

    @property
    def area(self):
        minx, miny, maxx, maxy = self._b
        return (maxx - minx) * (maxy - miny)

    @property
    def exterior(self):
        minx, miny, maxx, maxy = self._b

        class _E:
            xy = ([minx, maxx, maxx, minx, minx], [miny, miny, maxy, maxy, miny])
            coords = list(zip(*xy))
        return _E

    def __repr__(self):
        return "Rect(%r, %r, %r, %r)" % self._b

    def __eq__(self, o):
        return isinstance(o, Rect) and o._b == self._b

    @property
    def wkb(self) -> bytes:
        """Little-endian WKB of the rectangle as shapely.geometry.box builds it (densify:234-235):
        one ring of 5 vertices (maxx,miny) (maxx,maxy) (minx,maxy) (minx,miny) (maxx,miny)."""
        import struct
        minx, miny, maxx, maxy = self._b
        ring = [(maxx, miny), (maxx, maxy), (minx, maxy), (minx, miny), (maxx, miny)]
        return struct.pack("<BIII", 1, 3, 1, 5) + b"".join(struct.pack("<dd", x, y) for x, y in ring)

    def __reduce__(self):
        # shapely 2 pickles a geometry as shapely.from_wkb(<bytes>); dump_compat() rewrites the
        # callable's path to that, so the file loads as a REAL shapely Polygon where shapely exists
        # (consumers read `.origin_box.bounds` / shapely.vectorized.contains,
        # scene/ptgs/cut_partition.py:59-67, merge_partitions.py:10-57)
        return (_shapely_from_wkb, (self.wkb,))


def _shapely_from_wkb(data: bytes):
    """Rebuilds a pickled rectangle: a shapely Polygon when shapely is importable, else Rect."""
    try:
        import shapely
        return shapely.from_wkb(data)
    except Exception:
        return _rect_from_wkb(data)


def _rect_from_wkb(data: bytes):
    import struct
    order = "<" if data[0] == 1 else ">"
    gtype, nrings, npts = struct.unpack_from(order + "III", data, 1)
    if gtype != 3 or nrings != 1:
        raise ValueError("not a single-ring polygon WKB")
    xy = np.frombuffer(data, dtype=order + "f8", count=2 * npts, offset=13).reshape(-1, 2)
    return Rect(xy[:, 0].min(), xy[:, 1].min(), xy[:, 0].max(), xy[:, 1].max())


# Class paths the reference's pickles carry: the NamedTuples live in module
# `shen_partation_densify` (imported bare, utils-shen:7), cameras are scene.cameras.SimpleCamera,
# rectangles reduce to shapely.from_wkb.  dump_compat() writes protocol-3 pickles (text GLOBAL
# opcodes, no framing) and rewrites this package's paths to those.
_PICKLE_RENAMES = [
    (b"cptgs_b200.partition\n_shapely_from_wkb\n", b"cshapely\nfrom_wkb\n"),
    (b"cptgs_b200.partition\nSimpleCamera\n", b"cscene.cameras\nSimpleCamera\n"),
    (b"cptgs_b200.partition\nPartition\n", b"cshen_partation_densify\nPartition\n"),
    (b"cptgs_b200.partition\nBasicPointCloud\n", b"cshen_partation_densify\nBasicPointCloud\n"),
    (b"cptgs_b200.partition\nCameraPose\n", b"cshen_partation_densify\nCameraPose\n"),
]


def dumps_compat(obj) -> bytes:
    """pickle.dumps with the reference's class paths (see _PICKLE_RENAMES)."""
    data = pickle.dumps(obj, protocol=3)
    for a, b in _PICKLE_RENAMES:
        data = data.replace(a, b)
    return data


def dump_compat(obj, f) -> None:
    f.write(dumps_compat(obj))


class _CompatUnpickler(pickle.Unpickler):
    """Loads pickles written by dump_compat() or by the reference: the reference's own modules are
    used when importable, this package's equivalents otherwise."""
    _LOCAL = {"Partition", "BasicPointCloud", "CameraPose", "SimpleCamera"}

    def find_class(self, module, name):
        try:
            return super().find_class(module, name)
        except (ImportError, AttributeError):
            if module == "shapely" and name == "from_wkb":
                return _rect_from_wkb
            if name in self._LOCAL and module in ("shen_partation_densify", "scene.ptgs.shen_partation_densify",
                                                  "scene.cameras", "ptgs_b200.partition"):
                return globals()[name]
            raise


def load_compat(f):
    return _CompatUnpickler(f).load()


def box(minx, miny, maxx, maxy):
    """shapely.geometry.box (densify:234-235): a real shapely Polygon when
    shapely is importable, else Rect."""
    if HAVE_SHAPELY:
        return _shapely_box(minx, miny, maxx, maxy)
    return Rect(minx, miny, maxx, maxy)


# ---- the reference's record types (densify:21-35), field for field ------------------
class BasicPointCloud(NamedTuple):
    points: np.ndarray
    colors: np.ndarray
    normals: np.ndarray


class Partition(NamedTuple):
    partition_id: str
    camera: list
    point_cloud: Optional[BasicPointCloud]
    origin_box: object
    point_num: int
    extend_rate: float = 0
    extend_box: Optional[object] = None


class CameraPose(NamedTuple):
    camera: object
    pose: np.ndarray


def _world2view2(R, t, translate=np.array([.0, .0, .0]), scale=1.0):
    """utils/graphics_utils.py:38-50 (getWorld2View2)."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = np.asarray(R).transpose()
    Rt[:3, 3] = np.asarray(t).flatten()
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    cam_center = C2W[:3, 3]
    cam_center = (cam_center + translate) * scale
    C2W[:3, 3] = cam_center
    Rt = np.linalg.inv(C2W)
    return np.float32(Rt)


def _projection_matrix(znear, zfar, fovX, fovY):
    """utils/graphics_utils.py:52-74 (getProjectionMatrix)."""
    import math
    tanHalfFovY = math.tan((fovY / 2))
    tanHalfFovX = math.tan((fovX / 2))
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    z_sign = 1.0
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = z_sign
    P[2, 2] = z_sign * zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


class SimpleCamera(torch.nn.Module):
    """scene/cameras.py:18-71, attribute for attribute (an nn.Module like the reference's, so a
    pickled instance restores into scene.cameras.SimpleCamera with a complete state).
    camera_center is -R^T T evaluated by torch in float32 (:56-58).  data_device defaults to
    "cpu" (the partition script passes args.data_device = "cpu", utils-shen:110-114)."""
    def __init__(self, colmap_id, R, T, FoVx, FoVy, image_name, uid, width, height,
                 trans=np.array([0.0, 0.0, 0.0]), scale=1.0, data_device="cpu", points3D_ids=None,
                 point3D_ids=None):
        super().__init__()
        self.uid = uid
        self.colmap_id = colmap_id
        self.R = np.asarray(R, dtype=np.float64)
        self.T = np.asarray(T, dtype=np.float64)
        self.FoVx = FoVx
        self.FoVy = FoVy
        self.image_name = image_name
        self.image_width = width
        self.image_height = height
        self.trans = trans
        self.scale = scale
        if points3D_ids is None:
            points3D_ids = point3D_ids
        self.points3D_ids = points3D_ids if points3D_ids is not None else np.array([], dtype=np.int64)
        try:
            self.data_device = torch.device(data_device)
        except Exception as e:
            print(e)
            print(f"[Warning] Custom device {data_device} failed, fallback to default cuda device")
            self.data_device = torch.device("cuda")
        self.zfar = 100.0
        self.znear = 0.01
        W2C = _world2view2(self.R, self.T, trans, scale)
        self.world_view_transform = torch.tensor(W2C, dtype=torch.float32).to(self.data_device)
        Rt = torch.tensor(self.R, dtype=torch.float32).to(self.data_device)
        Tt = torch.tensor(self.T, dtype=torch.float32).to(self.data_device)
        self.camera_center = -torch.matmul(Rt.transpose(0, 1), Tt)
        self.projection_matrix = _projection_matrix(znear=self.znear, zfar=self.zfar, fovX=self.FoVx,
                                                    fovY=self.FoVy).transpose(0, 1).to(self.data_device)
        self.full_proj_transform = torch.matmul(self.world_view_transform, self.projection_matrix)

    @property
    def point3D_ids(self):          # round-1 spelling, kept as an alias
        return self.points3D_ids

    def get_camera_center(self):
        return self.camera_center


def _dev():
    if not torch.cuda.is_available():
        raise RuntimeError("ptgs_b200.partition: no CUDA device (libptgs has no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _bounds_of(poly):
    return tuple(float(v) for v in poly.bounds)


def _soa(points, dev):
    pts = np.ascontiguousarray(points)
    if pts.dtype not in (np.float32, np.float64):
        pts = pts.astype(np.float64)
    return D.aos_to_soa(torch.from_numpy(pts).to(dev))


def _center_of(cam):
    c = cam.camera_center
    return c.detach().cpu().numpy() if isinstance(c, torch.Tensor) else np.asarray(c)


def _camera_arrays(cameras):
    R = np.stack([np.asarray(c.R, dtype=np.float64) for c in cameras])
    T = np.stack([np.asarray(c.T, dtype=np.float64).reshape(3) for c in cameras])
    fx = np.array([c.FoVx for c in cameras], dtype=np.float64)
    fy = np.array([c.FoVy for c in cameras], dtype=np.float64)
    w = np.array([c.image_width for c in cameras])
    h = np.array([c.image_height for c in cameras])
    return R, T, fx, fy, w, h


# ---- A3 -------------------------------------------------------------------------------
def quad_tree_partition(points, bounds, threshold, depth=0) -> List[Partition]:
    """densify:37-49 -> density_partition.density_partition (dp:148-166):
    balanced binary (exact median) split of `points` ((N,3), only x,y are read)
    inside the rectangle `bounds`; leaves in DFS order with ids per dp:28/76/86,
    `origin_box` the closed leaf rectangle, `point_num` the closed-rectangle
    count, `camera=[]`, `point_cloud=None`.  `depth` is ignored like in the
    reference (always restarted at 0, densify:48)."""
    dev = _dev()
    pts = np.asarray(points)
    if len(pts) == 0:
        return []
    x, y, _ = _soa(pts, dev)
    minx, miny, maxx, maxy = _bounds_of(bounds)
    keep = ((x >= minx) & (x <= maxx) & (y >= miny) & (y <= maxy)).to(torch.uint8)   # dp:22-24
    fx, fy, nf = D.compact_xy(x, y, keep)
    root = torch.tensor([minx, miny, maxx, maxy], dtype=torch.float64, device=dev)
    lb, lp, ld, lc = D.balanced_split(fx, fy, nf, root, threshold)
    return [Partition(partition_id=D.leaf_id(lp[i], ld[i]), camera=[], point_cloud=None,
                      origin_box=box(*[float(v) for v in lb[i]]), point_num=int(lc[i]))
            for i in range(len(lb))]


# ---- A4 -------------------------------------------------------------------------------
def expand_partitions(filtered_partitions: List[Partition], point3D, expansion_distance: float = 80) \
        -> List[Partition]:
    """densify:182-252: every leaf grows by 5 % of its own width / height per
    side; its point cloud is the index-ordered subset of `point3D` inside the
    closed grown rectangle (bounds rounded to the cloud's dtype, :201-202)."""
    if not filtered_partitions:
        return []
    dev = _dev()
    lb = np.array([_bounds_of(p.origin_box) for p in filtered_partitions], dtype=np.float64).reshape(-1, 4)
    boxes = D.expanded_boxes(lb)
    pts = np.asarray(point3D.points)
    x, y, _ = _soa(pts, dev)
    off, idx = D.expand_select(x, y, boxes)
    off = off.cpu().numpy()
    idx = idx.cpu().numpy().astype(np.int64)
    out = []
    for b, p in enumerate(filtered_partitions):
        sel = idx[off[b]:off[b + 1]]
        pc = BasicPointCloud(points=pts[sel], colors=np.asarray(point3D.colors)[sel],
                             normals=np.asarray(point3D.normals)[sel])
        out.append(Partition(partition_id=p.partition_id, origin_box=p.origin_box, camera=[],
                             extend_rate=expansion_distance, extend_box=box(*[float(v) for v in boxes[b]]),
                             point_num=len(sel), point_cloud=pc))
    return out


# ---- A5 -------------------------------------------------------------------------------
def assign_cameras_to_partitions(partitions: List[Partition], cameras) -> List[Partition]:
    """densify:362-408: a camera joins every leaf whose `extend_box` strictly
    contains its centre in XY; `camera` becomes a list of CameraPose in camera
    order, everything else is carried over."""
    if not partitions:
        return []
    dev = _dev()
    boxes = np.array([_bounds_of(p.extend_box) for p in partitions], dtype=np.float64).reshape(-1, 4)
    centers = np.stack([_center_of(c) for c in cameras]) if len(cameras) else np.zeros((0, 3), np.float32)
    off, idx = D.assign_cameras(centers, boxes, dev)
    off = off.cpu().numpy()
    idx = idx.cpu().numpy()
    out = []
    for b, p in enumerate(partitions):
        poses = [CameraPose(camera=cameras[c], pose=centers[c]) for c in idx[off[b]:off[b + 1]]]
        out.append(p._replace(camera=poses))
    return out


# ---- A6 -------------------------------------------------------------------------------
def _attribute_means(attr, pts, voxel_size, first_h, normalise):
    """Per-voxel np.mean of a per-point attribute (densify:593-603), voxels in the order of
    `first_h` (ascending first-member index).  Singletons -- nearly all voxels at voxel 0.1 -- are
    the member's own value; the few multi-member voxels are averaged with the reference's own
    expression.  Host side: nothing downstream of the partitioner reads these (densify:928 uses
    `.points` only), they exist for API fidelity."""
    attr = np.asarray(attr)
    out = attr[first_h].astype(np.float64)
    vid = np.floor((pts - np.min(pts, axis=0)) / voxel_size).astype(np.int32).astype(np.int64)   # densify:572
    span = vid.max(axis=0) + 1
    key = (vid[:, 0] * span[1] + vid[:, 1]) * span[2] + vid[:, 2]
    order = np.argsort(key, kind="stable")                   # members of a voxel in index order
    ks = key[order]
    starts = np.flatnonzero(np.concatenate([[True], ks[1:] != ks[:-1]]))
    counts = np.diff(np.concatenate([starts, [len(ks)]]))
    for s0, c in zip(starts[counts > 1], counts[counts > 1]):
        idx_list = order[s0:s0 + c]
        row = int(np.searchsorted(first_h, idx_list[0]))
        out[row] = np.mean(attr[idx_list], axis=0)
    if normalise:                                            # densify:599-602
        # the reference normalises EVERY voxel's mean normal, single-member ones included
        ln = np.linalg.norm(out, axis=1)
        ok = ln > 1e-12
        out[ok] = out[ok] / ln[ok, None]
    return out.astype(np.float32)


def downsample_point_cloud(point_cloud: BasicPointCloud, voxel_size: float) -> BasicPointCloud:
    """densify:541-614: one point per occupied voxel (mean of its members,
    sequential sum in index order), voxels in first-occurrence order, float32; colours and
    normals are the per-voxel means too (normals re-normalised, :593-603)."""
    if voxel_size <= 0:
        raise ValueError("voxel_size must be positive.")
    dev = _dev()
    pts = np.asarray(point_cloud.points)
    n = len(pts)
    if n == 0:
        return point_cloud                                   # densify:560-561
    x, y, z = _soa(pts, dev)
    off = torch.tensor([0, n], dtype=torch.int64, device=dev)
    ds_off, dx, dy, dz, first = D.voxel_downsample(x, y, z, off, None, voxel_size, n_entries=n)
    order = torch.argsort(first.to(torch.int64), stable=True)
    out = torch.stack([dx[order], dy[order], dz[order]], dim=1).cpu().numpy()
    first_h = first[order].cpu().numpy().astype(np.int64)
    cols, nrm = point_cloud.colors, point_cloud.normals
    new_c = _attribute_means(cols, pts, voxel_size, first_h, False) if cols is not None and len(cols) == n else None
    new_n = _attribute_means(nrm, pts, voxel_size, first_h, True) if nrm is not None and len(nrm) == n else None
    return BasicPointCloud(points=out, colors=new_c, normals=new_n)


def plot_quad_tree(*args, **kwargs) -> None:
    """densify:52-110 and the other matplotlib figure writers of the module are not part of the
    result (out of scope); no-ops so that run_DataPartition's call sequence
    (utils-shen:82-92) runs unchanged against this module."""


plot_quad_tree_extend = plot_subregions = plot_partitions = plot_quad_tree


# ---- A10: per-block output (save_partition.py:19-158, densify:617-639) --------------------
def rotmat_to_colmap_qvec(R) -> np.ndarray:
    """save_partition.py:113-114: scipy Rotation.from_matrix(R).as_quat()
    reordered to COLMAP's [w, x, y, z]."""
    from scipy.spatial.transform import Rotation
    q = Rotation.from_matrix(np.asarray(R, dtype=np.float64)).as_quat()
    return np.array([q[3], q[0], q[1], q[2]], dtype=np.float64)


def save_partition_data(partition: Partition, base_dir: str, images_source_path: str) -> None:
    """save_partition.py:19-48 with vectorised writers (byte-identical files):
    <base_dir>/partition_<id>/{images/, sparse/0/images.bin, sparse/0/points3D.bin,
    sparse/0/cameras.bin}."""
    pdir = os.path.join(base_dir, "partition_%s" % partition.partition_id)
    images_dir = os.path.join(pdir, "images")
    sparse_dir = os.path.join(pdir, "sparse", "0")
    os.makedirs(images_dir, exist_ok=True)
    os.makedirs(sparse_dir, exist_ok=True)
    cams = [cp.camera for cp in partition.camera]
    # images dict is keyed by uid (save_partition.py:130): later duplicates win, order of first insertion
    by_uid = {}
    for c in cams:
        by_uid[c.uid] = c
    cams_u = list(by_uid.values())
    colmap_io.write_images_binary(
        [c.uid for c in cams_u],
        np.array([rotmat_to_colmap_qvec(c.R) for c in cams_u], dtype=np.float64).reshape(-1, 4),
        np.array([np.asarray(c.T, dtype=np.float64).reshape(3) for c in cams_u]).reshape(-1, 3),
        [c.colmap_id for c in cams_u], [c.image_name + ".jpg" for c in cams_u],
        os.path.join(sparse_dir, "images.bin"))
    pc = partition.point_cloud
    colmap_io.write_points3D_binary(np.asarray(pc.points, dtype=np.float64), colmap_io.colors_to_u8(pc.colors),
                                    os.path.join(sparse_dir, "points3D.bin"))
    for c in cams:                                                  # copy_images (:73-98)
        src = os.path.join(images_source_path, c.image_name + ".jpg")
        if os.path.exists(src):
            shutil.copy(src, os.path.join(images_dir, c.image_name + ".jpg"))
    src_cams = os.path.join(os.path.dirname(images_source_path), "sparse", "0", "cameras.bin")
    if os.path.exists(src_cams):                                    # copy_cameras_to_partitions (:50-71)
        shutil.copy(src_cams, os.path.join(sparse_dir, "cameras.bin"))


def save_partition_as_pkl(partition: Partition, base_directory: str) -> None:
    """densify:617-639: <base>/<id>/partition_<id>.pkl + save_partition_data."""
    pdir = os.path.join(base_directory, str(partition.partition_id))
    os.makedirs(pdir, exist_ok=True)
    model_path = os.path.dirname(os.path.dirname(base_directory))
    images_source_path = os.path.join(os.path.dirname(model_path), "images")
    save_partition_data(partition, pdir, images_source_path)
    try:
        with open(os.path.join(pdir, "partition_%s.pkl" % partition.partition_id), "wb") as f:
            dump_compat(partition, f)
    except Exception as e:  # the reference prints and continues (densify:638-639)
        print("saving partition %s failed: %s" % (partition.partition_id, e))


# ---- A7-A9 ----------------------------------------------------------------------------------
def visibility_based_camera_selection(partitions: List[Partition], plot_path: str, pcd=None,
                                      max_workers: int = 48, write_outputs: bool = True,
                                      visible_rate_threshold: float = D.VISIBLE_RATE,
                                      voxel_size: float = D.VOXEL_SIZE) -> List[Partition]:
    """densify:911-954.  Per leaf: voxel down-sample its cloud (A6), test every
    assigned camera (A7: projection, > 3 in-image points; A8: hull area /
    (W*H) >= 0.45), replace `camera` by the kept CameraPoses (ascending camera
    order -- the reference's order is the completion order of a thread pool),
    keep the point cloud, write the A10 outputs under `plot_path`.
    `max_workers` is accepted for signature compatibility (no thread pool)."""
    if not partitions:
        return []
    dev = _dev()
    cams, cam_slot = [], {}
    for p in partitions:
        for cp in p.camera:
            if id(cp.camera) not in cam_slot:
                cam_slot[id(cp.camera)] = len(cams)
                cams.append(cp.camera)
    if cams:
        recs = D.camera_records(*_camera_arrays(cams))
    else:
        recs = np.zeros((0, D.CAMERA_FLOATS), np.float32)
    sizes = [len(p.point_cloud.points) for p in partitions]
    blk_off = np.zeros(len(partitions) + 1, np.int64)
    blk_off[1:] = np.cumsum(sizes)
    dt = np.result_type(*[np.asarray(p.point_cloud.points).dtype for p in partitions])
    dt = np.float64 if dt == np.float64 else np.float32
    pts = np.concatenate([np.asarray(p.point_cloud.points, dtype=dt).reshape(-1, 3) for p in partitions]) \
        if blk_off[-1] else np.zeros((0, 3), dt)
    cam_off = np.zeros(len(partitions) + 1, np.int64)
    cam_off[1:] = np.cumsum([len(p.camera) for p in partitions])
    cam_idx = np.array([cam_slot[id(cp.camera)] for p in partitions for cp in p.camera], dtype=np.int32)
    out, _ = D.visibility_select_host(torch.from_numpy(np.ascontiguousarray(pts)), blk_off,
                                      torch.from_numpy(recs), cam_off, cam_idx, dev, voxel=voxel_size,
                                      threshold=visible_rate_threshold)
    keep = out[4].astype(bool)
    result = []
    for b, p in enumerate(partitions):
        kept = [cp for cp, k in zip(p.camera, keep[cam_off[b]:cam_off[b + 1]]) if k]
        q = p._replace(camera=kept)
        if write_outputs:
            save_partition_as_pkl(q, plot_path)
        result.append(q)
    return result


# ---- A12: COLMAP sparse model -> SceneInfo (shen_data_read.py:150-206) ------------------------
class CameraInfo(NamedTuple):           # shen_data_read.py:13-23
    uid: int
    R: np.ndarray
    T: np.ndarray
    FovY: float
    FovX: float
    image_path: str
    image_name: str
    width: int
    height: int
    point3D_ids: Optional[np.ndarray] = None


class SceneInfo(NamedTuple):            # shen_data_read.py:25-30
    point_cloud: BasicPointCloud
    train_cameras: list
    test_cameras: list
    nerf_normalization: dict
    ply_path: str


def _nerfpp_norm(cam_infos) -> dict:
    """shen_data_read.getNerfppNorm (:101-117)."""
    if not cam_infos:
        return {"translate": np.zeros(3), "radius": 0.0}
    centers = []
    for cam in cam_infos:
        W2C = _world2view2(cam.R, cam.T)
        centers.append(np.linalg.inv(W2C)[:3, 3:4])
    centers = np.hstack(centers)
    center = np.mean(centers, axis=1, keepdims=True)
    diagonal = np.max(np.linalg.norm(centers - center, axis=0, keepdims=True))
    return {"translate": -center.flatten(), "radius": diagonal * 1.1}


def read_scene(path: str, man_trans=None, eval: bool = False, llffhold: int = 83) -> SceneInfo:
    """shen_data_read.partition (:150-206): cameras from sparse/0/{images,cameras}.bin
    sorted by image name, R = qvec2rotmat(q) (world-to-camera, NOT transposed,
    :57), FoV = 2 atan(pixels / 2f) for PINHOLE / SIMPLE_PINHOLE (:74-85); the
    cloud from points3D.ply, created from points3D.bin on first use (:185-192).
    `man_trans` rotates the cloud only, in float64 (:126-130), like the reference.
    When the PLY does not exist yet the cloud is taken straight from the decoded points3D.bin
    (float32 coordinates, u8 colours / 255, zero normals -- exactly what the PLY round trip
    returns) and the PLY is still written for the next run and for `ply_path` consumers."""
    from .workload import qvec2rotmat, focal2fov
    sp = os.path.join(path, "sparse", "0")
    imgs = colmap_io.read_images_binary(os.path.join(sp, "images.bin"), with_point3D_ids=True)
    intr = colmap_io.read_cameras_binary(os.path.join(sp, "cameras.bin"))
    infos = []
    for i in range(len(imgs.ids)):
        cam = intr[int(imgs.camera_ids[i])]
        if cam.model == "SIMPLE_PINHOLE":
            fx = fy = cam.params[0]
        elif cam.model == "PINHOLE":
            fx, fy = cam.params[0], cam.params[1]
        else:
            raise ValueError("Colmap camera model not handled: only undistorted datasets "
                             "(PINHOLE or SIMPLE_PINHOLE cameras) supported!")
        name = os.path.basename(imgs.names[i]).split(".")[0]
        infos.append(CameraInfo(uid=cam.id, R=qvec2rotmat(imgs.qvecs[i]), T=np.array(imgs.tvecs[i]),
                                FovY=focal2fov(fy, cam.height), FovX=focal2fov(fx, cam.width),
                                image_path=os.path.join(path, "images", os.path.basename(imgs.names[i])),
                                image_name=name, width=cam.width, height=cam.height,
                                point3D_ids=imgs.point3D_ids[i] if imgs.point3D_ids is not None else None))
    infos.sort(key=lambda c: c.image_name)
    if eval:
        train = [c for i, c in enumerate(infos) if i % llffhold != 0]
        test = [c for i, c in enumerate(infos) if i % llffhold == 0]
    else:
        train, test = infos, []
    ply = os.path.join(sp, "points3D.ply")
    if not os.path.exists(ply):
        xyz, rgb, _ = colmap_io.read_points3D_binary(os.path.join(sp, "points3D.bin"))
        colmap_io.store_ply(ply, xyz, rgb)
        points = xyz.astype(np.float32)
        colors = rgb / 255.0
        normals = np.zeros_like(points)
    else:
        points, colors, normals = colmap_io.fetch_ply(ply)
    if man_trans is not None:
        mt = np.asarray(man_trans, dtype=np.float64)
        R_, T_ = mt[:3, :3], mt[:3, -1]
        new = np.dot(R_, points.transpose()) + np.repeat(T_, points.shape[0]).reshape(-1, points.shape[0])
        points = new.transpose()
    return SceneInfo(BasicPointCloud(points, colors, normals), train, test, _nerfpp_norm(train), ply)


def cameraList_from_camInfos_partition(cam_infos, args=None) -> list:
    """utils/camera_utils.py:130-148: uid = index in the name-sorted list,
    colmap_id = the COLMAP camera id, data_device = args.data_device ("cpu" for the script)."""
    dev = getattr(args, "data_device", "cpu") if args is not None else "cpu"
    return [SimpleCamera(colmap_id=c.uid, R=c.R, T=c.T, FoVx=c.FovX, FoVy=c.FovY, image_name=c.image_name,
                         uid=i, width=c.width, height=c.height, data_device=dev, points3D_ids=c.point3D_ids)
            for i, c in enumerate(cam_infos)]


def save_test_cameras(cameras, source_path: str) -> None:
    """save_partition.py:170-208: <sfm>/test/sparse/0/{images.bin, cameras.bin} and the test
    images copied to <sfm>/test/images (copy_images, :73-98: a missing image is reported and
    skipped); a missing <sfm>/sparse/0/cameras.bin raises FileNotFoundError like the reference."""
    test_path = os.path.join(source_path, "test")
    sp = os.path.join(test_path, "sparse", "0")
    images_src = os.path.join(source_path, "images")
    images_dst = os.path.join(test_path, "images")
    os.makedirs(sp, exist_ok=True)
    os.makedirs(images_src, exist_ok=True)
    os.makedirs(images_dst, exist_ok=True)
    for c in cameras:
        src = os.path.join(images_src, c.image_name + ".jpg")
        if os.path.exists(src):
            shutil.copy(src, os.path.join(images_dst, c.image_name + ".jpg"))
    # the reference keys the images by uid (save_partition.py:130): later duplicates win
    by_uid = {}
    for c in cameras:
        by_uid[c.uid] = c
    cams_u = list(by_uid.values())
    colmap_io.write_images_binary(
        [c.uid for c in cams_u],
        np.array([rotmat_to_colmap_qvec(c.R) for c in cams_u], dtype=np.float64).reshape(-1, 4),
        np.array([np.asarray(c.T, dtype=np.float64).reshape(3) for c in cams_u]).reshape(-1, 3),
        [c.colmap_id for c in cams_u], [c.image_name + ".jpg" for c in cams_u], os.path.join(sp, "images.bin"))
    src = os.path.join(source_path, "sparse", "0", "cameras.bin")
    if not os.path.exists(src):
        raise FileNotFoundError("cameras.bin not found: %s" % src)
    shutil.copy(src, os.path.join(sp, "cameras.bin"))


# ---- A0: the orchestrator (shen_partition_utils.py:13-101) ------------------------------------
class ProgressiveDataPartitioning:
    """Drop-in for shen_partition_utils.ProgressiveDataPartitioning: same
    constructor, same attributes (`partitions`, `partition_dir`,
    `partition_visible_dir`, `pcd`, `ply`, `train_cameras`, `threshold`, ...), same on-disk result.
    The whole of run_DataPartition (A1-A9) is ONE device pipeline (device.partition_device);
    only the per-block clouds / camera lists come back to the host for the A10
    writers.  Remaining deviations are listed in INTEGRATION.md."""

    def __init__(self, scene_info, train_cameras, threshold, model_path, write_outputs: bool = True):
        self.partition_scene = None
        self.ply = scene_info.ply_path                   # utils-shen:17
        self.ply_path = scene_info.ply_path              # round-1 spelling, kept
        self.pcd = scene_info.point_cloud
        self.train_cameras = train_cameras
        self.threshold = threshold
        self.model_path = model_path
        self.partition_dir = os.path.join(model_path, "split_result")
        self.partition_visible_dir = os.path.join(self.partition_dir, "visible")
        self.save_partition_data_dir = os.path.join(self.model_path, "partition_data.pkl")
        self.write_outputs = write_outputs
        self.device_result = None
        os.makedirs(self.partition_visible_dir, exist_ok=True)
        self.partitions = self.run_DataPartition()

    def run_DataPartition(self, train_cameras=None, threshold=None) -> List[Partition]:
        """utils-shen:57-93 (no arguments there; the optional ones override the attributes)."""
        train_cameras = self.train_cameras if train_cameras is None else train_cameras
        threshold = self.threshold if threshold is None else threshold
        dev = _dev()
        pts = np.ascontiguousarray(self.pcd.points)
        # the split reads the PLY (float32 coordinates, utils-shen:62); expansion and
        # visibility use self.pcd (float64 and rotated under man_trans) -- D6 in SURVEY.md
        split_xyz = None
        if pts.dtype != np.float32:
            ply_pts, _, _ = colmap_io.fetch_ply(self.ply)
            split_xyz = torch.from_numpy(np.ascontiguousarray(ply_pts, dtype=np.float32)).to(dev)
        recs = D.camera_records(*_camera_arrays(train_cameras))
        centers = np.stack([_center_of(c) for c in train_cameras]).astype(np.float32)
        # a float64 self.pcd (already rotated by fetchPly) runs the float64 kernels as it is
        res = D.partition_device(torch.from_numpy(pts).to(dev), recs, centers, threshold, split_xyz=split_xyz)
        self.device_result = res
        p_off = res.pt_offsets.cpu().numpy()
        p_idx = res.pt_index.cpu().numpy().astype(np.int64)
        k_off = res.kept_offsets.cpu().numpy()
        k_idx = res.kept_index.cpu().numpy()
        boxes = D.expanded_boxes(res.leaf_bounds)
        colors, normals = np.asarray(self.pcd.colors), np.asarray(self.pcd.normals)
        out = []
        for b, pid in enumerate(res.leaf_ids):
            sel = p_idx[p_off[b]:p_off[b + 1]]
            pc = BasicPointCloud(points=pts[sel], colors=colors[sel], normals=normals[sel])
            poses = [CameraPose(camera=train_cameras[c], pose=centers[c]) for c in k_idx[k_off[b]:k_off[b + 1]]]
            part = Partition(partition_id=pid, camera=poses, point_cloud=pc,
                             origin_box=box(*[float(v) for v in res.leaf_bounds[b]]), point_num=len(sel),
                             extend_rate=80, extend_box=box(*[float(v) for v in boxes[b]]))
            if self.write_outputs:
                save_partition_as_pkl(part, self.partition_visible_dir)
            out.append(part)
        return out

    def save_partition_data(self):                       # utils-shen:94-97
        with open(self.save_partition_data_dir, "wb") as f:
            dump_compat(self.partitions, f)

    def load_partition_data(self):                       # utils-shen:98-101 (assigns, returns None)
        with open(self.save_partition_data_dir, "rb") as f:
            self.partitions = load_compat(f)


def main(argv=None) -> "ProgressiveDataPartitioning":
    """The script body of shen_partition_utils.py (:102-126):
    `python shen_partition_utils.py <sfm_path>` with threshold_value = 500000."""
    argv = sys.argv[1:] if argv is None else list(argv)
    if not argv:
        raise SystemExit("usage: shen_partition_utils.py <sfm_path> [threshold_value]")
    path = argv[0]
    threshold_value = int(argv[1]) if len(argv) > 1 else 500000
    model_path = os.path.join(path, "model")
    scene_info = read_scene(path, None)

    class args:                                          # utils-shen:108-112
        scene_path = path
        partition_dir = model_path
        data_device = "cpu"
    args.model_path = model_path
    # train and test lists are enumerated separately (uid restarts at 0 for each, utils-shen:115-119)
    test_cameras = cameraList_from_camInfos_partition(scene_info.test_cameras, args)
    save_test_cameras(test_cameras, path)
    train_cameras = cameraList_from_camInfos_partition(scene_info.train_cameras, args)
    return ProgressiveDataPartitioning(scene_info, train_cameras, threshold_value, model_path)
