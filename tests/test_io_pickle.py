"""SURVEY.md "next" rows N2 / N3 on the CPU:

* N3 -- the COLMAP readers on real-shaped files: points3D.bin WITH tracks (variable-length records,
  scene/colmap_loader.py:195-224 unpacks them one by one) and images.bin with 2D observations
  (point3D ids forwarded like colmap_loader.py:276-277);
* N2 -- the pickles carry the reference's class paths (shen_partation_densify.Partition,
  scene.cameras.SimpleCamera, shapely.from_wkb rectangles) and load with PLAIN pickle.load where those
  modules exist -- the way cut_partition.py:121 / merge_partitions.py:116 load them -- as well as
  through load_compat() where they do not."""
import io
import os
import pickle
import struct
import sys
import types
from typing import NamedTuple, Optional

import numpy as np
import pytest
import torch

import ptgs_b200  # noqa: F401
from ptgs_b200 import colmap_io as IO, partition as P


def _cloud(n, seed=0):
    rng = np.random.default_rng(seed)
    return rng.normal(size=(n, 3)) * 50, rng.integers(0, 256, (n, 3), dtype=np.uint8), rng.integers(0, 9, n)


# ------------------------------------------------------------------------------------- N3
@pytest.mark.parametrize("n", [0, 1, 257, 20000])
def test_points3d_with_tracks_roundtrip(tmp_path, n):
    xyz, rgb, tl = _cloud(n, seed=n)
    path = str(tmp_path / "points3D.bin")
    IO.write_points3D_binary_with_tracks(xyz, rgb, tl, path, seed=3)
    assert os.path.getsize(path) == 8 + 51 * n + 8 * int(tl.sum())
    gx, gr, ge = IO.read_points3D_binary(path)
    assert gx.dtype == np.float64 and np.array_equal(gx, xyz) and np.array_equal(gr, rgb)
    assert np.array_equal(ge, np.random.default_rng(3).uniform(0, 2, n))
    # no tracks: the single-frombuffer path gives the same arrays
    IO.write_points3D_binary(xyz, rgb, path)
    hx, hr, he = IO.read_points3D_binary(path)
    assert np.array_equal(hx, xyz) and np.array_equal(hr, rgb) and not he.any()


def test_points3d_truncated_file_raises(tmp_path):
    xyz, rgb, tl = _cloud(100, seed=1)
    path = str(tmp_path / "points3D.bin")
    IO.write_points3D_binary_with_tracks(xyz, rgb, tl + 1, path)
    data = open(path, "rb").read()
    open(path, "wb").write(data[:-5])
    with pytest.raises(ValueError):
        IO.read_points3D_binary(path)


@pytest.mark.reference
def test_points3d_reader_equals_reference(tmp_path):
    import ref_harness as H
    ref = H.load()
    xyz, rgb, tl = _cloud(5000, seed=2)
    path = str(tmp_path / "points3D.bin")
    IO.write_points3D_binary_with_tracks(xyz, rgb, tl, path)
    wx, wr, we = ref.colmap_loader.read_points3D_binary(path)
    gx, gr, ge = IO.read_points3D_binary(path)
    assert np.array_equal(gx, wx) and np.array_equal(gr, wr) and np.array_equal(ge, np.asarray(we).reshape(-1))


def _write_images_with_points2d(path, ids, qvecs, tvecs, cam_ids, names, p3d_lists, seed=0):
    rng = np.random.default_rng(seed)
    out = bytearray(struct.pack("<Q", len(ids)))
    for i in range(len(ids)):
        out += struct.pack("<idddddddi", int(ids[i]), *map(float, qvecs[i]), *map(float, tvecs[i]), int(cam_ids[i]))
        out += names[i].encode() + b"\x00"
        out += struct.pack("<Q", len(p3d_lists[i]))
        for pid in p3d_lists[i]:
            out += struct.pack("<ddq", rng.uniform(0, 1600), rng.uniform(0, 1200), int(pid))
    open(path, "wb").write(bytes(out))


def test_images_with_observations(tmp_path):
    rng = np.random.default_rng(4)
    C = 9
    q = rng.normal(size=(C, 4))
    t = rng.normal(size=(C, 3))
    p3d = [rng.integers(-1, 1000, rng.integers(0, 40)) for _ in range(C)]
    names = ["sub/img_%03d.jpg" % i for i in range(C)]
    path = str(tmp_path / "images.bin")
    _write_images_with_points2d(path, np.arange(1, C + 1), q, t, np.ones(C, int), names, p3d)
    plain = IO.read_images_binary(path)
    assert plain.point3D_ids is None and plain.names == names
    assert np.array_equal(plain.qvecs, q) and np.array_equal(plain.tvecs, t)
    full = IO.read_images_binary(path, with_point3D_ids=True)
    assert all(np.array_equal(a, b) and a.dtype == np.int64 for a, b in zip(full.point3D_ids, p3d))


# ------------------------------------------------------------------------------------- N2
def _sample_partition():
    cam = P.SimpleCamera(colmap_id=7, R=np.eye(3), T=np.array([1.0, 2.0, 3.0]), FoVx=0.9, FoVy=0.7,
                         image_name="img_000004", uid=4, width=1600, height=1200,
                         points3D_ids=np.array([5, 9], dtype=np.int64))
    pts = np.arange(12, dtype=np.float32).reshape(4, 3)
    pc = P.BasicPointCloud(points=pts, colors=pts / 20.0, normals=np.zeros_like(pts))
    return P.Partition(partition_id="010", camera=[P.CameraPose(camera=cam, pose=np.array([-1, -2, -3], np.float32))],
                       point_cloud=pc, origin_box=P.Rect(-1.5, 2.0, 3.25, 8.0), point_num=4, extend_rate=80,
                       extend_box=P.Rect(-1.7375, 1.7, 3.4875, 8.3))


def test_rect_wkb_is_shapely_box_layout():
    r = P.Rect(-1.5, 2.0, 3.25, 8.0)
    w = r.wkb
    assert len(w) == 1 + 4 + 4 + 4 + 5 * 16 and w[0] == 1
    assert struct.unpack_from("<III", w, 1) == (3, 1, 5)
    ring = np.frombuffer(w, "<f8", 10, 13).reshape(5, 2)
    # shapely.geometry.box(minx, miny, maxx, maxy): (maxx,miny) (maxx,maxy) (minx,maxy) (minx,miny) closed
    assert ring.tolist() == [[3.25, 2.0], [3.25, 8.0], [-1.5, 8.0], [-1.5, 2.0], [3.25, 2.0]]
    assert P._rect_from_wkb(w) == r


def test_pickle_carries_reference_class_paths_and_loads_compat():
    part = _sample_partition()
    data = P.dumps_compat([part])
    for path in (b"cshen_partation_densify\nPartition\n", b"cshen_partation_densify\nBasicPointCloud\n",
                 b"cshen_partation_densify\nCameraPose\n", b"cscene.cameras\nSimpleCamera\n", b"cshapely\nfrom_wkb\n"):
        assert path in data
    assert b"ptgs_b200" not in data
    q = P.load_compat(io.BytesIO(data))[0]      # no shapely / scene / reference modules needed
    assert q.partition_id == "010" and q.origin_box.bounds == (-1.5, 2.0, 3.25, 8.0)
    assert np.array_equal(q.point_cloud.points, part.point_cloud.points)
    c = q.camera[0].camera
    assert c.uid == 4 and c.colmap_id == 7 and c.image_name == "img_000004"
    assert torch.equal(c.camera_center, part.camera[0].camera.camera_center)


@pytest.fixture
def consumer_modules():
    """A consumer machine: shapely (from_wkb), scene.cameras (SimpleCamera as an nn.Module) and
    shen_partation_densify (the NamedTuples) exist, this package's modules are not consulted."""
    class SimpleCamera(torch.nn.Module):              # scene/cameras.py:18 (state restored by pickle)
        def get_camera_center(self):
            return self.camera_center

    class Poly:
        def __init__(self, ring):
            self.ring = ring

        @property
        def bounds(self):
            xs, ys = [p[0] for p in self.ring], [p[1] for p in self.ring]
            return (min(xs), min(ys), max(xs), max(ys))

    def from_wkb(data):
        assert data[0] == 1 and struct.unpack_from("<III", data, 1)[:2] == (3, 1)
        n = struct.unpack_from("<I", data, 9)[0]
        xy = struct.unpack_from("<%dd" % (2 * n), data, 13)
        return Poly([(xy[2 * i], xy[2 * i + 1]) for i in range(n)])

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

    mods = {"shapely": types.ModuleType("shapely"), "scene": types.ModuleType("scene"),
            "scene.cameras": types.ModuleType("scene.cameras"),
            "shen_partation_densify": types.ModuleType("shen_partation_densify")}
    mods["shapely"].from_wkb = from_wkb
    mods["scene"].__path__ = []
    mods["scene.cameras"].SimpleCamera = SimpleCamera
    SimpleCamera.__module__, SimpleCamera.__qualname__ = "scene.cameras", "SimpleCamera"
    for t in (BasicPointCloud, Partition, CameraPose):
        t.__module__, t.__qualname__ = "shen_partation_densify", t.__name__
        setattr(mods["shen_partation_densify"], t.__name__, t)
    saved = {k: sys.modules.get(k) for k in mods}
    sys.modules.update(mods)
    try:
        yield mods
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_plain_pickle_load_on_a_consumer_machine(consumer_modules, tmp_path):
    """cut_partition.py:116-121 / merge_partitions.py:110-116: `partition = pickle.load(f)` then
    `partition.origin_box` / `.bounds`, `partition.point_cloud`, `partition.camera`."""
    part = _sample_partition()
    path = tmp_path / "partition_010.pkl"
    with open(path, "wb") as f:
        P.dump_compat(part, f)
    with open(path, "rb") as f:
        q = pickle.load(f)
    m = consumer_modules
    assert type(q) is m["shen_partation_densify"].Partition
    assert type(q.point_cloud) is m["shen_partation_densify"].BasicPointCloud
    assert q.origin_box.bounds == (-1.5, 2.0, 3.25, 8.0) and q.extend_box.bounds == (-1.7375, 1.7, 3.4875, 8.3)
    minx, miny, maxx, maxy = q.origin_box.bounds            # merge_partitions.py:16
    assert (minx, maxy) == (-1.5, 8.0)
    cam = q.camera[0].camera
    assert type(cam) is m["scene.cameras"].SimpleCamera and isinstance(cam, torch.nn.Module)
    for k, v in (("uid", 4), ("colmap_id", 7), ("image_name", "img_000004"), ("image_width", 1600),
                 ("image_height", 1200), ("zfar", 100.0), ("znear", 0.01), ("scale", 1.0)):
        assert getattr(cam, k) == v
    assert np.array_equal(cam.points3D_ids, [5, 9]) and cam.data_device == torch.device("cpu")
    assert torch.equal(cam.get_camera_center(), torch.tensor([-1.0, -2.0, -3.0]))
    assert cam.world_view_transform.shape == (4, 4) and cam.full_proj_transform.shape == (4, 4)
    assert list(cam.parameters()) == []                      # nn.Module state is complete


@pytest.mark.reference
def test_plain_pickle_load_with_the_reference_modules(tmp_path):
    """Same, with the reference's OWN shen_partation_densify and scene.cameras modules
    (oracle/ref_harness.py): the loaded camera equals one the reference constructs itself."""
    import ref_harness as H
    ref = H.load()
    part = _sample_partition()
    q = pickle.loads(P.dumps_compat(part))
    assert type(q) is ref.densify.Partition and type(q.camera[0]) is ref.densify.CameraPose
    cam = q.camera[0].camera
    assert type(cam) is ref.cameras.SimpleCamera
    want = H.make_simple_camera(ref, 4, 7, np.eye(3), np.array([1.0, 2.0, 3.0]), 0.9, 0.7, "img_000004", 1600, 1200)
    for k in ("uid", "colmap_id", "FoVx", "FoVy", "image_name", "image_width", "image_height", "zfar", "znear",
              "scale", "data_device"):
        assert getattr(cam, k) == getattr(want, k), k
    for k in ("world_view_transform", "projection_matrix", "full_proj_transform", "camera_center"):
        assert torch.equal(getattr(cam, k), getattr(want, k)), k
    assert set(vars(want)) <= set(vars(cam)) | {"points3D_ids"}
    assert tuple(q.origin_box.bounds) == (-1.5, 2.0, 3.25, 8.0)


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_points3d_to_device(tmp_path):
    xyz, rgb, tl = _cloud(30011, seed=8)
    path = str(tmp_path / "points3D.bin")
    IO.write_points3D_binary_with_tracks(xyz, rgb, tl, path)
    x, y, z, c = IO.read_points3D_to_device(path, "cuda:0")
    want = xyz.astype(np.float32)                            # storePly's cast
    got = np.stack([x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy()], 1)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)) and np.array_equal(c.cpu().numpy(), rgb)
    IO.write_points3D_binary(xyz, rgb, path)                  # no tracks
    x, y, z, c = IO.read_points3D_to_device(path, "cuda:0", want_rgb=False)
    got = np.stack([x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy()], 1)
    assert c is None and np.array_equal(got.view(np.uint32), want.view(np.uint32))
