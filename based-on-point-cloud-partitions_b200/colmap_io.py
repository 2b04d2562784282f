"""COLMAP binary / PLY readers and writers on the fast path (SURVEY.md rows
N1 and N3).  Byte-compatible with the reference's per-record struct loops:
  readers  scene/colmap_loader.py:195-311 (read_points3D_binary,
           read_extrinsics_binary, read_intrinsics_binary)
  writers  scene/ptgs/read_write_model.py:189-203 (cameras), :335-353 (images),
           :457-473 (points3D); PLY: scene/ptgs/shen_data_read.py:120-148.
Everything is NumPy structured arrays + one tofile()/frombuffer() where the
record size is fixed; variable-length records (tracks, 2D points) are walked
once with precomputed offsets."""
from __future__ import annotations

import os
import struct
from typing import NamedTuple

import numpy as np

CAMERA_MODEL_PARAMS = {0: 3, 1: 4, 2: 4, 3: 5, 4: 8, 5: 8, 6: 12, 7: 5, 8: 4, 9: 5, 10: 12}
CAMERA_MODEL_NAMES = {0: "SIMPLE_PINHOLE", 1: "PINHOLE", 2: "SIMPLE_RADIAL", 3: "RADIAL", 4: "OPENCV",
                      5: "OPENCV_FISHEYE", 6: "FULL_OPENCV", 7: "FOV", 8: "SIMPLE_RADIAL_FISHEYE",
                      9: "RADIAL_FISHEYE", 10: "THIN_PRISM_FISHEYE"}


class ColmapCamera(NamedTuple):
    id: int
    model: str
    width: int
    height: int
    params: np.ndarray


class ColmapImages(NamedTuple):
    """Columnar form of images.bin (the 2D observations are not needed by the
    partitioner and are skipped)."""
    ids: np.ndarray        # (C,) int32
    qvecs: np.ndarray      # (C,4) float64  [w,x,y,z]
    tvecs: np.ndarray      # (C,3) float64
    camera_ids: np.ndarray  # (C,) int32
    names: list            # str
    point3D_ids: list = None   # per image int64 array (colmap_loader.py:276-277), when requested


def read_cameras_binary(path) -> dict:
    """colmap_loader.read_intrinsics_binary (:285-311)."""
    cams = {}
    with open(path, "rb") as f:
        buf = f.read()
    n = struct.unpack_from("<Q", buf, 0)[0]
    off = 8
    for _ in range(n):
        cid, model, w, h = struct.unpack_from("<iiQQ", buf, off)
        off += 24
        k = CAMERA_MODEL_PARAMS[model]
        params = np.frombuffer(buf, dtype="<f8", count=k, offset=off).copy()
        off += 8 * k
        cams[cid] = ColmapCamera(cid, CAMERA_MODEL_NAMES[model], int(w), int(h), params)
    return cams


def read_images_binary(path, with_point3D_ids: bool = False) -> ColmapImages:
    """colmap_loader.read_extrinsics_binary (:250-282) without materialising
    the per-image 2D point lists; with_point3D_ids keeps the point3D id column of
    the 2D observations (one strided view per image, :276-277)."""
    with open(path, "rb") as f:
        buf = f.read()
    n = struct.unpack_from("<Q", buf, 0)[0]
    ids = np.empty(n, np.int32)
    q = np.empty((n, 4), np.float64)
    t = np.empty((n, 3), np.float64)
    cid = np.empty(n, np.int32)
    names = []
    p3d = [] if with_point3D_ids else None
    off = 8
    for i in range(n):
        rec = struct.unpack_from("<idddddddi", buf, off)
        ids[i] = rec[0]
        q[i] = rec[1:5]
        t[i] = rec[5:8]
        cid[i] = rec[8]
        off += 64
        end = buf.index(b"\x00", off)
        names.append(buf[off:end].decode("utf-8"))
        off = end + 1
        n2d = struct.unpack_from("<Q", buf, off)[0]
        if p3d is not None:
            p3d.append(np.frombuffer(buf, dtype=_P2D, count=n2d, offset=off + 8)["id"].copy())
        off += 8 + 24 * n2d
    return ColmapImages(ids, q, t, cid, names, p3d)


_P2D = np.dtype([("xy", "<f8", 2), ("id", "<i8")])
_P3D_HEAD = np.dtype([("id", "<u8"), ("xyz", "<f8", 3), ("rgb", "u1", 3), ("error", "<f8"), ("track", "<u8")])


def _record_offsets(buf, n):
    """Byte offset of every points3D record: libptgs' host walk (ptgs_points3d_index)."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    off = np.empty(n, dtype=np.int64)
    src = np.frombuffer(buf, dtype=np.uint8)
    got = lib.ptgs_points3d_index(C.c_void_p(src.ctypes.data), len(buf), C.c_void_p(off.ctypes.data), n)
    if got != n:
        raise ValueError("points3D.bin: %s" % lib.ptgs_last_error().decode("utf-8", "replace"))
    return off


def read_points3D_binary(path):
    """colmap_loader.read_points3D_binary (:195-224): (xyz f64 (N,3), rgb u8
    (N,3), error f64 (N,)).  Fixed 51-byte head + 8 bytes per track element.  Without tracks
    the whole file is one structured frombuffer; with tracks the record offsets come from one
    native walk over the file image and the heads are gathered with one fancy index -- no
    per-point Python loop either way (the reference unpacks record by record)."""
    with open(path, "rb") as f:
        buf = f.read()
    n = struct.unpack_from("<Q", buf, 0)[0]
    head = _P3D_HEAD.itemsize  # 51
    if n == 0:
        return np.zeros((0, 3)), np.zeros((0, 3), np.uint8), np.zeros(0)
    if len(buf) == 8 + n * head:  # no tracks at all
        a = np.frombuffer(buf, dtype=_P3D_HEAD, count=n, offset=8)
        return a["xyz"].copy(), a["rgb"].copy(), a["error"].copy()
    off = _record_offsets(buf, n)
    raw = np.frombuffer(buf, dtype=np.uint8)
    heads = raw[off[:, None] + np.arange(head, dtype=np.int64)[None, :]]        # (n, 43) gather
    a = heads.reshape(-1).view(_P3D_HEAD)
    return a["xyz"].copy(), a["rgb"].copy(), a["error"].copy()


def read_points3D_to_device(path, device, want_rgb: bool = True):
    """points3D.bin -> planar float32 device tensors without the PLY round trip
    (shen_data_read.py:185-193 converts to points3D.ply and reads it back): the file image goes
    to the device through pinned memory and ptgs_points3d_decode extracts xyz (with storePly's
    float32 cast), rgb and the error there.  Returns (x, y, z, rgb (N,3) u8 or None)."""
    import ctypes as C
    import torch
    from . import _lib
    lib = _lib.load()
    size = os.path.getsize(path)
    host = torch.empty(size, dtype=torch.uint8, pin_memory=True)
    with open(path, "rb") as f:
        f.readinto(memoryview(host.numpy()))
    buf = host.numpy()
    n = int(struct.unpack_from("<Q", buf, 0)[0])
    off = torch.empty(max(n, 1), dtype=torch.int64, pin_memory=True)
    got = lib.ptgs_points3d_index(C.c_void_p(buf.ctypes.data), size, C.c_void_p(off.numpy().ctypes.data), n)
    if got != n:
        raise ValueError("points3D.bin: %s" % lib.ptgs_last_error().decode("utf-8", "replace"))
    dev = torch.device(device)
    file_d = host.to(dev, non_blocking=True)
    off_d = off.to(dev, non_blocking=True)
    pitch = (n + 8 + 3) // 4 * 4
    planes = torch.zeros((3, pitch), dtype=torch.float32, device=dev)
    rgb = torch.empty((max(n, 1), 3), dtype=torch.uint8, device=dev) if want_rgb else None
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(lib.ptgs_points3d_decode(C.c_void_p(file_d.data_ptr()), C.c_void_p(off_d.data_ptr()), n,
                                        C.c_void_p(planes[0].data_ptr()), C.c_void_p(planes[1].data_ptr()),
                                        C.c_void_p(planes[2].data_ptr()),
                                        C.c_void_p(rgb.data_ptr()) if rgb is not None else C.c_void_p(0),
                                        C.c_void_p(0), st))
    return planes[0, :n], planes[1, :n], planes[2, :n], (rgb[:n] if rgb is not None else None)


def write_points3D_binary_with_tracks(xyz, rgb_u8, track_lengths, path, seed=0):
    """A points3D.bin WITH tracks (synthetic image ids / 2D indices), the shape real COLMAP
    reconstructions have: test and benchmark input for the readers above."""
    n = len(xyz)
    tl = np.asarray(track_lengths, dtype=np.int64)
    rec = 51 + 8 * tl
    off = np.concatenate([[0], np.cumsum(rec)])
    out = np.zeros(int(off[-1]), dtype=np.uint8)
    a = np.zeros(n, dtype=_P3D_HEAD)
    a["id"] = np.arange(1, n + 1, dtype=np.uint64)
    a["xyz"] = np.asarray(xyz, dtype=np.float64)
    a["rgb"] = rgb_u8
    a["error"] = np.random.default_rng(seed).uniform(0, 2, n)
    a["track"] = tl.astype(np.uint64)
    out[off[:-1, None] + np.arange(51)[None, :]] = a.view(np.uint8).reshape(n, 51)
    tracks = np.random.default_rng(seed + 1).integers(0, 1000, int(tl.sum()) * 2).astype("<i4")
    pos = np.repeat(off[:-1] + 51, tl) + 8 * (np.arange(int(tl.sum())) - np.repeat(np.cumsum(tl) - tl, tl))
    out[pos[:, None] + np.arange(8)[None, :]] = tracks.view(np.uint8).reshape(-1, 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", n))
        out.tofile(f)


# ---------------------------------------------------------------------------
# writers
# ---------------------------------------------------------------------------
def write_cameras_binary(cams: dict, path):
    """read_write_model.write_cameras_binary (:189-203)."""
    inv = {v: k for k, v in CAMERA_MODEL_NAMES.items()}
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(cams)))
        for c in cams.values():
            f.write(struct.pack("<iiQQ", c.id, inv[c.model], c.width, c.height))
            f.write(np.asarray(c.params, dtype="<f8").tobytes())


def write_images_binary(ids, qvecs, tvecs, camera_ids, names, path):
    """read_write_model.write_images_binary (:335-353) for images WITHOUT 2D
    points (what save_partition.simple_camera_to_images produces, :100-131):
    id i32, qvec 4xf64, tvec 3xf64, camera_id i32, name, NUL, u64 0."""
    out = bytearray(struct.pack("<Q", len(ids)))
    for i in range(len(ids)):
        out += struct.pack("<idddddddi", int(ids[i]), *[float(v) for v in qvecs[i]],
                           *[float(v) for v in tvecs[i]], int(camera_ids[i]))
        out += names[i].encode("utf-8") + b"\x00"
        out += struct.pack("<Q", 0)
    with open(path, "wb") as f:
        f.write(bytes(out))


def colors_to_u8(colors: np.ndarray) -> np.ndarray:
    """save_partition.convert_to_colmap_format (:139-158): when
    colors.max() <= 1.0 the float colours are scaled by 255 and TRUNCATED to
    uint8 (so u8/255*255 can land one below the original value); otherwise the
    per-point np.array(colors[idx], dtype=np.uint8) cast applies."""
    colors = np.asarray(colors)
    if colors.size and colors.max() <= 1.0:
        return (colors * 255).astype(np.uint8)
    return colors.astype(np.uint8)


def write_points3D_binary(xyz, rgb_u8, path):
    """read_write_model.write_points3D_binary (:457-473) for points without
    tracks: id = index (u64), xyz f64, rgb u8, error 0.0, track length 0 --
    one structured array, one tofile()."""
    n = len(xyz)
    a = np.zeros(n, dtype=_P3D_HEAD)
    a["id"] = np.arange(n, dtype=np.uint64)
    a["xyz"] = np.asarray(xyz, dtype=np.float64)
    a["rgb"] = rgb_u8
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", n))
        a.tofile(f)


# ---------------------------------------------------------------------------
# PLY (binary little endian; the layout plyfile writes for storePly)
# ---------------------------------------------------------------------------
_PLY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("nx", "<f4"), ("ny", "<f4"), ("nz", "<f4"),
                       ("red", "u1"), ("green", "u1"), ("blue", "u1")])


def store_ply(path, xyz, rgb):
    """shen_data_read.storePly (:135-148): float32 xyz, zero normals, u8 rgb."""
    n = len(xyz)
    a = np.zeros(n, dtype=_PLY_DTYPE)
    a["x"], a["y"], a["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    a["red"], a["green"], a["blue"] = rgb[:, 0], rgb[:, 1], rgb[:, 2]
    header = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\n"
              "property float x\nproperty float y\nproperty float z\n"
              "property float nx\nproperty float ny\nproperty float nz\n"
              "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n" % n)
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        a.tofile(f)


def fetch_ply(path):
    """shen_data_read.fetchPly (:120-133) without the Manhattan branch: points
    float32 (N,3), colors float64 = u8 / 255.0, normals float32.  Understands
    binary little-endian vertex elements with scalar properties."""
    with open(path, "rb") as f:
        buf = f.read()
    end = buf.index(b"end_header\n") + len(b"end_header\n")
    lines = buf[:end].decode("ascii").split("\n")
    if "format binary_little_endian 1.0" not in lines:
        raise ValueError("fetch_ply: only binary_little_endian PLY files are supported")
    tmap = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
            "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
            "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}
    n, fields, in_vertex = 0, [], False
    for ln in lines:
        tok = ln.split()
        if tok[:2] == ["element", "vertex"]:
            n, in_vertex = int(tok[2]), True
        elif tok[:1] == ["element"]:
            in_vertex = False
        elif tok[:1] == ["property"] and in_vertex:
            fields.append((tok[2], tmap[tok[1]]))
    a = np.frombuffer(buf, dtype=np.dtype(fields), count=n, offset=end)
    points = np.vstack([a["x"], a["y"], a["z"]]).T
    colors = np.vstack([a["red"], a["green"], a["blue"]]).T / 255.0
    normals = np.vstack([a["nx"], a["ny"], a["nz"]]).T
    return points, colors, normals


def write_synthetic_colmap(sc, root):
    """A SynthScene as a COLMAP sparse model under <root>/sparse/0 (+ empty
    <root>/images): cameras.bin (one PINHOLE camera), images.bin (no 2D
    points), points3D.bin (no tracks).  SURVEY.md section 8(d)."""
    sp = os.path.join(root, "sparse", "0")
    os.makedirs(sp, exist_ok=True)
    os.makedirs(os.path.join(root, "images"), exist_ok=True)
    cam = ColmapCamera(1, "PINHOLE", sc.width, sc.height, np.array([sc.fx, sc.fy, sc.cx, sc.cy]))
    write_cameras_binary({1: cam}, os.path.join(sp, "cameras.bin"))
    C = len(sc.qvec)
    write_images_binary(np.arange(1, C + 1), sc.qvec, sc.tvec, np.ones(C, np.int32), sc.names,
                        os.path.join(sp, "images.bin"))
    write_points3D_binary(sc.xyz.astype(np.float64), sc.rgb, os.path.join(sp, "points3D.bin"))
    return sp
