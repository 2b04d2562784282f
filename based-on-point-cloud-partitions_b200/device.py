"""Device pipeline of the point-count partitioner: thin Python over the C ABI
(include/ptgs.h).  torch is used for device memory and streams only.

Stages (reference: scene/ptgs/shen_partition_utils.py:57-93, run_DataPartition):
  A1 radius filter -> A2 bounds -> A3 balanced split -> A4 expansion ->
  A5 camera assignment -> A6 voxel down-sampling -> A7-A9 visibility/selection.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import os as _os

import numpy as np
import torch

from . import _lib
from ._lib import PTGS_F32, PTGS_F64, CAMERA_FLOATS, check

# constants hard-coded by the reference (file:line in DESIGN.md)
RADIUS, NB_POINTS = 1.0, 5            # shen_partition_utils.py:28-30
LEAF_FACTOR, MAX_DEPTH = 1.2, 10      # density_partition.py:32, :163
EXPAND_RATIO = 0.10                   # shen_partation_densify.py:222
VOXEL_SIZE = 0.1                      # :928
VISIBLE_RATE = 0.45                   # :930
MIN_HITS = 3                          # :898
NEAR_EPS = 1e-6
SAMPLE_STRIDE = 128                   # sweep-1 sample used for the interior regions (measured optimum at c2)
KRECT_MARGIN_PX = 1.0


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


def _dt(t):
    return PTGS_F64 if t.dtype == torch.float64 else PTGS_F32


@dataclass
class VisibilityResult:
    n_hit: torch.Tensor        # (P,) int32
    area: torch.Tensor         # (P,) float64
    ratio: torch.Tensor        # (P,) float64
    keep: torch.Tensor         # (P,) uint8
    flags: torch.Tensor        # (P,) uint8
    cand_count: torch.Tensor   # (P,) int32
    rounds: int = 1
    ratio_lo: torch.Tensor = None   # (P,) float64 lower / upper bound of the ratio
    ratio_hi: torch.Tensor = None
    n_exact: int = None             # pairs that went through the exact hull


@dataclass
class PartitionDeviceResult:
    keep_mask: torch.Tensor            # (N,) uint8 (A1)
    n_filtered: int
    root_bounds: tuple
    leaf_bounds: np.ndarray            # (L,4) float64 {minx,miny,maxx,maxy}
    leaf_path: np.ndarray
    leaf_depth: np.ndarray
    leaf_count: np.ndarray
    leaf_ids: list
    boxes: np.ndarray                  # (L,4) float64 expanded {minx,miny,maxx,maxy}
    pt_offsets: torch.Tensor           # (L+1,) int64, A4 CSR
    pt_index: torch.Tensor             # int32
    cam_offsets: torch.Tensor          # (L+1,) int64, A5 CSR
    cam_index: torch.Tensor
    ds_offsets: torch.Tensor = None    # A6
    ds_x: torch.Tensor = None
    ds_y: torch.Tensor = None
    ds_z: torch.Tensor = None
    ds_first: torch.Tensor = None
    vis: VisibilityResult = None
    kept_offsets: torch.Tensor = None  # A9 CSR
    kept_index: torch.Tensor = None
    timings_ms: dict = field(default_factory=dict)
    pt_counts_h: np.ndarray = None     # host copies of the L-sized tables (no extra synchronisation)
    cam_offsets_h: np.ndarray = None
    ds_block0: int = 0                 # first block of ds_* (range-sharded runs keep only their own blocks)
    pt_index_host: tuple = None        # (pinned int32 tensor, event): early device -> host copy of pt_index


def leaf_id(path: int, depth: int) -> str:
    """density_partition.py:28/:76/:86: root leaf -> "0"; otherwise the branch
    bits, left = "0", right = "1", most significant first."""
    if depth == 0:
        return "0"
    return format(int(path), "0%db" % depth)


def expanded_boxes(leaf_bounds: np.ndarray) -> np.ndarray:
    """shen_partation_densify.py:210-232 in Python floats (host, float64)."""
    out = np.empty((len(leaf_bounds), 4), dtype=np.float64)
    for i, (minx, miny, maxx, maxy) in enumerate(leaf_bounds.tolist()):
        ew = (maxx - minx) * EXPAND_RATIO / 2
        eh = (maxy - miny) * EXPAND_RATIO / 2
        out[i] = (minx - ew, miny - eh, maxx + ew, maxy + eh)
    return out


# --------------------------------------------------------------------------
# stage wrappers
# --------------------------------------------------------------------------
def aos_to_soa(xyz: torch.Tensor):
    """(N,3) float32 device tensor -> planar x, y, z (padded for bulk copies)."""
    lib = _lib.load()
    assert xyz.is_cuda and xyz.dtype in (torch.float32, torch.float64) and xyz.is_contiguous()
    n = xyz.shape[0]
    # rows must stay 16-byte aligned: pad the row pitch to a multiple of 4
    pitch = (n + 8 + 3) // 4 * 4
    buf = torch.zeros((3, pitch), dtype=xyz.dtype, device=xyz.device)
    x, y, z = buf[0, :n], buf[1, :n], buf[2, :n]
    fn = lib.ptgs_aos_to_soa_f32 if xyz.dtype == torch.float32 else lib.ptgs_aos_to_soa_f64
    check(fn(_ptr(xyz), n, _ptr(x), _ptr(y), _ptr(z), _stream()))
    return x, y, z


def manhattan_apply(x, y, z, man_trans: np.ndarray):
    lib = _lib.load()
    n = x.shape[0]
    m = np.ascontiguousarray(man_trans, dtype=np.float64)
    pitch = (n + 8 + 1) // 2 * 2
    buf = torch.zeros((3, pitch), dtype=torch.float64, device=x.device)
    ox, oy, oz = buf[0, :n], buf[1, :n], buf[2, :n]
    check(lib.ptgs_manhattan_apply(_ptr(x), _ptr(y), _ptr(z), n,
                                   m.ctypes.data_as(C.POINTER(C.c_double)), _ptr(ox), _ptr(oy),
                                   _ptr(oz), _stream()))
    return ox, oy, oz


def aos_to_soa_manhattan(xyz: torch.Tensor, man_trans: np.ndarray):
    """(N,3) float32 device tensor -> planar float64 x, y, z in the Manhattan frame, one pass
    (ptgs_aos_to_soa_manhattan == manhattan_apply(*aos_to_soa(xyz), man_trans))."""
    lib = _lib.load()
    assert xyz.is_cuda and xyz.dtype == torch.float32 and xyz.is_contiguous()
    n = xyz.shape[0]
    m = np.ascontiguousarray(man_trans, dtype=np.float64)
    pitch = (n + 8 + 1) // 2 * 2
    buf = torch.zeros((3, pitch), dtype=torch.float64, device=xyz.device)
    ox, oy, oz = buf[0, :n], buf[1, :n], buf[2, :n]
    check(lib.ptgs_aos_to_soa_manhattan(_ptr(xyz), n, m.ctypes.data_as(C.POINTER(C.c_double)), _ptr(ox), _ptr(oy),
                                        _ptr(oz), _stream()))
    return ox, oy, oz


def radius_outlier_mask(x, y, z, radius=RADIUS, nb_points=NB_POINTS, xyz_aos=None):
    """A1.  xyz_aos: the cloud as (N,3) rows of the planes' dtype (optional, cheaper gather)."""
    lib = _lib.load()
    n = x.shape[0]
    if xyz_aos is not None:
        assert xyz_aos.dtype == x.dtype and xyz_aos.is_contiguous() and xyz_aos.shape[0] == n
    keep = torch.empty(n, dtype=torch.uint8, device=x.device)
    n_keep = torch.zeros(1, dtype=torch.int64, device=x.device)
    ws = _ws(lib.ptgs_radius_outlier_workspace_bytes(n), x.device)
    check(lib.ptgs_radius_outlier_mask(_dt(x), _ptr(x), _ptr(y), _ptr(z), _ptr(xyz_aos), n, float(radius),
                                       int(nb_points), _ptr(keep), _ptr(n_keep), _ptr(ws),
                                       ws.numel(), _stream()))
    return keep, n_keep


def slab_bounds(x, world: int, sample: int = 65536) -> np.ndarray:
    """X boundaries of `world` slabs with about equal point counts: quantiles of an every-k-th
    sample of x (deterministic, identical on every rank; any boundaries give the same mask, only
    the balance depends on them).  bounds[0] = -inf, bounds[world] = +inf."""
    n = int(x.shape[0])
    b = np.full(world + 1, np.inf, dtype=np.float64)
    b[0] = -np.inf
    if n == 0 or world <= 1:
        return b
    step = max(1, n // sample)
    xs = torch.sort(x[::step].to(torch.float64)).values
    xs = xs[torch.isfinite(xs)]
    m = int(xs.shape[0])
    if m == 0:
        return b
    q = xs[[min(m - 1, (m * r) // world) for r in range(1, world)]].cpu().numpy()
    b[1:world] = np.maximum.accumulate(q)
    return b


def slab_select(x, y, z, lo: float, hi: float):
    """ptgs_slab_select: the points with lo <= x < hi, compacted in index order.  Returns
    (planes (3, pitch) of x's dtype, cloud indices int32, device count)."""
    lib = _lib.load()
    dev = x.device
    n = int(x.shape[0])
    pitch = (max(n, 1) + 3) // 4 * 4
    sub = torch.empty((3, pitch), dtype=x.dtype, device=dev)
    sidx = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    m_dev = torch.zeros(1, dtype=torch.int64, device=dev)
    ws = _ws(lib.ptgs_slab_select_workspace_bytes(n), dev)
    check(lib.ptgs_slab_select(_dt(x), _ptr(x), _ptr(y), _ptr(z), n, float(lo), float(hi), _ptr(sub[0]),
                               _ptr(sub[1]), _ptr(sub[2]), _ptr(sidx), _ptr(m_dev), _ptr(ws), ws.numel(), _stream()))
    return sub, sidx, m_dev


def radius_outlier_mask_sharded(x, y, z, rank: int, world: int, group=None, radius=RADIUS, nb_points=NB_POINTS,
                                bounds=None, reduce=True):
    """A1 over several GPUs: rank r filters the points of its X slab, with the points within one grid
    cell of the slab as halo (every neighbour within `radius` is inside it); the cloud-sized masks
    (zero outside the own slab) are combined with one MAX all-reduce.  Bit-identical to
    radius_outlier_mask.  reduce=False returns the rank's partial mask (tests)."""
    lib = _lib.load()
    if world <= 1:
        return radius_outlier_mask(x, y, z, radius, nb_points)
    dev = x.device
    n = int(x.shape[0])
    if bounds is None:
        bounds = slab_bounds(x, world)
    lo, hi = float(bounds[rank]), float(bounds[rank + 1])
    halo = float(radius) * (1.0 + 1.0 / 1024.0)
    keep = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
    sub, sidx, m_dev = slab_select(x, y, z, lo - halo, hi + halo)
    m = int(m_dev.item())
    if m > 0:
        skeep, _ = radius_outlier_mask(sub[0, :m], sub[1, :m], sub[2, :m], radius, nb_points)
        check(lib.ptgs_slab_keep(_dt(x), _ptr(sub[0]), _ptr(sidx), _ptr(skeep), m, lo, hi, _ptr(keep), _stream()))
    keep = keep[:n]
    if reduce:
        import torch.distributed as dist
        dist.all_reduce(keep, op=dist.ReduceOp.MAX, group=group)
    n_keep = keep.sum(dtype=torch.int64).reshape(1)
    return keep, n_keep


def compact_xy(x, y, keep):
    lib = _lib.load()
    n = x.shape[0]
    ox = torch.empty(n, dtype=x.dtype, device=x.device)
    oy = torch.empty(n, dtype=x.dtype, device=x.device)
    n_out = torch.zeros(1, dtype=torch.int64, device=x.device)
    ws = _ws(lib.ptgs_compact_xy_workspace_bytes(n), x.device)
    check(lib.ptgs_compact_xy(_dt(x), _ptr(x), _ptr(y), _ptr(keep), n, _ptr(ox), _ptr(oy),
                              _ptr(n_out), _ptr(ws), ws.numel(), _stream()))
    return ox, oy, n_out


def bounds_xy(x, y, n_dev):
    lib = _lib.load()
    b = torch.empty(4, dtype=torch.float64, device=x.device)
    check(lib.ptgs_bounds_xy(_dt(x), _ptr(x), _ptr(y), _ptr(n_dev), x.shape[0], _ptr(b), _stream()))
    return b


def balanced_split(x, y, n_dev, root_bounds, threshold, max_depth=MAX_DEPTH, capacity_factor=1.25):
    """A3.  Returns host arrays (bounds (L,4) f64, path, depth, count)."""
    lib = _lib.load()
    dev = x.device
    ncap = x.shape[0]
    max_leaves = 1 << max_depth
    while True:
        lb = torch.empty((max_leaves, 4), dtype=torch.float64, device=dev)
        lp = torch.empty(max_leaves, dtype=torch.int32, device=dev)
        ld = torch.empty(max_leaves, dtype=torch.int32, device=dev)
        lc = torch.empty(max_leaves, dtype=torch.int64, device=dev)
        nl = torch.zeros(1, dtype=torch.int32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        ws = _ws(lib.ptgs_balanced_split_workspace_bytes(_dt(x), ncap, capacity_factor, max_depth), dev)
        check(lib.ptgs_balanced_split(_dt(x), _ptr(x), _ptr(y), _ptr(n_dev), ncap, _ptr(root_bounds),
                                      float(threshold * LEAF_FACTOR), int(max_depth),
                                      float(capacity_factor), _ptr(lb), _ptr(lp), _ptr(ld), _ptr(lc),
                                      _ptr(nl), max_leaves, _ptr(status), _ptr(ws), ws.numel(),
                                      _stream()))
        st = int(status.item())
        if st == _lib.PTGS_E_CAPACITY and capacity_factor < 64:
            capacity_factor *= 2          # many points on split lines: duplicates grow the level
            continue
        if st != 0:
            raise _lib.PtgsError(st, "balanced_split device status")
        L = int(nl.item())
        return (lb[:L].cpu().numpy(), lp[:L].cpu().numpy().astype(np.uint32),
                ld[:L].cpu().numpy(), lc[:L].cpu().numpy())


GRID_MIN_BLOCKS = 33          # below this the lock-step kernel (every point x every rectangle) is as fast


def _cell_grid(rects: np.ndarray):
    """Uniform G x G cell grid over the rectangles {minx,maxx,miny,maxy} (float64 view of the values the
    kernel compares with) -> (grid[5], cell_off int32[G*G+1], cell_list int32) or None when the rectangles
    overlap so much that the grid would not prune.  The cell of a coordinate is floor((v - x0) * inv) in
    float64, evaluated identically on the device, and rounding is monotone, so a point inside a rectangle
    always falls into the rectangle's cell range; one extra cell on every side is kept anyway."""
    L = len(rects)
    r = rects.astype(np.float64)
    if not np.isfinite(r).all():
        return None
    x0, x1, y0, y1 = r[:, 0].min(), r[:, 1].max(), r[:, 2].min(), r[:, 3].max()
    if not (x1 > x0 and y1 > y0):
        return None
    G = int(min(256, max(16, 1 << int(np.ceil(np.log2(4.0 * np.sqrt(L)))))))
    inv_x, inv_y = G / (x1 - x0), G / (y1 - y0)
    cl = lambda v: np.clip(v, 0, G - 1).astype(np.int64)
    ix0, ix1 = cl(np.floor((r[:, 0] - x0) * inv_x) - 1), cl(np.floor((r[:, 1] - x0) * inv_x) + 1)
    iy0, iy1 = cl(np.floor((r[:, 2] - y0) * inv_y) - 1), cl(np.floor((r[:, 3] - y0) * inv_y) + 1)
    n_cells = (ix1 - ix0 + 1) * (iy1 - iy0 + 1)
    if int(n_cells.sum()) > 16 * G * G:
        return None
    # (cell, rectangle) pairs, vectorised: rectangle b covers w[b] x h[b] cells
    w = ix1 - ix0 + 1
    start = np.cumsum(n_cells) - n_cells
    owners = np.repeat(np.arange(L, dtype=np.int32), n_cells)
    local = np.arange(int(n_cells.sum()), dtype=np.int64) - np.repeat(start, n_cells)
    ww = np.repeat(w, n_cells)
    cells = (np.repeat(iy0, n_cells) + local // ww) * G + np.repeat(ix0, n_cells) + local % ww
    order = np.argsort(cells, kind="stable")          # rectangles stay ascending inside a cell
    cell_off = np.zeros(G * G + 1, np.int32)
    np.cumsum(np.bincount(cells, minlength=G * G), out=cell_off[1:])
    grid = np.array([x0, y0, inv_x, inv_y, float(G)], np.float64)
    return grid, cell_off, np.ascontiguousarray(owners[order])


def expand_select(x, y, boxes: np.ndarray, use_grid=None, want_counts=False):
    """A4.  boxes (L,4) float64 {minx,miny,maxx,maxy}; the bounds are rounded
    to the cloud's dtype exactly as NumPy does for `f32_array >= python_float`
    (shen_partation_densify.py:201-202).  With many blocks the selection goes through a
    cell grid (ptgs_expand_*_grid, identical results); use_grid forces one path."""
    import ctypes as C
    lib = _lib.load()
    dev = x.device
    n, L = x.shape[0], len(boxes)
    npdt = np.float64 if x.dtype == torch.float64 else np.float32
    rects = np.stack([boxes[:, 0], boxes[:, 2], boxes[:, 1], boxes[:, 3]], axis=1).astype(npdt)
    rects_d = torch.from_numpy(np.ascontiguousarray(rects)).to(dev)
    counts = torch.empty(L, dtype=torch.int64, device=dev)
    ws = _ws(lib.ptgs_expand_workspace_bytes(n, L), dev)
    cg = None
    if use_grid or (use_grid is None and L >= GRID_MIN_BLOCKS):
        cg = _cell_grid(rects)
    if cg is not None:
        grid, cell_off, cell_list = cg
        gp = grid.ctypes.data_as(C.POINTER(C.c_double))
        off_d = torch.from_numpy(cell_off).to(dev)
        lst_d = torch.from_numpy(cell_list if len(cell_list) else np.zeros(1, np.int32)).to(dev)
        check(lib.ptgs_expand_count_grid(_dt(x), _ptr(x), _ptr(y), n, _ptr(rects_d), L, gp, _ptr(off_d), _ptr(lst_d),
                                         _ptr(counts), _ptr(ws), ws.numel(), _stream()))
    else:
        check(lib.ptgs_expand_count(_dt(x), _ptr(x), _ptr(y), n, _ptr(rects_d), L, _ptr(counts),
                                    _ptr(ws), ws.numel(), _stream()))
    counts_h = counts.cpu().numpy()               # the one host read of the stage (L values)
    total = int(counts_h.sum())
    offsets = torch.empty(L + 1, dtype=torch.int64, device=dev)
    indices = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    if cg is not None:
        check(lib.ptgs_expand_fill_grid(_dt(x), _ptr(x), _ptr(y), n, _ptr(rects_d), L, gp, _ptr(off_d), _ptr(lst_d),
                                        _ptr(counts), _ptr(offsets), _ptr(indices), total, _ptr(ws), ws.numel(),
                                        _stream()))
    else:
        check(lib.ptgs_expand_fill(_dt(x), _ptr(x), _ptr(y), n, _ptr(rects_d), L, _ptr(counts),
                                   _ptr(offsets), _ptr(indices), total, _ptr(ws), ws.numel(), _stream()))
    if want_counts:
        return offsets, indices[:total], counts_h
    return offsets, indices[:total]


def assign_cameras(centers: np.ndarray, boxes: np.ndarray, device, want_host=False):
    """A5.  centers (C,3) (float32 values widened to float64), boxes (L,4) f64."""
    lib = _lib.load()
    Cn, L = len(centers), len(boxes)
    cen = torch.from_numpy(np.ascontiguousarray(centers, dtype=np.float64)).to(device)
    bx = torch.from_numpy(np.ascontiguousarray(boxes, dtype=np.float64)).to(device)
    offsets = torch.empty(L + 1, dtype=torch.int64, device=device)
    cap = max(Cn * L, 1)
    index = torch.empty(cap, dtype=torch.int32, device=device)
    check(lib.ptgs_assign_cameras(_ptr(cen), Cn, _ptr(bx), L, _ptr(offsets), _ptr(index), cap, _stream()))
    off_h = offsets.cpu().numpy()                 # the one host read of the stage (L + 1 values)
    total = int(off_h[L])
    if want_host:
        return offsets, index[:total].clone(), off_h
    return offsets, index[:total].clone()


def voxel_downsample(x, y, z, offsets, indices, voxel=VOXEL_SIZE, n_entries=None, sync=True, xyz_aos=None):
    """A6 over every block at once.  Voxels of a block come out in ascending
    key order; ds_first gives the reference's first-occurrence order.  xyz_aos: the cloud as (N,3)
    rows of the planes' dtype (optional; makes the entry gather of an index-list call cheaper).
    With xyz_aos the planes may be None (the rows are then the only source)."""
    lib = _lib.load()
    dev = offsets.device
    L = offsets.shape[0] - 1
    ne = indices.shape[0] if indices is not None else int(n_entries)
    pitch = (ne + 8 + 3) // 4 * 4
    buf = torch.zeros((3, pitch), dtype=torch.float32, device=dev)
    ds_first = torch.empty(max(ne, 1), dtype=torch.int32, device=dev)
    ds_off = torch.empty(L + 1, dtype=torch.int64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ws = _ws(lib.ptgs_voxel_downsample_workspace_bytes(ne, L), dev)
    if xyz_aos is not None:
        assert xyz_aos.is_cuda and xyz_aos.is_contiguous() and xyz_aos.shape[1] == 3
        assert x is None or (xyz_aos.dtype == x.dtype and xyz_aos.shape[0] == x.shape[0])
    check(lib.ptgs_voxel_downsample(_dt(x if x is not None else xyz_aos), _ptr(x), _ptr(y), _ptr(z), _ptr(xyz_aos), _ptr(offsets), _ptr(indices),
                                    ne, L, float(voxel), _ptr(buf[0]), _ptr(buf[1]), _ptr(buf[2]),
                                    _ptr(ds_first), _ptr(ds_off), _ptr(status), _ptr(ws), ws.numel(),
                                    _stream()))
    if not sync:
        # no host round trip: full-capacity planes (ds_off bounds every later read), status unchecked
        return ds_off, buf[0], buf[1], buf[2], ds_first
    m = int(ds_off[L].item())
    return ds_off, buf[0, :m], buf[1, :m], buf[2, :m], ds_first[:m]


def camera_records(R, T, FoVx, FoVy, width, height, front_axis=1) -> np.ndarray:
    """(C,20) float32 camera records (include/ptgs.h) from the SimpleCamera
    fields, using the same Python expressions as shen_partation_densify.py:
    433-434 (float32 casts of R, T) and :457-461 (fx, fy, cx, cy in Python
    floats, then float32)."""
    import math
    Cn = len(R)
    rec = np.zeros((Cn, CAMERA_FLOATS), dtype=np.float32)
    rec[:, 0:9] = np.asarray(R, dtype=np.float64).reshape(Cn, 9).astype(np.float32)
    rec[:, 9:12] = np.asarray(T, dtype=np.float64).reshape(Cn, 3).astype(np.float32)
    for i in range(Cn):
        w, h = int(width[i]), int(height[i])
        rec[i, 12] = np.float32(w / (2 * math.tan(float(FoVx[i]) / 2)))
        rec[i, 13] = np.float32(h / (2 * math.tan(float(FoVy[i]) / 2)))
        rec[i, 14] = np.float32(w / 2)
        rec[i, 15] = np.float32(h / 2)
        rec[i, 16] = np.float32(w)
        rec[i, 17] = np.float32(h)
    rec[:, 18] = np.float32(front_axis)
    return rec


def pick_cam_split(n_points_total: int, n_blocks: int) -> int:
    """Camera-list split factor of the sweep: enough work items for 148 SMs x 2
    CTAs when there are few point tiles.  Results do not depend on it."""
    import os
    if os.environ.get("PTGS_CAM_SPLIT"):          # diagnostics: results do not depend on it
        return int(os.environ["PTGS_CAM_SPLIT"])
    tiles = max(1, (int(n_points_total) + 2047) // 2048 + n_blocks)
    want = 148 * 2 * 4
    return max(1, min(64, (want + tiles - 1) // tiles))


def visibility(ds_x, ds_y, ds_z, ds_offsets, cameras_d, cam_offsets, cam_index,
               threshold=VISIBLE_RATE, min_hits=MIN_HITS, near_eps=NEAR_EPS,
               cand_capacity=2048, cam_split=None, timer=None, exact_ratio=False,
               cand_pool=None, interior=True, sample_stride=SAMPLE_STRIDE,
               defer_check=False, hull_prefilter=True) -> VisibilityResult:
    """A7-A9 for every (block, camera) pair.

    sweep 1 (hit counts, pivot, 8 extremes) -> area bounds + early decision ->
    for the undecided pairs only: sweep 2 (hull candidates) + exact hull area.
    exact_ratio=True sends every pair with > min_hits hits through the exact
    hull (ratio is then the reference's value for every pair; otherwise decided
    pairs carry the deciding bound and flags bit4).  One host synchronisation
    at the end (candidate-capacity check).

    interior=True first runs sweep 1 on a 1-in-`sample_stride` sample of every
    block and derives, per pair, a rectangle strictly inside the convex hull
    of the sample's hits; the full sweep only counts the points that project
    into it (no divisions, no extreme tracking).  Results are identical."""
    lib = _lib.load()
    dev = cameras_d.device
    L = ds_offsets.shape[0] - 1
    P = int(cam_index.shape[0])
    Pa = max(P, 1)
    n_hit = torch.empty(Pa, dtype=torch.int32, device=dev)
    area = torch.empty(Pa, dtype=torch.float64, device=dev)
    ratio = torch.empty(Pa, dtype=torch.float64, device=dev)
    keep = torch.empty(Pa, dtype=torch.uint8, device=dev)
    flags = torch.empty(Pa, dtype=torch.uint8, device=dev)
    cand_count = torch.empty(Pa, dtype=torch.int32, device=dev)
    if P == 0:
        return VisibilityResult(n_hit[:0], area[:0], ratio[:0], keep[:0], flags[:0], cand_count[:0], 0)
    pivot = torch.empty(P, dtype=torch.int64, device=dev)
    extreme = torch.empty(P * _lib.EXTREME_DIRS, dtype=torch.int64, device=dev)
    ratio_lo = torch.empty(P, dtype=torch.float64, device=dev)
    ratio_hi = torch.empty(P, dtype=torch.float64, device=dev)
    need = torch.empty(P, dtype=torch.uint8, device=dev)
    poly = torch.empty((P, _lib.POLY_FLOATS), dtype=torch.float32, device=dev)
    if cam_split is None:
        cam_split = pick_cam_split(int(ds_x.shape[0]), L)
    ws = _ws(lib.ptgs_visibility_workspace_bytes(L), dev)
    st = _stream()
    mark = timer.mark if timer is not None else (lambda name: None)
    check(lib.ptgs_visibility_init(P, _ptr(n_hit), _ptr(pivot), _ptr(extreme), _ptr(cand_count), st))
    mark("vis_init")
    krect = None
    if interior:
        n_pts = int(ds_x.shape[0])
        cap = (n_pts // int(sample_stride) + L + 8 + 3) & ~3   # >= sum ceil(M_b / stride); rows stay 16 B aligned
        smp = torch.empty((3, cap + 8), dtype=torch.float32, device=dev)
        s_off = torch.empty(L + 1, dtype=torch.int64, device=dev)
        n_hit_s = torch.empty(P, dtype=torch.int32, device=dev)
        pivot_s = torch.empty(P, dtype=torch.int64, device=dev)
        extreme_s = torch.empty(P * _lib.EXTREME_DIRS, dtype=torch.int64, device=dev)
        krect = torch.empty((P, 32), dtype=torch.float32, device=dev)     # PTGS_KRECT_FLOATS
        check(lib.ptgs_visibility_sample(_ptr(ds_x), _ptr(ds_y), _ptr(ds_z), _ptr(ds_offsets), L,
                                         int(sample_stride), _ptr(smp[0]), _ptr(smp[1]), _ptr(smp[2]),
                                         _ptr(s_off), cap, st))
        if _os.environ.get("PTGS_SAMPLE_PASS") == "sweep":      # A/B: the tile sweep on the sample
            check(lib.ptgs_visibility_init(P, _ptr(n_hit_s), _ptr(pivot_s), _ptr(extreme_s), None, st))
            check(lib.ptgs_visibility_count(_ptr(smp[0]), _ptr(smp[1]), _ptr(smp[2]), _ptr(s_off), L,
                                            _ptr(cameras_d), _ptr(cam_offsets), _ptr(cam_index),
                                            pick_cam_split(cap, L), _ptr(n_hit_s), _ptr(pivot_s),
                                            _ptr(extreme_s), None, _ptr(ws), ws.numel(), st))
        else:                                                    # one warp per pair, no atomics
            check(lib.ptgs_visibility_sample_count(_ptr(smp[0]), _ptr(smp[1]), _ptr(smp[2]), _ptr(s_off), L,
                                                   _ptr(cameras_d), _ptr(cam_offsets), _ptr(cam_index), P,
                                                   _ptr(n_hit_s), _ptr(pivot_s), _ptr(extreme_s), st))
        check(lib.ptgs_visibility_krect(_ptr(smp[0]), _ptr(smp[1]), _ptr(smp[2]), _ptr(s_off), L,
                                        _ptr(cameras_d), _ptr(cam_offsets), _ptr(cam_index),
                                        _ptr(n_hit_s), _ptr(extreme_s), P, float(KRECT_MARGIN_PX),
                                        _ptr(krect), st))
        mark("vis_interior_rects")
    check(lib.ptgs_visibility_count(_ptr(ds_x), _ptr(ds_y), _ptr(ds_z), _ptr(ds_offsets), L,
                                    _ptr(cameras_d), _ptr(cam_offsets), _ptr(cam_index), cam_split,
                                    _ptr(n_hit), _ptr(pivot), _ptr(extreme), _ptr(krect), _ptr(ws),
                                    ws.numel(), st))
    mark("vis_sweep1_count")
    check(lib.ptgs_visibility_bounds(_ptr(ds_x), _ptr(ds_y), _ptr(ds_z), _ptr(ds_offsets), L,
                                     _ptr(cameras_d), _ptr(cam_offsets), _ptr(cam_index), _ptr(n_hit),
                                     _ptr(pivot), _ptr(extreme), P, int(min_hits), float(threshold),
                                     float(near_eps), 1 if exact_ratio else 0, _ptr(poly), _ptr(ratio_lo),
                                     _ptr(ratio_hi), _ptr(need), _ptr(area), _ptr(ratio), _ptr(keep),
                                     _ptr(flags), st))
    mark("vis_bounds")
    sub_off = torch.empty(L + 1, dtype=torch.int64, device=dev)
    sub_idx = torch.empty(P, dtype=torch.int32, device=dev)
    sub_pos = torch.empty(P, dtype=torch.int32, device=dev)
    cand_off = torch.empty(P + 1, dtype=torch.int64, device=dev)
    hull_status = torch.empty(2, dtype=torch.int32, device=dev)
    hull_ws = _ws(lib.ptgs_visibility_hull_workspace_bytes(P), dev) if hull_prefilter else None
    caps_in = None
    rounds = 0
    n_exact = None
    while True:
        rounds += 1
        # sub-list of the pairs that need the exact hull (device side, no sync)
        check(lib.ptgs_select_cameras(_ptr(need), _ptr(cam_offsets), _ptr(cam_index), L, _ptr(sub_off),
                                      _ptr(sub_idx), _ptr(sub_pos), st))
        if caps_in is None:
            # fixed pool (no sync): pairs that do not fit overflow and are re-run
            # 8 bytes per slot; address space rather than traffic (only emitted candidates are touched)
            floor = (32 << 20) if int(ds_x.shape[0]) < (20 << 20) else (128 << 20)
            total = max(P * 256, floor) if cand_pool is None else int(cand_pool)
        else:
            total = int(caps_in[need.bool()].sum().item())
        check(lib.ptgs_visibility_plan(_ptr(n_hit), _ptr(need), _ptr(caps_in), _ptr(poly), _ptr(ratio_lo),
                                       _ptr(ratio_hi), P, int(cand_capacity), total, _ptr(cand_off), st))
        cand_uv = torch.empty((max(total, 1), 2), dtype=torch.float32, device=dev)
        check(lib.ptgs_visibility_init(P, None, None, None, _ptr(cand_count), st))
        mark("vis_plan")
        check(lib.ptgs_visibility_candidates(_ptr(ds_x), _ptr(ds_y), _ptr(ds_z), _ptr(ds_offsets), L,
                                             _ptr(cameras_d), _ptr(sub_off), _ptr(sub_idx), cam_split,
                                             _ptr(sub_pos), _ptr(poly), _ptr(cand_off), _ptr(cand_uv),
                                             _ptr(cand_count), _ptr(ws), ws.numel(), st))
        mark("vis_sweep2_candidates")
        check(lib.ptgs_visibility_hull(_ptr(cand_uv), _ptr(cand_off), _ptr(cand_count), _ptr(n_hit), P,
                                       _ptr(cam_index), _ptr(cameras_d), int(min_hits),
                                       float(threshold), float(near_eps), _ptr(area), _ptr(ratio),
                                       _ptr(keep), _ptr(flags), _ptr(sub_pos), _ptr(sub_off[L:]),
                                       _ptr(hull_status), _ptr(hull_ws), hull_ws.numel() if hull_ws is not None else 0,
                                       st))
        mark("vis_hull")
        # the one synchronisation: [overflowed pairs, hull budget exceeded, exact pairs]
        stat_d = torch.cat([hull_status.to(torch.int64), sub_off[L:L + 1]])
        if defer_check and rounds == 1:
            # the caller reads `pending` later (one synchronisation for several calls) and
            # re-runs this call without defer_check when pending[0] != 0
            res = VisibilityResult(n_hit, area, ratio, keep, flags, cand_count, rounds)
            res.ratio_lo, res.ratio_hi, res.n_exact, res.pending = ratio_lo, ratio_hi, None, stat_d
            mark("vis_check")
            return res
        stat = stat_d.cpu()
        if n_exact is None:
            n_exact = int(stat[2])
        if int(stat[1]):
            raise RuntimeError("visibility: convex hull exceeds the kernel's vertex budget")
        if int(stat[0]) == 0:
            break
        if rounds > 3:
            raise RuntimeError("visibility: candidate capacity did not converge")
        # re-run only the overflowed pairs with their exact candidate counts
        need = (need.bool() & ((flags & 4) != 0)).to(torch.uint8)
        caps_in = cand_count.clone()
    mark("vis_check")
    res = VisibilityResult(n_hit, area, ratio, keep, flags, cand_count, rounds)
    res.ratio_lo, res.ratio_hi, res.n_exact = ratio_lo, ratio_hi, n_exact
    return res


def select_cameras(keep, cam_offsets, cam_index, want_pos=False, sync=True):
    """A9: per block compaction of the kept cameras (device CSR)."""
    lib = _lib.load()
    dev = cam_offsets.device
    L = cam_offsets.shape[0] - 1
    kept_off = torch.empty(L + 1, dtype=torch.int64, device=dev)
    kept_idx = torch.empty(max(int(cam_index.shape[0]), 1), dtype=torch.int32, device=dev)
    kept_pos = torch.empty_like(kept_idx) if want_pos else None
    check(lib.ptgs_select_cameras(_ptr(keep), _ptr(cam_offsets), _ptr(cam_index), L, _ptr(kept_off),
                                  _ptr(kept_idx), _ptr(kept_pos), _stream()))
    if not sync:
        return (kept_off, kept_idx, kept_pos) if want_pos else (kept_off, kept_idx)
    n = int(kept_off[L].item())
    return (kept_off, kept_idx[:n], kept_pos[:n]) if want_pos else (kept_off, kept_idx[:n])


# per super-tile camera lists in ptgs_coverage_sweep (selected by the workspace size)
COVERAGE_SUPERTILES = True


def spatial_order(px, py, pz, xyz_aos=None):
    """Morton-ordered copy of a planar float32 cloud + the permutation
    (sorted[i] = input[perm[i]]); ptgs_spatial_order.  xyz_aos: the same cloud as (N,3) float32
    rows (optional, cheaper gather).  The sorted planes carry 16 bytes of tail padding: the sweep's
    bulk copies round the last tile's window up to 16 bytes."""
    lib = _lib.load()
    dev = px.device
    n = px.shape[0]
    if xyz_aos is not None:
        assert xyz_aos.dtype == torch.float32 and xyz_aos.is_contiguous() and xyz_aos.shape[0] == n
    pitch = (n + 8 + 3) // 4 * 4
    planes = torch.zeros((3, pitch), dtype=torch.float32, device=dev)
    sx, sy, sz = planes[0], planes[1], planes[2]
    perm = torch.empty(max(n, 1), dtype=torch.int32, device=dev)       # uint32 payload
    ws = _ws(lib.ptgs_spatial_order_workspace_bytes(n), dev)
    check(lib.ptgs_spatial_order(_ptr(px), _ptr(py), _ptr(pz), _ptr(xyz_aos), n, _ptr(sx), _ptr(sy), _ptr(sz), _ptr(perm),
                                 _ptr(ws), ws.numel(), _stream()))
    return sx[:n], sy[:n], sz[:n], perm[:n]


def coverage_sweep(px, py, pz, cameras_d, cam_first=0, cam_count=None, want_seen=False,
                   spatial_order_points=True, ordered=None):
    """Config 5 / Vast-style coverage (data_partition.py:385-424, :505): every point against
    cameras [cam_first, cam_first + cam_count).  The cloud is first put into Morton order
    (`spatial_order`) so that the sweep can cull cameras per tile; `ordered` takes the result of
    an earlier `spatial_order(px, py, pz)` call to reuse it across camera ranges.  Results do
    not depend on the order (seen flags come back in the caller's point order)."""
    lib = _lib.load()
    dev = px.device
    if cam_count is None:
        cam_count = cameras_d.shape[0] - cam_first
    n = px.shape[0]
    perm = None
    if ordered is not None:
        px, py, pz, perm = ordered
    elif spatial_order_points and n > 0:
        px, py, pz, perm = spatial_order(px, py, pz)
    n_hit = torch.empty(max(cam_count, 1), dtype=torch.int32, device=dev)
    seen = torch.empty(max(n, 1), dtype=torch.uint8, device=dev) if want_seen else None
    ws = _ws(lib.ptgs_coverage_workspace_bytes(n, int(cam_count)) if COVERAGE_SUPERTILES
             else lib.ptgs_visibility_workspace_bytes(1), dev)
    check(lib.ptgs_coverage_sweep(_ptr(px), _ptr(py), _ptr(pz), n, _ptr(cameras_d), int(cam_first),
                                  int(cam_count), _ptr(n_hit), _ptr(seen), _ptr(ws), ws.numel(), _stream()))
    if want_seen and perm is not None and n > 0:
        out = torch.empty_like(seen)
        check(lib.ptgs_unpermute_u8(_ptr(seen), _ptr(perm), n, _ptr(out), _stream()))
        seen = out
    return (n_hit[:cam_count], seen[:n]) if want_seen else n_hit[:cam_count]


def coverage_sweep_sharded(px, py, pz, cameras_d, rank: int, world: int, group=None, want_seen=False,
                           shard="points", ordered=None):
    """Config c5 over several GPUs.  Every rank holds the whole cloud and all cameras.
    shard="slabs": rank r orders and sweeps the X slab of the cloud between two quantiles of x
    (slab_bounds / ptgs_slab_select) against ALL cameras; counts are one int32 sum all-reduce, seen
    flags one MAX all-reduce.
    shard="points" (default): rank r orders and sweeps only its slice of the cloud
    (multigpu.point_range) against ALL cameras -- the Morton sort and the per-tile camera culls,
    which dominate after the sort, both divide by the number of ranks; per-camera counts are one
    int32 sum all-reduce, per-point seen flags an all-gather of the slices.
    shard="cameras": rank r sweeps the whole cloud against its camera range (SURVEY's layout; the
    sort is then repeated on every rank).  Returns the full result on every rank."""
    from . import multigpu as M
    C = cameras_d.shape[0]
    if shard == "cameras":
        first, cnt = M.camera_range(C, rank, world)
        out = coverage_sweep(px, py, pz, cameras_d, cam_first=first, cam_count=cnt, want_seen=want_seen,
                             ordered=ordered)
        n_hit, seen = out if want_seen else (out, None)
        n_hit = M.gather_camera_counts(n_hit, C, world, group)
        return (n_hit, M.merge_seen(seen, world, group)) if want_seen else n_hit
    n = px.shape[0]
    if shard == "slabs":
        # rank r takes the X slab [bounds[r], bounds[r+1]) of the cloud: a spatially compact n / world
        # points, so tiles and warp runs stay as small as on one GPU (a file-order slice is a random
        # subset of the whole scene: its tiles cover `world` times the area and are cut by more cameras)
        import torch.distributed as dist
        bounds = slab_bounds(px, world)
        sub, sidx, m_dev = slab_select(px, py, pz, bounds[rank], bounds[rank + 1])
        m = int(m_dev.item())
        if m > 0:
            out = coverage_sweep(sub[0, :m], sub[1, :m], sub[2, :m], cameras_d, want_seen=want_seen)
            n_hit, seen_l = out if want_seen else (out, None)
        else:
            n_hit, seen_l = torch.zeros(C, dtype=torch.int32, device=px.device), None
        n_hit = M.sum_camera_counts(n_hit.contiguous(), world, group)
        if not want_seen:
            return n_hit
        seen = torch.zeros(max(n, 1), dtype=torch.uint8, device=px.device)
        if m > 0:
            seen[sidx[:m].to(torch.int64)] = seen_l
        dist.all_reduce(seen, op=dist.ReduceOp.MAX, group=group)
        return n_hit, seen[:n]
    p0, p1 = M.point_range(n, rank, world)
    out = coverage_sweep(px[p0:p1], py[p0:p1], pz[p0:p1], cameras_d, want_seen=want_seen)
    n_hit, seen = out if want_seen else (out, None)
    if p1 <= p0:
        n_hit = torch.zeros(C, dtype=torch.int32, device=px.device)
    n_hit = M.sum_camera_counts(n_hit.contiguous(), world, group)
    return (n_hit, M.gather_seen(seen, n, world, group)) if want_seen else n_hit


# ---------------------------------------------------------------- A14: the camera-grid (Vast) path
VAST_CAMERA_DOUBLES = 20


def vast_camera_records(R, T, FoVx, FoVy, width, height) -> np.ndarray:
    """(C,20) float64 records for the ptgs_vast_* entry points, with the Python expressions of
    point_in_image (scene/ptgs/data_partition.py:388-407): w2c = R^T, fx = W / (2 tan(FoVx/2)),
    cx = W // 2, cy = H // 2."""
    import math
    Cn = len(R)
    rec = np.zeros((Cn, VAST_CAMERA_DOUBLES), dtype=np.float64)
    for i in range(Cn):
        w, h = int(width[i]), int(height[i])
        rec[i, 0:9] = np.transpose(np.asarray(R[i], dtype=np.float64)).reshape(9)
        rec[i, 9:12] = np.asarray(T[i], dtype=np.float64)
        rec[i, 12] = w / (2 * math.tan(float(FoVx[i]) / 2))
        rec[i, 13] = h / (2 * math.tan(float(FoVy[i]) / 2))
        rec[i, 14] = w // 2
        rec[i, 15] = h // 2
        rec[i, 16] = w
        rec[i, 17] = h
    return rec


def vast_point_range(x, y, z, offsets, indices) -> np.ndarray:
    """get_point_range (data_partition.py:285-293) of every block of a CSR -> (B,6) float64 on the host."""
    lib = _lib.load()
    dev = x.device
    B = offsets.shape[0] - 1
    out = torch.empty((B, 6), dtype=torch.float64, device=dev)
    ws = _ws(B * 48, dev)
    check(lib.ptgs_vast_point_range(_dt(x), _ptr(x), _ptr(y), _ptr(z), _ptr(offsets), _ptr(indices), B,
                                    _ptr(out), _ptr(ws), ws.numel(), _stream()))
    return out.cpu().numpy()


def vast_airspace(cams_d, boxes: np.ndarray, pair_block: np.ndarray, pair_cam: np.ndarray) -> np.ndarray:
    """8-corner visibility rate of every (block box, camera) pair (data_partition.py:478-488)."""
    lib = _lib.load()
    dev = cams_d.device
    P = len(pair_block)
    if P == 0:
        return np.zeros(0)
    boxes_d = torch.from_numpy(np.ascontiguousarray(boxes, dtype=np.float64)).to(dev)
    pb = torch.from_numpy(np.ascontiguousarray(pair_block, dtype=np.int32)).to(dev)
    pc = torch.from_numpy(np.ascontiguousarray(pair_cam, dtype=np.int32)).to(dev)
    rate = torch.empty(P, dtype=torch.float64, device=dev)
    check(lib.ptgs_vast_airspace(_ptr(cams_d), _ptr(boxes_d), _ptr(pb), _ptr(pc), P, _ptr(rate), _stream()))
    return rate.cpu().numpy()


def vast_block_cloud(x, y, z, cams_d, own_index, sources):
    """Coverage-based point selection + row dedup for one destination block
    (data_partition.py:441-557).  own_index: int32 cloud indices of the block's own (extended)
    cloud; sources: [(block_index int32 tensor, [camera ids accepted from that block, in order])]
    in the reference's block order.  Returns the cloud indices of the unique rows
    (lexicographic row order, first occurrence) as an int32 tensor."""
    lib = _lib.load()
    dev = x.device
    n = x.shape[0]
    first_pos = torch.empty(max(n, 1), dtype=torch.int64, device=dev)      # uint64 payload
    check(lib.ptgs_vast_begin_block(_ptr(first_pos), n, _stream()))
    seq = 0
    for idx_j, cam_ids in sources:
        k = len(cam_ids)
        if k == 0 or idx_j.shape[0] == 0:
            seq += k
            continue
        cl = torch.from_numpy(np.ascontiguousarray(cam_ids, dtype=np.int32)).to(dev)
        ws = _ws(lib.ptgs_vast_coverage_workspace_bytes(idx_j.shape[0], k), dev)
        check(lib.ptgs_vast_coverage(_dt(x), _ptr(x), _ptr(y), _ptr(z), _ptr(idx_j), idx_j.shape[0], _ptr(cams_d),
                                     _ptr(cl), k, seq, _ptr(first_pos), _ptr(ws), ws.numel(), _stream()))
        seq += k
    check(lib.ptgs_vast_mark_own(_ptr(own_index), own_index.shape[0], seq, _ptr(first_pos), _stream()))
    uniq = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    n_uniq = torch.zeros(1, dtype=torch.int64, device=dev)
    ws = _ws(lib.ptgs_vast_unique_workspace_bytes(n), dev)
    check(lib.ptgs_vast_unique(_dt(x), _ptr(x), _ptr(y), _ptr(z), n, _ptr(first_pos), _ptr(uniq), n,
                               _ptr(n_uniq), _ptr(ws), ws.numel(), _stream()))
    return uniq[:int(n_uniq.item())]


class _Timer:
    """CUDA-event stage timer on the current stream."""

    def __init__(self, enabled):
        self.enabled = enabled
        self.marks = []

    def mark(self, name):
        if self.enabled:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.marks.append((name, e))

    def result(self):
        if not self.enabled or len(self.marks) < 2:
            return {}
        torch.cuda.synchronize()
        out = {}
        for (n0, e0), (n1, e1) in zip(self.marks[:-1], self.marks[1:]):
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        out["total"] = self.marks[0][1].elapsed_time(self.marks[-1][1])
        return out


def partition_device(xyz: torch.Tensor, cam_records: np.ndarray, cam_centers: np.ndarray,
                     threshold: int, man_trans: np.ndarray | None = None, radius_filter: bool = True,
                     do_visibility: bool = True, time_stages: bool = False,
                     shard: tuple | None = None, split_xyz: torch.Tensor | None = None,
                     exact_ratio: bool = False, host_results: bool = False) -> PartitionDeviceResult:
    """run_DataPartition (shen_partition_utils.py:57-93) on the device.

    xyz         (N,3) float32 CUDA tensor: the scene cloud as fetchPly returns it.
    cam_records (C,20) float32 host array (camera_records()).
    cam_centers (C,3) host array, SimpleCamera.camera_center (float32 values).
    man_trans   optional 4x4: the cloud is rotated first (float64 from there on);
                the caller passes camera records / centres already in that frame.
    split_xyz   optional (N,3) float32 tensor used for A1-A3 instead of `xyz`
                (the reference's literal shen-path behaviour under man_trans:
                the split reads the un-rotated PLY, shen_partition_utils.py:62).
    shard       (rank, world[, group]): multi-GPU.  A1 is sharded by X slabs with a one-cell halo
                (one MAX all-reduce of the keep mask), A2-A5 are computed by every rank
                (deterministic, no communication); the visibility stage is range-sharded: rank r
                owns a contiguous range of the (block, camera) pair list of about 1 / world of the
                cost (multigpu.RangePlan on the expanded point counts), down-samples (A6) and
                sweeps (A7-A8) only the blocks that range touches, and the per-pair records are
                exchanged once (one padded all-gather).  Every rank returns the FULL per-pair
                result and kept-camera CSR; ds_* hold the rank's own blocks (ds_block0 = first).
    exact_ratio every pair's exact hull ratio (default: pairs decided by the
                area bounds skip the exact hull, see visibility()).
    host_results the caller will ask for partition_results_to_host(): the expanded index list (the
                large output, 4 bytes per expanded point) starts its device -> pinned host copy on the
                copy stream as soon as A4 has produced it, underneath A6-A9.
    """
    assert xyz.is_cuda
    dev = xyz.device
    tm = _Timer(time_stages)
    tm.mark("start")
    if man_trans is not None:
        x, y, z = aos_to_soa_manhattan(xyz, man_trans)
    else:
        x, y, z = aos_to_soa(xyz)
    tm.mark("layout")
    if split_xyz is not None:
        sx, sy, sz = aos_to_soa(split_xyz)
        s_aos = split_xyz if split_xyz.is_contiguous() else None
    else:
        sx, sy, sz = x, y, z
        s_aos = xyz if (man_trans is None and xyz.is_contiguous()) else None
    n = sx.shape[0]
    if radius_filter and shard is not None and int(shard[1]) > 1:
        keep_mask, n_keep = radius_outlier_mask_sharded(sx, sy, sz, int(shard[0]), int(shard[1]),
                                                        shard[2] if len(shard) > 2 else None)
    elif radius_filter:
        keep_mask, n_keep = radius_outlier_mask(sx, sy, sz, xyz_aos=s_aos)
    else:
        keep_mask = torch.ones(n, dtype=torch.uint8, device=dev)
        n_keep = torch.full((1,), n, dtype=torch.int64, device=dev)
    tm.mark("A1_radius")
    fx, fy, n_f = compact_xy(sx, sy, keep_mask)
    root = bounds_xy(fx, fy, n_f)
    tm.mark("A2_bounds")
    lb, lp, ld, lc = balanced_split(fx, fy, n_f, root, threshold)
    tm.mark("A3_split")
    del fx, fy
    boxes = expanded_boxes(lb)
    pt_off, pt_idx, pt_cnt_h = expand_select(x, y, boxes, want_counts=True)
    tm.mark("A4_expand")
    cam_off, cam_idx, cam_off_h = assign_cameras(cam_centers, boxes, dev, want_host=True)
    tm.mark("A5_assign")
    res = PartitionDeviceResult(
        keep_mask=keep_mask, n_filtered=int(n_f.item()), root_bounds=tuple(root.cpu().tolist()),
        leaf_bounds=lb, leaf_path=lp, leaf_depth=ld, leaf_count=lc,
        leaf_ids=[leaf_id(p, d) for p, d in zip(lp.tolist(), ld.tolist())], boxes=boxes,
        pt_offsets=pt_off, pt_index=pt_idx, cam_offsets=cam_off, cam_index=cam_idx)
    res.pt_counts_h, res.cam_offsets_h = pt_cnt_h, cam_off_h
    if host_results:
        main, copy = torch.cuda.current_stream(dev), _copy_stream(dev)
        copy.wait_stream(main)                                   # pt_idx is complete on the main stream
        with torch.cuda.stream(copy):
            idx_h = torch.empty(int(pt_idx.shape[0]), dtype=torch.int32, pin_memory=True)
            idx_h.copy_(pt_idx, non_blocking=True)
            pt_idx.record_stream(copy)
            done = torch.cuda.Event()
            done.record(copy)
        res.pt_index_host = (idx_h, done)
    if do_visibility:
        cams_d = torch.from_numpy(np.ascontiguousarray(cam_records, dtype=np.float32)).to(dev)
        # the input rows double as the gather source of A6 when the planes carry the same values
        aos = xyz if (man_trans is None and xyz.dtype == x.dtype and xyz.is_contiguous()) else None
        if shard is not None and int(shard[1]) > 1:
            from . import multigpu as M
            rank, world = int(shard[0]), int(shard[1])
            group = shard[2] if len(shard) > 2 else None
            plan = M.RangePlan(None, None, cam_off_h, pt_cnt_h, rank, world)
            b0, b1 = plan.bl0, plan.bl1
            pt_off_h = np.concatenate([[0], np.cumsum(pt_cnt_h)]).astype(np.int64)
            e0, e1 = int(pt_off_h[b0]), int(pt_off_h[b1])
            sub_off = torch.from_numpy(pt_off_h[b0:b1 + 1] - e0).to(dev)
            # sync=False: no host read-back of the voxel count here -- a small device -> host copy would
            # queue behind the index list's copy (host_results) on the copy engine; the planes keep their
            # full capacity until the exchange's own synchronisation below
            ds_off, dx, dy, dz, dfirst = voxel_downsample(x, y, z, sub_off, pt_idx[e0:e1], xyz_aos=aos, sync=False) \
                if b1 > b0 else \
                voxel_downsample(x, y, z, torch.zeros(2, dtype=torch.int64, device=dev), pt_idx[:0], sync=False)
            tm.mark("A6_voxel")
            loc = np.clip(cam_off_h[b0:b1 + 1], plan.lo, plan.hi) - plan.lo if b1 > b0 else np.zeros(2, np.int64)
            loc_off = torch.from_numpy(np.ascontiguousarray(loc, dtype=np.int64)).to(dev)
            nl = plan.hi - plan.lo
            v = visibility(dx, dy, dz, ds_off, cams_d, loc_off, cam_idx[plan.lo:plan.hi], exact_ratio=exact_ratio,
                           defer_check=True)
            tm.mark("A7_A8_visibility")
            local = (v.n_hit[:nl], v.ratio[:nl], v.area[:nl], v.keep[:nl], v.flags[:nl])
            n_hit, ratio, area, keep, flags = M.exchange_ranges(local, plan, pack_records_dev, unpack_records_dev,
                                                                group)
            bad = int((flags & 12).max().item()) if flags.numel() else 0     # same verdict on every rank
            if bad & 8:
                raise RuntimeError("visibility: convex hull exceeds the kernel's vertex budget")
            if bad & 4:       # candidate capacity overflow somewhere: every rank re-runs with the check inside
                v = visibility(dx, dy, dz, ds_off, cams_d, loc_off, cam_idx[plan.lo:plan.hi], exact_ratio=exact_ratio)
                local = (v.n_hit[:nl], v.ratio[:nl], v.area[:nl], v.keep[:nl], v.flags[:nl])
                n_hit, ratio, area, keep, flags = M.exchange_ranges(local, plan, pack_records_dev,
                                                                    unpack_records_dev, group)
            vis = VisibilityResult(n_hit, area, ratio, keep, flags, None, v.rounds)
            vis.n_exact = v.n_exact
            m_ds = int(ds_off[-1].item())
            dx, dy, dz, dfirst = dx[:m_ds], dy[:m_ds], dz[:m_ds], dfirst[:m_ds]
            tm.mark("exchange")
            res.ds_block0 = b0
        else:
            ds_off, dx, dy, dz, dfirst = voxel_downsample(x, y, z, pt_off, pt_idx, xyz_aos=aos)
            tm.mark("A6_voxel")
            vis = visibility(dx, dy, dz, ds_off, cams_d, cam_off, cam_idx, exact_ratio=exact_ratio)
            tm.mark("A7_A8_visibility")
            res.ds_block0 = 0
        k_off, k_idx = select_cameras(vis.keep, cam_off, cam_idx)
        tm.mark("A9_select")
        res.ds_offsets, res.ds_x, res.ds_y, res.ds_z, res.ds_first = ds_off, dx, dy, dz, dfirst
        res.vis, res.kept_offsets, res.kept_index = vis, k_off, k_idx
    res.timings_ms = tm.result()
    return res


def partition_results_to_host(res: PartitionDeviceResult) -> dict:
    """The result set SURVEY.md 8(d)(ii) ends the end-to-end partition on, as host arrays: leaf
    rectangles / ids / counts (already host side), the expanded index CSR, the block -> camera CSR,
    per-pair n_hit / ratio / keep / flags and the kept-camera CSR.  The index list is the large one
    (4 bytes per expanded point); it lands in pinned memory so the copy runs at PCIe speed."""
    out = {"leaf_bounds": res.leaf_bounds, "leaf_ids": res.leaf_ids, "leaf_count": res.leaf_count,
           "boxes": res.boxes}
    staged = getattr(res, "pt_index_host", None)
    if staged is not None:                    # partition_device(host_results=True) started the copy after A4
        idx_h, done = staged
        done.synchronize()
    else:
        idx_h = torch.empty(int(res.pt_index.shape[0]), dtype=torch.int32, pin_memory=True)
        idx_h.copy_(res.pt_index, non_blocking=True)
    out["pt_index"] = idx_h.numpy()
    small = {"pt_offsets": res.pt_offsets, "cam_offsets": res.cam_offsets, "cam_index": res.cam_index}
    if res.vis is not None:
        small.update({"n_hit": res.vis.n_hit, "ratio": res.vis.ratio, "keep": res.vis.keep, "flags": res.vis.flags,
                      "kept_offsets": res.kept_offsets, "kept_index": res.kept_index})
    for k, t in small.items():
        out[k] = t.cpu().numpy()
    torch.cuda.current_stream().synchronize()
    return out


def fma_probe(device, iters: int = 4096) -> float:
    """Measured FP32 FMA issue peak of the device in lane-ops/s (ptgs_debug_fma_probe: 8 independent
    FMA chains per thread, 148 x 8 CTAs of 256 threads).  Diagnostics for the roofline report."""
    lib = _lib.load()
    out = torch.empty(148 * 8 * 256, dtype=torch.float32, device=device)
    st = _stream()
    check(lib.ptgs_debug_fma_probe(_ptr(out), int(iters), st))
    torch.cuda.synchronize()
    best = None
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.ptgs_debug_fma_probe(_ptr(out), int(iters), st))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return float(148 * 8 * 256) * iters * 8 / (best * 1e-3)


def camera_centers(R, T) -> np.ndarray:
    """SimpleCamera.camera_center (scene/cameras.py:56-58): -R^T T evaluated by
    torch in float32 on the CPU, one camera at a time like the reference."""
    out = np.empty((len(R), 3), dtype=np.float32)
    for i in range(len(R)):
        Rt = torch.tensor(np.asarray(R[i]), dtype=torch.float32)
        Tt = torch.tensor(np.asarray(T[i]), dtype=torch.float32)
        out[i] = (-torch.matmul(Rt.transpose(0, 1), Tt)).numpy()
    return out


StageTimer = _Timer


def pack_records_dev(local, slot: int) -> torch.Tensor:
    """ptgs_pack_records: (n_hit, ratio, area, keep, flags) of this rank's range -> a (slot, 24) byte
    tensor whose first rows are the records (the padding rows are never read)."""
    lib = _lib.load()
    n_hit, ratio, area, keep, flags = local
    n = int(n_hit.shape[0])
    send = torch.empty((max(int(slot), 1), _lib.PAIR_RECORD_BYTES), dtype=torch.uint8, device=n_hit.device)
    check(lib.ptgs_pack_records(_ptr(n_hit), _ptr(ratio), _ptr(area), _ptr(keep), _ptr(flags), n, _ptr(send),
                                _stream()))
    return send


def unpack_records_dev(recv: torch.Tensor, plan):
    """ptgs_unpack_records: the gathered slots -> full-length (n_hit, ratio, area, keep, flags)."""
    lib = _lib.load()
    dev, P = recv.device, plan.P
    Pa = max(P, 1)
    n_hit = torch.empty(Pa, dtype=torch.int32, device=dev)
    ratio = torch.empty(Pa, dtype=torch.float64, device=dev)
    area = torch.empty(Pa, dtype=torch.float64, device=dev)
    keep = torch.empty(Pa, dtype=torch.uint8, device=dev)
    flags = torch.empty(Pa, dtype=torch.uint8, device=dev)
    lo = np.ascontiguousarray(plan.bounds, dtype=np.int64)
    base = np.ascontiguousarray(plan.base, dtype=np.int64)
    i64p = C.POINTER(C.c_int64)
    check(lib.ptgs_unpack_records(_ptr(recv), len(base), int(recv.shape[0]), lo.ctypes.data_as(i64p),
                                  base.ctypes.data_as(i64p), P, _ptr(n_hit), _ptr(ratio), _ptr(area), _ptr(keep),
                                  _ptr(flags), _stream()))
    return n_hit[:P], ratio[:P], area[:P], keep[:P], flags[:P]


def range_plan(ds_offsets, cam_offsets, cam_index, rank: int, world: int, cam_off_h=None, block_cost_h=None,
               chunks: int = 4):
    """multigpu.RangePlan of a device CSR (world * chunks interleaved ranges).  The L-sized host tables
    are taken from the caller when it has them (partition_device keeps them), otherwise copied from
    the device once."""
    from . import multigpu as M
    if cam_off_h is None:
        cam_off_h = cam_offsets.cpu().numpy()
    if block_cost_h is None:
        block_cost_h = (ds_offsets[1:] - ds_offsets[:-1]).cpu().numpy()
    return M.RangePlan(cam_offsets, cam_index, cam_off_h, block_cost_h, rank, world, chunks=chunks)


def visibility_sharded(ds_x, ds_y, ds_z, ds_offsets, cameras_d, cam_offsets, cam_index,
                       rank: int, world: int, group=None, timer=None, shard="ranges", plan=None,
                       **kw) -> VisibilityResult:
    """A7-A9 with the (block, camera) pairs sharded over the ranks (north star: only the
    visibility pass is partitioned; the per-pair result vectors are all-gathered).
    shard="ranges": contiguous equal-cost ranges of the pair list, a few per rank, interleaved
    (multigpu.RangePlan; pass a `plan` built once per partition result to keep its host arithmetic
    out of the step), records
    exchanged as one padded all-gather of the ranks' contiguous slices between a pack and an
    unpack kernel;
    shard="modulo": cameras c % world == rank of every block (multigpu.shard_csr + gather_records).
    Returns the FULL per-pair result on every rank."""
    from . import multigpu as M
    if world <= 1:
        return visibility(ds_x, ds_y, ds_z, ds_offsets, cameras_d, cam_offsets, cam_index,
                          timer=timer, **kw)
    if shard == "ranges":
        if plan is None:
            plan = range_plan(ds_offsets, cam_offsets, cam_index, rank, world)
        if timer is not None:
            timer.mark("vis_shard")
        # the candidate-capacity check of the pass is deferred past the exchange: pack / all-gather /
        # unpack are queued behind the hull kernel instead of waiting for a host round trip (the GPU
        # would idle for the host-side launch latency of the collective).  An overflow (rare) re-runs
        # the pass with the check inside and exchanges again -- on every rank, so that the collectives
        # stay matched (the overflow flag is part of the exchanged status below).
        n = plan.n_local
        v = visibility(ds_x, ds_y, ds_z, ds_offsets, cameras_d, plan.my_off, plan.my_idx, timer=timer,
                       defer_check=True, **kw)
        local = (v.n_hit[:n], v.ratio[:n], v.area[:n], v.keep[:n], v.flags[:n])
        n_hit, ratio, area, keep, flags = M.exchange_ranges(local, plan, pack_records_dev, unpack_records_dev, group)
        # flags bit2 (capacity overflow) / bit3 (vertex budget) travel with the records, so every rank
        # reaches the same verdict from the exchanged table (this is the step's one host round trip)
        bad = int((flags & 12).max().item()) if flags.numel() else 0
        if bad & 8:
            raise RuntimeError("visibility: convex hull exceeds the kernel's vertex budget")
        if bad & 4:
            v = visibility(ds_x, ds_y, ds_z, ds_offsets, cameras_d, plan.my_off, plan.my_idx, **kw)
            local = (v.n_hit[:n], v.ratio[:n], v.area[:n], v.keep[:n], v.flags[:n])
            n_hit, ratio, area, keep, flags = M.exchange_ranges(local, plan, pack_records_dev,
                                                                unpack_records_dev, group)
    else:
        my_off, my_idx, my_pos = M.shard_csr(cam_offsets, cam_index, rank, world)
        if timer is not None:
            timer.mark("vis_shard")
        v = visibility(ds_x, ds_y, ds_z, ds_offsets, cameras_d, my_off, my_idx, timer=timer, **kw)
        rec = M.pack_records(v.n_hit, v.ratio, v.area, v.keep, v.flags)
        full = M.gather_records(rec, cam_index, world, group, pos=my_pos)
        n_hit, ratio, area, keep, flags = M.unpack_records(full)
    if timer is not None:
        timer.mark("vis_allgather")
    return VisibilityResult(n_hit, area, ratio, keep, flags, None, v.rounds)


def upload_group_cuts(block_offsets: np.ndarray, bl0: int, bl1: int, groups) -> list:
    """Block boundaries of the upload groups of visibility_select_host: consecutive blocks of
    [bl0, bl1); `groups` = a count (~equal point counts) or the groups' shares of the points (a small
    first group starts the kernels early, a small last one shortens what is left to do once the last
    byte has landed).  Returns cuts with cuts[0] = bl0, cuts[-1] = bl1, non-decreasing (a group may
    be empty when a block is larger than a share)."""
    n_local = int(block_offsets[bl1] - block_offsets[bl0])
    if isinstance(groups, (list, tuple, np.ndarray)):
        shares = np.cumsum(np.asarray(groups, dtype=np.float64))
        shares = shares / shares[-1]
    else:
        G = max(1, int(groups))
        shares = np.arange(1, G + 1, dtype=np.float64) / G
    shares = shares[: max(bl1 - bl0, 1)]
    cuts = [bl0]
    for g in range(len(shares) - 1):
        b = int(np.searchsorted(block_offsets, block_offsets[bl0] + n_local * shares[g], side="left"))
        cuts.append(min(max(b, cuts[-1]), bl1))
    cuts.append(bl1)
    return cuts


def visibility_select_host(block_points: torch.Tensor, block_offsets: np.ndarray,
                           cam_records: torch.Tensor, cam_offsets: np.ndarray,
                           cam_index: np.ndarray, device, voxel=VOXEL_SIZE,
                           threshold=VISIBLE_RATE, rank=0, world=1, group=None, groups=2, exchange=None, streams=2,
                           trace=None):
    """visibility_based_camera_selection (shen_partation_densify.py:911-954)
    for HOST inputs: the expanded blocks' points back to back
    (`block_points` (sum N_b, 3) float32/float64, ideally pinned), the camera
    records and the block->camera CSR.  Uploads, down-samples (A6), runs the
    visibility pass (A7-A9) and returns host arrays
    (kept_offsets, kept_index, n_hit, ratio, keep, flags).

    With world > 1, rank r owns a contiguous range of the (block, camera) pair list of
    about 1/world of the cost (multigpu.range_bounds on the expanded point counts); it uploads,
    down-samples and sweeps only the blocks that range touches, and the per-pair records are
    exchanged once at the end (every rank returns the full result).

    The (local) blocks are cut into `groups` runs of consecutive blocks of about equal
    size; all uploads are queued on a copy stream up front and group g is
    processed on the current stream as soon as its points have landed, so the
    PCIe transfer of group g+1 overlaps the kernels of group g."""
    from . import multigpu as M
    L = len(block_offsets) - 1
    block_offsets = np.ascontiguousarray(block_offsets, dtype=np.int64)
    cam_offsets = np.ascontiguousarray(cam_offsets, dtype=np.int64)
    cam_index = np.ascontiguousarray(cam_index, dtype=np.int32)
    P = int(cam_offsets[L])
    # rank r owns the contiguous pair range [lo, hi) (cost of a pair = expanded points of its block,
    # known on the host) and uploads / down-samples only the blocks that range touches
    lo, hi = (0, P) if world <= 1 else M.range_bounds(cam_offsets, np.diff(block_offsets), rank, world)
    if hi > lo:
        bl0 = int(np.searchsorted(cam_offsets, lo, side="right")) - 1
        bl1 = int(np.searchsorted(cam_offsets, hi, side="left"))
    else:
        bl0 = bl1 = 0
    loc_off = np.clip(cam_offsets, lo, hi) - lo            # this rank's block -> camera CSR (global blocks)
    n_local = int(block_offsets[bl1] - block_offsets[bl0])
    cuts = upload_group_cuts(block_offsets, bl0, bl1, groups)
    G = len(cuts) - 1
    spans = [(cuts[i], cuts[i + 1]) for i in range(G) if cuts[i + 1] > cuts[i]]
    main = torch.cuda.current_stream(device)
    copy = _copy_stream(device)
    copy.wait_stream(main)
    # trace (diagnostics): a list that receives (label, ms since the call started on the device)
    tev = []
    def _tmark(label, stream):
        if trace is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            tev.append((label, e))
    _tmark("start", main)
    staged = []
    with torch.cuda.stream(copy):
        cams_d = cam_records.to(device, non_blocking=True)
        for (b0, b1) in spans:
            p0, p1 = int(block_offsets[b0]), int(block_offsets[b1])
            pts_d = block_points[p0:p1].to(device, non_blocking=True)
            off_d = torch.from_numpy(block_offsets[b0:b1 + 1] - p0).to(device, non_blocking=True)
            c0, c1 = int(loc_off[b0]), int(loc_off[b1])
            coff_d = torch.from_numpy(loc_off[b0:b1 + 1] - c0).to(device, non_blocking=True)
            cidx_d = torch.from_numpy(cam_index[lo + c0:lo + c1]).to(device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy)
            _tmark("upload%d" % len(staged), copy)
            staged.append((pts_d, off_d, coff_d, cidx_d, ev, p1 - p0, b1 - b0))
        # the whole block -> camera CSR: the final selection is ONE call on the concatenated results
        full_off_d = torch.from_numpy(cam_offsets).to(device, non_blocking=True)
        full_idx_d = torch.from_numpy(cam_index).to(device, non_blocking=True)
        ev_full = torch.cuda.Event()
        ev_full.record(copy)
    parts, n_ds = [], 0
    # groups alternate between the current stream and side streams: the latency-bound tail of one
    # group's visibility pass (sample pass, hull, small kernels) overlaps the next group's A6
    lanes = [main] + [_side_stream(device, i) for i in range(max(0, min(int(streams), len(staged)) - 1))]
    for ln in lanes[1:]:
        ln.wait_stream(main)
    for g, (pts_d, off_d, coff_d, cidx_d, ev, n_e, nb) in enumerate(staged):
        ln = lanes[g % len(lanes)]
        with torch.cuda.stream(ln):
            ln.wait_event(ev)
            for t in (pts_d, off_d, coff_d, cidx_d, cams_d):
                t.record_stream(ln)
            # the uploaded rows are read as they are (no plane layout pass); no host synchronisation
            # inside a group: the host queues group g+1 while group g runs
            _tmark("g%d_begin" % g, ln)
            ds_off, dx, dy, dz, _ = voxel_downsample(None, None, None, off_d, None, voxel, n_entries=n_e,
                                                     sync=False, xyz_aos=pts_d)
            _tmark("g%d_A6" % g, ln)
            args = (dx, dy, dz, ds_off, cams_d, coff_d, cidx_d)
            vis = visibility(*args, threshold=threshold, defer_check=True)
            _tmark("g%d_vis" % g, ln)
        parts.append([vis, ds_off, nb, args, coff_d, cidx_d])
    for ln in lanes[1:]:
        main.wait_stream(ln)
    for part in parts:                        # results are consumed on the current stream from here on
        v = part[0]
        for t in (v.n_hit, v.area, v.ratio, v.keep, v.flags, part[1]) + ((v.pending,) if getattr(v, "pending", None) is not None else ()):
            t.record_stream(main)
    # deferred candidate-capacity checks, ONE read-back for all groups (rare: re-run that group)
    pends = [getattr(part[0], "pending", None) for part in parts]
    if any(pd is not None for pd in pends):
        st_all = torch.stack([pd if pd is not None else torch.zeros(3, dtype=torch.int64, device=device)
                              for pd in pends]).cpu()
        for part, st_h in zip(parts, st_all):
            if int(st_h[1]):
                raise RuntimeError("visibility: convex hull exceeds the kernel's vertex budget")
            if int(st_h[0]):
                part[0] = visibility(*part[3], threshold=threshold)
    if world <= 1:
        # one selection over the concatenated keep flags and ONE round of device -> host copies
        # (per-group copies are synchronous round trips: 6 per group)
        main.wait_event(ev_full)
        full_off_d.record_stream(main); full_idx_d.record_stream(main)
        live = [p[0] for p in parts]
        cat = lambda f, dt: (torch.cat([getattr(v, f) for v in live]) if live
                             else torch.zeros(0, dtype=dt, device=device))
        n_hit, ratio, keep, flags = (cat("n_hit", torch.int32), cat("ratio", torch.float64),
                                     cat("keep", torch.uint8), cat("flags", torch.uint8))
        k_off, k_idx = select_cameras(keep, full_off_d, full_idx_d, sync=False)
        n_ds_d = torch.stack([p[1][p[2]] for p in parts]).sum() if parts else torch.zeros((), dtype=torch.int64, device=device)
        small = torch.cat([k_off, n_ds_d.reshape(1)]).cpu().numpy()          # synchronises
        ko, n_ds = small[:L + 1], int(small[L + 1])
        out = (ko, k_idx[: int(ko[L])].cpu().numpy(), n_hit.cpu().numpy(), ratio.cpu().numpy(),
               keep.cpu().numpy(), flags.cpu().numpy())
        if trace is not None:
            _tmark("end", main)
            torch.cuda.synchronize()
            trace.extend((lb, round(tev[0][1].elapsed_time(e), 3)) for lb, e in tev)
        return out, n_ds
    # world > 1: the groups' records are this rank's pairs lo..hi-1 in order; ONE exchange
    live = [p[0] for p in parts if p[0].n_hit.shape[0] > 0 and p[5].shape[0] > 0]
    main.wait_event(ev_full)
    full_off_d.record_stream(main); full_idx_d.record_stream(main)
    if exchange is not None:       # tests: stand-in for the collective
        recs = [M.pack_records(v.n_hit, v.ratio, v.area, v.keep, v.flags) for v in live]
        local = torch.cat(recs) if recs else torch.zeros((0, M.REC_COLS), dtype=torch.float64, device=device)
        n_hit, ratio, area, keep, flags = M.unpack_records(exchange(local[: hi - lo], torch.arange(lo, hi, device=device)))
    else:
        cat = lambda f, dt: (torch.cat([getattr(v, f) for v in live]) if live
                             else torch.zeros(0, dtype=dt, device=device))[: hi - lo]
        local = (cat("n_hit", torch.int32), cat("ratio", torch.float64), cat("area", torch.float64),
                 cat("keep", torch.uint8), cat("flags", torch.uint8))
        plan = M.RangePlan(None, None, cam_offsets, np.diff(block_offsets), rank, world)
        n_hit, ratio, area, keep, flags = M.exchange_ranges(local, plan, pack_records_dev, unpack_records_dev, group)
    k_off, k_idx = select_cameras(keep, full_off_d, full_idx_d, sync=False)
    ko = k_off.cpu().numpy()
    for p in parts:
        n_ds += int(p[1][p[2]].item())
    out = (ko, k_idx[: int(ko[L])].cpu().numpy(), n_hit.cpu().numpy(), ratio.cpu().numpy(),
           keep.cpu().numpy(), flags.cpu().numpy())
    return out, n_ds


_COPY_STREAMS = {}
_SIDE_STREAMS = {}


def _side_stream(device, i):
    d = torch.device(device)
    key = (d.index if d.index is not None else torch.cuda.current_device(), i)
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream(device=device)
    return _SIDE_STREAMS[key]


def _copy_stream(device):
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = torch.cuda.Stream(device=device)
    return _COPY_STREAMS[key]
