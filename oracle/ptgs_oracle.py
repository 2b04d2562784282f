"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy) of the reference's
point-count scene partitioner.  Nothing in the product path may import this
module; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs do, and only as the checker / timed CPU baseline.

Every function cites the reference lines it follows (paths relative to
/root/reference).  Abbreviations: densify = scene/ptgs/shen_partation_densify.py,
dp = scene/ptgs/density_partition.py, utils-shen = scene/ptgs/shen_partition_utils.py.

Pinning status (see DESIGN.md "Oracle"):
  * split / expand / assign / voxel down-sampling / projection / Graham scan /
    selection are pinned against the reference's OWN code executed in the build
    container through oracle/ref_harness.py (tests/test_oracle_vs_reference.py)
    and against the committed fixtures tests/golden/*.npz that the same code
    produced (oracle/make_golden.py).  The reference has no tests or golden
    vectors of its own.
  * radius_outlier_mask restates open3d's remove_radius_outlier (third party,
    un-pinned version, not vendored in the reference): PARITY UNPINNED.
"""
from __future__ import annotations

import math
from typing import NamedTuple

import numpy as np

# --------------------------------------------------------------------------
# constants hard-coded by the reference
# --------------------------------------------------------------------------
RADIUS = 1.0            # utils-shen:30
NB_POINTS = 5           # utils-shen:30
LEAF_FACTOR = 1.2       # dp:32
MAX_DEPTH = 10          # dp:163
EXPAND_RATIO = 0.10     # densify:222
EXTEND_RATE = 80        # densify:183 (label only)
VOXEL_SIZE = 0.1        # densify:928
VISIBLE_RATE = 0.45     # densify:930
MIN_HITS = 3            # densify:898 (keep needs len > 3)


class Leaf(NamedTuple):
    partition_id: str
    bounds: tuple        # (minx, miny, maxx, maxy) Python floats
    point_num: int


# --------------------------------------------------------------------------
# A1  radius outlier removal  (open3d, third party -- parity unpinned)
# --------------------------------------------------------------------------
def radius_outlier_mask(points, radius=RADIUS, nb_points=NB_POINTS):
    """utils-shen:50  pcd.remove_radius_outlier(nb_points, radius).

    open3d (geometry/PointCloud.cpp, RemoveRadiusOutliers) keeps point i iff
    the nanoflann radius search around it returns MORE than nb_points entries;
    the search keeps j when d2 < radius^2 (strict) with
    d2 = ((dx*dx) + dy*dy) + dz*dz in float64, and includes i itself.
    """
    from scipy.spatial import cKDTree
    p = np.ascontiguousarray(points, dtype=np.float64)
    n = len(p)
    if n == 0:
        return np.zeros(0, dtype=bool)
    tree = cKDTree(p)
    lo = tree.query_ball_point(p, radius * (1 - 1e-9), return_length=True, workers=-1)
    hi = tree.query_ball_point(p, radius * (1 + 1e-9), return_length=True, workers=-1)
    keep = lo > nb_points
    amb = np.nonzero((lo > nb_points) != (hi > nb_points))[0]
    r2 = radius * radius
    for i in amb:
        cand = tree.query_ball_point(p[i], radius * (1 + 1e-9))
        d = p[cand] - p[i]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        keep[i] = int(np.count_nonzero(d2 < r2)) > nb_points
    return keep


def radius_outlier_mask_bruteforce(points, radius=RADIUS, nb_points=NB_POINTS):
    """O(n^2) literal form of the same rule, for small cross-checks."""
    p = np.asarray(points, dtype=np.float64)
    keep = np.zeros(len(p), dtype=bool)
    r2 = radius * radius
    for i in range(len(p)):
        d = p - p[i]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        keep[i] = int(np.count_nonzero(d2 < r2)) > nb_points
    return keep


# --------------------------------------------------------------------------
# A2  XY bounds
# --------------------------------------------------------------------------
def bounds_xy(points):
    """utils-shen:66-70: min/max of the filtered cloud's x and y."""
    p = np.asarray(points)
    xmin, ymin = p[:, :2].min(axis=0)
    xmax, ymax = p[:, :2].max(axis=0)
    return (float(xmin), float(ymin), float(xmax), float(ymax))


# --------------------------------------------------------------------------
# A3  balanced binary (median) split
# --------------------------------------------------------------------------
def balanced_binary_partition(points, bounds, threshold, depth=0, max_depth=MAX_DEPTH,
                              index_prefix=""):
    """dp:9-99.  `points` (n,>=2) float64, `bounds` = (xmin,ymin,xmax,ymax).
    Returns the leaves depth-first, left child first."""
    xmin, ymin, xmax, ymax = bounds
    in_region = (points[:, 0] >= xmin) & (points[:, 0] <= xmax) & \
                (points[:, 1] >= ymin) & (points[:, 1] <= ymax)       # dp:22-23
    region = points[in_region]
    n = len(region)
    pid = index_prefix if index_prefix else "0"                        # dp:28
    if (n <= threshold * LEAF_FACTOR) or (depth >= max_depth):         # dp:32
        return [Leaf(pid, (float(xmin), float(ymin), float(xmax), float(ymax)), n)]
    axis = 0 if (xmax - xmin) >= (ymax - ymin) else 1                  # dp:45-50
    # dp:53-56  sorted(coord)[n//2]; the value does not depend on the sort's
    # tie order, so a selection is equivalent.
    split = float(np.partition(region[:, axis], n // 2)[n // 2])
    if axis == 0:                                                      # dp:59-64
        lb, rb = (xmin, ymin, split, ymax), (split, ymin, xmax, ymax)
    else:
        lb, rb = (xmin, ymin, xmax, split), (xmin, split, xmax, ymax)
    out = []
    out += balanced_binary_partition(region, lb, threshold, depth + 1, max_depth,
                                     f"{index_prefix}0" if index_prefix else "0")
    out += balanced_binary_partition(region, rb, threshold, depth + 1, max_depth,
                                     f"{index_prefix}1" if index_prefix else "1")
    return out


def quad_tree_partition(points, bounds, threshold):
    """densify:37-49 -> dp:148-166 (always depth=0, max_depth=10)."""
    return balanced_binary_partition(np.asarray(points), bounds, threshold)


# --------------------------------------------------------------------------
# A4  expansion
# --------------------------------------------------------------------------
def expanded_box(bounds):
    """densify:210-232: 5 % of the leaf's own width/height per side, evaluated
    in Python floats.  Returns (minx, miny, maxx, maxy)."""
    minx, miny, maxx, maxy = bounds
    width = maxx - minx
    height = maxy - miny
    ew = width * EXPAND_RATIO / 2
    eh = height * EXPAND_RATIO / 2
    return (minx - ew, miny - eh, maxx + ew, maxy + eh)


def expand_mask(points, box):
    """densify:197-202 (extract_point_cloud): closed mask; `points` keeps its
    own dtype so a float32 cloud is compared in float32 against the bound
    rounded to float32 (NumPy weak-scalar promotion)."""
    minx, miny, maxx, maxy = box
    return (points[:, 0] >= minx) & (points[:, 0] <= maxx) & \
           (points[:, 1] >= miny) & (points[:, 1] <= maxy)


# --------------------------------------------------------------------------
# A5  camera assignment
# --------------------------------------------------------------------------
def camera_centers(R, T):
    """scene/cameras.py:56-58: -R^T T in torch float32 on the CPU."""
    import torch
    out = np.empty((len(R), 3), dtype=np.float32)
    for i in range(len(R)):
        Rt = torch.tensor(np.asarray(R[i]), dtype=torch.float32)
        Tt = torch.tensor(np.asarray(T[i]), dtype=torch.float32)
        out[i] = (-torch.matmul(Rt.transpose(0, 1), Tt)).numpy()
    return out


def assign_cameras(boxes, centers):
    """densify:384-389: camera in leaf iff centre STRICTLY inside extend_box
    (shapely Polygon.contains(Point)), XY only.  Returns list of index arrays."""
    c = np.asarray(centers, dtype=np.float64)
    out = []
    for (minx, miny, maxx, maxy) in boxes:
        m = (c[:, 0] > minx) & (c[:, 0] < maxx) & (c[:, 1] > miny) & (c[:, 1] < maxy)
        out.append(np.nonzero(m)[0].astype(np.int32))
    return out


# --------------------------------------------------------------------------
# A6  voxel down-sampling
# --------------------------------------------------------------------------
def voxel_downsample(points, voxel_size=VOXEL_SIZE, return_first=False):
    """densify:541-614, points only (colours/normals are unused downstream).
    Vectorised but order-exact: voxels in first-occurrence order, each mean =
    sequential sum in index order (np.mean over axis 0 of a (k,3) array) then
    a divide, result cast to float32 (densify:605)."""
    points = np.asarray(points)
    n = len(points)
    if n == 0:
        return (points.astype(np.float32), np.zeros(0, np.int64)) if return_first \
            else points.astype(np.float32)
    min_bound = np.min(points, axis=0)
    vid = np.floor((points - min_bound) / voxel_size).astype(np.int32)   # densify:572
    v64 = vid.astype(np.int64)
    span = v64.max(axis=0) + 1
    key = (v64[:, 0] * span[1] + v64[:, 1]) * span[2] + v64[:, 2]
    uniq, first, inv, cnt = np.unique(key, return_index=True, return_inverse=True,
                                      return_counts=True)
    order = np.argsort(first, kind="stable")          # first-occurrence order
    rank_of = np.empty(len(uniq), dtype=np.int64)
    rank_of[order] = np.arange(len(uniq))
    g = rank_of[inv]                                  # output slot of each point
    cnt_o = cnt[order]
    # sequential accumulation: r-th member of each voxel added at step r
    srt = np.lexsort((np.arange(n), g))               # by voxel, then index
    gs = g[srt]
    start = np.zeros(len(uniq), dtype=np.int64)
    start[1:] = np.cumsum(cnt_o)[:-1]
    pos_in = np.arange(n) - start[gs]
    acc = np.zeros((len(uniq), 3), dtype=points.dtype)
    maxc = int(cnt_o.max())
    for r in range(maxc):
        sel = pos_in == r
        acc[gs[sel]] += points[srt[sel]]
    mean = acc / cnt_o[:, None].astype(points.dtype)
    out = mean.astype(np.float32)
    if return_first:
        return out, first[order]
    return out


def voxel_downsample_literal(points, voxel_size=VOXEL_SIZE):
    """The reference's own dict loop (densify:566-605), for small inputs."""
    points = np.asarray(points)
    if len(points) == 0:
        return points
    min_bound = np.min(points, axis=0)
    vid = np.floor((points - min_bound) / voxel_size).astype(np.int32)
    d = {}
    for i, v in enumerate(vid):
        d.setdefault((v[0], v[1], v[2]), []).append(i)
    return np.array([np.mean(points[idx], axis=0) for idx in d.values()], dtype=np.float32)


# --------------------------------------------------------------------------
# A7  projection
# --------------------------------------------------------------------------
class CameraSet(NamedTuple):
    """Per camera fields of SimpleCamera used by the path (scene/cameras.py:18-40)."""
    R: np.ndarray        # (C,3,3) float64, world->camera
    T: np.ndarray        # (C,3) float64
    FoVx: np.ndarray     # (C,) float64
    FoVy: np.ndarray
    width: np.ndarray    # (C,) int
    height: np.ndarray


def focal2fov(focal, pixels):
    """utils/graphics_utils.py:76."""
    return 2 * math.atan(pixels / (2 * focal))


def qvec2rotmat(q):
    """scene/colmap_loader.py:43-53."""
    return np.array([
        [1 - 2 * q[2] ** 2 - 2 * q[3] ** 2, 2 * q[1] * q[2] - 2 * q[0] * q[3], 2 * q[3] * q[1] + 2 * q[0] * q[2]],
        [2 * q[1] * q[2] + 2 * q[0] * q[3], 1 - 2 * q[1] ** 2 - 2 * q[3] ** 2, 2 * q[2] * q[3] - 2 * q[0] * q[1]],
        [2 * q[3] * q[1] - 2 * q[0] * q[2], 2 * q[2] * q[3] + 2 * q[0] * q[1], 1 - 2 * q[1] ** 2 - 2 * q[2] ** 2]])


def camera_record(R, T, FoVx, FoVy, width, height, front_axis=1):
    """The 20-float camera record of include/ptgs.h built from the same Python
    expressions as densify:433-463."""
    rec = np.zeros(20, dtype=np.float32)
    rec[0:9] = np.asarray(R, dtype=np.float64).reshape(9).astype(np.float32)   # :433
    rec[9:12] = np.asarray(T, dtype=np.float64).reshape(3).astype(np.float32)  # :434
    fx = width / (2 * math.tan(FoVx / 2))                                       # :457
    fy = height / (2 * math.tan(FoVy / 2))                                      # :458
    rec[12] = np.float32(fx)
    rec[13] = np.float32(fy)
    rec[14] = np.float32(width / 2)                                             # :460
    rec[15] = np.float32(height / 2)                                            # :461
    rec[16] = np.float32(width)
    rec[17] = np.float32(height)
    rec[18] = np.float32(front_axis)
    return rec


def _fma32(a, b, c):
    """float32 fused multiply-add emulated through float64 (the product of two
    float32 is exact in float64; the single extra rounding of the sum is the
    documented 2^-29-probability difference from a hardware FMA)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def project_port(points, rec):
    """densify:429-471 in the arithmetic libptgs fixes (include/ptgs.h,
    "One test"): float32, camera space by the SGEMM's FMA chain, IEEE division.
    Returns (hit mask over ALL points, u, v) -- u,v valid where hit."""
    p = np.asarray(points, dtype=np.float32)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    r = rec
    f = np.float32

    def row(i):
        # k-ordered accumulation of the 4x4 @ 4xN SGEMM (densify:444): r0*x,
        # fma r1*y, fma r2*z, then + t (fma(t, 1, acc) is a plain add).  Checked
        # bit-identical to torch's CPU matmul (tests/test_oracle_vs_reference.py).
        acc = np.full_like(x, r[3 * i]) * x
        acc = _fma32(np.full_like(x, r[3 * i + 1]), y, acc)
        acc = _fma32(np.full_like(x, r[3 * i + 2]), z, acc)
        return acc + np.full_like(x, r[9 + i])
    X, Y, Z = row(0), row(1), row(2)
    front = (Y > 0) if r[18] == f(1.0) else (Z > 0)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        uh = _fma32(np.full_like(x, r[14]), Z, (np.full_like(x, r[12]) * X))
        vh = _fma32(np.full_like(x, r[15]), Z, (np.full_like(x, r[13]) * Y))
        u = uh / Z
        v = vh / Z
        hit = front & (u >= 0) & (u < r[16]) & (v >= 0) & (v < r[17])
    return hit, u, v


def project_points_to_camera_ref(points, R, T, FoVx, FoVy, width, height, device="cpu"):
    """densify:410-477 restated call for call with torch (device='cpu' by default; 'cuda' is the
    reference's own default, densify:410: one host->device copy of the points per camera, two
    K = 4 matmuls, a device->host copy).  Returns the (M,2) float32 array of in-image points (the
    reference returns its .tolist())."""
    import torch
    points = np.asarray(points, dtype=np.float32)
    points_t = torch.tensor(points, device=device)
    R_wc = torch.tensor(np.asarray(R), dtype=torch.float32, device=device)
    T_wc = torch.tensor(np.asarray(T), dtype=torch.float32, device=device)
    W2C = torch.zeros((4, 4), dtype=torch.float32, device=device)
    W2C[:3, :3] = R_wc
    W2C[:3, -1] = T_wc
    W2C[3, 3] = 1.0
    N = points_t.shape[0]
    ones = torch.ones((N, 1), dtype=torch.float32, device=device)
    homog = torch.cat((points_t, ones), dim=1)
    pc_h = (W2C @ homog.transpose(0, 1)).transpose(0, 1)
    pc = pc_h[:, :3] / pc_h[:, 3, None]
    front = pc[:, 1] > 0                                                    # :447
    if torch.sum(front) == 0:
        return np.zeros((0, 2), dtype=np.float32)
    pcf = pc[front]
    fx = width / (2 * math.tan(FoVx / 2))
    fy = height / (2 * math.tan(FoVy / 2))
    K = torch.tensor([[fx, 0, width / 2], [0, fy, height / 2], [0, 0, 1]],
                     dtype=torch.float32, device=device)
    ih = (K @ pcf.transpose(0, 1)).transpose(0, 1)
    img = ih[:, :2] / ih[:, 2, None]
    ok = (img[:, 0] >= 0) & (img[:, 0] < width) & (img[:, 1] >= 0) & (img[:, 1] < height)
    return img[ok].cpu().numpy()


# --------------------------------------------------------------------------
# A8  Graham scan + area ratio
# --------------------------------------------------------------------------
def run_graham_scan(image_points, image_width, image_height):
    """densify:481-539, literal (including its tie handling: the first element
    of the sorted list is skipped, equal-angle points are visited farthest
    first and popped on cross <= 0)."""
    points = np.array(image_points, dtype=float)
    if len(points) < 3:
        return 0.0
    min_y = np.min(points[:, 1])
    cand = points[points[:, 1] == min_y]
    p0 = cand[np.argmin(cand[:, 0])]                                        # :509-511
    ang = np.arctan2(points[:, 1] - p0[1], points[:, 0] - p0[0])
    d2 = (points[:, 0] - p0[0]) ** 2 + (points[:, 1] - p0[1]) ** 2
    order = np.lexsort((-d2, ang))                # sorted(key=(angle, -dist^2)), stable
    sp = points[order]
    hx = [float(p0[0])]
    hy = [float(p0[1])]
    xs = sp[:, 0].tolist()
    ys = sp[:, 1].tolist()
    for i in range(1, len(xs)):                                             # :516-527
        px, py = xs[i], ys[i]
        while len(hx) > 1:
            cross = ((hx[-1] - hx[-2]) * (py - hy[-2]) - (hy[-1] - hy[-2]) * (px - hx[-2]))
            if cross <= 0:
                hx.pop()
                hy.pop()
            else:
                break
        hx.append(px)
        hy.append(py)
    if len(hx) < 3:
        return 0.0
    x = np.array(hx)
    y = np.array(hy)
    area = 0.5 * np.abs(np.dot(x, np.roll(y, 1)) - np.dot(y, np.roll(x, 1)))  # :535
    return float(area / (image_width * image_height))


def first_ray_collinear(image_points):
    """True when >= 2 distinct points with y > ymin are exactly collinear with
    the pivot on the minimum-angle ray -- the one configuration in which the
    reference's result depends on whether libm's atan2 returns bit-identical
    angles for them (tests report such pairs separately)."""
    pts = np.unique(np.asarray(image_points, dtype=np.float64), axis=0)
    if len(pts) < 3:
        return False
    ymin = pts[:, 1].min()
    p0 = pts[pts[:, 1] == ymin][0]   # unique() sorted by x then y -> min x first
    up = pts[pts[:, 1] > ymin]
    if len(up) < 2:
        return False
    d = up - p0
    best = d[0]
    for q in d[1:]:
        if best[0] * q[1] - best[1] * q[0] < 0:
            best = q
    col = np.count_nonzero(best[0] * d[:, 1] - best[1] * d[:, 0] == 0)
    return col >= 2


# --------------------------------------------------------------------------
# A9  per-camera decision and the whole path
# --------------------------------------------------------------------------
def camera_visibility(ds_points, rec, threshold=VISIBLE_RATE, project=project_port):
    """densify:890-910.  Returns (n_hit, ratio, keep)."""
    hit, u, v = project(ds_points, rec)
    n_hit = int(np.count_nonzero(hit))
    if n_hit <= MIN_HITS:                                                   # :898
        return n_hit, 0.0, False
    uv = np.stack([u[hit], v[hit]], axis=1).astype(np.float64)
    ratio = run_graham_scan(uv, float(rec[16]), float(rec[17]))
    return n_hit, ratio, bool(ratio >= threshold)                           # :904


class PartitionResult(NamedTuple):
    keep_mask: np.ndarray          # A1, over the unfiltered cloud
    root_bounds: tuple
    leaves: list                   # [Leaf]
    boxes: list                    # expanded boxes (minx,miny,maxx,maxy)
    block_index: list              # per leaf: ascending indices into the unfiltered cloud
    cam_index: list                # per leaf: ascending camera indices (A5)
    ds_points: list                # per leaf (M_b,3) float32, first-occurrence order
    n_hit: list                    # per leaf arrays aligned with cam_index
    ratio: list
    keep: list
    kept_cameras: list             # per leaf: ascending camera indices kept (A9)


def run_partition(points, cams: CameraSet, threshold, centers=None,
                  radius_filter=True, visibility=True, front_axis=1, split_points=None):
    """utils-shen:57-93 (run_DataPartition) without plotting and file output.
    `points` is the scene cloud as fetchPly returns it (float32, or float64
    under a Manhattan transform) = self.pcd, which A4 (utils-shen:86) and A6-A9
    (:91) read.  `split_points` (default: `points`) is what A1-A3 read: the
    reference re-reads the PLY with open3d (utils-shen:62, self.ply), which is
    the UN-rotated float32 cloud even when fetchPly rotated self.pcd
    (shen_data_read.py:126-130) -- the literal shen-path behaviour under
    man_trans."""
    points = np.asarray(points)
    p64 = np.asarray(points if split_points is None else split_points).astype(np.float64)   # o3d.io.read_point_cloud -> f64
    keep_mask = radius_outlier_mask(p64) if radius_filter else np.ones(len(p64), bool)
    filt = p64[keep_mask]
    root = bounds_xy(filt)
    leaves = quad_tree_partition(filt, root, threshold)
    boxes = [expanded_box(l.bounds) for l in leaves]
    block_index = [np.nonzero(expand_mask(points, b))[0] for b in boxes]
    if centers is None:
        centers = camera_centers(cams.R, cams.T)
    cam_index = assign_cameras(boxes, centers)
    ds, nh, rt, kp, kept = [], [], [], [], []
    if visibility:
        recs = [camera_record(cams.R[c], cams.T[c], cams.FoVx[c], cams.FoVy[c],
                              int(cams.width[c]), int(cams.height[c]), front_axis)
                for c in range(len(cams.R))]
        for b, idx in enumerate(block_index):
            d = voxel_downsample(points[idx])
            ds.append(d)
            r = [camera_visibility(d, recs[c]) for c in cam_index[b]]
            nh.append(np.array([q[0] for q in r], dtype=np.int32))
            rt.append(np.array([q[1] for q in r], dtype=np.float64))
            kp.append(np.array([q[2] for q in r], dtype=bool))
            kept.append(cam_index[b][kp[-1]] if len(r) else np.zeros(0, np.int32))
    return PartitionResult(keep_mask, root, leaves, boxes, block_index, cam_index,
                           ds, nh, rt, kp, kept)


# --------------------------------------------------------------------------
# A13  Manhattan transform of the cloud
# --------------------------------------------------------------------------
def manhattan_points(points_f32, man_trans):
    """scene/ptgs/shen_data_read.py:126-130 (fetchPly), same NumPy calls."""
    positions = np.asarray(points_f32)
    R = man_trans[:3, :3]
    T = man_trans[:3, -1]
    new = np.dot(R, positions.transpose()) + np.repeat(T, positions.shape[0]).reshape(-1, positions.shape[0])
    return new.transpose()


def manhattan_cameras(R, T, man_trans):
    """scene/dataset_readers.py:96-104: W2C @ inv(man_trans)."""
    inv = np.linalg.inv(man_trans)
    Ro = np.empty_like(np.asarray(R, dtype=np.float64))
    To = np.empty_like(np.asarray(T, dtype=np.float64))
    for i in range(len(R)):
        W2C = np.zeros((4, 4))
        W2C[:3, :3] = R[i]
        W2C[:3, -1] = T[i]
        W2C[3, 3] = 1.0
        W2nC = W2C @ inv
        Ro[i] = W2nC[:3, :3]
        To[i] = W2nC[:3, -1]
    return Ro, To


# --------------------------------------------------------------------------
# c5 / Vast-style coverage: every point against every camera, hit counts
# --------------------------------------------------------------------------
def coverage_counts(points, recs, want_seen=False):
    n_hit = np.zeros(len(recs), dtype=np.int32)
    seen = np.zeros(len(points), dtype=bool)
    for c, rec in enumerate(recs):
        hit, _, _ = project_port(points, rec)
        n_hit[c] = np.count_nonzero(hit)
        if want_seen:
            seen |= hit
    return (n_hit, seen) if want_seen else n_hit
