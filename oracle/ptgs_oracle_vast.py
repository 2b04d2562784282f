"""TEST INFRASTRUCTURE ONLY -- CPU (NumPy) restatement of the reference's second entry
point, the camera-grid ("Vast") partitioner: scene/ptgs/data_partition.py and
scene/ptgs/graham_scan.py:65-105.  Only tests/, smoke() and bench.py's
CPU legs may import this file; the product path never does.

Pinned against the reference's own code by tests/test_oracle_vs_reference.py (run under
oracle/ref_harness.py) and through tests/golden/vast.npz (oracle/make_golden.py).
Third-party pieces restated here because they are not in /root/reference: scipy's
ConvexHull (Qhull) -> Andrew monotone chain, shapely's polygon & box -> Sutherland-Hodgman;
both only enter through an area, compared with 1e-9 relative tolerance, and through the
`>= visible_rate` decision (pairs within 1e-9 of the threshold are reported, not compared).
"""
from __future__ import annotations

import math
from typing import Dict, List, NamedTuple

import numpy as np


class VastCameras(NamedTuple):
    """The fields data_partition.py reads from each camera (scene/cameras.py:18-71)."""
    center: np.ndarray      # (C,3) float32  camera.camera_center
    R: np.ndarray           # (C,3,3) float64 camera.R (this reader stores C2W: w2c = R^T, :389-392)
    T: np.ndarray           # (C,3) float64
    FoVx: np.ndarray        # (C,) float64
    FoVy: np.ndarray
    width: np.ndarray       # (C,) int
    height: np.ndarray
    names: List[str]


class VastPartition(NamedTuple):
    partition_id: str
    cameras: np.ndarray             # camera indices, in the reference's list order
    point_index: np.ndarray         # indices into the scene cloud (extended box), ascending
    ori_camera_bbox: list
    extend_camera_bbox: list
    ori_point_bbox: list
    extend_point_bbox: list


# ----------------------------------------------------------------- region division
def region_division(center: np.ndarray, m: int, n: int):
    """Camera_position_based_region_division (data_partition.py:220-274).  Returns
    {id: {"camera_list": [camera indices], "x_mid_camera": idx|None, "z_mid_camera": idx|None}}
    in the reference's dict insertion order.  sorted() is stable, keys are float32 scalars."""
    C = center.shape[0]
    by_x = sorted(range(C), key=lambda i: center[i, 0])
    per_m = C // m
    m_dict = {}
    for i in range(m):
        d = {"camera_list": by_x[i * per_m:(i + 1) * per_m]}
        d["x_mid_camera"] = by_x[(i + 1) * per_m - 1] if i != m - 1 else None
        m_dict[str(i + 1)] = d
    if C % m != 0:
        m_dict[str(m)]["camera_list"].extend(by_x[m * per_m:])
    out = {}
    for pid, cams in m_dict.items():
        tot = len(cams["camera_list"])
        per_n = tot // n
        by_z = sorted(cams["camera_list"], key=lambda i: center[i, 2])
        for i in range(n):
            d = {"camera_list": by_z[i * per_n:(i + 1) * per_n], "x_mid_camera": cams["x_mid_camera"]}
            d["z_mid_camera"] = by_z[(i + 1) * per_n - 1] if i != n - 1 else None
            out["%s_%d" % (pid, i + 1)] = d
        if tot % n != 0:
            out["%s_%d" % (pid, n)]["camera_list"].extend(by_z[n * per_n:])
    return out


def refine_ori_bbox(pdict, center: np.ndarray, m: int, n: int):
    """refine_ori_bbox (data_partition.py:179-218): camera extents, then the z seams
    (z_mid_camera of the lower block) and the x seams (x_mid_camera of the left block)."""
    bbox = {}
    for pid, d in pdict.items():
        cl = d["camera_list"]
        xs = [center[i, 0] for i in cl]
        zs = [center[i, 2] for i in cl]
        bbox[pid] = [min(xs), max(xs), min(zs), max(zs)]
    for mm in range(1, m + 1):
        for nn in range(1, n + 1):
            if nn + 1 == n + 1:
                break
            up, lo = "%d_%d" % (mm, nn + 1), "%d_%d" % (mm, nn)
            a, b = bbox[up], bbox[lo]
            mid_z = center[pdict[lo]["z_mid_camera"], 2]
            bbox[up] = [a[0], a[1], mid_z, a[3]]
            bbox[lo] = [b[0], b[1], b[2], mid_z]
    for nn in range(1, n + 1):
        for mm in range(1, m + 1):
            if mm + 1 == m + 1:
                break
            le, ri = "%d_%d" % (mm, nn), "%d_%d" % (mm + 1, nn)
            a, b = bbox[le], bbox[ri]
            mid_x = center[pdict[le]["x_mid_camera"], 0]
            bbox[le] = [a[0], mid_x, a[2], a[3]]
            bbox[ri] = [mid_x, b[1], b[2], b[3]]
    return {pid: d["camera_list"] for pid, d in pdict.items()}, bbox


def xz_mask(points: np.ndarray, bbox) -> np.ndarray:
    """extract_point_cloud (data_partition.py:276-283): closed box on x and z."""
    return (points[:, 0] >= bbox[0]) & (points[:, 0] <= bbox[1]) & \
           (points[:, 2] >= bbox[2]) & (points[:, 2] <= bbox[3])


def point_range(p: np.ndarray) -> list:
    """get_point_range (data_partition.py:285-293)."""
    return [p[:, 0].min(), p[:, 0].max(), p[:, 1].min(), p[:, 1].max(), p[:, 2].min(), p[:, 2].max()]


def position_based_selection(points: np.ndarray, cam_lists, bbox, center: np.ndarray,
                             extend_rate: float) -> List[VastPartition]:
    """Position_based_data_selection (data_partition.py:295-365)."""
    out = []
    for pid in cam_lists:
        min_x, max_x, min_z, max_z = bbox[pid]
        ori = [min_x, max_x, min_z, max_z]
        ext = [min_x - extend_rate * (max_x - min_x), max_x + extend_rate * (max_x - min_x),
               min_z - extend_rate * (max_z - min_z), max_z + extend_rate * (max_z - min_z)]
        cams = []
        for _pid, cl in cam_lists.items():         # every camera of every block, dict order (:330-335)
            for c in cl:
                if ext[0] <= center[c, 0] <= ext[1] and ext[2] <= center[c, 2] <= ext[3]:
                    cams.append(c)
        mo = xz_mask(points, ori)
        me = xz_mask(points, ext)
        out.append(VastPartition(pid, np.asarray(cams, np.int64), np.nonzero(me)[0], ori, ext,
                                 point_range(points[mo]), point_range(points[me])))
    return out


# ----------------------------------------------------------------- projection
def corner_points(b) -> np.ndarray:
    """get_8_corner_points (data_partition.py:368-382), dict insertion order."""
    x0, x1, y0, y1, z0, z1 = b
    return np.array([[x0, y0, z0], [x0, y0, z1], [x0, y1, z0], [x0, y1, z1],
                     [x1, y0, z0], [x1, y0, z1], [x1, y1, z0], [x1, y1, z1]])


def point_in_image(cams: VastCameras, c: int, points: np.ndarray):
    """point_in_image (data_partition.py:385-424).  Returns (points_image of the z>0 points,
    mask = indices INTO THAT FILTERED ARRAY of the points inside the image).  Quirks kept:
    cx = W//2, cy = H//2; x is compared with H and y with W (:417-422)."""
    w2c_R = np.transpose(cams.R[c])
    T = cams.T[c]
    W, H = int(cams.width[c]), int(cams.height[c])
    fx = W / (2 * math.tan(cams.FoVx[c] / 2))
    fy = H / (2 * math.tan(cams.FoVy[c] / 2))
    K = np.array([[fx, 0, W // 2], [0, fy, H // 2], [0, 0, 1]])
    pc = (np.dot(w2c_R, points.T) + T.reshape(3, 1)).T
    pc = pc[np.where(pc[:, 2] > 0)]
    pi = np.dot(K, pc.T)
    pi = (pi[:2, :] / pi[2, :]).T
    mask = np.where(np.logical_and.reduce((pi[:, 0] >= 0, pi[:, 0] < H, pi[:, 1] >= 0, pi[:, 1] < W)))[0]
    return pi, mask


# ----------------------------------------------------------------- hull & image box
def convex_hull(pts: np.ndarray) -> np.ndarray:
    """Andrew monotone chain (counter-clockwise, no collinear vertices); stands in for
    scipy ConvexHull(points).vertices (graham_scan.py:85-86)."""
    p = sorted(set((float(a), float(b)) for a, b in pts))
    if len(p) < 3:
        return np.zeros((0, 2))

    def cross(o, a, b):
        return (a[0] - o[0]) * (b[1] - o[1]) - (a[1] - o[1]) * (b[0] - o[0])
    lo, up = [], []
    for q in p:
        while len(lo) >= 2 and cross(lo[-2], lo[-1], q) <= 0:
            lo.pop()
        lo.append(q)
    for q in reversed(p):
        while len(up) >= 2 and cross(up[-2], up[-1], q) <= 0:
            up.pop()
        up.append(q)
    h = lo[:-1] + up[:-1]
    return np.array(h) if len(h) >= 3 else np.zeros((0, 2))


def clip_to_image(ring: np.ndarray, W: float, H: float) -> np.ndarray:
    """Convex ring & box(0, 0, W, H) (graham_scan.py:88-90), Sutherland-Hodgman."""
    r = [tuple(v) for v in ring]
    for axis, bound, le in ((0, 0.0, False), (0, float(W), True), (1, 0.0, False), (1, float(H), True)):
        out = []
        for i in range(len(r)):
            p, q = r[i], r[(i + 1) % len(r)]
            fp = (p[axis] - bound) if le else (bound - p[axis])
            fq = (q[axis] - bound) if le else (bound - q[axis])
            if fp <= 0.0:
                out.append(p)
            if (fp <= 0.0) != (fq <= 0.0):
                t = fp / (fp - fq)
                out.append((p[0] + t * (q[0] - p[0]), p[1] + t * (q[1] - p[1])))
        r = out
        if not r:
            break
    return np.array(r) if r else np.zeros((0, 2))


def shoelace(r: np.ndarray) -> float:
    if len(r) < 3:
        return 0.0
    x, y = r[:, 0], r[:, 1]
    return 0.5 * abs(float(np.sum(x * np.roll(y, -1) - np.roll(x, -1) * y)))


def airspace_rate(cams: VastCameras, c: int, corners: np.ndarray) -> float:
    """The 8-corner visibility of data_partition.py:478-488 + graham_scan.run_graham_scan:
    corners behind the camera are left out; fewer than 3 projected corners or a degenerate
    hull -> 0."""
    proj = []
    for k in range(8):
        pi, _ = point_in_image(cams, c, corners[k:k + 1])
        if len(pi):
            proj.append(pi[0])
    if len(proj) < 3:
        return 0.0
    W, H = int(cams.width[c]), int(cams.height[c])
    hull = convex_hull(np.array(proj))
    if len(hull) < 3:
        return 0.0
    return shoelace(clip_to_image(hull, W, H)) / (W * H)


# ----------------------------------------------------------------- camera / point selection
def visibility_selection(points: np.ndarray, cams: VastCameras, parts: List[VastPartition],
                         visible_rate: float):
    """Visibility_based_camera_selection (data_partition.py:427-557).  Per block i returns
    (camera index list, unique point rows (lexicographic, np.unique axis=0), for each row the
    scene-cloud index of its first occurrence in the reference's concatenation order,
    [(j, camera, rate)] of the cameras that were appended).  The mask quirk of :505-506 is
    kept: the in-image indices refer to the z>0-filtered array but select from the
    unfiltered block cloud."""
    res = []
    for i, pi_ in enumerate(parts):
        cam_list = list(pi_.cameras)
        names = set(cams.names[c] for c in cam_list)
        corners = corner_points(pi_.extend_point_bbox)
        sel = []                                    # scene-cloud indices, concatenation order
        added = []
        for pj in parts:
            if pj.partition_id == pi_.partition_id:
                continue
            pts_j = points[pj.point_index]
            for c in pj.cameras:
                rate = airspace_rate(cams, int(c), corners)
                if rate >= visible_rate:
                    if cams.names[c] in names:
                        continue
                    names.add(cams.names[c])
                    cam_list.append(int(c))
                    _, mask = point_in_image(cams, int(c), pts_j)
                    sel.append(pj.point_index[mask])
                    added.append((pj.partition_id, int(c), rate))
        sel.append(pi_.point_index)
        cat = np.concatenate(sel)
        uniq, first = np.unique(points[cat], return_index=True, axis=0)
        res.append((np.asarray(cam_list, np.int64), uniq, cat[first], added))
    return res


def run_vast(points: np.ndarray, cams: VastCameras, m: int, n: int, extend_rate: float,
             visible_rate: float):
    """ProgressiveDataPartitioning.run_DataPartition (data_partition.py:91-99)."""
    pdict = region_division(cams.center, m, n)
    cam_lists, bbox = refine_ori_bbox(pdict, cams.center, m, n)
    parts = position_based_selection(points, cam_lists, bbox, cams.center, extend_rate)
    return parts, visibility_selection(points, cams, parts, visible_rate)
