"""Drop-in mirror of the reference's second entry point, the camera-grid ("Vast")
partitioner: scene/ptgs/data_partition.py (`ProgressiveDataPartitioning`, `CameraPose`,
`CameraPartition`) and utils/partition_utils.py (`data_partition`, `read_camList`) --
SURVEY.md row A14.

Same names, arguments, NamedTuples and output files as the reference; the device work
(XZ box selection, point ranges, 8-corner airspace visibility, coverage with the
reference's index quirk, row dedup) runs through libptgs (`ptgs_expand_*` on (x, z),
`ptgs_vast_*`).  Camera-count-sized host logic (sorting cameras into the m x n grid, the
seam refinement, the accepted-camera bookkeeping) stays in Python with the reference's
own expressions so that float32 / float64 promotions match the installed NumPy.
There is no CPU fallback: without a CUDA device the constructor raises."""
from __future__ import annotations

import os
import pickle
from typing import NamedTuple

import numpy as np
import torch

from . import colmap_io, device as D
from .partition import BasicPointCloud, _dev


class CameraPose(NamedTuple):           # data_partition.py:17-19
    camera: object
    pose: np.ndarray


class CameraPartition(NamedTuple):      # data_partition.py:22-32
    partition_id: str
    cameras: list
    point_cloud: BasicPointCloud
    ori_camera_bbox: list
    extend_camera_bbox: list
    extend_rate: float
    ori_point_bbox: list
    extend_point_bbox: list


def _pose_of(cam) -> np.ndarray:
    c = cam.camera_center
    return c.detach().cpu().numpy().copy() if isinstance(c, torch.Tensor) else np.array(c)


class ProgressiveDataPartitioning:
    """data_partition.py:35-557.  `partition_scene` holds the final list of CameraPartition.
    write_outputs=False skips the PLY / log / pickle files (the arithmetic is the same);
    `plots` (pcd.png, camera_on_pcd.png, regions.png) are matplotlib output and not produced."""

    def __init__(self, scene_info, train_cameras, model_path, m_region=2, n_region=4, extend_rate=0.2,
                 visible_rate=0.25, write_outputs=True):
        self.partition_scene = None
        self.pcd = scene_info.point_cloud
        self.model_path = model_path
        self.partition_dir = os.path.join(model_path, "partition_point_cloud")
        self.partition_ori_dir = os.path.join(self.partition_dir, "ori")
        self.partition_extend_dir = os.path.join(self.partition_dir, "extend")
        self.partition_visible_dir = os.path.join(self.partition_dir, "visible")
        self.save_partition_data_dir = os.path.join(self.model_path, "partition_data.pkl")
        self.m_region, self.n_region = m_region, n_region
        self.extend_rate, self.visible_rate = extend_rate, visible_rate
        self.write_outputs = write_outputs
        self.appended = []          # (partition_i, partition_j, camera name, rate) of every appended camera
        if write_outputs:
            for d in (self.partition_ori_dir, self.partition_extend_dir, self.partition_visible_dir):
                os.makedirs(d, exist_ok=True)
        self.run_DataPartition(train_cameras)

    # ------------------------------------------------------------------ orchestration
    def run_DataPartition(self, train_cameras):           # :91-99
        if self.write_outputs and os.path.exists(self.save_partition_data_dir):
            self.partition_scene = self.load_partition_data()
            return
        self._upload()
        partition_dict = self.Camera_position_based_region_division(train_cameras)
        partition_dict, refined_ori_bbox = self.refine_ori_bbox(partition_dict)
        partition_list = self.Position_based_data_selection(partition_dict, refined_ori_bbox)
        self.partition_scene = self.Visibility_based_camera_selection(partition_list)
        if self.write_outputs:
            self.save_partition_data()

    def save_partition_data(self):                        # :102-105
        with open(self.save_partition_data_dir, "wb") as f:
            pickle.dump(self.partition_scene, f)

    def load_partition_data(self):                        # :107-111
        with open(self.save_partition_data_dir, "rb") as f:
            return pickle.load(f)

    def _upload(self):
        dev = _dev()
        pts = np.ascontiguousarray(self.pcd.points)
        if pts.dtype not in (np.float32, np.float64):
            pts = pts.astype(np.float64)
        self._pts_d = torch.from_numpy(pts).to(dev)
        self._x, self._y, self._z = D.aos_to_soa(self._pts_d)
        # colours / normals ride along so that block clouds are gathered on the device
        # (a NumPy fancy index of ~10^7 rows in coordinate order costs seconds per block)
        self._col_d = torch.from_numpy(np.ascontiguousarray(self.pcd.colors)).to(dev)
        self._nrm_d = torch.from_numpy(np.ascontiguousarray(self.pcd.normals)).to(dev)
        self._dev = dev

    def _gather(self, idx_d):
        """points / colors / normals rows `idx_d` (device int tensor) -> host arrays in the cloud's dtypes."""
        i = idx_d.to(torch.int64)
        return (self._pts_d.index_select(0, i).cpu().numpy(), self._col_d.index_select(0, i).cpu().numpy(),
                self._nrm_d.index_select(0, i).cpu().numpy())

    def _store_ply(self, path, xyz, rgb):
        if self.write_outputs:
            colmap_io.store_ply(path, np.asarray(xyz).reshape(-1, 3), np.asarray(rgb).reshape(-1, 3))

    # ------------------------------------------------------------------ 1. camera grid
    def Camera_position_based_region_division(self, train_cameras):      # :220-274
        m, n = self.m_region, self.n_region
        poses = [CameraPose(camera=c, pose=_pose_of(c)) for c in train_cameras]
        self._store_ply(os.path.join(self.partition_dir, "camera_centers.ply"),
                        np.array([p.pose for p in poses]), np.zeros((len(poses), 3)))
        m_dict = {}
        total = len(poses)
        per_m = total // m
        by_x = sorted(poses, key=lambda p: p.pose[0])
        for i in range(m):
            d = {"camera_list": by_x[i * per_m:(i + 1) * per_m]}
            d["x_mid_camera"] = by_x[(i + 1) * per_m - 1] if i != m - 1 else None
            m_dict[str(i + 1)] = d
        if total % m != 0:
            m_dict[str(m)]["camera_list"].extend(by_x[m * per_m:])
        out = {}
        for pid, cams in m_dict.items():
            tot = len(cams["camera_list"])
            per_n = tot // n
            by_z = sorted(cams["camera_list"], key=lambda p: p.pose[2])
            for i in range(n):
                d = {"camera_list": by_z[i * per_n:(i + 1) * per_n], "x_mid_camera": cams["x_mid_camera"]}
                d["z_mid_camera"] = by_z[(i + 1) * per_n - 1] if i != n - 1 else None
                out["%s_%d" % (pid, i + 1)] = d
            if tot % n != 0:
                out["%s_%d" % (pid, n)]["camera_list"].extend(by_z[n * per_n:])
        return out

    def refine_ori_bbox(self, partition_dict):                            # :179-218
        bbox = {}
        for pid, d in partition_dict.items():
            cl = d["camera_list"]
            bbox[pid] = [min(c.pose[0] for c in cl), max(c.pose[0] for c in cl),
                         min(c.pose[2] for c in cl), max(c.pose[2] for c in cl)]
        for m in range(1, self.m_region + 1):
            for n in range(1, self.n_region):
                up, lo = "%d_%d" % (m, n + 1), "%d_%d" % (m, n)
                a, b = bbox[up], bbox[lo]
                mid_z = partition_dict[lo]["z_mid_camera"].pose[2]
                bbox[up] = [a[0], a[1], mid_z, a[3]]
                bbox[lo] = [b[0], b[1], b[2], mid_z]
        for n in range(1, self.n_region + 1):
            for m in range(1, self.m_region):
                le, ri = "%d_%d" % (m, n), "%d_%d" % (m + 1, n)
                a, b = bbox[le], bbox[ri]
                mid_x = partition_dict[le]["x_mid_camera"].pose[0]
                bbox[le] = [a[0], mid_x, a[2], a[3]]
                bbox[ri] = [mid_x, b[1], b[2], b[3]]
        return {pid: d["camera_list"] for pid, d in partition_dict.items()}, bbox

    # ------------------------------------------------------------------ 2. position-based selection
    def extract_point_cloud(self, pcd, bbox):                             # :276-283 (one box)
        off, idx = D.expand_select(self._x, self._z, np.array([[bbox[0], bbox[2], bbox[1], bbox[3]]], np.float64))
        i = idx.cpu().numpy()
        return pcd.points[i], pcd.colors[i], pcd.normals[i]

    def Position_based_data_selection(self, partition_dict, refined_ori_bbox):   # :295-365
        pcd = self.pcd
        ids = list(partition_dict)
        er = self.extend_rate
        ori_boxes, ext_boxes, cam_lists = [], [], []
        all_poses = [cp for cl in partition_dict.values() for cp in cl]
        # float64 on both sides: exact for float32 poses against float32 or float64 box values
        px = np.array([cp.pose[0] for cp in all_poses], dtype=np.float64)
        pz = np.array([cp.pose[2] for cp in all_poses], dtype=np.float64)
        for pid in ids:
            min_x, max_x, min_z, max_z = refined_ori_bbox[pid]
            ori = [min_x, max_x, min_z, max_z]
            ext = [min_x - er * (max_x - min_x), max_x + er * (max_x - min_x),
                   min_z - er * (max_z - min_z), max_z + er * (max_z - min_z)]
            print("Partition", pid, "ori_camera_bbox", ori, "\textend_camera_bbox", ext)
            inside = (float(ext[0]) <= px) & (px <= float(ext[1])) & (float(ext[2]) <= pz) & (pz <= float(ext[3]))
            cam_lists.append([all_poses[k] for k in np.nonzero(inside)[0]])
            ori_boxes.append(ori)
            ext_boxes.append(ext)
        # closed XZ boxes over the whole cloud (expand_select takes {min a, min b, max a, max b})
        as_rect = lambda b: [float(b[0]), float(b[2]), float(b[1]), float(b[3])]
        o_off, o_idx = D.expand_select(self._x, self._z, np.array([as_rect(b) for b in ori_boxes], np.float64))
        e_off, e_idx = D.expand_select(self._x, self._z, np.array([as_rect(b) for b in ext_boxes], np.float64))
        o_rng = D.vast_point_range(self._x, self._y, self._z, o_off, o_idx)
        e_rng = D.vast_point_range(self._x, self._y, self._z, e_off, e_idx)
        o_off_h, e_off_h = o_off.cpu().numpy(), e_off.cpu().numpy()
        self._ext_index = [e_idx[e_off_h[b]:e_off_h[b + 1]] for b in range(len(ids))]
        sc = pcd.points.dtype.type
        out = []
        n_ori = n_ext = 0
        for b, pid in enumerate(ids):
            n_o, n_e = int(o_off_h[b + 1] - o_off_h[b]), int(e_off_h[b + 1] - e_off_h[b])
            if n_o == 0 or n_e == 0:
                raise ValueError("min() arg is an empty sequence")       # get_point_range of an empty block (:289)
            self._store_ply(os.path.join(self.partition_ori_dir, "%s_camera_centers.ply" % pid),
                            np.array([cp.pose for cp in partition_dict[pid]]), np.zeros((len(partition_dict[pid]), 3)))
            self._store_ply(os.path.join(self.partition_extend_dir, "%s_camera_centers.ply" % pid),
                            np.array([cp.pose for cp in cam_lists[b]]), np.zeros((len(cam_lists[b]), 3)))
            pts_e, col_e, nrm_e = self._gather(self._ext_index[b])
            out.append(CameraPartition(partition_id=pid, cameras=cam_lists[b],
                                       point_cloud=BasicPointCloud(pts_e, col_e, nrm_e),
                                       ori_camera_bbox=ori_boxes[b], extend_camera_bbox=ext_boxes[b],
                                       extend_rate=self.extend_rate,
                                       ori_point_bbox=[sc(v) for v in o_rng[b]],
                                       extend_point_bbox=[sc(v) for v in e_rng[b]]))
            n_ori += n_o
            n_ext += n_e
            if self.write_outputs:
                pts_o, col_o, _ = self._gather(o_idx[o_off_h[b]:o_off_h[b + 1]])
                self._store_ply(os.path.join(self.partition_ori_dir, "%s.ply" % pid), pts_o, col_o)
                self._store_ply(os.path.join(self.partition_extend_dir, "%s_extend.ply" % pid), pts_e, col_e)
        print("Total ori point number: %d\n" % pcd.points.shape[0], "Total before extend point number: %d\n" % n_ori,
              "Total extend point number: %d\n" % n_ext)
        return out

    # ------------------------------------------------------------------ 3. visibility + coverage
    def get_8_corner_points(self, bbox):                                   # :368-382
        x0, x1, y0, y1, z0, z1 = bbox
        return {"minx_miny_minz": [x0, y0, z0], "minx_miny_maxz": [x0, y0, z1],
                "minx_maxy_minz": [x0, y1, z0], "minx_maxy_maxz": [x0, y1, z1],
                "maxx_miny_minz": [x1, y0, z0], "maxx_miny_maxz": [x1, y0, z1],
                "maxx_maxy_minz": [x1, y1, z0], "maxx_maxy_maxz": [x1, y1, z1]}

    def Visibility_based_camera_selection(self, partition_list):          # :427-557
        pcd = self.pcd
        B = len(partition_list)
        # distinct camera objects -> device records
        cam_of, cams = {}, []
        for p in partition_list:
            for cp in p.cameras:
                if id(cp.camera) not in cam_of:
                    cam_of[id(cp.camera)] = len(cams)
                    cams.append(cp.camera)
        recs = D.vast_camera_records([c.R for c in cams], [c.T for c in cams], [c.FoVx for c in cams],
                                     [c.FoVy for c in cams], [c.image_width for c in cams],
                                     [c.image_height for c in cams])
        cams_d = torch.from_numpy(recs).to(self._dev)
        # every (block i, camera of block j != i) pair at once
        boxes = np.array([[float(v) for v in p.extend_point_bbox] for p in partition_list], np.float64)
        pb, pc, span = [], [], {}
        for i in range(B):
            for j in range(B):
                if partition_list[i].partition_id == partition_list[j].partition_id:
                    continue
                ids = [cam_of[id(cp.camera)] for cp in partition_list[j].cameras]
                span[(i, j)] = (len(pb), len(ids))
                pb.extend([i] * len(ids))
                pc.extend(ids)
        rate = D.vast_airspace(cams_d, boxes, np.asarray(pb, np.int32), np.asarray(pc, np.int32))

        out = []
        for i, part_i in enumerate(partition_list):
            pid_i = part_i.partition_id
            if self.write_outputs:
                corners = list(self.get_8_corner_points(part_i.extend_point_bbox).values())
                self._store_ply(os.path.join(self.partition_extend_dir, "%s_corner_points.ply" % pid_i),
                                np.array(corners), np.zeros((8, 3)))
            cameras = list(part_i.cameras)
            names = set(cp.camera.image_name for cp in cameras)
            sources = []
            total_added = 0
            for j, part_j in enumerate(partition_list):
                if part_j.partition_id == pid_i:
                    continue
                print("Now processing partition i:%s and j:%s" % (pid_i, part_j.partition_id))
                s0, cnt = span[(i, j)]
                accepted, added = [], 0
                for k, cp in enumerate(part_j.cameras):
                    r = rate[s0 + k]
                    if r >= self.visible_rate:
                        if cp.camera.image_name in names:
                            continue
                        names.add(cp.camera.image_name)
                        added += 1
                        cameras.append(cp)
                        accepted.append(cam_of[id(cp.camera)])
                        self.appended.append((pid_i, part_j.partition_id, cp.camera.image_name, float(r)))
                        if self.write_outputs:
                            W, H = cp.camera.image_width, cp.camera.image_height
                            with open(os.path.join(self.model_path, "graham_scan"), "a") as f:
                                f.write("intersection_area:%s\timage_area:%s\tintersection_rate:%s\tpartition_i:%s\t"
                                        "partition_j:%s\tappend_camera_id:%s\tappend_camera_count:%d\n"
                                        % (float(r) * (W * H), W * H, float(r), pid_i, part_j.partition_id,
                                           cp.camera.image_name, added))
                sources.append((self._ext_index[j], accepted))
                total_added += added
            if self.write_outputs:
                with open(os.path.join(self.model_path, "partition_cameras"), "a") as f:
                    f.write("partition_id:%s\ttotal_append_camera_count:%d\ttotal_camera:%d\n"
                            % (pid_i, total_added, len(cameras)))
                self._store_ply(os.path.join(self.partition_visible_dir, "%s_camera_centers.ply" % pid_i),
                                np.array([cp.pose for cp in cameras]), np.zeros((len(cameras), 3)))
            uniq = D.vast_block_cloud(self._x, self._y, self._z, cams_d, self._ext_index[i], sources)
            new_points, new_colors, new_normals = self._gather(uniq)
            out.append(part_i._replace(cameras=cameras,
                                       point_cloud=BasicPointCloud(new_points, new_colors, new_normals)))
            self._store_ply(os.path.join(self.partition_visible_dir, "%s_visible.ply" % pid_i), new_points, new_colors)
        return out


def read_camList(path):                                   # utils/partition_utils.py:54-61
    with open(path, "r") as f:
        return [line.replace("\n", "") for line in f.readlines()]


def data_partition(lp, scene_info=None, all_cameras=None):
    """utils/partition_utils.py:13-51.  `lp` carries source_path, images, man_trans, eval, llffhold,
    model_path, m_region, n_region, extend_rate, visible_rate.  The reference reads the scene through
    sceneLoadTypeCallbacks["Partition"] (scene/dataset_readers.py) -- pass `scene_info` (with
    train_cameras / test_cameras / point_cloud) and `all_cameras` built by that reader; without them
    the COLMAP reader of this package is used (every camera is a training camera)."""
    if scene_info is None:
        from .partition import read_scene, SimpleCamera
        scene_info = read_scene(lp.source_path)
        # this reader stores the transposed rotation (scene/dataset_readers.py:90)
        all_cameras = [SimpleCamera(colmap_id=c.uid, R=np.transpose(c.R), T=c.T, FoVx=c.FovX, FoVy=c.FovY,
                                    image_name=c.image_name, uid=i, width=c.width, height=c.height)
                       for i, c in enumerate(scene_info.train_cameras + list(scene_info.test_cameras))]
    os.makedirs(lp.model_path, exist_ok=True)
    with open(os.path.join(lp.model_path, "train_cameras.txt"), "w") as f:
        for cam in scene_info.train_cameras:
            f.write("%s\n" % cam.image_name)
    with open(os.path.join(lp.model_path, "test_cameras.txt"), "w") as f:
        for cam in scene_info.test_cameras:
            f.write("%s\n" % cam.image_name)
    dp = ProgressiveDataPartitioning(scene_info, all_cameras, lp.model_path, lp.m_region, lp.n_region,
                                     lp.extend_rate, lp.visible_rate)
    client = 0
    partition_id_list = []
    for partition in dp.partition_scene:
        partition_id_list.append(partition.partition_id)
        txt = "%s/partition_point_cloud/visible/%s_camera.txt" % (lp.model_path, partition.partition_id)
        with open(txt, "w") as f:
            for cp in partition.cameras:
                f.write("%s\n" % (cp.camera.image_name + ".jpg"))
        client += 1
    return client, partition_id_list
