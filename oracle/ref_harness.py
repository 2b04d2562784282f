"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference modules from
/root/reference under sys.modules stubs so their live code can be executed in
this container (which lacks shapely / open3d / plyfile / matplotlib).

Used by  oracle/make_golden.py  (fixture generation) and by the container-only
tests that validate the restatement in oracle/ptgs_oracle.py against the real
reference.  /root/reference does not exist on the GPU box: nothing that runs
there may import this module (``available()`` gates every use).

Only one stub carries semantics: an axis-aligned rectangle standing in for
shapely's Polygon/box/Point (the reference builds rectangles only:
density_partition.py:60-64, shen_partation_densify.py:234-235, :388).
    .bounds          -> (minx, miny, maxx, maxy) as Python floats
    .contains(Point) -> strict interior (shapely semantics for a polygon)
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REF_ROOT = "/root/reference"


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "scene", "ptgs"))


class _Point:
    def __init__(self, *a):
        if len(a) == 1:
            a = tuple(a[0])
        self.x = float(a[0])
        self.y = float(a[1])
        self.coords = [tuple(float(v) for v in a)]


class _Exterior:
    def __init__(self, pts):
        self.coords = list(pts) + [pts[0]]

    @property
    def xy(self):
        xs = [p[0] for p in self.coords]
        ys = [p[1] for p in self.coords]
        return xs, ys


class _Polygon:
    """Axis-aligned rectangle given as 4 corner vertices (any order)."""
    geom_type = "Polygon"

    def __init__(self, shell=None, holes=None):
        pts = [(float(p[0]), float(p[1])) for p in (shell if shell is not None else [])]
        self._pts = pts
        self.exterior = _Exterior(pts) if pts else None

    @property
    def bounds(self):
        xs = [p[0] for p in self._pts]
        ys = [p[1] for p in self._pts]
        return (min(xs), min(ys), max(xs), max(ys))

    def contains(self, pt):
        minx, miny, maxx, maxy = self.bounds
        return (minx < pt.x < maxx) and (miny < pt.y < maxy)

    @property
    def area(self):
        if getattr(self, "_general", False):      # ring of a convex polygon: shoelace
            a = 0.0
            n = len(self._pts)
            for i in range(n):
                x0, y0 = self._pts[i]
                x1, y1 = self._pts[(i + 1) % n]
                a += x0 * y1 - x1 * y0
            return abs(a) * 0.5
        minx, miny, maxx, maxy = self.bounds
        return (maxx - minx) * (maxy - miny)

    def intersection(self, other):
        """Convex ring (self) clipped by the bounds rectangle of `other` -- the only use the
        Vast path makes of shapely (graham_scan.py:92-94: hull polygon & box(0, 0, W, H)).
        Third-party restatement (shapely is not in the reference tree): Sutherland-Hodgman."""
        minx, miny, maxx, maxy = other.bounds
        ring = list(self._pts)
        for axis, bound, keep_le in ((0, minx, False), (0, maxx, True), (1, miny, False), (1, maxy, True)):
            out = []
            n = len(ring)
            for i in range(n):
                p, q = ring[i], ring[(i + 1) % n]
                fp = (p[axis] - bound) if keep_le else (bound - p[axis])
                fq = (q[axis] - bound) if keep_le else (bound - q[axis])
                if fp <= 0.0:
                    out.append(p)
                if (fp <= 0.0) != (fq <= 0.0):
                    t = fp / (fp - fq)
                    out.append((p[0] + t * (q[0] - p[0]), p[1] + t * (q[1] - p[1])))
            ring = out
            if not ring:
                break
        r = _Polygon(ring)
        r._general = True
        return r

    def __repr__(self):
        return "RectPolygon(%r)" % (self.bounds,)


def _box(minx, miny, maxx, maxy, ccw=True):
    # shapely.geometry.box vertex order: (maxx,miny),(maxx,maxy),(minx,maxy),(minx,miny)
    return _Polygon([(maxx, miny), (maxx, maxy), (minx, maxy), (minx, miny)])


def _from_wkb(data):
    """shapely.from_wkb for the single-ring polygons the partitioner pickles (shapely 2 reduces a
    geometry to from_wkb(bytes)); third-party restatement, shapely is absent here."""
    import struct
    order = "<" if data[0] == 1 else ">"
    gtype, nrings, npts = struct.unpack_from(order + "III", data, 1)
    assert gtype == 3 and nrings == 1
    xy = struct.unpack_from(order + "%dd" % (2 * npts), data, 13)
    return _Polygon([(xy[2 * i], xy[2 * i + 1]) for i in range(npts - 1)])


class _Anything:
    """Attribute sink for plotting / unused third-party APIs."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()

    def __iter__(self):
        return iter(())


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    def _ga(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        return _Anything()
    m.__getattr__ = _ga  # type: ignore[attr-defined]
    sys.modules[name] = m
    return m


_loaded = {}


def load():
    """Returns a namespace of reference modules (cached)."""
    if _loaded:
        return types.SimpleNamespace(**_loaded)
    if not available():
        raise RuntimeError("reference tree not present (GPU box?)")
    sys.dont_write_bytecode = True  # /root/reference is read-only
    import torch  # noqa: F401  (real torch must be imported before the stubs exist)
    import scipy.spatial.transform  # noqa: F401

    shp = _mod("shapely", from_wkb=_from_wkb)
    geo = _mod("shapely.geometry", Polygon=_Polygon, Point=_Point, box=_box,
               LineString=_Anything, MultiPolygon=_Anything)
    _mod("shapely.geometry.geo", box=_box)
    _mod("shapely.geometry.polygon", Polygon=_Polygon)
    _mod("shapely.vectorized", contains=_Anything())
    _mod("shapely.strtree", STRtree=_Anything)
    _mod("shapely.ops")
    shp.geometry = geo
    plt = _mod("matplotlib.pyplot", subplots=lambda *a, **k: (_Anything(), _Anything()))
    mpl = _mod("matplotlib", pyplot=plt)
    _mod("matplotlib.patches")
    _mod("matplotlib.font_manager")
    mpl.use = lambda *a, **k: None
    _mod("open3d")
    _mod("plyfile", PlyData=_Anything(), PlyElement=_Anything())
    _mod("simple_knn")
    _mod("simple_knn._C", distCUDA2=_Anything())
    _mod("diff_gaussian_rasterization")
    _mod("trimesh")
    _mod("trimesh.path")
    _mod("trimesh.path.packing")

    for p in (os.path.join(REF_ROOT, "scene", "ptgs"), REF_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    # `import scene` must not run scene/__init__.py (pulls the 3DGS trainer);
    # register bare namespace packages instead.
    for pkg, path in (("scene", os.path.join(REF_ROOT, "scene")),
                      ("scene.ptgs", os.path.join(REF_ROOT, "scene", "ptgs")),
                      ("utils", os.path.join(REF_ROOT, "utils"))):
        m = types.ModuleType(pkg)
        m.__path__ = [path]
        sys.modules[pkg] = m
    gm = types.ModuleType("scene.gaussian_model")
    from typing import NamedTuple
    import numpy as np

    class BasicPointCloud(NamedTuple):  # utils/graphics_utils.py:17-20
        points: np.array
        colors: np.array
        normals: np.array
    gm.BasicPointCloud = BasicPointCloud
    sys.modules["scene.gaussian_model"] = gm

    # a drop-in module of the same name (compat/) must not stand in for the reference
    ref_ptgs = os.path.join(REF_ROOT, "scene", "ptgs")
    for name in ("shen_partation_densify", "shen_partition_utils", "density_partition", "shen_data_read"):
        m = sys.modules.get(name)
        if m is not None and not os.path.abspath(getattr(m, "__file__", "") or "").startswith(ref_ptgs):
            del sys.modules[name]
    densify = importlib.import_module("shen_partation_densify")
    if not os.path.abspath(densify.__file__).startswith(ref_ptgs):
        raise ImportError("shen_partation_densify resolved to %s, not the reference" % densify.__file__)
    _loaded.update(
        densify=densify,
        density_partition=importlib.import_module("density_partition"),
        vast_data_partition=importlib.import_module("scene.ptgs.data_partition"),
        vast_graham_scan=importlib.import_module("scene.ptgs.graham_scan"),
        save_partition=importlib.import_module("scene.ptgs.save_partition"),
        read_write_model=importlib.import_module("scene.ptgs.read_write_model"),
        cameras=importlib.import_module("scene.cameras"),
        colmap_loader=importlib.import_module("scene.colmap_loader"),
        shen_data_read=importlib.import_module("shen_data_read"),
        camera_utils=None,
        manhattan_utils=importlib.import_module("utils.manhattan_utils"),
        graphics_utils=importlib.import_module("utils.graphics_utils"),
        Polygon=_Polygon, Point=_Point, box=_box,
        BasicPointCloud=densify.BasicPointCloud,
    )
    # densify.visibility_based_camera_selection uses `concurrent.futures`
    # without importing it (it leaks in through torch); make sure it is there.
    import concurrent.futures  # noqa: F401
    if not hasattr(densify, "concurrent"):
        import concurrent
        densify.concurrent = concurrent
    return types.SimpleNamespace(**_loaded)


def make_simple_camera(ref, uid, colmap_id, R, T, FoVx, FoVy, name, width, height):
    """utils/camera_utils.py:130-135 (loadCamPartition) with data_device='cpu'."""
    return ref.cameras.SimpleCamera(
        colmap_id=colmap_id, R=R, T=T, FoVx=FoVx, FoVy=FoVy, image_name=name,
        uid=uid, width=width, height=height, data_device="cpu",
        points3D_ids=None)
