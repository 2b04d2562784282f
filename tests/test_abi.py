"""C-ABI checks that need no GPU: libptgs.so loads, exports every function
include/ptgs.h declares (and nothing is declared that the ctypes table
lacks), the signatures carry no torch types, and the product path refuses to
run without CUDA instead of falling back to anything on the CPU."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

import ptgs_b200  # noqa: F401
from ptgs_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ptgs.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    return sorted(set(re.findall(r"\b(ptgs_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_functions():
    names = _declared()
    assert len(names) >= 25
    for must in ("ptgs_radius_outlier_mask", "ptgs_balanced_split", "ptgs_expand_fill", "ptgs_assign_cameras",
                 "ptgs_voxel_downsample", "ptgs_visibility_count", "ptgs_visibility_hull", "ptgs_select_cameras",
                 "ptgs_coverage_sweep", "ptgs_manhattan_apply", "ptgs_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = _lib.load(build_if_missing=True)
    for name in _declared():
        assert hasattr(lib, name), "libptgs.so does not export %s" % name


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _declared()


def test_header_is_plain_c():
    """The header must compile as C (no torch / C++ types in the signatures)."""
    r = subprocess.run(["gcc", "-std=c99", "-fsyntax-only", "-x", "c", "-I/usr/local/cuda/include", HEADER],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    src = open(HEADER).read()
    assert "at::Tensor" not in src and "std::" not in src


def test_version_and_error_string_without_gpu():
    lib = _lib.load()
    assert lib.ptgs_version() >= 1
    msg = lib.ptgs_last_error()
    assert isinstance(msg, (bytes, type(None)))


def test_no_cpu_fallback_in_product_path():
    """Nothing under the package imports the oracle, and smoke() refuses to run
    without a device."""
    pkg = os.path.join(ROOT, "based-on-point-cloud-partitions_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "ptgs_oracle" not in src and "ref_harness" not in src and "cpu_arm" not in src, fn
    import torch
    if not torch.cuda.is_available():
        sys.path.insert(0, ROOT)
        import __graft_entry__ as G
        with pytest.raises(RuntimeError):
            G.smoke()
