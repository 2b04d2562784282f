"""ptgs_compact_xy_static: the allocation-free variant of ptgs_compact_xy (scratch in a file-scope
`static __device__` array of the library instead of a caller workspace).  Called through the C ABI
(ctypes) and compared bit for bit with the existing entry point, on two calls in a row that share the
array.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from ptgs_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _p(t):
    return C.c_void_p(t.data_ptr())


def _existing(lib, x, y, keep, st):
    n = x.shape[0]
    ox, oy = torch.empty_like(x), torch.empty_like(y)
    n_out = torch.zeros(1, dtype=torch.int64, device=DEV)
    ws = torch.empty(max(int(lib.ptgs_compact_xy_workspace_bytes(n)), 256), dtype=torch.uint8, device=DEV)
    dt = _lib.PTGS_F64 if x.dtype == torch.float64 else _lib.PTGS_F32
    rc = lib.ptgs_compact_xy(dt, _p(x), _p(y), _p(keep), n, _p(ox), _p(oy), _p(n_out), _p(ws), ws.numel(), st)
    assert rc == 0, lib.ptgs_last_error()
    torch.cuda.synchronize()
    m = int(n_out.item())
    return ox[:m].clone(), oy[:m].clone(), m


def test_compact_xy_static_equals_workspace_entry_on_consecutive_calls():
    lib = _lib.load()
    torch.cuda.set_device(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rng = np.random.default_rng(101)
    # two calls in a row on the shared array: a large float32 call, then a smaller float64 one whose
    # tiles are a prefix of the first call's (stale entries beyond them must not matter); then the
    # first one again
    cases = [(300_001, torch.float32, 0.7), (70_003, torch.float64, 0.31), (300_001, torch.float32, 0.05)]
    for k, (n, dtype, frac) in enumerate(cases):
        x = torch.from_numpy(rng.standard_normal(n)).to(DEV, dtype)
        y = torch.from_numpy(rng.standard_normal(n)).to(DEV, dtype)
        keep = torch.from_numpy((rng.uniform(size=n) < frac).astype(np.uint8)).to(DEV)
        # every buffer of the call exists before it: the entry point itself allocates nothing
        ox = torch.full_like(x, float("nan"))
        oy = torch.full_like(y, float("nan"))
        n_out = torch.full((1,), -1, dtype=torch.int64, device=DEV)
        dt = _lib.PTGS_F64 if dtype == torch.float64 else _lib.PTGS_F32
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = lib.ptgs_compact_xy_static(dt, _p(x), _p(y), _p(keep), n, _p(ox), _p(oy), _p(n_out), st)
        assert rc == 0, lib.ptgs_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        wx, wy, m = _existing(lib, x, y, keep, st)
        assert int(n_out.item()) == m == int(keep.sum().item()), k
        assert m > 0
        # bit-exact (a copy of selected values, index order kept), as the other compaction tests ask
        assert torch.equal(ox[:m], wx) and torch.equal(oy[:m], wy), k
        assert torch.equal(ox[:m], x[keep.bool()]), k
        assert bool(torch.isnan(ox[m:]).all()) and bool(torch.isnan(oy[m:]).all())   # nothing written past the count


def test_compact_xy_static_empty_and_limit():
    lib = _lib.load()
    torch.cuda.set_device(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    n_out = torch.full((1,), -1, dtype=torch.int64, device=DEV)
    rc = lib.ptgs_compact_xy_static(_lib.PTGS_F32, None, None, None, 0, None, None, _p(n_out), st)
    assert rc == 0 and int(n_out.item()) == 0
    # beyond the built-in scratch: rejected before anything is launched or read
    x = torch.zeros(8, dtype=torch.float32, device=DEV)
    keep = torch.ones(8, dtype=torch.uint8, device=DEV)
    n_out.fill_(-1)
    too_many = _lib.COMPACT_XY_STATIC_MAX_POINTS + 1
    rc = lib.ptgs_compact_xy_static(_lib.PTGS_F32, _p(x), _p(x), _p(keep), too_many, _p(x), _p(x), _p(n_out), st)
    torch.cuda.synchronize()
    assert rc == -5                                   # PTGS_E_WORKSPACE
    msg = lib.ptgs_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "PTGS_COMPACT_XY_STATIC_MAX_POINTS" in msg and str(1 << 26) in msg
    assert int(n_out.item()) == -1
