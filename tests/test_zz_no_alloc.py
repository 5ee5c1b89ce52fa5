# -*- coding: utf-8 -*-
"""
``pyodesys_cuda_compact_rows_noalloc`` (include/pyodesys_cuda.h): the compaction stage of the adaptive path on
caller-owned device buffers, scan scratch in a static ``__device__`` array - for callers that cannot call cudaMalloc.
No counterpart in the reference; the checker is the existing entry
point ``pyodesys_cuda_integrate_adaptive_compact`` + ``pyodesys_cuda_fetch_compact_rows`` on the same batch.
Integer offsets and copied rows: bit-exact.  Collected last on purpose.
"""
import ctypes

import numpy as np
import pytest

from pyodesys_b200 import systems
from systems_for_tests import full_params

pytestmark = pytest.mark.gpu


def _existing_entry_point(mod, b, x0, x1, n):
    pieces, status, counters, _t = mod.integrate_adaptive_compact(
        b['y0'], x0, x1, full_params(b['y0'], b['params']), b['atol'], b['rtol'], mxsteps=b['nsteps'], method='dopri5')
    assert len(pieces) == 1 and (status == 0).all()
    first, rows, offsets = pieces[0]
    assert first == 0 and offsets.shape == (n + 1,)
    return np.array(rows), np.array(offsets)


def test_compact_rows_noalloc_matches_compact_entry_point_twice(cuda_device):
    import torch
    dev = torch.device('cuda', 0)
    vd = systems.vdp()
    mod = vd._native.mod
    lib = mod.lib
    ny = 2
    outs = []
    # two calls in a row with different batches (the second SMALLER: whatever the first left in the static scratch
    # beyond the part the second uses must not matter, and the part it uses is cleared by the call)
    for n, seed, max_rows in ((6000, 11, 1001), (2500, 12, 1001)):
        b = systems.batch_vdp(n=n, seed=seed)
        x0, x1 = np.zeros(n), np.full(n, 10.0)
        rows_ref, off_ref = _existing_entry_point(mod, b, x0, x1, n)
        # device-resident adaptive call (ptr_kind = 1): bounded buffers xyout[n, max_rows, 1 + ny] + nrows[n]
        d_y0 = torch.from_numpy(b['y0']).to(dev)
        d_p = torch.from_numpy(full_params(b['y0'], b['params'])).to(dev)
        d_x0, d_x1 = torch.from_numpy(x0).to(dev), torch.from_numpy(x1).to(dev)
        d_xy = torch.full((n, max_rows, 1 + ny), float('nan'), dtype=torch.float64, device=dev)
        d_nrows = torch.empty(n, dtype=torch.int32, device=dev)
        d_status = torch.empty(n, dtype=torch.int32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        mod.integrate_adaptive_device(n, d_y0.data_ptr(), d_x0.data_ptr(), d_x1.data_ptr(), d_p.data_ptr(), b['atol'],
                                      b['rtol'], max_rows, d_xy.data_ptr(), d_nrows.data_ptr(), d_status.data_ptr(),
                                      mxsteps=b['nsteps'], method='dopri5', device=0, stream=stream)
        assert int(d_status.abs().sum()) == 0
        total = int(off_ref[-1])
        # every buffer of the new entry point exists BEFORE the call: it allocates nothing
        d_off = torch.full((n + 1,), -1, dtype=torch.int64, device=dev)
        d_rows = torch.full((total, 1 + ny), float('nan'), dtype=torch.float64, device=dev)
        torch.cuda.synchronize(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        # straight through the C ABI
        rc = lib.pyodesys_cuda_compact_rows_noalloc(
            0, n, max_rows, ctypes.cast(ctypes.c_void_p(d_xy.data_ptr()), ctypes.POINTER(ctypes.c_double)),
            ctypes.cast(ctypes.c_void_p(d_nrows.data_ptr()), ctypes.POINTER(ctypes.c_int32)),
            ctypes.cast(ctypes.c_void_p(d_off.data_ptr()), ctypes.POINTER(ctypes.c_int64)),
            ctypes.cast(ctypes.c_void_p(d_rows.data_ptr()), ctypes.POINTER(ctypes.c_double)), total,
            ctypes.c_void_p(stream))
        assert rc == 0, lib.pyodesys_cuda_last_error()
        torch.cuda.synchronize(dev)
        if outs:  # second call: every kernel of the path is loaded by now, the call must not take device memory
            assert torch.cuda.mem_get_info(dev)[0] >= free_before
        off = d_off.cpu().numpy()
        rows = d_rows.cpu().numpy()
        # tolerance: none - integer offsets and copied doubles are bit-exact against the existing entry point
        assert np.array_equal(off, off_ref)
        assert rows.shape == rows_ref.shape and np.array_equal(rows, rows_ref)
        assert np.array_equal(np.diff(off), d_nrows.cpu().numpy())
        outs.append(total)
    assert outs[0] != outs[1]


def test_compact_rows_noalloc_limits(cuda_device):
    import torch
    dev = torch.device('cuda', 0)
    mod = systems.vdp()._native.mod
    nmax = int(8192*2048)
    # beyond the static scratch: refused with a message, before anything is touched
    with pytest.raises(RuntimeError, match='at most'):
        mod.compact_rows_noalloc(nmax + 1, 4, 1, 1, 1, 1, 0)
    # a capacity that is too small drops the rows beyond it and still reports the full count in offsets[n]
    n, max_rows = 5, 4
    d_xy = torch.arange(n*max_rows*3, dtype=torch.float64, device=dev).reshape(n, max_rows, 3)
    d_nrows = torch.tensor([1, 4, 0, 2, 3], dtype=torch.int32, device=dev)
    d_off = torch.empty(n + 1, dtype=torch.int64, device=dev)
    d_rows = torch.full((6, 3), -1.0, dtype=torch.float64, device=dev)
    mod.compact_rows_noalloc(n, max_rows, d_xy.data_ptr(), d_nrows.data_ptr(), d_off.data_ptr(), d_rows.data_ptr(), 6)
    torch.cuda.synchronize(dev)
    assert d_off.cpu().tolist() == [0, 1, 5, 5, 7, 10]
    xy = d_xy.cpu().numpy()
    want = np.concatenate([xy[0, :1], xy[1, :4], xy[3, :1]])
    assert np.array_equal(d_rows.cpu().numpy(), want)
    # offsets only (rows_capacity = 0, no rows buffer), n = 0
    d_off.fill_(-7)
    mod.compact_rows_noalloc(n, max_rows, d_xy.data_ptr(), d_nrows.data_ptr(), d_off.data_ptr(), None, 0)
    torch.cuda.synchronize(dev)
    assert d_off.cpu().tolist() == [0, 1, 5, 5, 7, 10]
    mod.compact_rows_noalloc(0, max_rows, None, None, d_off.data_ptr(), None, 0)
    torch.cuda.synchronize(dev)
    assert int(d_off[0]) == 0
