# -*- coding: utf-8 -*-
"""
The ``cuda`` native back end: code generation, on-the-fly compilation for sm_100a and the
ctypes binding of the C-ABI shim.

It sits where ``pyodesys/native/{cvode,odeint,gsl}.py`` sit in the reference:

* :class:`NativeCudaCode`  <->  ``_NativeCodeBase`` (``pyodesys/native/_base.py:54-281``): renders
  the system specific source, compiles it once (cached by content hash) and exposes the
  loaded module as ``.mod`` with ``integrate_adaptive`` / ``integrate_predefined``;
* :class:`NativeCudaSys`   <->  ``_NativeSysBase`` (``_base.py:284-342``): a ``SymbolicSys`` whose
  ``integrate`` is forced onto ``_integrate_native``.

There is no CPU fallback: without a CUDA device the integrate calls raise.
"""
from __future__ import annotations

import ctypes
import hashlib
import logging
import os
import shutil
import subprocess
import threading
import time

import numpy as np

from ..core import BatchInfo
from ..symbolic import SymbolicSys
from . import _codegen

logger = logging.getLogger(__name__)

_here = os.path.dirname(os.path.abspath(__file__))
_pkg = os.path.dirname(_here)
_csrc = os.path.join(_pkg, 'csrc')
_include = os.path.join(os.path.dirname(_pkg), 'include')

METHODS = {'rk4': 0, 'dopri5': 1, 'rosenbrock4': 2, 'ros4': 2, 'rodas': 2, 'bdf': 3}
STATUS_TEXT = {0: 'success', 1: 'too many steps (nsteps/mxsteps exceeded)', 2: 'step size underflow',
               3: 'rhs/jac error or non-finite state', 4: 'adaptive output buffer exhausted'}
NCOUNTERS = 6
NTIMINGS = 4


class Options(ctypes.Structure):
    """ ``pyodesys_cuda_options`` of include/pyodesys_cuda.h """
    _fields_ = [('get_dx_max_factor', ctypes.c_double),
                ('max_invariant_violation', ctypes.c_double),
                ('error_outside_bounds', ctypes.c_int32),
                ('xout_shared', ctypes.c_int32),
                ('schedule', ctypes.c_int32),
                ('block_size', ctypes.c_int32),
                ('blocks_per_sm', ctypes.c_int32),
                ('yout_soa', ctypes.c_int32),
                ('stream', ctypes.c_void_p),
                ('iv_from_y0', ctypes.c_int32),
                ('params_ld', ctypes.c_int32),
                ('kernel', ctypes.c_int32),
                ('return_on_root', ctypes.c_int32),
                ('max_roots', ctypes.c_int32),
                ('linear_solver', ctypes.c_int32),
                ('root_out', ctypes.c_void_p),
                ('root_indices', ctypes.c_void_p),
                ('nroots_found', ctypes.c_void_p),
                ('counters_ld', ctypes.c_int64),
                ('ctx_lane', ctypes.c_int32),
                ('reserved4', ctypes.c_int32)]


def _linear_solver_id(linear_solver='default', with_jtimes=False):
    key = str(linear_solver).lower()
    table = {'default': 0, 'dense': 0, 'klu': 0, 'lapack': 0, 'gmres': 1}
    if key not in table:
        raise ValueError("Unknown linear_solver: %s (dense, klu, gmres)" % linear_solver)
    if table[key] == 1 and not with_jtimes:
        raise ValueError("linear_solver='gmres' is matrix-free here: pass with_jtimes=True (system with jtimes=True)")
    return table[key]


def cache_dir():
    return os.environ.get('PYODESYS_CUDA_CACHE', os.path.join(_pkg, '_build'))


def _csrc_digest():
    h = hashlib.sha256()
    for fn in ('integrators.cuh', 'bdf_integrator.cuh', 'group_integrators.cuh', 'coop_integrators.cuh', 'coop_bdf.cuh', 'smem_integrators.cuh', 'pyodesys_cuda_shim.cu', 'odesys_prelude.h'):
        path = os.path.join(_csrc, fn)
        if os.path.exists(path):
            with open(path, 'rb') as fh:
                h.update(fh.read())
    with open(os.path.join(_include, 'pyodesys_cuda.h'), 'rb') as fh:
        h.update(fh.read())
    return h.hexdigest()[:12]


def nvcc_command(src_dir, out, arch=None):
    arch = arch or os.environ.get('PYODESYS_CUDA_ARCH', 'sm_100a')
    comp = arch.replace('sm_', 'compute_')
    nvcc = os.environ.get('NVCC', shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc')
    extra = os.environ.get('PYODESYS_CUDA_NVCC_FLAGS', '').split()
    return [nvcc, '-gencode', 'arch=%s,code=%s' % (comp, arch), '-O3', '-lineinfo', '-std=c++17',
            '--shared', '-Xcompiler', '-fPIC', '-Xptxas', '-v'] + extra + [
            '-I', src_dir, '-I', _csrc, '-I', _include,
            os.path.join(_csrc, 'pyodesys_cuda_shim.cu'), '-o', out]


_d = ctypes.POINTER(ctypes.c_double)
_i32 = ctypes.POINTER(ctypes.c_int32)


def _ptr(arr, tp=_d):
    if arr is None:
        return None
    if isinstance(arr, int):
        return ctypes.cast(ctypes.c_void_p(arr), tp)
    return arr.ctypes.data_as(tp)


class _PinnedBlock:
    """ Owner of one page-locked block; hands it back to its pool when the last numpy view dies. """

    def __init__(self, pool, ptr, nbytes):
        self.pool, self.ptr, self.nbytes = pool, ptr, nbytes

    def __del__(self):
        try:
            self.pool._release(self.ptr, self.nbytes)
        except Exception:  # interpreter shutdown
            pass


class PinnedPool:
    """ numpy arrays in page-locked host memory (``pyodesys_cuda_host_alloc``), recycled by size.

    Falls back to ordinary numpy memory when pinning fails (e.g. no CUDA device: the integrate call
    that follows reports the real error).  Small arrays are not worth pinning.
    """
    min_bytes = 1 << 16
    # blocks handed back are kept for reuse up to this many bytes (page-locking costs ~0.2 s per GB: the 12 GB result
    # of a 1e8-system sweep took 2.4 s of a 3.3 s call when the pool would not keep it)
    max_cached_bytes = int(float(os.environ.get('PYODESYS_CUDA_PINNED_CACHE_GB', '32'))*(1 << 30))

    def __init__(self, lib):
        self.lib = lib
        self.free = {}
        self.cached = 0
        self.lock = threading.Lock()

    def empty(self, shape, dtype):
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape, dtype=np.int64))*dtype.itemsize
        if nbytes < self.min_bytes:
            return np.empty(shape, dtype=dtype)
        size = 1 << (nbytes - 1).bit_length() if nbytes < (1 << 26) else -(-nbytes // (1 << 26))*(1 << 26)
        with self.lock:
            lst = self.free.get(size)
            ptr = lst.pop() if lst else None
            if ptr is not None:
                self.cached -= size
        if ptr is None:
            ptr = self.lib.pyodesys_cuda_host_alloc(size)
            if not ptr:
                return np.empty(shape, dtype=dtype)
        buf = (ctypes.c_char*nbytes).from_address(ptr)
        buf._owner = _PinnedBlock(self, ptr, size)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def _release(self, ptr, size):
        with self.lock:
            if self.cached + size <= self.max_cached_bytes:
                self.free.setdefault(size, []).append(ptr)
                self.cached += size
                return
        self.lib.pyodesys_cuda_host_free(ptr)


class CudaModule:
    """ The loaded shared library of one system; stands in for the compiled Cython module.

    ``integrate_predefined`` / ``integrate_adaptive`` keep the wrappers' Python signatures
    (``_odeint_wrapper.pyx:37-46,111-119``); the ``*_arrays`` variants return plain arrays and
    are what :class:`NativeCudaSys` uses for large batches.
    """

    def __init__(self, path, gen):
        self.path = path
        self.gen = gen
        self.lib = lib = ctypes.CDLL(path)
        lib.pyodesys_cuda_last_error.restype = ctypes.c_char_p
        lib.pyodesys_cuda_query.restype = ctypes.c_int64
        lib.pyodesys_cuda_query.argtypes = [ctypes.c_char_p]
        lib.pyodesys_cuda_launch_count.restype = ctypes.c_int64
        lib.pyodesys_cuda_integrate_predefined.restype = ctypes.c_int
        lib.pyodesys_cuda_integrate_predefined.argtypes = [
            ctypes.c_int, ctypes.c_int64, ctypes.c_int, _d, _d, _d, _d, ctypes.c_int, ctypes.c_double,
            _d, _d, _d, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, _d, _i32, _i32, _i32, _d,
            ctypes.c_int, ctypes.POINTER(Options)]
        lib.pyodesys_cuda_integrate_adaptive.restype = ctypes.c_int
        lib.pyodesys_cuda_integrate_adaptive.argtypes = [
            ctypes.c_int, ctypes.c_int64, _d, _d, _d, _d, _d, ctypes.c_int, ctypes.c_double,
            _d, _d, _d, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
            _d, _i32, _i32, _i32, _d, ctypes.c_int, ctypes.POINTER(Options)]
        _i64 = ctypes.POINTER(ctypes.c_int64)
        lib.pyodesys_cuda_integrate_adaptive_compact.restype = ctypes.c_int
        lib.pyodesys_cuda_integrate_adaptive_compact.argtypes = [
            ctypes.c_int, ctypes.c_int64, _d, _d, _d, _d, _d, ctypes.c_int, ctypes.c_double,
            _d, _d, _d, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
            _i64, _i32, _i32, _d, ctypes.POINTER(Options)]
        lib.pyodesys_cuda_fetch_compact_rows.restype = ctypes.c_int
        lib.pyodesys_cuda_fetch_compact_rows.argtypes = [ctypes.c_int, ctypes.c_int, _d, ctypes.c_int64]
        lib.pyodesys_cuda_eval.restype = ctypes.c_int
        lib.pyodesys_cuda_eval.argtypes = [ctypes.c_int, ctypes.c_int64, _d, _d, _d, _d, _d, _d]
        lib.pyodesys_cuda_eval_jtimes.restype = ctypes.c_int
        lib.pyodesys_cuda_eval_jtimes.argtypes = [ctypes.c_int, ctypes.c_int64, _d, _d, _d, _d, _d]
        lib.pyodesys_cuda_eval_sparse_jac.restype = ctypes.c_int
        lib.pyodesys_cuda_eval_sparse_jac.argtypes = [ctypes.c_int, ctypes.c_int64, _d, _d, _d, _d, _i32, _i32]
        lib.pyodesys_cuda_lu_solve.restype = ctypes.c_int
        lib.pyodesys_cuda_lu_solve.argtypes = [ctypes.c_int, ctypes.c_int64, _d, _d, _d, _i32]
        lib.pyodesys_cuda_host_alloc.restype = ctypes.c_void_p
        lib.pyodesys_cuda_host_alloc.argtypes = [ctypes.c_size_t]
        lib.pyodesys_cuda_host_free.restype = None
        lib.pyodesys_cuda_host_free.argtypes = [ctypes.c_void_p]
        lib.pyodesys_cuda_device_pci_bus_id.restype = ctypes.c_int
        lib.pyodesys_cuda_device_pci_bus_id.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
        lib.pyodesys_cuda_device_count.restype = ctypes.c_int
        lib.pyodesys_cuda_integrate_predefined_multi.restype = ctypes.c_int
        lib.pyodesys_cuda_integrate_predefined_multi.argtypes = [
            ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_int,
            _d, _d, _d, _d, ctypes.c_int, ctypes.c_double, _d, _d, _d, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, _d, _i32, _i32, _i32, _d, _i32, ctypes.POINTER(ctypes.c_int64), ctypes.c_void_p]
        lib.pyodesys_cuda_compact_rows_noalloc.restype = ctypes.c_int
        lib.pyodesys_cuda_compact_rows_noalloc.argtypes = [
            ctypes.c_int, ctypes.c_int64, ctypes.c_int64, _d, _i32, ctypes.POINTER(ctypes.c_int64), _d,
            ctypes.c_int64, ctypes.c_void_p]
        self.pinned = PinnedPool(lib)
        self._lane_locks = {}  # (device, ctx_lane) -> lock: a compacting call and its fetch form one transaction
        self._lane_locks_guard = threading.Lock()
        self.nparams = int(self.query('nparams'))
        self.ny = int(self.query('ny'))
        self.nroots = max(0, int(self.query('nroots')))
        self.ninvar = max(0, int(self.query('ninvar')))
        self.np_tot = int(self.query('np_tot'))
        self.has_jac = bool(self.query('has_jac'))
        self.nnz = int(self.query('nnz'))  # -1: dense storage (template:131-133)
        self.has_jtimes = bool(self.query('has_jtimes'))

    def query(self, key):
        return self.lib.pyodesys_cuda_query(key.encode())

    def launch_count(self):
        return int(self.lib.pyodesys_cuda_launch_count())

    def release(self, device=0):
        """ free the device buffers / streams the library keeps for ``device`` (re-created on the next call) """
        self._check(self.lib.pyodesys_cuda_release(int(device)))

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError("pyodesys_cuda: " + self.lib.pyodesys_cuda_last_error().decode())

    def default_method(self):
        return 'rosenbrock4' if self.has_jac else 'dopri5'  # _odeint_wrapper.pyx:43

    # -- argument validation shared by both modes (``_odeint_wrapper.pyx:58-82``) -------------
    def _prep(self, y0, params, atol, dx0, dx_min, dx_max, method, iv_from_y0=False):
        y0 = np.ascontiguousarray(y0, dtype=np.float64)
        if y0.ndim != 2 or y0.shape[1] != self.ny:
            raise ValueError("y0 needs to be of shape (N, %d)" % self.ny)
        n = y0.shape[0]
        # "NaN found in y0" (_odeint_wrapper.pyx:58-59): small batches are screened here; for large ones a host pass
        # over y0 costs as much as 10 % of the whole call (2 ms per 3e6 doubles), so the kernels' own non-finite test
        # does the screening (status 3 before the first step) and _check_nan_y0 turns it into the same exception
        if y0.size <= (1 << 16) and np.isnan(y0).any():
            raise ValueError("NaN found in y0")
        params = np.ascontiguousarray(params, dtype=np.float64)
        pcols = self.nparams if iv_from_y0 else self.np_tot
        if params.ndim != 2 or params.shape != (n, pcols):
            raise ValueError("params needs to be of shape (N, %d), got %s" % (pcols, params.shape))
        atol = np.ascontiguousarray(np.atleast_1d(atol), dtype=np.float64)
        if atol.size != 1 and atol.size != self.ny:
            raise ValueError("atol needs to be of length 1 or %d" % self.ny)

        def bcast(v, name, skip):
            if v is None:
                return None
            v = np.ascontiguousarray(v, dtype=np.float64)
            if v.size == 1:
                if skip(float(v.reshape(-1)[0])):
                    return None
                v = v.reshape(-1)[0]*np.ones(n)
            if v.size < n:
                raise ValueError('%s too short' % name)
            return np.ascontiguousarray(v.reshape(-1)[:n])
        dx0 = bcast(dx0, 'dx0', lambda x: x == 0.0)
        dx_min = bcast(dx_min, 'dx_min', lambda x: x == 0.0)
        dx_max = bcast(dx_max, 'dx_max', lambda x: x == 0.0 or np.isinf(x))
        if method is None:
            method = self.default_method()
        try:
            mid = METHODS[method.lower()]
        except KeyError:
            raise ValueError("Unknown method: %s (choose from %s)" % (method, ', '.join(sorted(METHODS))))
        return y0, params, atol, dx0, dx_min, dx_max, mid, n

    @staticmethod
    def _options(get_dx_max_factor=0.0, error_outside_bounds=False, xout_shared=False, schedule=0,
                 blocks_per_sm=0, max_invariant_violation=0.0, iv_from_y0=False, kernel=0,
                 linear_solver='default', with_jacobian=None, with_jtimes=False, ctx_lane=0, **unknown):
        if unknown:
            raise TypeError("Unknown keyword arguments: %s" % ', '.join(sorted(unknown)))
        # cvode's names (_cvode_wrapper.pyx:88-92): 'dense' / 'klu' factorise the ny x ny iteration matrix (here: in
        # registers; for a sparse=True system it is assembled from the CSC entries), 'gmres' is matrix-free on jtimes
        linsol = _linear_solver_id(linear_solver, with_jtimes)
        sched = {'auto': 0, 'static': 1, 'dynamic': 2}.get(schedule, schedule)
        return Options(float(get_dx_max_factor or 0.0), float(max_invariant_violation or 0.0),
                       int(bool(error_outside_bounds)), int(bool(xout_shared)), int(sched), 0,
                       int(blocks_per_sm), 0, None, int(bool(iv_from_y0)), 0,
                       int({'auto': 0, 'thread': 1, 'group': 2, 'thread_smem': 3, 'coop': 4}.get(kernel, kernel)), 0, 0, linsol,
                       None, None, None, 0, int(ctx_lane), 0)

    def _root_buffers(self, opts, n, kw_roots):
        """ Event functions (``roots=`` of the system): attach the three output arrays of ABI v3 to ``opts``.
        ``kw_roots`` = (roots_sink, return_on_root, max_roots, roots_out); the arrays are handed back through the
        sink.  ``roots_out``: caller-provided (root_out, root_indices, nroots_found) - row ranges of batch-wide arrays
        (:func:`pyodesys_b200.multi.integrate_sharded`). """
        sink, return_on_root, max_roots, given = kw_roots
        if sink is None or self.nroots == 0 or max_roots <= 0:
            return None
        if given is None:
            root_out = self.pinned.empty((n, max_roots, 1 + self.ny), np.float64)
            root_indices = self.pinned.empty((n, max_roots), np.int32)
            nfound = self.pinned.empty((n,), np.int32)
        else:
            root_out, root_indices, nfound = given
        opts.return_on_root = int(bool(return_on_root))
        opts.max_roots = int(max_roots)
        opts.root_out = root_out.ctypes.data
        opts.root_indices = root_indices.ctypes.data
        opts.nroots_found = nfound.ctypes.data
        sink.update(root_out=root_out, root_indices=root_indices, nroots_found=nfound, max_roots=int(max_roots))
        return sink

    @staticmethod
    def _pop_roots(kw):
        return (kw.pop('roots_sink', None), kw.pop('return_on_root', False), int(kw.pop('max_roots', 8)),
                kw.pop('roots_out', None))

    def integrate_predefined_arrays(self, y0, xout, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                                    mxsteps=0, method=None, autorestart=0, return_on_error=False,
                                    device=0, out=None, out_aux=None, special_settings=None, **kw):
        """ ``out`` / ``out_aux = (nreached, status, counters, counters_ld)``: caller-provided (ideally page-locked)
        destinations - row ranges of arrays that span a larger batch (:func:`pyodesys_b200.multi.integrate_sharded`);
        ``counters`` is then the column window ``[:, lo:hi]`` of a ``[NCOUNTERS, counters_ld]`` array. """
        y0, params, atol, dx0, dx_min, dx_max, mid, n = self._prep(y0, params, atol, dx0, dx_min, dx_max, method,
                                                                   kw.get('iv_from_y0', False))
        xout = np.ascontiguousarray(xout, dtype=np.float64)
        shared = xout.ndim == 1
        if not shared and xout.shape[0] != n:
            raise ValueError("xout needs to be of shape (N, nt)")
        nt = xout.shape[-1]
        kw_roots = self._pop_roots(kw)
        opts = self._options(xout_shared=shared, **kw)
        self._root_buffers(opts, n, kw_roots)
        # outputs live in page-locked memory (D2H at PCIe speed); the blocks return to the pool when the
        # arrays are garbage collected
        yout = self.pinned.empty((n, nt, self.ny), np.float64) if out is None else out
        if out_aux is None:
            nreached = self.pinned.empty((n,), np.int32)
            status = self.pinned.empty((n,), np.int32)
            counters = self.pinned.empty((NCOUNTERS, n), np.int32)
        else:
            nreached, status, counters, opts.counters_ld = out_aux
        timings = np.zeros(NTIMINGS)
        rc = self.lib.pyodesys_cuda_integrate_predefined(
            device, n, nt, _ptr(y0), _ptr(xout), _ptr(params), _ptr(atol), atol.size, float(rtol),
            _ptr(dx0), _ptr(dx_min), _ptr(dx_max), int(mxsteps), mid, int(autorestart), int(return_on_error),
            _ptr(yout), _ptr(nreached, _i32), _ptr(status, _i32), _ptr(counters, _i32), _ptr(timings), 0,
            ctypes.byref(opts))
        self._check(rc)
        nbad = int(self.lib.pyodesys_cuda_query(b'last_failed_systems'))  # counted while the call waited (or -1)
        _check_nan_y0(y0, status, nbad=nbad if nbad >= 0 else None)
        return yout, nreached, status, counters, timings

    def integrate_predefined_multi(self, devices, lanes, chunk, y0, xout, params, atol, rtol, dx0=None, dx_min=None,
                                   dx_max=None, mxsteps=0, method=None, autorestart=0, return_on_error=False,
                                   special_settings=None, out=None, **kw):
        """ ``pyodesys_cuda_integrate_predefined_multi``: the whole batch in ONE call, scheduled natively over
        ``devices`` (``lanes`` host threads per device draw chunks of ``chunk`` systems from an atomic counter).
        Returns ``(yout, nreached, status, counters, busy_s[ndev], chunks_taken[ndev])``; all outputs page-locked. """
        y0, params, atol, dx0, dx_min, dx_max, mid, n = self._prep(y0, params, atol, dx0, dx_min, dx_max, method,
                                                                   kw.get('iv_from_y0', False))
        xout = np.ascontiguousarray(xout, dtype=np.float64)
        shared = xout.ndim == 1
        if not shared and xout.shape[0] != n:
            raise ValueError("xout needs to be of shape (N, nt)")
        nt = xout.shape[-1]
        opts = self._options(xout_shared=shared, **kw)
        if out is None:
            out = (self.pinned.empty((n, nt, self.ny), np.float64), self.pinned.empty((n,), np.int32),
                   self.pinned.empty((n,), np.int32), self.pinned.empty((NCOUNTERS, n), np.int32))
        yout, nreached, status, counters = out
        devs = (ctypes.c_int*len(devices))(*[int(d) for d in devices])
        busy = np.zeros(len(devices))
        taken = np.zeros(len(devices), dtype=np.int32)
        nbad = ctypes.c_int64(0)
        rc = self.lib.pyodesys_cuda_integrate_predefined_multi(
            devs, len(devices), int(lanes), int(chunk), n, nt, _ptr(y0), _ptr(xout), _ptr(params), _ptr(atol), atol.size,
            float(rtol), _ptr(dx0), _ptr(dx_min), _ptr(dx_max), int(mxsteps), mid, int(autorestart), int(return_on_error),
            _ptr(yout), _ptr(nreached, _i32), _ptr(status, _i32), _ptr(counters, _i32), _ptr(busy), _ptr(taken, _i32),
            ctypes.byref(nbad), ctypes.cast(ctypes.byref(opts), ctypes.c_void_p))
        self._check(rc)
        _check_nan_y0(y0, status, nbad=int(nbad.value))
        return yout, nreached, status, counters, busy, taken

    def integrate_adaptive_arrays(self, y0, x0, xend, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                                  mxsteps=0, method=None, autorestart=0, return_on_error=False,
                                  device=0, max_rows=None, out=None, out_aux=None, special_settings=None, **kw):
        y0, params, atol, dx0, dx_min, dx_max, mid, n = self._prep(y0, params, atol, dx0, dx_min, dx_max, method,
                                                                   kw.get('iv_from_y0', False))
        x0 = np.ascontiguousarray(x0, dtype=np.float64).reshape(-1)
        xend = np.ascontiguousarray(xend, dtype=np.float64).reshape(-1)
        if x0.size != n or xend.size != n:
            raise ValueError("x0/xend need to be of length N")
        kw_roots = self._pop_roots(kw)
        opts = self._options(**kw)
        has_roots = self._root_buffers(opts, n, kw_roots) is not None
        if max_rows is None:
            # a system can never produce more rows than steps it is allowed to take (+ the start)
            steps = int(mxsteps) if mxsteps else 500
            max_rows = steps*(1 + int(autorestart)) + 1
            if has_roots:
                max_rows += kw_roots[2]  # every stored root is one more row
        xyout = self.pinned.empty((n, max_rows, 1 + self.ny), np.float64) if out is None else out
        if out_aux is None:
            nrows = self.pinned.empty((n,), np.int32)
            status = self.pinned.empty((n,), np.int32)
            counters = self.pinned.empty((NCOUNTERS, n), np.int32)
        else:
            nrows, status, counters, opts.counters_ld = out_aux
        timings = np.zeros(NTIMINGS)
        rc = self.lib.pyodesys_cuda_integrate_adaptive(
            device, n, _ptr(y0), _ptr(x0), _ptr(xend), _ptr(params), _ptr(atol), atol.size, float(rtol),
            _ptr(dx0), _ptr(dx_min), _ptr(dx_max), int(mxsteps), mid, int(autorestart), int(return_on_error),
            int(max_rows), _ptr(xyout), _ptr(nrows, _i32), _ptr(status, _i32), _ptr(counters, _i32),
            _ptr(timings), 0, ctypes.byref(opts))
        self._check(rc)
        _check_nan_y0(y0, status)
        return xyout, nrows, status, counters, timings

    # device scratch of the bounded row buffers of one compacting call (systems beyond it go into further calls)
    adaptive_scratch_bytes = 4 << 30

    def integrate_adaptive_compact(self, y0, x0, xend, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                                   mxsteps=0, method=None, autorestart=0, return_on_error=False, device=0,
                                   max_rows=None, out_aux=None, special_settings=None, **kw):
        """ Adaptive mode, compacted output: ``(pieces, status, counters, timings)`` with ``pieces`` a list of
        ``(first_system, rows[total, 1 + ny], offsets[m + 1])`` - exactly the rows that were written
        (``_cvode_wrapper.pyx:198-205``), in page-locked memory.  One piece per device call; a batch is cut into
        several calls when its bounded device scratch ``n * max_rows * (1 + ny) * 8`` would exceed
        :attr:`adaptive_scratch_bytes` (round-1 advisor finding: that product used to be allocated - and copied over
        PCIe - in one piece, 16 GB for 1e6 three-component systems at the default ``nsteps``). """
        y0, params, atol, dx0, dx_min, dx_max, mid, n = self._prep(y0, params, atol, dx0, dx_min, dx_max, method,
                                                                   kw.get('iv_from_y0', False))
        x0 = np.ascontiguousarray(x0, dtype=np.float64).reshape(-1)
        xend = np.ascontiguousarray(xend, dtype=np.float64).reshape(-1)
        if x0.size != n or xend.size != n:
            raise ValueError("x0/xend need to be of length N")
        kw_roots = self._pop_roots(kw)
        opts = self._options(**kw)
        has_roots = self._root_buffers(opts, n, kw_roots) is not None
        if max_rows is None:
            steps = int(mxsteps) if mxsteps else 500
            max_rows = steps*(1 + int(autorestart)) + 1 + (kw_roots[2] if has_roots else 0)
        if out_aux is None:
            status = self.pinned.empty((n,), np.int32)
            counters = self.pinned.empty((NCOUNTERS, n), np.int32)
            opts.counters_ld = n
        else:  # windows of batch-wide arrays (multi.integrate_sharded)
            status, counters, opts.counters_ld = out_aux
        per = max(1, int(self.adaptive_scratch_bytes // (max_rows*(1 + self.ny)*8)))
        timings = np.zeros(NTIMINGS)
        pieces = []
        root_ptrs = (opts.root_out, opts.root_indices, opts.nroots_found) if has_roots else None
        for lo in range(0, max(n, 1), per):
            hi = min(n, lo + per)
            m = hi - lo
            if m <= 0:
                break
            offsets = np.empty(m + 1, dtype=np.int64)
            t1 = np.zeros(NTIMINGS)
            if has_roots:  # row windows of the three root arrays
                mr = kw_roots[2]
                opts.root_out = root_ptrs[0] + lo*mr*(1 + self.ny)*8
                opts.root_indices = root_ptrs[1] + lo*mr*4
                opts.nroots_found = root_ptrs[2] + lo*4
            sl = slice(lo, hi)
            with self._lane_locks_guard:
                lane_lock = self._lane_locks.setdefault((int(device), int(opts.ctx_lane)), threading.Lock())
            lane_lock.acquire()
            try:
                rc = self._compact_call(device, m, y0, x0, xend, params, atol, rtol, dx0, dx_min, dx_max, mxsteps, mid,
                                        autorestart, return_on_error, max_rows, offsets, status, counters, lo, sl, t1,
                                        opts)
                self._check(rc)
                total = int(offsets[m])
                rows = self.pinned.empty((total, 1 + self.ny), np.float64)
                self._check(self.lib.pyodesys_cuda_fetch_compact_rows(device, int(opts.ctx_lane), _ptr(rows), total))
            finally:
                lane_lock.release()
            pieces.append((lo, rows, offsets))
            timings += t1
        _check_nan_y0(y0, status)
        return pieces, status, counters, timings

    def _compact_call(self, device, m, y0, x0, xend, params, atol, rtol, dx0, dx_min, dx_max, mxsteps, mid,
                      autorestart, return_on_error, max_rows, offsets, status, counters, lo, sl, t1, opts):
        return self.lib.pyodesys_cuda_integrate_adaptive_compact(
            device, m, _ptr(y0[sl]), _ptr(x0[sl]), _ptr(xend[sl]), _ptr(params[sl]), _ptr(atol), atol.size,
            float(rtol), _ptr(None if dx0 is None else dx0[sl]), _ptr(None if dx_min is None else dx_min[sl]),
            _ptr(None if dx_max is None else dx_max[sl]), int(mxsteps), mid, int(autorestart),
            int(return_on_error), int(max_rows), offsets.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
            _ptr(status[sl], _i32), ctypes.cast(ctypes.c_void_p(counters.ctypes.data + lo*4), _i32), _ptr(t1),
            ctypes.byref(opts))

    def integrate_predefined_device(self, n, nt, y0, xout, params, atol, rtol, yout, nreached, status,
                                    counters=None, dx0=None, dx_min=None, dx_max=None, mxsteps=0,
                                    method=None, autorestart=0, device=0, stream=None, xout_shared=False,
                                    yout_soa=False, debug_buffer=None, **kw):
        """ Device-resident call (``ptr_kind=1``): every array argument is a device address (int,
        e.g. ``tensor.data_ptr()``) on ``device``; nothing is copied.  Returns the timings array. """
        atol = np.ascontiguousarray(np.atleast_1d(atol), dtype=np.float64)
        if atol.size != 1 and atol.size != self.ny:
            raise ValueError("atol needs to be of length 1 or %d" % self.ny)
        mid = METHODS[(method or self.default_method()).lower()]
        opts = self._options(xout_shared=xout_shared, **kw)
        opts.yout_soa = int(bool(yout_soa))
        opts.stream = stream
        if debug_buffer is not None:  # profiling builds (-DPYODESYS_COOP_PROFILE) write their phase counters here
            opts.root_out = debug_buffer
        timings = np.zeros(NTIMINGS)
        rc = self.lib.pyodesys_cuda_integrate_predefined(
            device, n, nt, _ptr(y0), _ptr(xout), _ptr(params), _ptr(atol), atol.size, float(rtol),
            _ptr(dx0), _ptr(dx_min), _ptr(dx_max), int(mxsteps), mid, int(autorestart), 0,
            _ptr(yout), _ptr(nreached, _i32), _ptr(status, _i32), _ptr(counters, _i32), _ptr(timings), 1,
            ctypes.byref(opts))
        self._check(rc)
        return timings

    def integrate_adaptive_device(self, n, y0, x0, xend, params, atol, rtol, max_rows, xyout, nrows, status,
                                  counters=None, dx0=None, dx_min=None, dx_max=None, mxsteps=0, method=None,
                                  autorestart=0, device=0, stream=None, **kw):
        """ Device-resident adaptive call (``ptr_kind=1``): ``xyout[n, max_rows, 1 + ny]`` + ``nrows[n]`` stay on
        the device.  Returns the timings array. """
        atol = np.ascontiguousarray(np.atleast_1d(atol), dtype=np.float64)
        if atol.size != 1 and atol.size != self.ny:
            raise ValueError("atol needs to be of length 1 or %d" % self.ny)
        mid = METHODS[(method or self.default_method()).lower()]
        opts = self._options(**kw)
        opts.stream = stream
        timings = np.zeros(NTIMINGS)
        rc = self.lib.pyodesys_cuda_integrate_adaptive(
            device, n, _ptr(y0), _ptr(x0), _ptr(xend), _ptr(params), _ptr(atol), atol.size, float(rtol),
            _ptr(dx0), _ptr(dx_min), _ptr(dx_max), int(mxsteps), mid, int(autorestart), 0, int(max_rows),
            _ptr(xyout), _ptr(nrows, _i32), _ptr(status, _i32), _ptr(counters, _i32), _ptr(timings), 1,
            ctypes.byref(opts))
        self._check(rc)
        return timings

    def compact_rows_noalloc(self, n, max_rows, xyout, nrows, offsets, rows, rows_capacity, device=0, stream=None):
        """ ``pyodesys_cuda_compact_rows_noalloc`` (include/pyodesys_cuda.h): compaction of a device-resident
        adaptive result on caller-owned device buffers, scan scratch in a static device array - no allocation, no
        synchronisation; at most ``PYODESYS_CUDA_NOALLOC_MAX_SYSTEMS`` systems per call.  Arguments are device
        pointers as ints (``tensor.data_ptr()``). """
        rc = self.lib.pyodesys_cuda_compact_rows_noalloc(
            int(device), int(n), int(max_rows), _ptr(xyout), _ptr(nrows, _i32),
            _ptr(offsets, ctypes.POINTER(ctypes.c_int64)), _ptr(rows), int(rows_capacity), stream)
        self._check(rc)

    # -- reference-shaped entry points --------------------------------------------------------
    @staticmethod
    def _nfo(i, n, status, counters, timings, mode, nreached=None, roots=None, krylov=False):
        d = {k: int(counters[j, i]) for j, k in enumerate(BatchInfo.counter_names)}
        if krylov:  # matrix-free Newton iteration: the factorisation slot counts the J v products
            d['njvev'], d['nlu'] = d['nlu'], 0
        d['n_rhs_evals'] = d['nfev']
        d['dense_n_dls_jac_evals'] = d['njev']
        d['time_wall'] = float(timings[0])/n
        d['time_cpu'] = float(timings[3])/n
        d['mode'] = mode
        d['success'] = bool(status[i] == 0)
        if nreached is not None:
            d['nreached'] = int(nreached[i])
        if roots:
            d.update(root_info(roots, i, mode))
        return d

    def integrate_predefined(self, y0, xout, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                             mxsteps=0, method=None, autorestart=0, return_on_error=False, **kw):
        roots = {}
        yout, nreached, status, counters, timings = self.integrate_predefined_arrays(
            y0, xout, params, atol, rtol, dx0, dx_min, dx_max, mxsteps, method, autorestart,
            return_on_error, roots_sink=roots, **kw)
        _raise_on_failure(status, return_on_error)
        n = yout.shape[0]
        kry = _linear_solver_id(kw.get('linear_solver', 'default'), kw.get('with_jtimes', False)) == 1
        return yout, [self._nfo(i, n, status, counters, timings, 'predefined', nreached, roots, kry) for i in range(n)]

    def integrate_adaptive(self, y0, x0, xend, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                           mxsteps=0, method=None, autorestart=0, return_on_error=False, **kw):
        roots = {}
        xyout, nrows, status, counters, timings = self.integrate_adaptive_arrays(
            y0, x0, xend, params, atol, rtol, dx0, dx_min, dx_max, mxsteps, method, autorestart,
            return_on_error, roots_sink=roots, **kw)
        _raise_on_failure(status, return_on_error)
        n = xyout.shape[0]
        xs = [xyout[i, :nrows[i], 0] for i in range(n)]
        ys = [xyout[i, :nrows[i], 1:] for i in range(n)]
        kry = _linear_solver_id(kw.get('linear_solver', 'default'), kw.get('with_jtimes', False)) == 1
        return xs, ys, [self._nfo(i, n, status, counters, timings, 'adaptive', None, roots, kry) for i in range(n)]

    # -- verification hooks -----------------------------------------------------------------
    def eval_rhs_jac(self, x, y, params, device=0, jac=True):
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        f = np.empty((n, self.ny))
        J = np.empty((n, self.ny, self.ny)) if jac and self.has_jac else None
        dfdt = np.empty((n, self.ny)) if J is not None else None
        self._check(self.lib.pyodesys_cuda_eval(device, n, _ptr(x), _ptr(y), _ptr(params), _ptr(f),
                                                _ptr(J), _ptr(dfdt)))
        return f, J, dfdt

    def eval_jtimes(self, x, y, params, v, device=0):
        """ ``J(x, y) v`` for a batch by the generated ``OdeSys::jtimes`` """
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        v = np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.float64), (n, self.ny)))
        out = np.empty((n, self.ny))
        self._check(self.lib.pyodesys_cuda_eval_jtimes(device, n, _ptr(x), _ptr(y), _ptr(params), _ptr(v), _ptr(out)))
        return out

    def eval_sparse_jac(self, x, y, params, device=0):
        """ ``(data[n, nnz], colptrs[ny + 1], rowvals[nnz])`` by the generated ``OdeSys::sparse_jac_csc`` """
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        data = np.empty((n, max(self.nnz, 0)))
        colptrs = np.zeros(self.ny + 1, dtype=np.int32)
        rowvals = np.zeros(max(self.nnz, 0), dtype=np.int32)
        self._check(self.lib.pyodesys_cuda_eval_sparse_jac(device, n, _ptr(x), _ptr(y), _ptr(params), _ptr(data),
                                                           _ptr(colptrs, _i32), _ptr(rowvals, _i32)))
        return data, colptrs, rowvals

    def lu_solve(self, A, b, device=0):
        A = np.ascontiguousarray(A, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        n = A.shape[0]
        x = np.empty((n, self.ny))
        ok = np.empty(n, dtype=np.int32)
        self._check(self.lib.pyodesys_cuda_lu_solve(device, n, _ptr(A), _ptr(b), _ptr(x), _ptr(ok, _i32)))
        return x, ok


def root_info(roots, i, mode):
    """ ``root_indices`` (and ``root_out`` in predefined mode) of system ``i`` as the cvode wrapper reports them
    (``_cvode_wrapper.pyx:63-70,206-219,354-355``) """
    k = min(int(roots['nroots_found'][i]), roots['max_roots'])
    d = {'root_indices': [int(v) for v in roots['root_indices'][i, :k]]}
    if mode == 'predefined':
        d['root_out'] = np.array(roots['root_out'][i, :k])
    d['n_root_evals'] = int(roots['nroots_found'][i])
    return d


_screen_memo = threading.local()


def _check_nan_y0(y0, status, nbad=None):
    """ deferred NaN screen of large batches: only rows the kernel rejected are looked at.  ``nbad``: the number of
    failed systems where the library counted them already (multi-device entry point) """
    if nbad is None:
        nbad = int(np.count_nonzero(status))  # one SIMD pass, 0.25 ms per 1e6 systems
    _screen_memo.last = (status, nbad)
    if y0.size <= (1 << 16) or not nbad:
        return
    bad = np.flatnonzero(status == 3)
    if bad.size and np.isnan(y0[bad]).any():
        raise ValueError("NaN found in y0")


def _raise_on_failure(status, return_on_error):
    last = getattr(_screen_memo, 'last', None)  # counted a moment ago by _check_nan_y0 (same live array)?
    nbad = last[1] if last is not None and last[0] is status else None
    _screen_memo.last = None
    if return_on_error or not (np.count_nonzero(status) if nbad is None else nbad):
        return
    bad = np.flatnonzero(status != 0)
    if bad.size:
        i = int(bad[0])
        raise RuntimeError("Integration failed for %d of %d systems (first: index %d: %s). "
                           "Pass return_on_error=True to obtain partial results." %
                           (bad.size, status.size, i, STATUS_TEXT.get(int(status[i]), 'unknown')))


_build_locks = {}
_build_locks_guard = threading.Lock()


def _lock_for(tag):
    with _build_locks_guard:
        return _build_locks.setdefault(tag, threading.Lock())


class NativeCudaCode:
    """ Generated + compiled code of one system (``_NativeCodeBase`` counterpart). """

    def __init__(self, odesys, namespace_override=None, namespace_extend=None, save_temp=False,
                 compile=True):
        self.odesys = odesys
        self.gen = _codegen.generate(odesys, namespace_override, namespace_extend)
        tag = hashlib.sha256((self.gen.digest + _csrc_digest() +
                              os.environ.get('PYODESYS_CUDA_ARCH', 'sm_100a') +
                              os.environ.get('PYODESYS_CUDA_NVCC_FLAGS', '')).encode()).hexdigest()[:16]
        self.tag = tag
        self.build_dir = os.path.join(cache_dir(), tag)
        self.header = os.path.join(self.build_dir, 'odesys_generated.h')
        self.so_file = os.path.join(self.build_dir, 'libodesys_%s.so' % tag)
        self._mod = None
        if compile:
            self.ensure_built()

    def write_source(self):
        os.makedirs(self.build_dir, exist_ok=True)
        if not os.path.exists(self.header):
            tmp = self.header + '.%d.tmp' % os.getpid()
            with open(tmp, 'w') as fh:
                fh.write(self.gen.text)
            os.replace(tmp, self.header)

    def ensure_built(self):
        if os.path.exists(self.so_file):
            return self.so_file
        with _lock_for(self.tag):
            if os.path.exists(self.so_file):
                return self.so_file
            self.write_source()
            tmp_out = self.so_file + '.%d.tmp' % os.getpid()
            cmd = nvcc_command(self.build_dir, tmp_out)
            t0 = time.time()
            proc = subprocess.run(cmd, capture_output=True, text=True)
            with open(os.path.join(self.build_dir, 'nvcc.log'), 'w') as fh:
                fh.write(' '.join(cmd) + '\n' + proc.stdout + proc.stderr)
            if proc.returncode != 0:
                raise RuntimeError("nvcc failed (%s):\n%s" % (' '.join(cmd), proc.stderr[-4000:]))
            os.replace(tmp_out, self.so_file)
            logger.info("compiled %s in %.1f s", self.so_file, time.time() - t0)
        return self.so_file

    @property
    def mod(self):
        if self._mod is None:
            self._mod = CudaModule(self.ensure_built(), self.gen)
        return self._mod


class NativeCudaSys(SymbolicSys):
    """ ``SymbolicSys`` integrated by the CUDA back end (``_NativeSysBase`` counterpart).

    ``integrate(x, y0, params, atol=, rtol=, nsteps=, first_step=, force_predefined=,
    method=, return_on_error=, dx_max=, autorestart=, device=, ...)``.
    """

    _NativeCode = NativeCudaCode
    _native_name = 'cuda'
    _append_iv_on_device = True

    def __init__(self, *args, **kwargs):
        namespace_override = kwargs.pop('namespace_override', {})
        namespace_extend = kwargs.pop('namespace_extend', {})
        save_temp = kwargs.pop('save_temp', False)
        compile_ = kwargs.pop('compile', True)
        if 'init_indep' not in kwargs:  # _base.py:293-295: triggers append_iv
            kwargs['init_indep'] = True
            kwargs['init_dep'] = True
        super().__init__(*args, **kwargs)
        self._native = self._NativeCode(self, save_temp=save_temp, namespace_override=namespace_override,
                                        namespace_extend=namespace_extend, compile=compile_)

    def integrate(self, *args, **kwargs):
        integrator = kwargs.pop('integrator', 'native')
        if integrator not in ('native', self._native_name):
            raise ValueError("Got incompatible kwargs integrator=%s" % integrator)
        kwargs['integrator'] = 'native'
        return super().integrate(*args, **kwargs)

    def _integrate_native(self, intern_x, intern_y0, intern_p, force_predefined=False,
                          atol=1e-8, rtol=1e-8, nsteps=500, first_step=0.0, **kwargs):
        """ ``pyodesys/native/_base.py:310-342`` with array-backed info instead of N dicts. """
        atol = np.atleast_1d(atol)
        if atol.size != 1 and atol.size != self.ny:
            raise ValueError("atol needs to be of length 1 or %d" % self.ny)
        return_on_error = kwargs.get('return_on_error', False)
        devices = kwargs.pop('devices', None)
        iv = intern_y0 if kwargs.get('iv_from_y0', False) else None
        mod = self._native.mod
        method = kwargs.get('method') or mod.default_method()
        y0 = np.ascontiguousarray(intern_y0, dtype=np.float64)
        params = np.ascontiguousarray(intern_p, dtype=np.float64)
        t0 = time.time()
        roots = {}
        if mod.nroots:
            kwargs['roots_sink'] = roots
        if devices is not None and len(devices) > 1:
            from ..multi import integrate_sharded
            binfo = integrate_sharded(mod, devices, intern_x, y0, params, force_predefined, atol, rtol,
                                      nsteps, first_step, kwargs)
        elif intern_x.shape[-1] == 2 and not force_predefined:
            pieces, status, counters, timings = mod.integrate_adaptive_compact(
                y0, intern_x[:, 0], intern_x[:, 1], params, atol, rtol, dx0=first_step, mxsteps=nsteps, **kwargs)
            xs = _RaggedPieces(pieces, len(status), 0)
            ys = _RaggedPieces(pieces, len(status), 1)
            binfo = BatchInfo('adaptive', status, None, counters, 0.0, 0.0, xs, ys, intern_p,
                              return_on_error, timings=_tdict(timings), method=method, iv=iv)
            binfo.d2h_bytes = sum(r.nbytes + o.nbytes for _lo, r, o in pieces) + status.nbytes + counters.nbytes
        else:
            if intern_x.shape[0] > 1 and intern_x.strides[0] == 0:
                # one broadcast row (core.to_arrays): ship it once (options.xout_shared)
                xout_arg = np.ascontiguousarray(intern_x[0], dtype=np.float64)
            else:
                xout_arg = np.ascontiguousarray(intern_x, dtype=np.float64)
            yout, nreached, status, counters, timings = mod.integrate_predefined_arrays(
                y0, xout_arg, params, atol, rtol, dx0=first_step, mxsteps=nsteps, **kwargs)
            binfo = BatchInfo('predefined', status, nreached, counters, 0.0, 0.0, intern_x, yout, intern_p,
                              return_on_error, nt=intern_x.shape[-1], timings=_tdict(timings), method=method, iv=iv)
        n = len(binfo)
        binfo.roots = roots or None
        binfo.krylov = _linear_solver_id(kwargs.get('linear_solver', 'default'), kwargs.get('with_jtimes', False)) == 1
        binfo.time_wall = binfo.timings.get('kernel', 0.0)/max(n, 1)
        binfo.time_cpu = (time.time() - t0)/max(n, 1)
        _raise_on_failure(binfo.status, return_on_error)
        return binfo


def _tdict(t):
    return {'kernel': float(t[0]), 'h2d': float(t[1]), 'd2h': float(t[2]), 'total': float(t[3])}


class _Ragged:
    """ View of the adaptive output buffer [N, max_rows, 1+ny] as N ragged arrays. """

    def __init__(self, xyout, nrows, kind):
        self.xyout, self.nrows, self.kind = xyout, nrows, kind

    def __len__(self):
        return self.xyout.shape[0]

    def __getitem__(self, i):
        if self.kind == 0:
            return self.xyout[i, :self.nrows[i], 0]
        return self.xyout[i, :self.nrows[i], 1:]


class _RaggedPieces:
    """ N ragged arrays backed by compacted pieces ``(first_system, rows[total, 1 + ny], offsets[m + 1])``. """

    def __init__(self, pieces, n, kind):
        self.pieces, self.n, self.kind = pieces, n, kind
        self.starts = np.array([p[0] for p in pieces], dtype=np.int64)

    def __len__(self):
        return self.n

    def nrows(self):
        """ rows per system, [N] """
        out = np.empty(self.n, dtype=np.int64)
        for lo, _rows, off in self.pieces:
            out[lo:lo + off.size - 1] = np.diff(off)
        return out

    def __getitem__(self, i):
        if i < 0:
            i += self.n
        if not 0 <= i < self.n:
            raise IndexError(i)
        k = int(np.searchsorted(self.starts, i, side='right')) - 1
        lo, rows, off = self.pieces[k]
        a, b = int(off[i - lo]), int(off[i - lo + 1])
        return rows[a:b, 0] if self.kind == 0 else rows[a:b, 1:]
