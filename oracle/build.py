# -*- coding: utf-8 -*-
"""
ORACLE - TEST INFRASTRUCTURE ONLY.

Build recipes:

* :func:`build_oracle` - ``g++ -O3 -march=native -fopenmp`` of ``oracle/oracle.cpp`` against the generated header
  of one ODE system -> ``oracle/_build/<tag>/liboracle.so`` (+ ctypes wrapper
  :class:`OracleModule`);
* :func:`build_ref` - compiles the reference's own LAPACK-free LU
  (``external/anyode/include/anyode/anyode_blasless.hpp`` under ``/root/reference``) from the
  sources where they lie into ``oracle/_ref/libanyode_blasless.so``.  Only possible where
  ``/root/reference`` exists (this container); the GPU box uses the prebuilt file.
"""
import ctypes
import hashlib
import os
import subprocess
import threading

import numpy as np

_here = os.path.dirname(os.path.abspath(__file__))
_repo = os.path.dirname(_here)
_csrc = os.path.join(_repo, 'pyodesys_b200', 'csrc')
REFERENCE = os.environ.get('PYODESYS_REFERENCE_ROOT', '/root/reference')
REF_DIR = os.path.join(_here, '_ref')
REF_LU = os.path.join(REF_DIR, 'libanyode_blasless.so')

_build_lock = threading.Lock()
_d = ctypes.POINTER(ctypes.c_double)
_i32 = ctypes.POINTER(ctypes.c_int32)


def _p(a, tp=_d):
    return None if a is None else a.ctypes.data_as(tp)


def _oracle_digest():
    h = hashlib.sha256()
    for path in (os.path.join(_here, 'oracle.cpp'), os.path.join(_csrc, 'odesys_prelude.h')):
        with open(path, 'rb') as fh:
            h.update(fh.read())
    return h.hexdigest()[:12]


def _cpu_signature():
    """ ``-march=native`` binaries only run on CPUs with the same instruction sets: the build tag carries the
    host's flag list, so a box with another CPU rebuilds (g++ is part of the image) instead of loading them """
    try:
        with open('/proc/cpuinfo') as fh:
            for line in fh:
                if line.startswith('flags'):
                    return hashlib.sha256(line.encode()).hexdigest()[:8]
    except OSError:
        pass
    return 'generic'


# The reference compiles its generated native code with pycompilation's 'fast' option
# (pyodesys/native/_base.py:41-42: -O3 -march=native -ffast-math -funroll-loops).  The parity oracle keeps IEEE
# semantics (its NaN / inf tests must work; no FMA contraction, so its LU stays bit-equal to oracle/_ref) and uses
# -O3 -march=native -ffp-contract=off; the TIMING build (bench.py's cpu_baseline and --impl reference) uses the
# reference's full flag set.
FLAGS_PARITY = ['-O3', '-march=native', '-ffp-contract=off']
FLAGS_TIMING = ['-O3', '-march=native', '-ffast-math', '-funroll-loops']
# the BDF stepper (in the reference: the separately compiled CVODE library, not 'fast' code) does not survive
# -fassociative-math: its difference-table updates get re-associated and the Newton iteration stops converging
FLAGS_TIMING_BDF = ['-O3', '-march=native', '-ffast-math', '-fno-associative-math', '-funroll-loops']


def build_oracle(native_code, timing=False, method=None):
    """ ``native_code``: a ``pyodesys_b200.native.cuda.NativeCudaCode`` (only its generated header is used).
    ``timing=True``: the reference's 'fast' flags (results equal to ~1e-14; NaN screens are not reliable). """
    native_code.write_source()
    flags = (FLAGS_TIMING_BDF if method == 'bdf' else FLAGS_TIMING) if timing else FLAGS_PARITY
    tag = hashlib.sha256((native_code.gen.digest + _oracle_digest() + _cpu_signature() +
                          ' '.join(flags)).encode()).hexdigest()[:16]
    out_dir = os.path.join(_here, '_build', tag)
    so = os.path.join(out_dir, 'liboracle.so')
    with _build_lock:
        exists = os.path.exists(so)
    if not exists:
        os.makedirs(out_dir, exist_ok=True)
        tmp = so + '.%d.%d.tmp' % (os.getpid(), threading.get_ident())
        cmd = ['g++'] + flags + ['-std=c++17', '-fopenmp', '-shared', '-fPIC', '-I', native_code.build_dir,
                                 '-I', _csrc, os.path.join(_here, 'oracle.cpp'), '-o', tmp]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("g++ failed: %s\n%s" % (' '.join(cmd), proc.stderr[-4000:]))
        os.replace(tmp, so)
    mod = OracleModule(so)
    mod.flags = ' '.join(flags)
    return mod


def build_ref(force=False):
    """ Returns the path of ``oracle/_ref/libanyode_blasless.so`` (building it when possible). """
    if os.path.exists(REF_LU) and not force:
        return REF_LU
    inc = os.path.join(REFERENCE, 'external', 'anyode', 'include')
    if not os.path.exists(os.path.join(inc, 'anyode', 'anyode_blasless.hpp')):
        return None
    os.makedirs(REF_DIR, exist_ok=True)
    tmp = REF_LU + '.%d.tmp' % os.getpid()
    cmd = ['g++', '-O2', '-std=c++17', '-shared', '-fPIC', '-DANYODE_NO_LAPACK=1', '-I', inc,
           os.path.join(_here, 'ref_anyode_lu.cpp'), '-o', tmp]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("g++ failed: %s\n%s" % (' '.join(cmd), proc.stderr[-4000:]))
    os.replace(tmp, REF_LU)
    return REF_LU


class RefAnyodeLU:
    """ The reference's own getrf/getrs (AnyODE) behind ctypes. """

    def __init__(self, path=None):
        path = path or build_ref()
        if path is None:
            raise RuntimeError("oracle/_ref/libanyode_blasless.so is missing and /root/reference is absent")
        self.lib = ctypes.CDLL(path)
        self.lib.ref_anyode_lu_solve.restype = ctypes.c_int
        self.lib.ref_anyode_lu_solve.argtypes = [ctypes.c_int64, ctypes.c_int, _d, _d, _d, _i32]

    def solve(self, A, b):
        A = np.ascontiguousarray(A, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        nsys, n = b.shape
        x = np.empty_like(b)
        info = np.empty(nsys, dtype=np.int32)
        self.lib.ref_anyode_lu_solve(nsys, n, _p(A), _p(b), _p(x), _p(info, _i32))
        return x, info


METHODS = {'rk4': 0, 'dopri5': 1, 'rosenbrock4': 2, 'bdf': 3}


class OracleModule:
    def __init__(self, path):
        self.path = path
        self.lib = lib = ctypes.CDLL(path)
        ny, npt = ctypes.c_int(), ctypes.c_int()
        lib.oracle_info(ctypes.byref(ny), ctypes.byref(npt))
        self.ny, self.np_tot = ny.value, npt.value
        lib.oracle_integrate_predefined.restype = ctypes.c_int
        lib.oracle_integrate_predefined.argtypes = [
            ctypes.c_int64, ctypes.c_int, _d, _d, ctypes.c_int, _d, _d, ctypes.c_int, ctypes.c_double,
            _d, _d, _d, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int,
            _d, _i32, _i32, _i32, ctypes.c_int]
        lib.oracle_integrate_adaptive.restype = ctypes.c_int
        lib.oracle_integrate_adaptive.argtypes = [
            ctypes.c_int64, _d, _d, _d, _d, _d, ctypes.c_int, ctypes.c_double, _d, _d, _d, ctypes.c_int64,
            ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int64, _d, _i32, _i32, _i32,
            ctypes.c_int]
        lib.oracle_eval.argtypes = [ctypes.c_int64, _d, _d, _d, _d, _d, _d]
        lib.oracle_set_extras.argtypes = [ctypes.c_double, ctypes.c_int, ctypes.c_int, _d, _i32, _i32]
        lib.oracle_lu_solve.argtypes = [ctypes.c_int64, _d, _d, _d, _i32]
        lib.oracle_set_linear_solver.argtypes = [ctypes.c_int]
        lib.oracle_eval_jtimes.argtypes = [ctypes.c_int64, _d, _d, _d, _d, _d]
        lib.oracle_eval_sparse_jac.argtypes = [ctypes.c_int64, _d, _d, _d, _d, _i32, _i32]

    def _extras(self, n, max_invariant_violation, return_on_root, max_roots, sink):
        """ invariant guard / event functions for the next call; ``sink`` (a dict) receives the root arrays """
        if sink is None or max_roots <= 0:
            self.lib.oracle_set_extras(float(max_invariant_violation), 0, 0, None, None, None)
            return
        root_out = np.full((n, max_roots, 1 + self.ny), np.nan)
        root_indices = np.zeros((n, max_roots), dtype=np.int32)
        nfound = np.zeros(n, dtype=np.int32)
        self.lib.oracle_set_extras(float(max_invariant_violation), int(bool(return_on_root)), int(max_roots),
                                   _p(root_out), _p(root_indices, _i32), _p(nfound, _i32))
        sink.update(root_out=root_out, root_indices=root_indices, nroots_found=nfound, max_roots=int(max_roots))

    def _reset_extras(self):
        self.lib.oracle_set_extras(0.0, 0, 0, None, None, None)
        self.lib.oracle_set_linear_solver(0)

    def set_linear_solver(self, which):
        """ 'gmres': the next integrate call (method='bdf') solves its Newton corrections matrix-free on jtimes """
        if self.lib.oracle_set_linear_solver({'dense': 0, 'gmres': 1}[which]) != 0:
            raise RuntimeError("oracle: system generated without jtimes")

    def max_threads(self):
        return int(self.lib.oracle_max_threads())

    @staticmethod
    def _opt(v, n):
        if v is None:
            return None
        v = np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.float64), (n,)))
        return v

    def integrate_predefined(self, y0, xout, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                             mxsteps=0, method='rosenbrock4', autorestart=0, get_dx_max_factor=0.0,
                             error_outside_bounds=False, nthreads=0, max_invariant_violation=0.0,
                             max_roots=0, roots_sink=None):
        y0 = np.ascontiguousarray(y0, dtype=np.float64)
        n = y0.shape[0]
        self._extras(n, max_invariant_violation, False, max_roots, roots_sink)
        xout = np.ascontiguousarray(xout, dtype=np.float64)
        shared = xout.ndim == 1
        nt = xout.shape[-1]
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        atol = np.ascontiguousarray(np.atleast_1d(atol), dtype=np.float64)
        if dx0 is not None and np.all(np.asarray(dx0) == 0):
            dx0 = None
        dx0, dx_min, dx_max = (self._opt(v, n) for v in (dx0, dx_min, dx_max))
        yout = np.empty((n, nt, self.ny))
        nreached = np.empty(n, dtype=np.int32)
        status = np.empty(n, dtype=np.int32)
        counters = np.zeros((6, n), dtype=np.int32)
        rc = self.lib.oracle_integrate_predefined(
            n, nt, _p(y0), _p(xout), int(shared), _p(params), _p(atol), atol.size, float(rtol), _p(dx0),
            _p(dx_min), _p(dx_max), int(mxsteps), METHODS[method], int(autorestart), float(get_dx_max_factor),
            int(error_outside_bounds), _p(yout), _p(nreached, _i32), _p(status, _i32), _p(counters, _i32),
            int(nthreads))
        self._reset_extras()
        if rc != 0:
            raise RuntimeError("oracle: unsupported method")
        return yout, nreached, status, counters

    def integrate_adaptive(self, y0, x0, xend, params, atol, rtol, dx0=None, dx_min=None, dx_max=None,
                           mxsteps=0, method='rosenbrock4', autorestart=0, get_dx_max_factor=0.0,
                           error_outside_bounds=False, max_rows=None, nthreads=0, max_invariant_violation=0.0,
                           return_on_root=False, max_roots=0, roots_sink=None):
        y0 = np.ascontiguousarray(y0, dtype=np.float64)
        n = y0.shape[0]
        self._extras(n, max_invariant_violation, return_on_root, max_roots, roots_sink)
        x0 = np.ascontiguousarray(np.broadcast_to(np.asarray(x0, dtype=np.float64), (n,)))
        xend = np.ascontiguousarray(np.broadcast_to(np.asarray(xend, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        atol = np.ascontiguousarray(np.atleast_1d(atol), dtype=np.float64)
        if dx0 is not None and np.all(np.asarray(dx0) == 0):
            dx0 = None
        dx0, dx_min, dx_max = (self._opt(v, n) for v in (dx0, dx_min, dx_max))
        if max_rows is None:
            max_rows = (int(mxsteps) if mxsteps else 500)*(1 + int(autorestart)) + 1
            if roots_sink is not None:
                max_rows += max_roots
        xyout = np.empty((n, max_rows, 1 + self.ny))
        nrows = np.empty(n, dtype=np.int32)
        status = np.empty(n, dtype=np.int32)
        counters = np.zeros((6, n), dtype=np.int32)
        rc = self.lib.oracle_integrate_adaptive(
            n, _p(y0), _p(x0), _p(xend), _p(params), _p(atol), atol.size, float(rtol), _p(dx0), _p(dx_min),
            _p(dx_max), int(mxsteps), METHODS[method], int(autorestart), float(get_dx_max_factor),
            int(error_outside_bounds), int(max_rows), _p(xyout), _p(nrows, _i32), _p(status, _i32),
            _p(counters, _i32), int(nthreads))
        self._reset_extras()
        if rc != 0:
            raise RuntimeError("oracle: unsupported method")
        return xyout, nrows, status, counters

    def eval_rhs_jac(self, x, y, params, jac=True):
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        f = np.empty((n, self.ny))
        J = np.empty((n, self.ny, self.ny)) if jac else None
        dfdt = np.empty((n, self.ny)) if jac else None
        self.lib.oracle_eval(n, _p(x), _p(y), _p(params), _p(f), _p(J), _p(dfdt))
        return f, J, dfdt

    def eval_jtimes(self, x, y, params, v):
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        v = np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.float64), (n, self.ny)))
        out = np.empty((n, self.ny))
        if self.lib.oracle_eval_jtimes(n, _p(x), _p(y), _p(params), _p(v), _p(out)) != 0:
            raise RuntimeError("oracle: system generated without jtimes")
        return out

    def eval_sparse_jac(self, x, y, params, nnz):
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = y.shape[0]
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, dtype=np.float64), (n,)))
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, self.np_tot)
        data = np.empty((n, nnz))
        colptrs = np.zeros(self.ny + 1, dtype=np.int32)
        rowvals = np.zeros(nnz, dtype=np.int32)
        if self.lib.oracle_eval_sparse_jac(n, _p(x), _p(y), _p(params), _p(data), _p(colptrs, _i32),
                                           _p(rowvals, _i32)) != 0:
            raise RuntimeError("oracle: system generated without sparse=True")
        return data, colptrs, rowvals

    def lu_solve(self, A, b):
        A = np.ascontiguousarray(A, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        n = b.shape[0]
        x = np.empty((n, self.ny))
        ok = np.empty(n, dtype=np.int32)
        self.lib.oracle_lu_solve(n, _p(A), _p(b), _p(x), _p(ok, _i32))
        return x, ok
