# -*- coding: utf-8 -*-
"""
CPU tests (``-m "not gpu"``) of the host side: code generation, argument shaping, the C-ABI library (loads and
exports every symbol ``include/pyodesys_cuda.h`` declares - no compute calls without a GPU), sharding and the
world_size-2 gloo gather.  The reference behaviours they mirror are cited per test.
"""
import ctypes
import os
import re
import socket
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from pyodesys_b200 import systems, native_sys
from pyodesys_b200.core import ODESys, BatchInfo, BatchResult
from pyodesys_b200.multi import shard_range
from pyodesys_b200.native import _codegen

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    with open(os.path.join(_root, 'include', 'pyodesys_cuda.h')) as fh:
        text = fh.read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(pyodesys_cuda_[a-z0-9_]+)\s*\(', text)))


def test_c_abi_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert {'pyodesys_cuda_integrate_predefined', 'pyodesys_cuda_integrate_adaptive', 'pyodesys_cuda_info',
            'pyodesys_cuda_query', 'pyodesys_cuda_last_error', 'pyodesys_cuda_eval', 'pyodesys_cuda_lu_solve',
            'pyodesys_cuda_eval_jtimes', 'pyodesys_cuda_eval_sparse_jac',
            'pyodesys_cuda_launch_count', 'pyodesys_cuda_host_alloc', 'pyodesys_cuda_host_free',
            'pyodesys_cuda_release', 'pyodesys_cuda_device_pci_bus_id', 'pyodesys_cuda_device_count'} <= set(names)
    for sysobj in (systems.robertson(), systems.vdp(), systems.radiolysis()):
        lib = ctypes.CDLL(sysobj._native.so_file)
        for nm in names:
            assert hasattr(lib, nm), nm
        lib.pyodesys_cuda_query.restype = ctypes.c_int64
        lib.pyodesys_cuda_query.argtypes = [ctypes.c_char_p]
        assert lib.pyodesys_cuda_query(b'abi_version') == 4
        assert lib.pyodesys_cuda_query(b'ny') == sysobj.ny
        assert lib.pyodesys_cuda_query(b'np_tot') == len(sysobj.params) + sysobj.ny  # append_iv, core.py:410-411
        assert lib.pyodesys_cuda_query(b'no such key') == -1
        ny, npt = ctypes.c_int(), ctypes.c_int()
        assert lib.pyodesys_cuda_info(ctypes.byref(ny), ctypes.byref(npt), None, None, None) == 0
        assert (ny.value, npt.value) == (sysobj.ny, len(sysobj.params) + sysobj.ny)
    lib = ctypes.CDLL(systems.radiolysis()._native.so_file)
    lib.pyodesys_cuda_query.restype = ctypes.c_int64
    assert lib.pyodesys_cuda_query(b'has_group_kernel') == 1 and lib.pyodesys_cuda_query(b'group_size') == 8


def test_product_path_has_no_cpu_fallback_and_never_imports_the_oracle():
    """ without a CUDA device the integrate call must fail loudly; the package must not reference oracle/ """
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="pyodesys_cuda"):
        systems.decay().integrate([0, 1], [1.0, 0.0], [1.0])
    pkg = os.path.join(_root, 'pyodesys_b200')
    for dirpath, _d, files in os.walk(pkg):
        if '_build' in dirpath:
            continue
        for fn in files:
            if fn.endswith(('.py', '.cu', '.cuh', '.h')):
                with open(os.path.join(dirpath, fn)) as fh:
                    src = fh.read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), fn
                assert 'liboracle' not in src, fn


def test_native_sys_registry_and_constructors():
    """ ``native_sys`` (pyodesys/native/__init__.py:18-22) gains 'cuda'; from_callback / from_other as symbolic.py:384-508 """
    cls = native_sys['cuda']
    a = cls.from_callback(systems.vdp_f, 2, 1, compile=False)
    b = cls.from_other(a, compile=False)
    assert a.ny == b.ny == 2 and a._native.gen.digest == b._native.gen.digest
    assert a.append_iv and a._native.gen.np_tot == 3
    with pytest.raises(ValueError, match="incompatible kwargs integrator"):
        a.integrate([0, 1], [1, 0], [1.0], integrator='cvode')


def test_codegen_two_level_cse_and_flop_counts():
    """ ``_NativeCodeBase.variables`` (native/_base.py:126-281): parameter-only CSE goes to m_p_cse and is evaluated
    once per system (template:104-110); PYODESYS_NATIVE_CSE=0 disables CSE (_base.py:175-179) """
    def f(x, y, p, be):
        k = be.exp(p[0])/(1 + p[1]**2)
        return [-k*y[0], k*y[0] - p[1]*y[1]]
    s = native_sys['cuda'].from_callback(f, 2, 2, compile=False)
    gen = s._native.gen
    assert gen.npcse >= 1 and 'm_p_cse[0] =' in gen.text and 'exp(' in gen.text.split('ODESYS_FN void setup()')[1].split('}')[0]
    rhs_body = gen.text.split('ODESYS_FN AnyODE::Status rhs(')[1].split('this->nfev++')[0]
    assert 'exp(' not in rhs_body  # hoisted out of the per-step code
    assert gen.flops['rhs'] > 0 and gen.flops['jac'] > 0 and gen.has_jac and gen.autonomous
    os.environ['PYODESYS_NATIVE_CSE'] = '0'
    try:
        s0 = native_sys['cuda'].from_callback(f, 2, 2, compile=False)
        assert s0._native.gen.npcse == 0 and s0._native.gen.digest != gen.digest
    finally:
        del os.environ['PYODESYS_NATIVE_CSE']
    rob = systems.robertson(compile=False)._native.gen
    assert (rob.ny, rob.nparams, rob.np_tot, rob.np_used, rob.jac_nnz) == (3, 3, 6, 3, 7)
    with pytest.raises(NotImplementedError):  # _base.py:128-129
        s.band = (1, 1)
        _codegen.generate(s)
    # the kernels load only the first np_used parameter columns: an invariant that reads a parameter the right-hand
    # side does not use must still find it there (round-1 advisor finding)
    import sympy
    y0_, y1_, k0_, k1_ = sympy.symbols('y0 y1 k0 k1', real=True)
    exprs = [(y0_, -k0_*y0_), (y1_, k0_*y0_)]
    base = native_sys['cuda'](exprs, params=(k0_, k1_), compile=False)
    assert base._native.gen.np_used == 1
    sinv = native_sys['cuda'](exprs, params=(k0_, k1_), nonlinear_invariants=[k1_*(y0_ + y1_)], compile=False)
    assert sinv._native.gen.np_used == 2 and 'm_p[1]' in sinv._native.gen.text.split('void invariants(')[1]


def test_generated_header_compiles_as_plain_cxx_and_matches_sympy():
    """ the same generated text feeds the CUDA kernels and the oracle; checked here against lambdified sympy """
    from oracle import build_oracle
    rad = systems.radiolysis(compile=False)
    orc = build_oracle(rad._native)
    rng = np.random.default_rng(0)
    y = rng.uniform(0.1, 2, size=(5, 14))
    p = rng.uniform(0.1, 2, size=(5, 54))
    pf = np.concatenate((p, y), axis=1)
    f, J, dfdt = orc.eval_rhs_jac(0.0, y, pf)
    assert np.allclose(f, rad.f_numeric(0.0, y, pf), rtol=1e-12, atol=1e-12)
    assert np.allclose(J, rad.j_numeric(0.0, y, pf), rtol=1e-12, atol=1e-12)
    assert (dfdt == 0).all()


def test_to_arrays_broadcasting_table():
    """ pyodesys/tests/test_core.py:80-98 """
    odes = ODESys()
    x, y, p = odes.to_arrays(3.0, [1, 2], [5])
    assert x.tolist() == [0, 3] and y.tolist() == [1, 2] and p.tolist() == [5]
    x, y, p = odes.to_arrays([1, 2, 3], [[1, 2], [3, 4]], [5])
    assert x.shape == (2, 3) and y.shape == (2, 2) and p.shape == (2, 1)
    x, y, p = odes.to_arrays([[0, 1], [0, 2], [0, 3]], [1, 2], [[5], [6], [7]])
    assert x.shape == (3, 2) and y.shape == (3, 2) and p[:, 0].tolist() == [5, 6, 7]
    with pytest.raises(ValueError):
        odes.to_arrays([1, 2], [[1, 2], [3, 4]], [[5], [6], [7]])
    named = ODESys(names=('a', 'b'), param_names=('k',), dep_by_name=True, par_by_name=True)
    x, y, p = named.to_arrays(1.0, {'a': [1, 2, 3], 'b': 4}, {'k': 2})
    assert y.tolist() == [[1, 4], [2, 4], [3, 4]] and p.shape == (3, 1)


def test_batch_result_is_lazy_and_list_like():
    n, nt, ny = 5, 3, 2
    counters = np.arange(6*n, dtype=np.int32).reshape(6, n)
    info = BatchInfo('predefined', np.array([0, 0, 1, 0, 0], dtype=np.int32), np.full(n, nt, dtype=np.int32), counters,
                     1e-6, 2e-6, np.tile(np.linspace(0, 1, nt), (n, 1)), np.zeros((n, nt, ny)), np.ones((n, 1)), True,
                     nt=nt, iv=np.full((n, ny), 7.0))
    res = BatchResult(info, ODESys())
    assert len(res) == n and len(list(res)) == n and len(res[1:3]) == 2
    xout, yout, nfo = res[2]
    assert nfo['success'] is False and nfo['nreached'] == nt and nfo['mode'] == 'predefined'
    assert nfo['nfev'] == counters[1, 2] and nfo['njev'] == counters[2, 2] and nfo['n_steps'] == counters[0, 2]
    assert nfo['internal_params'].tolist() == [1.0, 7.0, 7.0]  # lazily appended initial values (core.py:410-411)
    assert res[-1].info['nfev'] == counters[1, 4]
    with pytest.raises(IndexError):
        info[n]


def test_shard_ranges_cover_the_batch():
    for n in (0, 1, 7, 8, 1000003):
        for g in (1, 2, 3, 8):
            spans = [shard_range(n, g, k) for k in range(g)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(hi - lo for lo, hi in spans) <= -(-n // g)
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def test_two_rank_gloo_sharding_and_gather():
    """ N>1 path on CPU: two processes, gloo, each takes its contiguous shard (here integrated by the CPU oracle, which
    stands in for the device), rank 0 gathers; result must equal the single-process run """
    script = textwrap.dedent('''
        import os, sys
        sys.path.insert(0, %r)
        import numpy as np, torch.distributed as dist
        from pyodesys_b200 import systems
        from pyodesys_b200.multi import shard_range, gather_to_rank0
        from oracle import build_oracle
        dist.init_process_group('gloo')
        rank, world = dist.get_rank(), dist.get_world_size()
        n = 101
        b = systems.batch_robertson(n=n, seed=3, tend=10.0)
        lo, hi = shard_range(n, world, rank)
        orc = build_oracle(systems.robertson(compile=False)._native)
        p = np.concatenate((b['params'], b['y0']), axis=1)
        y, nr, st, cn = orc.integrate_predefined(b['y0'][lo:hi], np.array([0.0, 10.0]), p[lo:hi], b['atol'], b['rtol'],
                                                 mxsteps=b['nsteps'], method='rosenbrock4', nthreads=1)
        full = gather_to_rank0(y[:, -1, :], n)
        steps = gather_to_rank0(cn[0].astype(np.int64)[:, None], n)
        if rank == 0:
            yref, _, _, cref = orc.integrate_predefined(b['y0'], np.array([0.0, 10.0]), p, b['atol'], b['rtol'],
                                                        mxsteps=b['nsteps'], method='rosenbrock4', nthreads=1)
            assert full.shape == (n, 3) and np.array_equal(full, yref[:, -1, :])
            assert np.array_equal(steps[:, 0], cref[0])
            print('GATHER_OK', world)
        else:
            assert full is None
        dist.destroy_process_group()
    ''') % _root
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2', '--master-addr', '127.0.0.1',
           '--master-port', str(_free_port()), '-c', script]
    path = os.path.join(_root, 'tests', '_gloo_worker.py')
    with open(path, 'w') as fh:
        fh.write(script)
    try:
        cmd = cmd[:-2] + [path]
        proc = subprocess.run(cmd, capture_output=True, text=True, timeout=300,
                              env=dict(os.environ, OMP_NUM_THREADS='1', MASTER_ADDR='127.0.0.1'))
    finally:
        os.remove(path)
    assert proc.returncode == 0, proc.stderr[-3000:]
    assert 'GATHER_OK 2' in proc.stdout


def test_band_is_not_implemented_like_the_reference():
    """ ``_test_NativeSys__band`` is xfail in the reference: ``native/_base.py:128-129`` raises for banded Jacobians """
    import systems_for_tests as tsys
    with pytest.raises(NotImplementedError, match="Banded"):
        tsys.band_chain3(compile=False)


def test_bench_reference_arm_contract():
    """ ``bench.py --impl reference`` (the CPU arm the driver times beside the CUDA arm): runs without a GPU,
    prints ONE JSON line on stdout with the metric / config of the CUDA arm and the keys of the contract """
    import json
    out = subprocess.run([sys.executable, os.path.join(_root, 'bench.py'), '--impl', 'reference', '--steps', '1',
                          '--warmup', '1', '--cpu-seconds', '1'], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['unit'] == 'systems/s' and d['higher_is_better'] is True
    assert d['config']['workload'] == 'robertson_1e6_rtol1e-8' and d['dtype'] == 'f64'
    assert d['value'] > 0 and d['steps'] == 1 and d['n_gpus'] == 1
    cb = d['cpu_baseline']
    assert cb['kind'] in ('port', 'reference') and cb['cores'] >= 1 and cb['value'] == d['value'] and cb['sample']
    assert d['e2e'] == {"value": d['value'], "unit": d['unit'], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_jtimes_and_sparse_jacobian_codegen_against_sympy():
    """ ``OdeSys::jtimes`` / ``sparse_jac_csc`` / ``get_nnz`` (odesys_anyode_template.cpp:214-243,311-344,131-133)
    generated from ``SymbolicSys.get_jtimes`` / ``be.sparse_jacobian_csc`` (symbolic.py:652-665,641): the oracle
    compiles the same generated text as plain C++; checked against lambdified sympy J and J v """
    import systems_for_tests as tsys
    from oracle import build_oracle
    s = tsys.robertson_jtimes_sparse(compile=False)
    gen = s._native.gen
    assert gen.has_jtimes and gen.nnz == 7 and s.nnz == 7
    assert list(s._colptrs) == [0, 2, 5, 7] and list(s._rowvals) == [0, 1, 0, 1, 2, 0, 1]
    plain = systems.robertson(compile=False)._native.gen
    assert plain.nnz == -1 and not plain.has_jtimes
    orc = build_oracle(s._native)
    rng = np.random.default_rng(11)
    n = 33
    y = rng.uniform(0.1, 2, size=(n, 3))
    p = rng.uniform(0.1, 2, size=(n, 3)) * [0.04, 1e4, 3e7]
    v = rng.standard_normal((n, 3))
    pf = np.concatenate((p, y), axis=1)
    J = s.j_numeric(0.0, y, pf)
    Jv = orc.eval_jtimes(0.0, y, pf, v)
    ref = np.einsum('nij,nj->ni', J, v)
    assert np.allclose(Jv, ref, rtol=1e-13, atol=1e-13*np.abs(ref).max())
    data, colptrs, rowvals = orc.eval_sparse_jac(0.0, y, pf, s.nnz)
    assert colptrs.tolist() == list(s._colptrs) and rowvals.tolist() == list(s._rowvals)
    dense = np.zeros_like(J)
    for c in range(3):
        for k in range(colptrs[c], colptrs[c + 1]):
            dense[:, rowvals[k], c] = data[:, k]
    assert np.allclose(dense, J, rtol=1e-14, atol=0)
    # the dense Jacobian the kernels factorise is assembled from the same CSC entries
    _f, Jd, _ = orc.eval_rhs_jac(0.0, y, pf)
    assert np.array_equal(Jd, dense)
    # the damped oscillator of native/tests/test_cvode.py:198-238: constant Jacobian [[0, 1], [-w0^2, -nu]]
    osc = tsys.osc_jtimes(compile=False)
    yo = rng.standard_normal((4, 2))
    vo = rng.standard_normal((4, 2))
    Jvo = build_oracle(osc._native).eval_jtimes(0.0, yo, yo, vo)
    assert np.allclose(Jvo, np.stack((vo[:, 1], -vo[:, 0] - 0.01*vo[:, 1]), axis=1), rtol=1e-15, atol=1e-15)
    sp_ = tsys.osc_sparse(compile=False)
    assert sp_.nnz == 3 and list(sp_._colptrs) == [0, 1, 3] and list(sp_._rowvals) == [1, 0, 1]
    with pytest.raises(RuntimeError):
        build_oracle(systems.robertson(compile=False)._native).eval_jtimes(0.0, y, pf, v)


def test_linear_solver_names_of_the_cvode_wrapper():
    """ ``linear_solver`` / ``with_jtimes`` (_cvode_wrapper.pyx:88-92): dense and klu map to the register LU, gmres to
    the matrix-free iteration and needs the generated product """
    from pyodesys_b200.native.cuda import _linear_solver_id, NativeCudaCode, Options
    assert _linear_solver_id() == 0 and _linear_solver_id('dense') == 0 and _linear_solver_id('klu') == 0
    assert _linear_solver_id('gmres', with_jtimes=True) == 1
    with pytest.raises(ValueError, match='with_jtimes'):
        _linear_solver_id('gmres')
    with pytest.raises(ValueError, match='Unknown linear_solver'):
        _linear_solver_id('bicgstab', with_jtimes=True)
    # the slot that was ``reserved`` in ABI v3 keeps its offset (old callers passed 0 = dense)
    assert Options.linear_solver.offset == Options.max_roots.offset + 4
    assert Options.root_out.offset == Options.linear_solver.offset + 4


def test_eval_parts_cover_every_row_and_column_once():
    """ the four evaluation pieces of the block-cooperative kernel (``_codegen._balanced_parts``): every row of f and
    every column of J is set by exactly one piece, the largest piece is well below the whole """
    import re
    from pyodesys_b200 import systems
    from pyodesys_b200.native import _codegen
    gen = systems.radiolysis(compile=False)._native.gen
    txt = gen.text
    assert '#define ODESYS_EVAL_PARTS 4' in txt
    ny = gen.ny
    body = txt[txt.index('rhs_sink_part'):txt.index('dense_jac_sink_part')]
    rows = [int(m) for m in re.findall(r's\.set\((\d+),', body)]
    assert sorted(rows) == list(range(ny))
    jbody = txt[txt.index('dense_jac_sink_part'):]
    jbody = jbody[:jbody.index('#endif')]
    ent = re.findall(r's\.setj\((\d+), (\d+),', jbody)
    assert len(ent) == ny*ny and len(set(ent)) == ny*ny
    parts = [jbody.split('if constexpr (P == %d)' % q)[1].split('if constexpr')[0] for q in range(4)]
    cols = [set(int(j) for _i, j in re.findall(r's\.setj\((\d+), (\d+),', pt)) for pt in parts]
    assert all(cols[a].isdisjoint(cols[b]) for a in range(4) for b in range(a)) and set.union(*cols) == set(range(ny))
    # small systems keep the single-piece evaluation
    assert '#define ODESYS_EVAL_PARTS 0' in systems.robertson(compile=False)._native.gen.text
    # the partition helper itself: additive costs -> a balanced split
    pts, cost = _codegen._balanced_parts(8, lambda items: len(items), [1]*8)
    assert sorted(i for pt in pts for i in pt) == list(range(8)) and max(len(pt) for pt in pts) == 2
