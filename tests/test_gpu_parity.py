# -*- coding: utf-8 -*-
"""
GPU parity tests (``-m gpu``): the CUDA path, called through the C ABI, against

* the reference's own known answers (``tests/golden/golden.json``: G1-G9 of SURVEY.md §8c),
* the CPU oracle (``oracle/``) on the same seeded inputs,
* size independent properties (mass conservation, idempotence under tolerance tightening).

Tolerance rule of BASELINE.json ``north_star``: final and tout states within
``max(10*rtol*|y|, 10*atol)`` of the reference path; 1e-12 for the fixed-step RK4 mode.
"""
import numpy as np
import pytest

from pyodesys_b200 import systems
from oracle import build_oracle, RefAnyodeLU

import systems_for_tests as tsys
from systems_for_tests import full_params

pytestmark = pytest.mark.gpu


def bateman_full(y0, lmbd, t):
    """ analytic linear decay chain (restated from the Bateman equation; distinct rates) """
    n = len(lmbd)
    t = np.asarray(t, dtype=float)
    N = [np.zeros_like(t) for _ in range(n)]
    for i0, y00 in enumerate(y0):
        if y00 == 0:
            continue
        lm = lmbd[i0:]
        for i in range(len(lm)):
            prod = np.prod(lm[:i]) if i > 0 else 1.0
            s = np.zeros_like(t)
            for k in range(i + 1):
                den = 1.0
                for l in range(i + 1):
                    if l != k:
                        den *= lm[l] - lm[k]
                s += np.exp(-lm[k]*t)/den
            N[i0 + i] = N[i0 + i] + y00*prod*s
    return np.array(N).T


def within(y, ref, rtol, atol, factor=10):
    return np.abs(y - ref) <= np.maximum(factor*rtol*np.abs(ref), factor*atol)


# ---- G1 --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4', 'bdf'])
def test_vdp_blessed_values(cuda_device, golden, method):
    """ ``_test_NativeSys`` (pyodesys/native/tests/_tests.py:20-29) """
    native = systems.vdp()
    assert native.ny == 2 and len(native.params) == 1
    xout, yout, info = native.integrate([0, 1, 2], [1, 0], params=[2.0], method=method)
    assert np.allclose(yout, golden['known_answers']['vdp_mu2_t012'])
    assert info['nfev'] > 0 and info['success'] and info['mode'] == 'predefined' and info['nreached'] == 3


# ---- G2 --------------------------------------------------------------------------------------------------
def test_rk4_fixed_step_matches_reference_integrator(cuda_device, golden):
    """ 1e-12 against pyodesys.integrators.RK4_example_integrator (test_core.py:149-166) """
    g = golden['rk4_predefined']
    native = systems.vdp()
    res = native.integrate(np.array(g['xout']), [1, 0], params=[2.0], method='rk4')
    assert np.abs(res.yout - np.array(g['yout'])).max() < 1e-12
    assert res.info['nfev'] == g['nfev'] == 4*149
    assert np.allclose(res.yout[-1], [-1.89021896, -0.71633577])
    # the first_step=1e-2 variant through the adaptive API (the reference repeats the end point once)
    g = golden['rk4_adaptive']
    res = native.integrate([0, 2], [1, 0], params=[2.0], method='rk4', first_step=1e-2)
    nrow = res.xout.size
    assert nrow == 201
    assert np.abs(res.xout - np.array(g['xout'])[:nrow]).max() < 1e-12
    assert np.abs(res.yout - np.array(g['yout'])[:nrow]).max() < 1e-12
    assert res.info['nfev'] == g['nfev'] == 800


# ---- G3 / G4 ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4', 'bdf'])
def test_multiple_predefined(cuda_device, method):
    """ ``_test_integrate_multiple_predefined`` (test_core.py:182-193), atol=rtol=1e-10 (test_odeint.py:44-46) """
    odes = systems.decay()
    _xout = np.array([[0, 1, 2], [0, 1, 2], [0, 1, 2]])
    _y0 = np.array([[1, 2], [2, 3], [3, 4]])
    _params = np.array([[5], [6], [7]])
    results = odes.integrate(_xout, _y0, params=_params, atol=1e-10, rtol=1e-10, method=method, nsteps=20000)
    assert len(results) == 3
    for idx in range(3):
        xout, yout, info = results[idx]
        ref = _y0[idx, 0]*np.exp(-_params[idx, 0]*xout)
        assert np.allclose(yout[:, 0], ref)
        assert np.allclose(yout[:, 1], _y0[idx, 0] - ref + _y0[idx, 1])
        assert info['nfev'] > 0


@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4', 'bdf'])
def test_multiple_adaptive(cuda_device, method):
    """ ``_test_integrate_multiple_adaptive`` (test_core.py:281-297): per-system spans """
    odes = systems.sine()
    _xout = np.array([[0, 1], [1, 2], [1, 7]])
    Ak = [[2, 3], [3, 4], [4, 5]]
    _y0 = [[0., A*k] for A, k in Ak]
    _params = [[k] for A, k in Ak]
    results = odes.integrate(_xout, _y0, _params, method=method, nsteps=5000, atol=1e-8, rtol=1e-8)
    for idx in range(3):
        xout, yout, info = results[idx]
        A, k = Ak[idx]
        assert xout[0] == _xout[idx, 0] and xout[-1] == _xout[idx, 1] and xout.size > 5
        assert np.allclose(yout[:, 0], A*np.sin(k*(xout - xout[0])), atol=1e-5, rtol=1e-5)
        assert np.allclose(yout[:, 1], A*np.cos(k*(xout - xout[0]))*k, atol=1e-5, rtol=1e-5)
        assert info['nfev'] > 0 and info['mode'] == 'adaptive' and info['success']


# ---- G5 --------------------------------------------------------------------------------------------------
def test_two_systems_alive_and_bateman(cuda_device, golden):
    """ ``_test_NativeSys_two`` (_tests.py:32-50) """
    native1 = systems.vdp()
    native2 = tsys.decay_k43()
    tol = 1e-11
    xout1, yout1, info1 = native1.integrate([0, 1, 2], [1, 0], params=[2.0], nsteps=500)
    xout2, yout2, info2 = native2.integrate(2, (5, 4, 2), atol=tol, rtol=tol, nsteps=5000, method='dopri5')
    assert np.allclose(yout1, golden['known_answers']['vdp_mu2_t012'])
    ref2 = bateman_full((5, 4, 2), [4, 3, 0], xout2 - xout2[0])
    assert np.allclose(yout2, ref2, rtol=150*tol, atol=150*tol)
    g = golden['bateman']['k43']
    res = native2.integrate(np.array(g['t']), g['y0'], atol=tol, rtol=tol, nsteps=5000)
    assert np.allclose(res.yout, np.array(g['y']), rtol=150*tol, atol=150*tol)


# ---- G6 --------------------------------------------------------------------------------------------------
def test_robertson_1e11_batched(cuda_device, golden):
    """ N=3 identical Robertson systems to t=1e11 (_test_robertson_native.py:52-71; test_robertson.py:15) """
    rob = systems.robertson()
    yref = golden['known_answers']['robertson_1e11']
    res = rob.integrate([(0, 1e11)]*3, [[1, 0, 0]]*3, [[0.04, 1e4, 3e7]]*3, atol=1e-16, rtol=1e-8,
                        nsteps=5000, method='rosenbrock4')
    for r in res:
        assert r.info['success'] and r.info['nfev'] > 100 and r.info['njev'] > 10
        assert np.allclose(yref, r.yout[-1], atol=1e-16, rtol=0.02)
        assert abs(r.yout[-1].sum() - 1) < 1e-9


# ---- G8 --------------------------------------------------------------------------------------------------
def test_return_on_error(cuda_device):
    """ ``_test_return_on_error_success`` (_tests.py:321-340) """
    k, y0 = [4, 3], (5, 4, 2)
    native = tsys.decay_err()
    xout = [0, 0.2, 0.4, 0.6, 0.8, 1.0]
    result = native.integrate(xout, y0, k, atol=1e-11, rtol=1e-11, return_on_error=True, dx_max=.05,
                              nsteps=2000, method='dopri5')
    nreached = result.info['nreached']
    assert nreached == 3
    ref = bateman_full(y0, k + [0], result.xout[:nreached] - xout[0])
    assert result.info['success'] is False
    assert np.allclose(result.yout[:nreached, :], ref, rtol=1e-8, atol=1e-8)
    with pytest.raises(RuntimeError):
        native.integrate(xout, y0, k, atol=1e-11, rtol=1e-11, dx_max=.05, nsteps=2000, method='dopri5')


# ---- G9 --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('method', ['rosenbrock4', 'dopri5'])
def test_exp_y(cuda_device, method):
    """ dy/dx = e^y, y(0)=1 -> y(0.1) = -ln(1/e - 0.1) to 4e-8 (test_symbolic.py:274-286) """
    odesys = tsys.expy()
    res = odesys.integrate(0.1, [1.0], atol=1e-10, rtol=1e-10, method=method, nsteps=1000)
    assert abs(res.yout[-1, 0] - (-np.log(1/np.e - 0.1))) < 4e-8


# ---- first step / dx_max overrides ---------------------------------------------------------------------------
def test_first_step_expr(cuda_device):
    """ ``_test_NativeSys__first_step_cb`` (_tests.py:230-240) """
    odesys = tsys.decay3_first_step()
    y0, k = [.7, 0, 0], [1e23, 2, 3.]
    res = odesys.integrate(5, y0, k, atol=1e-8, rtol=1e-8, nsteps=5000)
    ref = bateman_full(y0, k, res.xout - res.xout[0])
    assert res.info['success'] and res.info['nfev'] > 10
    assert np.allclose(res.yout, ref, atol=20e-8, rtol=20e-8)


def test_get_dx_max_source_code(cuda_device):
    """ ``_test_NativeSys__get_dx_max_source_code`` (_tests.py:275-292) """
    odesys = tsys.decay3_dxmax()
    y0, k = [.7, 0, 0], [7., 2, 3.]
    res = odesys.integrate(1, y0, k, get_dx_max_factor=1.0, atol=1e-8, rtol=1e-8, nsteps=1000, method='dopri5')
    ref = bateman_full(y0, k, res.xout - res.xout[0])
    assert np.allclose(res.yout, ref, atol=20e-8, rtol=20e-8)
    assert res.info['success'] and res.info['nfev'] > 10
    assert 750 < res.info['n_steps'] <= 1000


def test_dep_by_name_single_varied(cuda_device):
    """ ``_test_NativeSys__dep_by_name__single_varied`` (_tests.py:305-318): N=7 by broadcasting """
    tend, kf, y0 = 2, [4, 3], {'a': (5, 3, 7, 9, 1, 6, 11), 'b': 4, 'c': 2}
    odesys = tsys.decay_named()
    results = odesys.integrate(tend, y0)
    assert len(results) == 7
    for idx in range(7):
        xout, yout, info = results[idx]
        assert info['success']
        assert xout.size == yout.shape[0] and yout.shape[1] == 3
        ref = bateman_full([y0[k][idx] if k == 'a' else y0[k] for k in odesys.names], kf + [0], xout - xout[0])
        assert np.allclose(yout, ref)


# ---- generated device functions and the register LU -------------------------------------------------------------
@pytest.mark.parametrize('name', ['robertson', 'vdp', 'lorenz'])
def test_generated_rhs_and_jacobian(cuda_device, name):
    sysobj = getattr(systems, name)()
    mod = sysobj._native.mod
    rng = np.random.default_rng(5)
    n = 257
    y = rng.uniform(0.1, 2.0, size=(n, sysobj.ny))
    p = rng.uniform(0.5, 3.0, size=(n, len(sysobj.params)))
    pf = full_params(y, p)
    f, J, dfdt = mod.eval_rhs_jac(0.3, y, pf)
    f_ref = sysobj.f_numeric(0.3, y, pf)
    J_ref = sysobj.j_numeric(0.3, y, pf)
    assert np.allclose(f, f_ref, rtol=1e-14, atol=1e-14)
    assert np.allclose(J, J_ref, rtol=1e-14, atol=1e-14)
    assert np.all(dfdt == 0)
    fo, Jo, _ = build_oracle(sysobj._native).eval_rhs_jac(0.3, y, pf)
    # same generated text on both sides; nvcc contracts a*b+c into FMA, g++ (x86-64, no -mfma) does not,
    # so agreement is to a few ulp of the largest term rather than bit-exact
    assert np.allclose(f, fo, rtol=1e-13, atol=4e-16*np.abs(fo).max())
    assert np.allclose(J, Jo, rtol=1e-13, atol=4e-16*np.abs(Jo).max())


def test_register_lu_against_anyode(cuda_device):
    """ the reference's own getrf/getrs (oracle/_ref) and its 6x6 known system (test_decomp.cpp:19-45 style) """
    mod = tsys.lu6()._native.mod
    rng = np.random.default_rng(6)
    n = 1000
    A = rng.standard_normal((n, 6, 6))
    A[::7, 0, 0] = 0.0  # force pivoting
    b = rng.standard_normal((n, 6))
    x, ok = mod.lu_solve(A, b)
    xr, info = RefAnyodeLU().solve(A, b)
    assert ok.all() and (info == 0).all()
    scale = np.abs(xr).max(axis=1, keepdims=True)
    assert (np.abs(x - xr) <= 1e-9*scale).all()
    assert np.allclose(np.einsum('nij,nj->ni', A, x), b, atol=1e-8)
    x3, ok3 = systems.robertson()._native.mod.lu_solve(A[:, :3, :3], b[:, :3])
    x3r, _ = RefAnyodeLU().solve(A[:, :3, :3], b[:, :3])
    assert (np.abs(x3 - x3r) <= 1e-9*np.abs(x3r).max(axis=1, keepdims=True)).all()


# ---- CUDA path vs CPU oracle on seeded batches ---------------------------------------------------------------------
def test_robertson_batch_vs_oracle_and_lsoda(cuda_device, golden):
    rob = systems.robertson()
    orc = build_oracle(rob._native)
    n = 4096
    b = systems.batch_robertson(n=n, seed=3)
    res = rob.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                        method='rosenbrock4', force_predefined=True)
    info = res.info
    assert (info.status == 0).all() and (info.nreached == 2).all()
    y_gpu = res.yout[:, -1, :]
    y_cpu, nr, st, cn = orc.integrate_predefined(b['y0'], np.array([0.0, 1e5]), full_params(b['y0'], b['params']),
                                                 b['atol'], b['rtol'], mxsteps=b['nsteps'], method='rosenbrock4')
    assert (st == 0).all()
    ok = within(y_gpu, y_cpu[:, -1, :], b['rtol'], b['atol'])
    assert ok.all(), "worst err/tol %g" % (np.abs(y_gpu - y_cpu[:, -1]) /
                                          np.maximum(10*b['rtol']*np.abs(y_cpu[:, -1]), 10*b['atol'])).max()
    # same algorithm -> nearly the same amount of work
    assert abs(info.counters[0].sum() - cn[0].sum()) <= 0.02*cn[0].sum()
    assert np.abs(y_gpu.sum(axis=1) - 1).max() < 1e-9  # linear invariant A+B+C
    # tight LSODA runs of the reference (first 48 systems of the same seeded batch)
    g = golden['robertson_1e5']
    ref = np.array(g['yout'])[:, -1, :]
    assert np.allclose(np.array(g['y0']), b['y0'][:48]) and np.allclose(np.array(g['k']), b['params'][:48])
    assert within(y_gpu[:48], ref, b['rtol'], b['atol']).all()


def test_robertson_dense_output_vs_lsoda(cuda_device, golden):
    g = golden['robertson_1e5']
    rob = systems.robertson()
    y0, k = np.array(g['y0']), np.array(g['k'])
    # C3's tolerances (rtol 1e-8, atol 1e-12) at every output time, transient points included
    res = rob.integrate(np.array(g['xout']), y0, k, atol=1e-12, rtol=1e-8, nsteps=50000, method='rosenbrock4')
    ref = np.array(g['yout'])
    assert within(res.yout, ref, 1e-8, 1e-12).all()
    orc = build_oracle(rob._native)
    y_cpu, _, st, _ = orc.integrate_predefined(y0, np.array(g['xout']), full_params(y0, k), 1e-12, 1e-8,
                                               mxsteps=50000, method='rosenbrock4')
    assert (st == 0).all() and within(res.yout, y_cpu, 1e-8, 1e-12).all()


def test_vdp_batch_vs_oracle_and_lsoda(cuda_device, golden):
    vd = systems.vdp()
    orc = build_oracle(vd._native)
    g = golden['vdp_lsoda']
    y0 = np.array(g['y0'])
    xo = np.array(g['xout'])
    res = vd.integrate(xo, y0, [[1.0]]*len(y0), atol=1e-10, rtol=1e-10, nsteps=5000, method='dopri5')
    assert within(res.yout, np.array(g['yout']), 1e-10, 1e-10).all()
    y_cpu, nr, st, cn = orc.integrate_predefined(y0, xo, full_params(y0, [[1.0]]), 1e-10, 1e-10, mxsteps=5000,
                                                 method='dopri5')
    assert within(res.yout, y_cpu, 1e-10, 1e-10).all()
    # C1 as specified: adaptive mode, 1e4 initial conditions, t=10
    b = systems.batch_vdp(n=10000)
    res = vd.integrate([b['x']]*10000, b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                       method='dopri5')
    assert (res.info.status == 0).all()
    xy, nrows, st, cn = orc.integrate_adaptive(b['y0'][:512], 0.0, 10.0, full_params(b['y0'][:512], b['params'][:512]),
                                               b['atol'], b['rtol'], mxsteps=b['nsteps'], method='dopri5')
    for i in range(0, 512, 17):
        yo = xy[i, nrows[i] - 1, 1:]
        assert within(res[i].yout[-1], yo, b['rtol'], b['atol']).all()
        assert res[i].xout[-1] == 10.0


def test_lorenz_dense_output(cuda_device, golden):
    g = golden['lorenz_lsoda']
    lo = systems.lorenz()
    res = lo.integrate(np.array(g['xout']), [g['y0']]*4, [g['p']]*4, atol=1e-11, rtol=1e-11, nsteps=5000,
                       method='dopri5')
    assert within(res.yout[0], np.array(g['yout']), 1e-11, 1e-11).all()
    assert np.array_equal(res.yout[0], res.yout[3])  # identical systems give identical bits
    b = systems.batch_lorenz(n=20000, nt=100)
    r = lo.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='dopri5')
    assert (r.info.status == 0).all() and np.isfinite(r.yout).all()
    orc = build_oracle(lo._native)
    y_cpu, _, _, _ = orc.integrate_predefined(b['y0'][:256], b['x'], full_params(b['y0'][:256], b['params'][:256]),
                                              b['atol'], b['rtol'], mxsteps=b['nsteps'], method='dopri5')
    short = b['x'] <= 2.0  # chaotic: only the short horizon is comparable
    assert within(r.yout[:256][:, short], y_cpu[:, short], 1e-8, 1e-8).all()


def test_non_autonomous_rosenbrock(cuda_device):
    fo = tsys.forced()
    orc = build_oracle(fo._native)
    y0 = np.array([[1.0, 0.5], [0.2, 0.1]])
    p = np.array([[50.0, 3.0], [5.0, 1.0]])
    xo = np.linspace(0, 3, 7)
    res = fo.integrate(xo, y0, p, atol=1e-10, rtol=1e-10, nsteps=20000, method='rosenbrock4')
    ref = fo.integrate(xo, y0, p, atol=1e-12, rtol=1e-12, nsteps=20000, method='dopri5')
    assert within(res.yout, ref.yout, 1e-10, 1e-10).all()
    y_cpu, _, st, _ = orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-10, 1e-10, mxsteps=20000,
                                               method='rosenbrock4')
    assert within(res.yout, y_cpu, 1e-10, 1e-10).all()


# ---- error conventions and edge cases ---------------------------------------------------------------------------------
def test_error_conventions(cuda_device):
    odes = systems.decay()
    with pytest.raises(ValueError, match="NaN found in y0"):
        odes.integrate([0, 1], [np.nan, 1], [1.0])
    big = np.ones((40000, 2))  # large batches are screened by the kernel (status 3 before the first step)
    big[31234, 1] = np.nan
    with pytest.raises(ValueError, match="NaN found in y0"):
        odes.integrate([0, 1], big, [[1.0]]*40000, force_predefined=True)
    with pytest.raises(ValueError, match="atol needs to be of length 1 or 2"):
        odes.integrate([0, 1], [1, 1], [1.0], atol=[1e-8, 1e-8, 1e-8])
    with pytest.raises(ValueError, match="incompatible kwargs integrator"):
        odes.integrate([0, 1], [1, 1], [1.0], integrator='cvode')
    with pytest.raises(ValueError, match="dx_max too short"):
        odes.integrate([[0, 1]]*3, [[1, 1]]*3, [[1.0]]*3, dx_max=[0.1, 0.1])
    with pytest.raises(RuntimeError):
        tsys.nojac().integrate([0, 1], [1, 1], [1.0], method='rosenbrock4')
    # vector atol is accepted
    r = odes.integrate([0, 1, 2], [1, 1], [1.0], atol=[1e-9, 1e-8], rtol=1e-9, force_predefined=True)
    assert np.allclose(r.yout[:, 0], np.exp(-r.xout))
    # too few steps -> failure raised / reported
    rob = systems.robertson()
    with pytest.raises(RuntimeError, match="too many steps"):
        rob.integrate([0, 1e5], [1, 0, 0], [0.04, 1e4, 3e7], nsteps=5, force_predefined=True)
    r = rob.integrate([0, 1e5], [1, 0, 0], [0.04, 1e4, 3e7], nsteps=5, force_predefined=True, return_on_error=True)
    assert r.info['success'] is False and r.info['nreached'] == 1 and np.isnan(r.yout[-1]).all()
    # autorestart buys more budget
    r = rob.integrate([0, 1e5], [1, 0, 0], [0.04, 1e4, 3e7], nsteps=50, autorestart=40, force_predefined=True)
    assert r.info['success']


def test_edge_cases(cuda_device):
    odes = systems.decay()
    # single output point, zero-length span, backward integration, ragged x handled per system
    r = odes.integrate([[0.0], [1.0]], [[1, 2], [3, 4]], [[1.0], [2.0]], force_predefined=True)
    assert np.array_equal(r.yout[:, 0, :], [[1, 2], [3, 4]])
    r = odes.integrate([[0.0, 0.0]], [[1, 2]], [[1.0]], force_predefined=True)
    assert np.array_equal(r.yout[0], [[1, 2], [1, 2]])
    r = odes.integrate([0, -1, -2], [1, 0], [1.0], atol=1e-10, rtol=1e-10, method='dopri5')
    assert np.allclose(r.yout[:, 0], np.exp(-r.xout))
    # many more systems than threads in flight, odd sizes
    for n in (1, 31, 33, 129, 70001):
        y0 = np.ones((n, 2))
        p = np.linspace(0.5, 2.0, n)[:, None]
        rr = odes.integrate([0.0, 1.0], y0, p, force_predefined=True, atol=1e-10, rtol=1e-10)
        assert np.allclose(rr.yout[:, -1, 0], np.exp(-p[:, 0]), rtol=1e-8)
    # static and dynamic schedules give identical bits
    b = systems.batch_robertson(n=5000, seed=9, tend=1e2)
    rob = systems.robertson()
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], force_predefined=True)
    r1 = rob.integrate(b['x'], b['y0'], b['params'], schedule='static', **kw)
    r2 = rob.integrate(b['x'], b['y0'], b['params'], schedule='dynamic', **kw)
    assert np.array_equal(r1.yout, r2.yout) and np.array_equal(r1.info.counters, r2.info.counters)


# ---- larger systems: the three Rosenbrock kernel families ------------------------------------------------------------
@pytest.mark.parametrize('kernel', ['thread', 'group', 'thread_smem', 'coop'])
def test_radiolysis_kernel_families_vs_oracle(cuda_device, kernel):
    """ C4 (radiolysis network of examples/_chem_kinet.py:57-85, ny=14, 54 rate parameters): register, lane-group and
    shared-memory kernels against the CPU oracle on the same seeded batch, dense output included """
    rad = systems.radiolysis()
    assert rad.ny == 14 and len(rad.params) == 54
    orc = build_oracle(rad._native)
    n = 777
    b = systems.batch_radiolysis(n=n, seed=4)
    xo = np.array([0.0, 1e-3, 1.0, 40.0, 1e3])
    res = rad.integrate(xo, b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                        method='rosenbrock4', kernel=kernel)
    assert (res.info.status == 0).all() and (res.info.nreached == xo.size).all()
    y_cpu, nr, st, cn = orc.integrate_predefined(b['y0'], xo, full_params(b['y0'], b['params']), b['atol'], b['rtol'],
                                                 mxsteps=b['nsteps'], method='rosenbrock4')
    assert (st == 0).all()
    assert within(res.yout, y_cpu, b['rtol'], b['atol']).all()
    assert abs(int(res.info.counters[0].sum()) - int(cn[0].sum())) <= 0.02*cn[0].sum()
    # adaptive mode: every accepted step, last row = final state of the predefined run
    m = 40
    ra = rad.integrate([(0.0, 1e3)]*m, b['y0'][:m], b['params'][:m], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                       method='rosenbrock4', kernel=kernel)
    for i in range(0, m, 7):
        assert ra[i].xout[0] == 0.0 and ra[i].xout[-1] == 1e3 and ra[i].info['success']
        assert ra[i].xout.size == ra[i].info['n_steps'] - ra[i].info['n_err_test_fails'] + 1
        assert within(ra[i].yout[-1], y_cpu[i, -1], b['rtol'], b['atol']).all()


def test_lane_group_kernel_small_system_and_failures(cuda_device):
    """ ny=6 decay chain on 4-lane groups (two rows per lane, padding rows): analytic Bateman solution, and the
    per-system failure semantics of the group kernel """
    ch = tsys.lu6()
    k = np.array([7.0, 5e3, 3.0, 2e5, 1.0])
    y0 = np.array([1.0, 0, 0, 0, 0, 0])
    xo = np.linspace(0, 2, 6)
    n = 37
    scale = np.linspace(0.5, 2.0, n)[:, None]
    res = ch.integrate(xo, np.tile(y0, (n, 1)), k*scale, atol=1e-12, rtol=1e-10, nsteps=20000, method='rosenbrock4',
                       kernel='group')
    thr = ch.integrate(xo, np.tile(y0, (n, 1)), k*scale, atol=1e-12, rtol=1e-10, nsteps=20000, method='rosenbrock4',
                       kernel='thread')
    for i in (0, 11, 36):
        ref = bateman_full(y0, list(k*scale[i, 0]) + [0], xo)
        assert np.allclose(res.yout[i], ref, rtol=1e-7, atol=1e-10)
        assert within(res.yout[i], thr.yout[i], 1e-10, 1e-12).all()
    assert np.abs(res.yout.sum(axis=2) - 1).max() < 1e-9
    r = ch.integrate(xo, [y0, y0], [k, k], atol=1e-12, rtol=1e-10, nsteps=3, method='rosenbrock4', kernel='group',
                     return_on_error=True)
    assert (r.info.status == 1).all() and (r.info.nreached == 1).all() and np.isnan(r.yout[:, 1:]).all()
    with pytest.raises(RuntimeError, match="not available"):
        systems.robertson().integrate([0, 1], [1, 0, 0], [0.04, 1e4, 3e7], kernel='group', force_predefined=True)


def test_coupled_network_ny28_auto_selects_group_kernel(cuda_device):
    """ C4(ii): 28 species (two coupled copies of the reference network), 32-lane groups, vs the CPU oracle """
    net = systems.radiolysis2()
    assert net.ny == 28
    orc = build_oracle(net._native)
    n = 300
    b = systems.batch_radiolysis2(n=n)
    res = net.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                        method='rosenbrock4', force_predefined=True)
    assert (res.info.status == 0).all()
    y_cpu, nr, st, cn = orc.integrate_predefined(b['y0'], np.array(b['x']), full_params(b['y0'], b['params']),
                                                 b['atol'], b['rtol'], mxsteps=b['nsteps'], method='rosenbrock4')
    assert within(res.yout, y_cpu, b['rtol'], b['atol']).all()
    with pytest.raises(RuntimeError, match="ny <= 16"):
        net.integrate(b['x'], b['y0'][:2], b['params'][:2], kernel='thread', force_predefined=True, method='rosenbrock4')


# ---- BDF (cvode ``method='BDF'`` role): CUDA stepper vs oracle restatement and the reference's known answers ------------
def test_bdf_robertson_batch_vs_oracle_and_lsoda(cuda_device, golden):
    rob = systems.robertson()
    orc = build_oracle(rob._native)
    n = 4096
    b = systems.batch_robertson(n=n, seed=3)
    res = rob.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                        method='bdf', force_predefined=True)
    info = res.info
    assert (info.status == 0).all() and (info.nreached == 2).all()
    y_gpu = res.yout[:, -1, :]
    y_cpu, nr, st, cn = orc.integrate_predefined(b['y0'], np.array([0.0, 1e5]), full_params(b['y0'], b['params']),
                                                 b['atol'], b['rtol'], mxsteps=b['nsteps'], method='bdf')
    assert (st == 0).all()
    ok = within(y_gpu, y_cpu[:, -1, :], b['rtol'], b['atol'])
    assert ok.all(), "worst err/tol %g" % (np.abs(y_gpu - y_cpu[:, -1]) /
                                          np.maximum(10*b['rtol']*np.abs(y_cpu[:, -1]), 10*b['atol'])).max()
    # same algorithm -> about the same amount of work (the step-size factors use fp32 powers on the device)
    assert abs(info.counters[0].sum() - cn[0].sum()) <= 0.05*cn[0].sum()
    # the Jacobian is reused across the steps of an epoch (odesys_prelude.h: one evaluation per order / step selection)
    assert info.counters[2].sum() < 0.25*info.counters[0].sum()
    assert abs(info.counters[2].sum() - cn[2].sum()) <= 0.05*cn[2].sum()
    assert np.abs(y_gpu.sum(axis=1) - 1).max() < 1e-9
    g = golden['robertson_1e5']
    ref = np.array(g['yout'])[:, -1, :]
    assert within(y_gpu[:48], ref, b['rtol'], b['atol'], factor=10).all()
    # dense output rows against the tight LSODA runs of the reference
    y0, k = np.array(g['y0']), np.array(g['k'])
    res = rob.integrate(np.array(g['xout']), y0, k, atol=1e-14, rtol=1e-9, nsteps=50000, method='bdf')
    assert within(res.yout, np.array(g['yout']), 1e-9, 1e-14).all()


def test_bdf_robertson_1e11_batched(cuda_device, golden):
    """ G6 with the BDF stepper: tolerances and assertions of _test_robertson_native.py:52-71 """
    rob = systems.robertson()
    res = rob.integrate([0, 1e11], [[1.0, 0, 0]]*3, [[0.04, 1e4, 3e7]]*3, atol=1e-10, rtol=1e-14, nsteps=5000,
                        method='bdf', force_predefined=True)
    for r in res:
        assert r.info['success'] and r.info['nfev'] > 100 and r.info['njev'] > 10
        assert np.allclose(golden['known_answers']['robertson_1e11'], r.yout[-1], atol=1e-16, rtol=0.02)


def test_bdf_failure_modes(cuda_device):
    rob = systems.robertson()
    b = systems.batch_robertson(n=64, seed=5)
    res = rob.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=20, method='bdf',
                        force_predefined=True, return_on_error=True)
    assert (res.info.status == 1).all() and (res.info.nreached == 1).all() and np.isnan(res.yout[:, 1]).all()
    # ny = 14: BDF runs on the block-cooperative kernel (round 1: "limited to ny <= 8"); parity at the band in
    # tests/test_parity_r2.py, here the failure semantics
    b = systems.batch_radiolysis(n=40)
    rad = systems.radiolysis()
    r = rad.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='bdf',
                      force_predefined=True)
    # the Jacobian is kept over several steps (re-evaluated after convergence failures only)
    assert (r.info.status == 0).all() and (r.info.counters[2] > 0).all()
    assert r.info.counters[2].sum() < 0.5*r.info.counters[0].sum()
    r = rad.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=15, method='bdf',
                      force_predefined=True, return_on_error=True)
    assert (r.info.status == 1).all() and (r.info.nreached == 1).all()


# ---- transformed systems + formulation switching ----------------------------------------------
def test_symmetricsys_native(cuda_device):
    """ ``_test_symmetricsys_nativesys`` (pyodesys/native/tests/_tests.py:75-100): log-log transformed chain """
    from pyodesys_b200 import integrate_chained
    ts = tsys.loglog_decay4()
    k = [7., 3, 2]
    y0 = [1e-20]*4
    y0[0] = 1
    rtol = atol = 1e-7
    xout, yout, info = ts.integrate([1e-12, 1], y0, k, atol=atol, rtol=rtol, first_step=1e-4, nsteps=800)
    ref = bateman_full(y0, k + [0], xout - xout[0])
    assert info['success'] and abs(xout[0]/1e-12 - 1) < 1e-13 and abs(xout[-1] - 1) < 1e-13  # exp(log(x)) round trip
    assert np.allclose(yout, ref, rtol=rtol*150, atol=atol*150)
    # a batch goes through the vectorised processors: same numbers as one system at a time
    res = ts.integrate(np.array([[1e-12, .25, .5, 1.0]]*3), [y0]*3, [k, [5., 3, 2], [7., 1, 2]], atol=atol, rtol=rtol,
                       first_step=1e-4, nsteps=800)
    xb, yb, pb = res.processed()
    assert np.allclose(xb, [[1e-12, .25, .5, 1.0]]*3, rtol=1e-13, atol=0)
    for i, kk in enumerate([k, [5., 3, 2], [7., 1, 2]]):
        assert np.allclose(yb[i], bateman_full(y0, list(kk) + [0], xb[i] - xb[i, 0]), rtol=rtol*150, atol=atol*150)
        assert np.array_equal(res[i].yout, yb[i])


def test_multiple_adaptive_chained(cuda_device):
    """ ``_test_multiple_adaptive_chained`` (pyodesys/native/tests/_tests.py:198-227) with the tolerances the
    odeint variant of the reference runs it at (test_odeint.py: rosenbrock4, atol 1e-8 / rtol 1e-10 scale) """
    from pyodesys_b200 import integrate_chained
    rtol, atol = 1e-10, 1e-10
    ks = [[7e13, 3, 2], [2e5, 3e4, 12.7]]
    y0s = [[1.0, 3.0, 2.0, 5.0], [2.0, 1.0, 3.0, 4.0]]
    t0, tend = 1e-16, 7
    touts = [(t0, tend)]*2
    tsys_, osys = tsys.loglog_decay4(), tsys.lin_decay4()
    comb = integrate_chained([tsys_, osys], {'nsteps': [30, 5000], 'return_on_error': [True, False]}, touts, y0s, ks,
                             atol=atol, rtol=rtol, first_step=1e-14)
    for y0, k, res in zip(y0s, ks, comb):
        xout, yout = res.xout, res.yout
        assert abs(xout[0]/t0 - 1) < 1e-13 and abs(xout[-1] - tend) < 1e-12 and np.all(np.diff(xout) >= 0)
        ref = bateman_full(y0, k + [0], xout - xout[0])
        assert np.allclose(yout, ref, rtol=rtol*1000, atol=atol*1000)
        assert 0 <= res.info['time_cpu'] < 100 and 0 <= res.info['time_wall'] < 100
        assert res.info['success'] is True and res.info['nsys'] == 2 and res.info['nfev'] > 10


def test_robertson_chained_log_then_linear(cuda_device, golden):
    """ ``_test_chained_multi_native`` (pyodesys/native/tests/_test_robertson_native.py:12-71), logc = logt = True:
    N = 3 systems start in log(c), log(t), switch to the linear formulation, reach t = 1e11 """
    from pyodesys_b200 import integrate_chained
    k = (.04, 1e4, 3e7)
    init = (1, 1e-20, 1e-20)
    for odes, kw in [([tsys.loglog_robertson(), systems.robertson()],
                      {'nsteps': [100, 5000], 'return_on_error': [True, False]}),
                     ([systems.robertson()], {'nsteps': [5000]})]:
        results = integrate_chained(odes, kw, [(1e-10, 1e11)]*3, [init]*3, [k]*3, atol=1e-10, rtol=1e-14)
        for res in results:
            x, y, nfo = res
            assert np.allclose(golden['known_answers']['robertson_1e11'], y[-1, :], atol=1e-16, rtol=0.02)
            assert nfo['success'] is True and nfo['nfev'] > 100 and nfo['njev'] > 10 and nfo['nsys'] in (1, 2)


def test_chunk_pipeline_is_invisible(cuda_device):
    """ output-heavy batches are cut into chunks that alternate between two streams (H2D / kernel / D2H overlap);
    systems are independent, so every row must equal the row of an un-chunked call bit for bit """
    lo = systems.lorenz()
    n = 100000  # 100 samples x 3 doubles per system = 250 MB -> 3 chunks of 33792 rows
    b = systems.batch_lorenz(n=n, nt=100)
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='dopri5')
    full = lo.integrate(b['x'], b['y0'], b['params'], **kw)
    assert (full.info.status == 0).all() and (full.info.nreached == 100).all()
    for sl in (slice(0, 5000), slice(33792 - 100, 33792 + 100), slice(n - 3000, n)):
        part = lo.integrate(b['x'], b['y0'][sl], b['params'][sl], **kw)
        assert np.array_equal(part.yout, full.yout[sl])
        assert np.array_equal(part.info.counters, full.info.counters[:, sl])
    # adaptive mode through the same pipeline (rows of (x, y...) per accepted step)
    x2 = np.array([[0.0, 0.5]]*40000)
    ad = lo.integrate(x2, b['y0'][:40000], b['params'][:40000], atol=1e-6, rtol=1e-6, nsteps=400, method='dopri5')
    ad1 = lo.integrate(x2[:64], b['y0'][39000:39064], b['params'][39000:39064], atol=1e-6, rtol=1e-6, nsteps=400,
                       method='dopri5')
    for i in range(64):
        assert np.array_equal(ad[39000 + i].yout, ad1[i].yout) and ad[39000 + i].xout[-1] == 0.5


def test_streaming_call_is_invisible(cuda_device, monkeypatch):
    """ large host-pointer batches run as ONE persistent launch with sliced copies underneath (inputs published
    through a ready word, finished output slices copied back while the kernel is still running): every array must
    equal the chunked-launch path (PYODESYS_CUDA_STREAMING=0) bit for bit, for final-state and dense-output calls,
    with per-system step-size arrays, failures included """
    rob = systems.robertson()
    n = 300000
    b = systems.batch_robertson(n=n)
    b['params'][1234, 0] = np.nan              # a system that fails at once (non-finite rhs)
    b['params'][777, 2] = 3e12                 # and one that runs out of steps
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=3000, method='rosenbrock4', force_predefined=True,
              return_on_error=True, first_step=np.linspace(1e-6, 1e-5, n))
    out = {}
    lib = rob._native.mod.lib
    for flag in ('1', '0'):
        monkeypatch.setenv('PYODESYS_CUDA_STREAMING', flag)
        r1 = rob.integrate(b['x'], b['y0'], b['params'], **kw)
        # the streaming call screens the status words while it waits for the device; the chunked path leaves that
        # to the caller (-1)
        counted = int(lib.pyodesys_cuda_query(b'last_failed_systems'))
        assert counted == (int((np.array(r1.info.status) != 0).sum()) if flag == '1' else -1), (flag, counted)
        r2 = rob.integrate(np.array([0.0, 1.0, 1e2, 1e4, 1e5]), b['y0'], b['params'], **kw)
        out[flag] = (np.array(r1.yout), np.array(r1.info.status), np.array(r1.info.counters), np.array(r1.info.nreached),
                     np.array(r2.yout), np.array(r2.info.nreached))
    for u, v in zip(out['1'], out['0']):
        assert np.array_equal(u, v, equal_nan=True)
    st = out['1'][1]
    assert st[1234] != 0, st[1234]
    assert (st != 0).sum() < n//100, ((st != 0).sum(), st[777])
    assert np.abs(out['1'][0][st == 0, -1].sum(axis=1) - 1).max() < 1e-9


@pytest.mark.parametrize('multiple', [False, True])
def test_partially_solved_symmetric_native_multi(cuda_device, multiple):
    """ G7: ``_test_PartiallySolved_symmetric_native(_multi)`` (pyodesys/native/tests/_tests.py:156-185): decay3 with
    analytic y0 (and y2), log-log transformed, N = 4 per-system (y0, k); Bateman + mass conservation """
    trnsfsys = tsys.partsolved_loglog_decay3(multiple)
    y0s = [[3., 2., 1.], [3.1, 2.1, 1.1], [3.2, 2.3, 1.2], [3.6, 2.4, 1.3]]
    ks = [[3.5, 2.5, 0], [3.3, 2.4, 0], [3.2, 2.1, 0], [3.3, 2.4, 0]]
    kw = dict(atol=1e-8, rtol=1e-8, nsteps=2000)
    results = trnsfsys.integrate([(1e-10, 1)]*len(ks), y0s, ks, **kw)
    for i, (y0, k) in enumerate(zip(y0s, ks)):
        xout, yout, info = results[i]
        ref = bateman_full(y0, k, xout - xout[0])
        assert info['success'] and info['nfev'] > 10 and info['time_cpu'] < 100
        assert yout.shape[1] == 3 and np.allclose(yout, ref, atol=2e-7, rtol=2e-7)
        assert np.allclose(np.sum(yout, axis=1), sum(y0), atol=2e-7, rtol=2e-7)
    xout, yout, info = trnsfsys.integrate([1e-10, 1], y0s[0], ks[0], **kw)
    assert info['success'] and np.allclose(yout, bateman_full(y0s[0], ks[0], xout - xout[0]), atol=2e-7, rtol=2e-7)


def test_scaledsys_native(cuda_device):
    """ ``_test_ScaledSys_NativeSys`` (pyodesys/native/tests/_tests.py:53-72): dep_scaling = 1e8, tol 1e-12,
    nsteps = 4000 as for odeint's rosenbrock4 (native/tests/test_odeint.py:29-30) """
    ss = tsys.scaled_chain4()
    k = [7., 3, 2]
    y0 = [1, 0, 0, 0]
    xout, yout, info = ss.integrate([1e-12, 1], y0, atol=1e-12, rtol=1e-12, nsteps=4000)
    ref = bateman_full(y0, k + [0], xout - xout[0])
    assert info['success'] and np.allclose(yout, ref, rtol=2e-11, atol=2e-11)


def test_decay_nonnegative(cuda_device):
    """ ``_test_Decay_nonnegative`` (_tests.py:98-104) """
    odesys = tsys.decay3_nonneg()
    y0, k = [3., 2., 1.], [3.5, 2.5, 0]
    xout, yout, info = odesys.integrate([1e-10, 1], y0, k, integrator='native')
    ref = bateman_full(y0, k, xout - xout[0])
    assert info['success'] and info['nfev'] > 10 and info['time_cpu'] < 100
    assert np.allclose(yout, ref) and np.allclose(np.sum(yout, axis=1), sum(y0))


def test_partially_solved_system_native_family(cuda_device):
    """ ``_test_PartiallySolvedSystem_Native`` (_tests.py:107-134): the same decay through every native
    formulation - plain, scaled by 42, reduced by the linear invariant, and the log-log transforms of both """
    fam = tsys.partsolved_family()
    assert fam['odesys'].linear_invariants == fam['n_sys'].linear_invariants
    assert len(fam['nll_scaledsys'].nonlinear_invariants) > 0
    y0, k = [3.3, 2.4, 1.5], [3.5, 2.5, 0]
    for name in ('n_sys', 'ns_sys', 'np_sys', 'nll_scaledsys', 'nll_partsys'):
        result = fam[name].integrate([0, .3, .5, .7, .9, 1.3], y0, k, integrator='native', atol=1e-9, rtol=1e-9)
        ref = bateman_full(y0, k, result.xout - result.xout[0])
        assert result.info['success'], name
        assert result.yout.shape == ref.shape and np.allclose(result.yout, ref), name
    # the same batched: per-system (y0, k), N = 5
    rng = np.random.default_rng(8)
    Y0 = np.array(y0)*(1 + 0.2*rng.uniform(size=(5, 3)))
    K = np.array(k)*(1 + 0.2*rng.uniform(size=(5, 3)))
    for name in ('np_sys', 'nll_partsys'):
        results = fam[name].integrate([0, .3, .5, .7, .9, 1.3], Y0, K, atol=1e-9, rtol=1e-9)
        for i in range(5):
            assert np.allclose(results[i].yout, bateman_full(Y0[i], K[i], results[i].xout - results[i].xout[0])), name


def test_chained_parameter_variation(cuda_device):
    """ ``_test_chained_parameter_variation`` (pyodesys/tests/test_symbolic.py:1602-1660, run natively by
    native/tests/test_cvode.py:192-194): exponent e of the rate t**e switched after the first duration; then the
    same for a batch, where every duration is one batched call.  BDF as in the reference (cvode): the symbolic
    d/dt of t**e is e*t**e/t, not finite at t = 0, and only Rosenbrock methods evaluate it """
    from pyodesys_b200 import chained_parameter_variation
    odesys = tsys.power_decay()
    durs = dur1, dur2 = 1.0, 3.0
    e2 = 0

    def analytic(tout, a0, b0, e):
        a = a0*np.exp(-tout**(e + 1)/(e + 1))
        return np.array([a, a0 - a + b0]).T

    def check(result, e1):
        assert result.info['success']
        m1, m2 = result.xout <= dur1, result.xout >= dur1
        y1, y2 = result.yout[m1], result.yout[m2]
        assert np.allclose(y1, analytic(result.xout[m1], y1[0, 0], y1[0, 1], e1), atol=1e-6)
        assert np.allclose(y2, analytic(result.xout[m2] - dur1, y2[0, 0], y2[0, 1], e2), atol=1e-6)
        assert np.allclose(result.calc_invariant_violations(), 0)

    for e1 in (2, 3):
        res = chained_parameter_variation(odesys, durs, {'a': 2, 'b': 1}, {'e': [e1, e2]},
                                          integrate_kwargs=dict(integrator='native', method='bdf'))
        assert res.params[0] == e1  # only the initial parameter is stored (test_symbolic.py:1624)
        check(res, e1)
        res = chained_parameter_variation(odesys, durs, {'a': 2, 'b': 1}, {'e': [e1, e2]}, npoints=5,
                                          integrate_kwargs=dict(integrator='native', method='bdf', nsteps=2000))
        assert res.xout.size == 11 and res.xout[5] == dur1 and res.xout[-1] == dur1 + dur2
        check(res, e1)
    # batch of initial states: one call per duration
    A0 = np.array([[2.0, 1.0], [1.5, 0.5], [3.0, 0.0], [0.7, 0.2]])
    launches0 = odesys._native.mod.launch_count()
    results = chained_parameter_variation(odesys, durs, A0, {'e': [2, e2]}, npoints=3,
                                          integrate_kwargs=dict(method='bdf'))
    assert odesys._native.mod.launch_count() - launches0 == len(durs)
    assert len(results) == 4
    for i, r in enumerate(results):
        assert r.yout[0].tolist() == A0[i].tolist() and r.xout.size == 7
        check(r, 2)


def test_release_and_failed_call_leave_the_library_usable(cuda_device):
    """ ``pyodesys_cuda_release`` frees the per-device state (buffers are re-created lazily); a call that fails on
    its arguments must not poison the next one """
    dec = systems.decay()
    mod = dec._native.mod
    y0 = np.tile([1.0, 0.0], (1000, 1))
    ref = dec.integrate([0, 1.0], y0, np.full((1000, 1), 2.0))
    mod.release(cuda_device)
    mod.release(cuda_device)  # idempotent
    again = dec.integrate([0, 1.0], y0, np.full((1000, 1), 2.0))
    assert np.array_equal(np.array(ref.yout), np.array(again.yout))
    with pytest.raises(RuntimeError, match="pyodesys_cuda"):
        mod.integrate_predefined_arrays(y0, np.array([0.0, 1.0]), np.full((1000, 3), 2.0), 1e-8, 1e-8, kernel=3)
    third = dec.integrate([0, 1.0], y0, np.full((1000, 1), 2.0))
    assert np.array_equal(np.array(ref.yout), np.array(third.yout))


# ---- the alternative Jacobian forms: jtimes (matrix-free product) and sparse_jac_csc ----------------------------
def test_generated_jtimes_and_sparse_jacobian(cuda_device):
    """ device ``OdeSys::jtimes`` / ``sparse_jac_csc`` (template:214-243,311-344) against lambdified sympy and the
    oracle's compilation of the same generated text """
    s = tsys.robertson_jtimes_sparse()
    mod = s._native.mod
    assert mod.has_jtimes and mod.nnz == 7
    rng = np.random.default_rng(12)
    n = 1025
    y = rng.uniform(0.1, 2, size=(n, 3))
    p = rng.uniform(0.1, 2, size=(n, 3)) * [0.04, 1e4, 3e7]
    v = rng.standard_normal((n, 3))
    pf = full_params(y, p)
    J = s.j_numeric(0.0, y, pf)
    ref = np.einsum('nij,nj->ni', J, v)
    Jv = mod.eval_jtimes(0.0, y, pf, v)
    assert np.allclose(Jv, ref, rtol=1e-13, atol=1e-13*np.abs(ref).max())
    orc = build_oracle(s._native)
    Jvo = orc.eval_jtimes(0.0, y, pf, v)
    assert np.allclose(Jv, Jvo, rtol=1e-13, atol=4e-16*np.abs(Jvo).max())
    data, colptrs, rowvals = mod.eval_sparse_jac(0.0, y, pf)
    do, co, ro = orc.eval_sparse_jac(0.0, y, pf, 7)
    assert np.array_equal(colptrs, co) and np.array_equal(rowvals, ro)  # index work: bit-exact
    assert colptrs.tolist() == list(s._colptrs) and rowvals.tolist() == list(s._rowvals)
    assert np.allclose(data, do, rtol=1e-14, atol=0)
    dense = np.zeros_like(J)
    for c in range(3):
        for k in range(colptrs[c], colptrs[c + 1]):
            dense[:, rowvals[k], c] = data[:, k]
    assert np.allclose(dense, J, rtol=1e-14, atol=0)
    _f, Jd, _ = mod.eval_rhs_jac(0.0, y, pf)
    assert np.array_equal(Jd, dense)  # the kernels' dense matrix is assembled from the CSC entries
    # systems generated without the options say so instead of returning garbage
    plain = systems.robertson()._native.mod
    assert plain.nnz == -1 and not plain.has_jtimes
    with pytest.raises(RuntimeError, match='jtimes'):
        plain.eval_jtimes(0.0, y, pf, v)
    with pytest.raises(RuntimeError, match='sparse'):
        plain.eval_sparse_jac(0.0, y, pf)


@pytest.mark.parametrize('which', ['sparse', 'jtimes'])
def test_damped_oscillator_sparse_and_jtimes_systems(cuda_device, which):
    """ ``test_sparse_jac_native_cvode`` / ``test_jtimes_native_cvode`` (native/tests/test_cvode.py:198-238):
    damped oscillator to t = 100 at atol = rtol = 1e-12 against the closed form, N = 64 copies batched """
    nu, x0 = 0.01, 1.0
    odesys = tsys.osc_sparse() if which == 'sparse' else tsys.osc_jtimes()
    n = 64
    xout = np.tile(np.linspace(0, 100, 41), (n, 1))
    y0 = np.tile([x0, 0.0], (n, 1))
    if which == 'sparse':
        kw = dict(linear_solver='klu', with_jacobian=True)
    else:
        kw = dict(linear_solver='gmres', with_jtimes=True, with_jacobian=False)
    res = odesys.integrate(xout, y0, atol=1e-12, rtol=1e-12, nsteps=20000, method='bdf', **kw)
    w = (1.0 - nu**2/4.0)**0.5
    a = (x0**2 + (nu*x0/2)**2/w**2)**0.5
    phi = np.arctan(nu*x0/(2*x0*w))
    for r in (res[0], res[n - 1]):
        ref = a*np.exp(-nu*r.xout/2)*np.cos(w*r.xout - phi)
        assert r.info['success']
        if which == 'sparse':
            assert r.info['njev'] > 0
        else:  # test_cvode.py:215-216
            assert r.info['njvev'] > 0
            assert r.info['njev'] == 0
        assert np.allclose(r.yout[:, 0], ref, rtol=1e-8, atol=1e-8)
    # the reference's own call shape: one system, adaptive mode (every accepted step is a row)
    tout, yout, info = odesys.integrate(100, [x0, 0], atol=1e-12, rtol=1e-12, nsteps=20000, method='bdf', **kw)
    ref = a*np.exp(-nu*tout/2)*np.cos(w*tout - phi)
    assert info['success'] and tout[-1] == 100 and len(tout) > 100
    assert (info['njvev'] > 0 and info['njev'] == 0) if which == 'jtimes' else info['njev'] > 0
    assert np.allclose(yout[:, 0], ref, rtol=1e-8, atol=1e-8)


def test_sparse_decay_chain_against_bateman(cuda_device):
    """ ``test_SymbolicSys_jacobian_sparse`` (pyodesys/tests/test_symbolic.py:145-153) """
    k, y0 = (4, 3), (5, 4, 2)
    odesys = tsys.decay_k43_sparse()
    assert odesys.nnz == 4
    xout, yout, info = odesys.integrate(np.linspace(0, 2, 20), y0, atol=1e-12, rtol=1e-12, nsteps=1000)
    ref = bateman_full(y0, list(k) + [0], xout - xout[0])
    assert info['success']
    assert np.allclose(yout, ref, rtol=1e-10, atol=1e-10)


def test_matrix_free_newton_krylov_bdf_on_robertson(cuda_device):
    """ ``linear_solver='gmres', with_jtimes=True``: stiff Robertson to t = 1e5 with the Newton corrections solved by
    GMRES on ``OdeSys::jtimes``; every trajectory within max(10 rtol |y|, 10 atol) of the dense-LU BDF kernel and of
    the oracle's BDF (tolerance of BASELINE.json's north_star), no Jacobian evaluated """
    s = tsys.robertson_jtimes_sparse()
    rng = np.random.default_rng(21)
    n = 777
    k = np.array([0.04, 1e4, 3e7])*10**rng.uniform(-0.5, 0.5, size=(n, 3))
    y0 = np.tile([1.0, 0.0, 0.0], (n, 1))
    rtol, atol = 1e-8, 1e-12
    xout = np.array([0.0, 1.0, 1e3, 1e5])
    kw = dict(atol=atol, rtol=rtol, nsteps=50000, method='bdf', force_predefined=True)
    res_k = s.integrate(xout, y0, k, linear_solver='gmres', with_jtimes=True, **kw)
    res_d = s.integrate(xout, y0, k, **kw)
    yk = np.array([r.yout for r in res_k])
    yd = np.array([r.yout for r in res_d])
    # both are BDF solutions at rtol: they differ by their (different) accepted global errors
    ratio = (np.abs(yk - yd)/np.maximum(10*rtol*np.abs(yd), 10*atol)).max()
    print('gmres vs dense BDF: max err/tol = %.3g' % ratio)
    assert ratio <= 1.0
    assert np.abs(yk.sum(axis=2) - 1).max() < 1e-6  # global error of two BDF runs, mass is only conserved to that
    info = res_k[0].info
    assert info['success'] and info['njev'] == 0 and info['njvev'] > 0 and info['n_lu'] == 0
    assert res_d[0].info['njev'] > 0 and 'njvev' not in res_d[0].info
    # the oracle's restatement of the same matrix-free iteration (same tolerances), and as an independent referee
    # its dense BDF at a 100 x tighter tolerance
    orc = build_oracle(s._native)
    m = 64
    pf = full_params(y0[:m], k[:m])
    orc.set_linear_solver('gmres')
    yo, _nr, status, co = orc.integrate_predefined(y0[:m], np.tile(xout, (m, 1)), pf, [atol], rtol, mxsteps=50000,
                                                   method='bdf')
    assert (status == 0).all() and (co[2] == 0).all() and (co[3] > 0).all()
    ratio_m = (np.abs(yk[:m] - yo)/np.maximum(10*rtol*np.abs(yo), 10*atol)).max()
    print('gmres BDF kernel vs oracle gmres BDF: max err/tol = %.3g' % ratio_m)
    assert ratio_m <= 1.0
    steps_k = np.array([res_k[i].info['n_steps'] for i in range(m)])
    assert np.abs(steps_k - co[0]).max() <= 0.05*co[0].max()  # same controller: step counts agree to a few percent
    yt, _nr, status, _c = orc.integrate_predefined(y0[:m], np.tile(xout, (m, 1)), pf, [atol*1e-2], rtol*1e-2,
                                                   mxsteps=500000, method='bdf')
    assert (status == 0).all()
    ratio_o = (np.abs(yk[:m] - yt)/np.maximum(10*rtol*np.abs(yt), 10*atol)).max()
    print('gmres BDF vs tight oracle BDF: max err/tol = %.3g' % ratio_o)
    assert ratio_o <= 1.0
    with pytest.raises((RuntimeError, ValueError), match='jtimes'):
        systems.robertson().integrate(xout, y0[:2], k[:2], linear_solver='gmres', with_jtimes=True, **kw)


# ---- adaptive mode: compacted output, bounded device scratch, restart of the unfinished subset ----------
def test_adaptive_output_is_compacted_on_the_device(cuda_device):
    """ C1 as BASELINE states it (van der Pol, 1e4 initial conditions, adaptive mode, nsteps = 1000): only the rows that
    were written cross PCIe (round 1 shipped the whole [N, nsteps + 1, 1 + ny] buffer: 240 MB for ~40 MB of rows), and
    they are the rows of the bounded-buffer call bit for bit - also when the batch does not fit the device scratch in
    one piece """
    vd = systems.vdp()
    mod = vd._native.mod
    b = systems.batch_vdp(n=10000)
    x = np.tile([0.0, 10.0], (10000, 1))
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='dopri5')
    res = vd.integrate(x, b['y0'], b['params'], **kw)
    assert (res.info.status == 0).all()
    nrows = res.info.xout.nrows()
    rowbytes = 3*8
    assert res.info.d2h_bytes <= nrows.sum()*rowbytes + 10000*(8 + 4 + 6*4) + 64
    assert res.info.d2h_bytes < 0.5*10000*1001*rowbytes  # the bounded buffers of round 1: 240 MB
    xy, nr, st, cn, _t = mod.integrate_adaptive_arrays(b['y0'], x[:, 0], x[:, 1], full_params(b['y0'], b['params']),
                                                       b['atol'], b['rtol'], mxsteps=b['nsteps'], method='dopri5')
    assert np.array_equal(nr, nrows) and np.array_equal(cn, res.info.counters)
    for i in (0, 1, 4999, 9999):
        assert np.array_equal(res[i].xout, xy[i, :nr[i], 0]) and np.array_equal(res[i].yout, xy[i, :nr[i], 1:])
        assert res[i].xout[-1] == 10.0 and res[i].info['n_steps'] >= nr[i] - 1
    # scratch limit: 10000 systems x 1001 rows x 24 B = 240 MB -> pieces of <= 1747 systems
    old = mod.adaptive_scratch_bytes
    try:
        mod.adaptive_scratch_bytes = 42 << 20
        res2 = vd.integrate(x, b['y0'], b['params'], **kw)
    finally:
        mod.adaptive_scratch_bytes = old
    assert len(res2.info.xout.pieces) == 6
    for i in (0, 1747, 1748, 5000, 9999):
        assert np.array_equal(res[i].xout, res2[i].xout) and np.array_equal(res[i].yout, res2[i].yout)
    assert np.array_equal(res.info.counters, res2.info.counters)


def test_auto_switch_restarts_only_what_did_not_finish(cuda_device):
    """ ``integrate_auto_switch`` (pyodesys/core.py:865-908): the second formulation sees the unfinished systems only """
    from pyodesys_b200 import integrate_chained
    rtol, atol = 1e-10, 1e-10
    tsys_, osys = tsys.loglog_decay4(), tsys.lin_decay4()
    n = 12
    ks = np.tile([7e13, 3, 2], (n, 1))
    y0s = np.tile([1.0, 3.0, 2.0, 5.0], (n, 1))
    # every fourth system has the full span (needs far more than the 30 steps of the first formulation), the others
    # a span that the step-size ramp from first_step covers in about ten steps (the oracle takes 9)
    spans = [(1e-16, 7.0) if i % 4 == 0 else (1e-16, 1e-16*(1 + 1e-8)) for i in range(n)]
    seen = []

    class Spy:
        def __init__(self, inner):
            self.inner = inner

        def __getattr__(self, name):
            return getattr(self.inner, name)

        def integrate(self, x, y0, params, **kw):
            seen.append(np.atleast_2d(y0).shape[0])
            return self.inner.integrate(x, y0, params, **kw)

    comb = integrate_chained([Spy(tsys_), Spy(osys)], {'nsteps': [30, 5000], 'return_on_error': [True, False]},
                             spans, y0s, ks, atol=atol, rtol=rtol, first_step=1e-14)
    assert seen == [n, n//4], seen
    for i, res in enumerate(comb):
        ref = bateman_full(y0s[i], list(ks[i]) + [0], res.xout - res.xout[0])
        assert res.info['success'] and abs(res.xout[-1]/spans[i][1] - 1) < 1e-12 and np.all(np.diff(res.xout) >= 0)
        assert np.allclose(res.yout, ref, rtol=rtol*1000, atol=atol*1000)
