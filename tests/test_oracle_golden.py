# -*- coding: utf-8 -*-
"""
Pins the CPU oracle (``oracle/``) before anything is compared against it:

* every known answer the reference's tests hold for the batched native path (G1-G9, G11 of
  SURVEY.md §8c / BASELINE.md §2),
* fixtures generated in the build container by importing the Python reference
  (``tests/golden/make_golden.py``: RK4_example_integrator, ODESys + scipy LSODA, bateman.py),
* the reference's own LU (``oracle/_ref``: AnyODE ``anyode_blasless.hpp`` compiled from
  ``/root/reference``).

CPU only - these run in the build container with ``-m "not gpu"``.
"""
import os

import numpy as np
import pytest

from pyodesys_b200 import systems
from oracle import build_oracle, build_ref, RefAnyodeLU
from oracle import check_ros4_coeffs

import systems_for_tests as tsys
from systems_for_tests import full_params
from test_gpu_parity import bateman_full, within


def _orc(sysobj):
    return build_oracle(sysobj._native)


@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4'])
def test_g1_vdp_blessed(golden, method):
    orc = _orc(systems.vdp(compile=False))
    y, nr, st, cn = orc.integrate_predefined([[1.0, 0]], [[0, 1., 2]], [[2.0, 1, 0]], 1e-8, 1e-8, mxsteps=500,
                                             method=method)
    assert st[0] == 0 and nr[0] == 3
    assert np.allclose(y[0], golden['known_answers']['vdp_mu2_t012'])


def test_g2_rk4_bit_level(golden):
    orc = _orc(systems.vdp(compile=False))
    g = golden['rk4_predefined']
    y, nr, st, cn = orc.integrate_predefined([[1.0, 0]], [g['xout']], [[2.0, 1, 0]], 1e-8, 1e-8, method='rk4')
    assert np.abs(y[0] - np.array(g['yout'])).max() < 1e-12
    assert cn[1, 0] == g['nfev'] == 596
    g = golden['rk4_adaptive']
    xy, nrows, st, cn = orc.integrate_adaptive([[1.0, 0]], 0.0, 2.0, [[2.0, 1, 0]], 1e-8, 1e-8, dx0=1e-2,
                                               method='rk4', mxsteps=500)
    n = nrows[0]
    assert n == 201 and cn[1, 0] == g['nfev'] == 800
    assert np.abs(xy[0, :n, 0] - np.array(g['xout'])[:n]).max() < 1e-12
    assert np.abs(xy[0, :n, 1:] - np.array(g['yout'])[:n]).max() < 1e-12


@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4'])
def test_g3_g4_batched(method):
    orc = _orc(systems.decay(compile=False))
    xout = np.array([[0, 1, 2.]]*3)
    y0 = np.array([[1., 2], [2, 3], [3, 4]])
    p = np.array([[5.], [6], [7]])
    y, nr, st, cn = orc.integrate_predefined(y0, xout, full_params(y0, p), 1e-10, 1e-10, mxsteps=2000, method=method)
    for i in range(3):
        ref = y0[i, 0]*np.exp(-p[i, 0]*xout[i])
        assert np.allclose(y[i, :, 0], ref) and np.allclose(y[i, :, 1], y0[i, 0] - ref + y0[i, 1])
    orc = _orc(systems.sine(compile=False))
    Ak = [[2, 3], [3, 4], [4, 5]]
    y0 = np.array([[0., A*k] for A, k in Ak])
    p = np.array([[k] for A, k in Ak], dtype=float)
    xy, nrows, st, cn = orc.integrate_adaptive(y0, [0., 1, 1], [1., 2, 7], full_params(y0, p), 1e-8, 1e-8,
                                               mxsteps=5000, method=method)
    for i, (A, k) in enumerate(Ak):
        x = xy[i, :nrows[i], 0]
        assert x[-1] == [1., 2, 7][i] and st[i] == 0
        assert np.allclose(xy[i, :nrows[i], 1], A*np.sin(k*(x - x[0])), atol=1e-5, rtol=1e-5)


def test_g5_bateman(golden):
    orc = _orc(tsys.decay_k43(compile=False))
    g = golden['bateman']['k43']
    tol = 1e-11
    for method in ('dopri5', 'rosenbrock4'):
        y, nr, st, cn = orc.integrate_predefined([g['y0']], [g['t']], [g['y0']], tol, tol, mxsteps=20000, method=method)
        assert st[0] == 0
        assert np.allclose(y[0], np.array(g['y']), rtol=150*tol, atol=150*tol)
    assert np.allclose(bateman_full(g['y0'], [4, 3, 0], g['t']), np.array(g['y']), rtol=1e-13)


def test_g6_robertson_1e11(golden):
    orc = _orc(systems.robertson(compile=False))
    y0 = np.array([[1., 0, 0]]*3)
    k = np.array([[0.04, 1e4, 3e7]]*3)
    y, nr, st, cn = orc.integrate_predefined(y0, [0, 1e11], full_params(y0, k), 1e-16, 1e-8, mxsteps=5000,
                                             method='rosenbrock4')
    assert (st == 0).all() and (cn[1] > 100).all() and (cn[2] > 10).all()
    assert np.allclose(golden['known_answers']['robertson_1e11'], y[0, -1], atol=1e-16, rtol=0.02)
    assert np.allclose(golden['robertson_std_1e11_lsoda'], y[0, -1], atol=1e-16, rtol=1e-4)


def test_g8_return_on_error():
    orc = _orc(tsys.decay_err(compile=False))
    y0 = [[5., 4, 2]]
    y, nr, st, cn = orc.integrate_predefined(y0, [[0, 0.2, 0.4, 0.6, 0.8, 1.0]], full_params(y0, [[4., 3]]), 1e-11,
                                             1e-11, dx_max=0.05, mxsteps=2000, method='dopri5')
    assert nr[0] == 3 and st[0] != 0
    ref = bateman_full(y0[0], [4, 3, 0], np.array([0, 0.2, 0.4]))
    assert np.allclose(y[0, :3], ref, rtol=1e-8, atol=1e-8) and np.isnan(y[0, 3:]).all()


@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4'])
def test_g9_exp_y(method):
    orc = _orc(tsys.expy(compile=False))
    y, nr, st, cn = orc.integrate_predefined([[1.0]], [[0, 0.1]], [[1.0]], 1e-10, 1e-10, mxsteps=1000, method=method)
    assert abs(y[0, -1, 0] - (-np.log(1/np.e - 0.1))) < 4e-8


def test_g11_lu_against_reference_anyode():
    if build_ref() is None:
        pytest.skip("oracle/_ref not built and /root/reference absent")
    ref = RefAnyodeLU()
    orc = _orc(tsys.lu6(compile=False))
    # the 6x6 system style of external/anyode/tests/decomp/test_decomp.cpp:19-45: diagonally dominant band
    A = np.zeros((1, 6, 6))
    for i in range(6):
        A[0, i, i] = 5 + i
        if i > 0:
            A[0, i, i - 1] = 2
        if i < 5:
            A[0, i, i + 1] = 3
    xtrue = np.arange(1., 7)[None, :]
    b = np.einsum('nij,nj->ni', A, xtrue)
    x, ok = orc.lu_solve(A, b)
    xr, info = ref.solve(A, b)
    assert np.abs(x - xtrue).max() < 2e-13 and np.abs(xr - xtrue).max() < 2e-13
    rng = np.random.default_rng(0)
    A = rng.standard_normal((500, 6, 6))
    A[::5, 0, 0] = 0
    b = rng.standard_normal((500, 6))
    x, ok = orc.lu_solve(A, b)
    xr, info = ref.solve(A, b)
    assert np.array_equal(x, xr)  # same algorithm, same operation order -> same bits


def test_robertson_1e5_against_reference_lsoda(golden):
    """ the restated Rosenbrock-4 vs tight LSODA runs of the reference ODESys (C3 inputs) """
    g = golden['robertson_1e5']
    orc = _orc(systems.robertson(compile=False))
    y0, k, ref = np.array(g['y0']), np.array(g['k']), np.array(g['yout'])
    y, nr, st, cn = orc.integrate_predefined(y0, g['xout'], full_params(y0, k), 1e-12, 1e-8, mxsteps=50000,
                                             method='rosenbrock4')
    assert (st == 0).all()
    assert within(y, ref, 1e-8, 1e-12, factor=10).all()
    assert np.abs(y.sum(axis=2) - 1).max() < 1e-10


def test_vdp_and_lorenz_against_reference_lsoda(golden):
    g = golden['vdp_lsoda']
    orc = _orc(systems.vdp(compile=False))
    y0 = np.array(g['y0'])
    y, nr, st, cn = orc.integrate_predefined(y0, g['xout'], full_params(y0, [[1.0]]), 1e-10, 1e-10, mxsteps=5000,
                                             method='dopri5')
    assert within(y, np.array(g['yout']), 1e-10, 1e-10, factor=10).all()
    g = golden['lorenz_lsoda']
    orc = _orc(systems.lorenz(compile=False))
    y, nr, st, cn = orc.integrate_predefined([g['y0']], g['xout'], full_params([g['y0']], [g['p']]), 1e-11, 1e-11,
                                             mxsteps=5000, method='dopri5')
    assert within(y[0], np.array(g['yout']), 1e-11, 1e-11, factor=10).all()


def test_dense_output_orders():
    """ the continuous extensions hit the step end exactly and converge with the expected order """
    orc = _orc(systems.sine(compile=False))
    y0 = [[0.0, 1.0]]
    for method, tol in (('dopri5', 1e-9), ('rosenbrock4', 1e-9)):
        xo = np.linspace(0, 3, 301)
        y, nr, st, cn = orc.integrate_predefined(y0, [xo], full_params(y0, [[1.0]]), tol, tol, mxsteps=100000,
                                                 method=method)
        assert cn[0, 0] < 300  # far fewer steps than output points: most values are interpolated
        assert np.abs(y[0, :, 0] - np.sin(xo)).max() < 2e-6


def test_ros4_coefficients_satisfy_order_conditions():
    assert check_ros4_coeffs.main() < 5e-13


def test_first_step_and_dx_max_hooks():
    orc = _orc(tsys.decay3_dxmax(compile=False))
    y0, k = [[.7, 0, 0]], [[7., 2, 3.]]
    y, nr, st, cn = orc.integrate_predefined(y0, [[0, 1.0]], full_params(y0, k), 1e-8, 1e-8, mxsteps=1000,
                                             method='dopri5', get_dx_max_factor=1.0)
    assert st[0] == 0 and 750 < cn[0, 0] <= 1000
    orc = _orc(tsys.decay3_first_step(compile=False))
    y0, k = [[.7, 0, 0]], [[1e23, 2, 3.]]
    xy, nrows, st, cn = orc.integrate_adaptive(y0, 0.0, 5.0, full_params(y0, k), 1e-8, 1e-8, mxsteps=5000,
                                               method='rosenbrock4')
    assert st[0] == 0
    assert abs(xy[0, 1, 0] - 0.7e-30) < 1e-40  # the first step is the generated expression y0*1e-30
    ref = bateman_full(y0[0], k[0], xy[0, :nrows[0], 0])
    assert np.allclose(xy[0, :nrows[0], 1:], ref, atol=20e-8, rtol=20e-8)


def test_oracle_header_says_test_only():
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, '..', 'oracle', 'oracle.cpp')) as fh:
        assert 'TEST INFRASTRUCTURE ONLY' in fh.read(300)
    # the product package never imports the oracle
    pkg = os.path.join(here, '..', 'pyodesys_b200')
    for dirpath, _dirs, files in os.walk(pkg):
        for fn in files:
            if fn.endswith('.py'):
                with open(os.path.join(dirpath, fn)) as fh:
                    src = fh.read()
                assert 'import oracle' not in src and 'from oracle' not in src, fn


# ---- BDF (cvode ``method='BDF'`` role) ---------------------------------------------------------------------
def test_bdf_oracle_against_known_answers(golden):
    """ restated variable-order BDF/NDF vs the reference's known answers: G1 (vdP blessed values), G3 (batched
    decay), G6 (Robertson t=1e11, tolerances of _test_robertson_native.py:52-71), tight LSODA fixtures at t=1e5 """
    orc = _orc(systems.vdp(compile=False))
    y, nr, st, cn = orc.integrate_predefined([[1.0, 0]], [[0, 1., 2]], [[2.0, 1, 0]], 1e-8, 1e-8, mxsteps=5000,
                                             method='bdf')
    assert st[0] == 0 and nr[0] == 3
    assert np.allclose(y[0], golden['known_answers']['vdp_mu2_t012'])
    orc = _orc(systems.decay(compile=False))
    xout = np.array([[0, 1, 2.]]*3)
    y0 = np.array([[1., 2], [2, 3], [3, 4]])
    p = np.array([[5.], [6], [7]])
    y, nr, st, cn = orc.integrate_predefined(y0, xout, full_params(y0, p), 1e-10, 1e-10, mxsteps=20000, method='bdf')
    for i in range(3):
        ref = y0[i, 0]*np.exp(-p[i, 0]*xout[i])
        assert np.allclose(y[i, :, 0], ref) and np.allclose(y[i, :, 1], y0[i, 0] - ref + y0[i, 1])
    orc = _orc(systems.robertson(compile=False))
    y0 = np.array([[1.0, 0, 0]]*3)
    k = np.array([[0.04, 1e4, 3e7]]*3)
    y, nr, st, cn = orc.integrate_predefined(y0, [0, 1e11], full_params(y0, k), 1e-10, 1e-14, mxsteps=5000, method='bdf')
    assert (st == 0).all()
    assert np.allclose(golden['known_answers']['robertson_1e11'], y[0, -1], atol=1e-16, rtol=0.02)
    assert cn[1, 0] > 100 and cn[2, 0] > 10  # nfev, njev as asserted by the reference test
    g = golden['robertson_1e5']
    y0, k, ref = np.array(g['y0']), np.array(g['k']), np.array(g['yout'])
    y, nr, st, cn = orc.integrate_predefined(y0, g['xout'], full_params(y0, k), 1e-12, 1e-8, mxsteps=50000, method='bdf')
    assert (st == 0).all()
    assert within(y, ref, 1e-8, 1e-12, factor=10).all()
    assert np.abs(y.sum(axis=2) - 1).max() < 1e-10
    # the two stiff methods agree with each other far inside the band when both run tight
    ya, _, sa, _ = orc.integrate_predefined(y0, g['xout'], full_params(y0, k), 1e-15, 1e-11, mxsteps=500000, method='bdf')
    yb, _, sb, _ = orc.integrate_predefined(y0, g['xout'], full_params(y0, k), 1e-15, 1e-11, mxsteps=500000,
                                            method='rosenbrock4')
    assert (sa == 0).all() and (sb == 0).all()
    assert within(ya, yb, 1e-9, 1e-14, factor=10).all()


def test_bdf_oracle_adaptive_rows_and_dense_output():
    """ adaptive mode returns every accepted step; interpolated rows equal landing on the points """
    orc = _orc(systems.sine(compile=False))
    y0 = np.array([[0., 6.0]])
    p = np.array([[3.0]])
    xy, nrows, st, cn = orc.integrate_adaptive(y0, [0.], [2.], full_params(y0, p), 1e-9, 1e-9, mxsteps=50000,
                                               method='bdf', max_rows=50001)
    n = nrows[0]
    x = xy[0, :n, 0]
    assert st[0] == 0 and x[0] == 0 and x[-1] == 2.0 and np.all(np.diff(x) > 0) and n == cn[0, 0] - cn[5, 0] + 1
    assert np.allclose(xy[0, :n, 1], 2*np.sin(3*x), atol=2e-6, rtol=0)
    xo = np.linspace(0, 2, 41)
    y, nr, st, cn = orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-9, 1e-9, mxsteps=50000, method='bdf')
    assert st[0] == 0 and nr[0] == 41
    assert np.allclose(y[0, :, 0], 2*np.sin(3*xo), atol=2e-6, rtol=0)


def test_oracle_matrix_free_bdf_against_closed_form_and_dense():
    """ the oracle's Newton-Krylov BDF (GMRES on the generated ``jtimes``): the reference's damped oscillator
    (native/tests/test_cvode.py:198-217: closed form, njvev > 0, njev == 0) and Robertson against the oracle's own
    dense-LU BDF within max(10 rtol |y|, 10 atol) """
    import systems_for_tests as tsys
    from systems_for_tests import full_params
    from oracle import build_oracle
    nu, x0 = 0.01, 1.0
    osc = build_oracle(tsys.osc_jtimes(compile=False)._native)
    xout = np.linspace(0, 100, 41)[None, :]
    y0 = np.array([[x0, 0.0]])
    osc.set_linear_solver('gmres')
    yout, nreached, status, c = osc.integrate_predefined(y0, xout, y0, [1e-12], 1e-12, mxsteps=20000, method='bdf')
    assert status[0] == 0 and c[2, 0] == 0 and c[3, 0] > 0
    w = (1.0 - nu**2/4.0)**0.5
    a = (x0**2 + (nu*x0/2)**2/w**2)**0.5
    phi = np.arctan(nu*x0/(2*x0*w))
    ref = a*np.exp(-nu*xout[0]/2)*np.cos(w*xout[0] - phi)
    assert np.allclose(yout[0, :, 0], ref, rtol=1e-8, atol=1e-8)
    rob = build_oracle(tsys.robertson_jtimes_sparse(compile=False)._native)
    rng = np.random.default_rng(21)
    n = 8
    k = np.array([0.04, 1e4, 3e7])*10**rng.uniform(-0.5, 0.5, size=(n, 3))
    y0 = np.tile([1.0, 0.0, 0.0], (n, 1))
    xo = np.tile([0.0, 1.0, 1e3, 1e5], (n, 1))
    rob.set_linear_solver('gmres')
    yk, _n, st, ck = rob.integrate_predefined(y0, xo, full_params(y0, k), [1e-12], 1e-8, mxsteps=50000, method='bdf')
    yd, _n, st2, cd = rob.integrate_predefined(y0, xo, full_params(y0, k), [1e-12], 1e-8, mxsteps=50000, method='bdf')
    assert (st == 0).all() and (st2 == 0).all()
    assert (ck[2] == 0).all() and (cd[2] > 0).all()  # the setting does not outlive the call
    assert (np.abs(yk - yd) <= np.maximum(1e-7*np.abs(yd), 1e-11)).all()
    with pytest.raises(RuntimeError):
        build_oracle(tsys.decay3(compile=False)._native).set_linear_solver('gmres')
