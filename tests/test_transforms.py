# -*- coding: utf-8 -*-
"""
Host-side logic of the transformed-system front end and of ``integrate_auto_switch`` (CPU only).

Known answers: the doctest examples of the reference (``pyodesys/symbolic.py:985-995`` symmetricsys,
``:1049-1058,1108-1113`` ScaledSys) and the transformation rules of ``pyodesys/util.py:126-187``.
"""
import numpy as np
import pytest
import sympy as sp

from pyodesys_b200 import (PartiallySolvedSystem, Result, ScaledSys, SymbolicSys, TransformedSys, get_logexp,
                           integrate_auto_switch, integrate_chained, symmetricsys)

import systems_for_tests as tsys

logexp = (sp.log, sp.exp)


def test_symmetricsys_doc_example():
    def psimp(exprs):
        return [sp.powsimp(e.expand(), force=True) for e in exprs]
    LogLogSys = symmetricsys(logexp, logexp, exprs_process_cb=psimp)
    mysys = LogLogSys.from_callback(lambda x, y, p: [-y[0], y[0] - y[1]], 2, 0)
    x, (y0, y1) = mysys.indep, mysys.dep
    assert mysys.exprs == (-sp.exp(x), -sp.exp(x) + sp.exp(x + y0 - y1))


def test_scaledsys_doc_examples():
    x = sp.Symbol('x')
    assert ScaledSys([(x, x*x)], dep_scaling=1000).exprs == (x**2/1000,)
    assert ScaledSys([(x, x**3)], dep_scaling=1000).exprs == (x**3/1000000,)
    odesys = ScaledSys.from_callback(lambda x, y, p: [p[0]*y[0]**2], 1, 1, dep_scaling=10)
    (y0,), (p0,) = odesys.dep, odesys.params
    assert odesys.exprs == (p0*y0**2/10,)


def test_incorrect_transform_is_rejected():
    x, y = sp.Symbol('x', real=True), sp.Symbol('y', real=True)
    with pytest.raises(ValueError):
        TransformedSys([(y, -y)], x, dep_transf=[(sp.log(y), sp.exp(2*y))])
    with pytest.raises(ValueError):
        symmetricsys((sp.log, 3))


def test_vectorised_processors_roundtrip():
    """ forward / backward transformation of a whole batch in one call; equals the per-system evaluation """
    ts = tsys.loglog_decay4(compile=False)
    rng = np.random.default_rng(0)
    n, nt = 5, 7
    x = np.sort(rng.uniform(1e-3, 5, (n, nt)), axis=1)
    y0 = rng.uniform(0.1, 3, (n, 4))
    p = rng.uniform(0.5, 2, (n, 3))
    xi, yi, pi = ts.pre_process(x, y0, p)
    assert np.allclose(xi, np.log(x)) and np.allclose(yi, np.log(y0)) and np.array_equal(pi, p)
    for i in range(n):  # one system at a time gives the same numbers
        xs, ys, ps = ts.pre_process(x[i], y0[i], p[i])
        assert np.array_equal(xs, xi[i]) and np.array_equal(ys, yi[i])
    yint = rng.normal(size=(n, nt, 4))
    pint = np.concatenate((p, yi), axis=1)  # internal parameter rows carry the appended initial values
    xo, yo, po = ts.post_process(xi, yint, pint)
    assert np.allclose(xo, x) and np.allclose(yo, np.exp(yint)) and po is pint
    xo1, yo1, _ = ts.post_process(xi[2], yint[2], pint[2])
    assert np.array_equal(xo1, xo[2]) and np.array_equal(yo1, yo[2])
    # ragged (adaptive) output: list of per-system arrays
    xr, yr, _ = ts.post_process([xi[0, :3], xi[1]], [yint[0, :3], yint[1]], pint[:2])
    assert np.allclose(xr[0], x[0, :3]) and np.allclose(yr[1], np.exp(yint[1]))
    # the transformed right-hand side is d(log y)/d(log t) = t f(y)/y
    f = ts.f_numeric(xi[0, 0], yi[0], pint[0])
    k = p[0]
    fy = np.array([-k[0]*y0[0, 0], k[0]*y0[0, 0] - k[1]*y0[0, 1], k[1]*y0[0, 1] - k[2]*y0[0, 2], k[2]*y0[0, 2]])
    assert np.allclose(f, x[0, 0]*fy/y0[0])


class _FakeSys:
    """ stands in for an ODE system: 'integrates' dy/dx = 1 but gives up after ``nsteps`` unit steps """

    def __init__(self):
        self.calls = []

    def integrate(self, x, y0, params=(), nsteps=3, return_on_error=False, **kw):
        x, y0 = np.asarray(x, dtype=float), np.asarray(y0, dtype=float)
        self.calls.append((x.copy(), y0.copy(), nsteps))
        if x.ndim == 2:
            return [self.integrate(x[i], y0[i], params, nsteps, return_on_error) for i in range(len(x))]
        npts = int(min(nsteps, np.ceil(x[1] - x[0])))
        xs = np.minimum(x[0] + np.arange(npts + 1), x[1])
        ok = xs[-1] == x[1]
        if not ok and not return_on_error:
            raise RuntimeError("too many steps")
        return Result(xs, y0 + (xs - x[0])[:, None], np.asarray(params), dict(success=ok, nfev=npts, njev=1,
                                                                             time_cpu=0.5, time_wall=0.25), self)


def test_integrate_auto_switch_logic():
    a, b = _FakeSys(), _FakeSys()
    kw = {'nsteps': [3, 100], 'return_on_error': [True, False]}
    res = integrate_auto_switch([a, b], kw, [(0., 2.), (1., 9.)], [[0.], [10.]], [[1.], [2.]])
    assert integrate_chained is integrate_auto_switch and len(res) == 2
    # first system finished with the first formulation, the second needed the restart
    assert np.array_equal(res[0].xout, [0, 1, 2]) and np.array_equal(res[0].yout[:, 0], [0, 1, 2])
    assert np.array_equal(res[1].xout, np.arange(1, 10)) and np.array_equal(res[1].yout[:, 0], 10 + np.arange(9))
    assert res[1].info == dict(nsys=2, nfev=8, njev=2, time_cpu=1.0, time_wall=0.5, success=True)
    assert res[0].info['nsys'] == 2 and res[0].info['nfev'] == 2
    x2, y2, _ = b.calls[0]  # the restart is handed the unfinished system only
    assert np.array_equal(x2, [[4, 9]]) and np.array_equal(y2, [[13.]])
    # single system, success with the first formulation -> the second is never called
    c, d = _FakeSys(), _FakeSys()
    r = integrate_auto_switch([c, d], kw, (0., 2.), [5.], [1.])
    assert r.info['nsys'] == 1 and not d.calls and np.array_equal(r.yout[:, 0], [5, 6, 7])
    with pytest.raises(NotImplementedError):
        integrate_auto_switch([c], {}, [0., 1., 2.], [5.], [1.])


def test_partially_solved_system_processors():
    """ PartiallySolvedSystem (pyodesys/symbolic.py:1134-1358): eliminated variables, appended initial values,
    vectorised reconstruction; then stacked under a log-log transformation like the reference's native test """
    odesys = SymbolicSys.from_callback(tsys.decay3_rhs, 3, 3)
    x, (y0s, y1s, y2s), (k0, k1, k2) = odesys.indep, odesys.dep, odesys.params

    def subs(x0, y0, p0, be):
        a0 = y0[0]*be.exp(-p0[0]*(x - x0))
        return {y0s: a0, y2s: y0[0] + y0[1] + y0[2] - a0 - y1s}
    ps = PartiallySolvedSystem(odesys, subs)
    assert ps.dep == (y1s,) and len(ps.params) == 3 + 1 + 3 and ps.ny == 1
    px0, py0 = ps.params[3], ps.params[4]
    assert sp.simplify(ps.exprs[0] - (k0*py0*sp.exp(-k0*(x - px0)) - k1*y1s)) == 0
    X = np.array([[0., 1, 2], [1., 2, 3]])
    Y0 = np.array([[3., 2, 1], [4., 5, 6]])
    P = np.array([[3.5, 2.5, 0], [1., 2, 0]])
    xi, yi, pi = ps.pre_process(X, Y0, P)
    assert np.array_equal(yi, Y0[:, 1:2]) and np.array_equal(pi, np.concatenate((P, X[:, :1], Y0), axis=1))
    yint = np.full((2, 3, 1), 1.5)
    xo, yo, po = ps.post_process(xi, yint, np.concatenate((pi, yi), axis=1))  # internal rows carry appended iv
    a0 = Y0[:, :1]*np.exp(-P[:, :1]*(X - X[:, :1]))
    assert np.allclose(yo[..., 0], a0) and np.allclose(yo[..., 1], 1.5)
    assert np.allclose(yo.sum(axis=-1), Y0.sum(axis=1, keepdims=True)) and np.array_equal(po, P)
    xo1, yo1, po1 = ps.post_process(xi[1], yint[1], np.concatenate((pi, yi), axis=1)[1])
    assert np.array_equal(yo1, yo[1]) and np.array_equal(po1, P[1])
    # per-component atol refers to the original variables
    assert ps._keep_idx == [1] and ps._analytic_idx == [0, 2]
    ts = tsys.partsolved_loglog_decay3(True, compile=False)
    assert ts.ny == 1 and len(ts.pre_processors) == 2 and len(ts.post_processors) == 2
    xi, yi, pi = ts.pre_process(np.array([[1e-10, 1.]]*2), Y0, P)
    assert np.allclose(xi, np.log([[1e-10, 1.]]*2)) and np.allclose(yi, np.log(Y0[:, 1:2]))
    assert np.allclose(pi, np.concatenate((P, [[1e-10]]*2, Y0), axis=1))
    with pytest.raises(ValueError):
        PartiallySolvedSystem(odesys, lambda x0, y0, p0: {})


def test_from_linear_invariants():
    """ mass conservation A + B + C = const removes one variable (pyodesys/symbolic.py:1275-1336) """
    from pyodesys_b200 import systems
    rob = SymbolicSys.from_callback(systems.robertson_f, 3, 3, linear_invariants=[[1, 1, 1]])
    red = PartiallySolvedSystem.from_linear_invariants(rob, preferred=[2])
    A, B, C = rob.dep
    assert red.dep == (A, B) and red.linear_invariants is None
    iv = red.params[4:]
    assert sp.simplify(red.analytic_exprs[C] - (iv[2] - (A - iv[0]) - (B - iv[1]))) == 0
