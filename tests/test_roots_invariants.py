# -*- coding: utf-8 -*-
"""
Event functions (``roots``) and the invariant guard (``max_invariant_violation``) - SURVEY.md §8f-4.

Reference behaviour that is pinned here:

* ``_test_NativeSys__roots`` / ``test_NativeSys__roots`` (``pyodesys/native/tests/_tests.py:262-273``,
  ``native/tests/test_cvode.py:135-150``): y' = y, root of ``y - e``: one entry in ``info['root_indices']``,
  a point of ``xout`` within 1e-11 of x = 1, ``success`` with and without ``return_on_root``;
* ``test_NativeSys__PartiallySolvedSystem__roots`` (``native/tests/test_cvode.py:153-176``): crossings of the
  A -> B -> C chain at x = 0.11299628093544488, 0.20674119231833346, 0.3541828705348678 (1e-7), also for the
  partially solved system, ``return_on_root=True`` ends the integration there;
* the violation test of the generated ``rhs`` (``odesys_anyode_template.cpp:112-123,191-198``).

CPU part: the oracle against those known answers.  GPU part (``-m gpu``): the CUDA kernels against the oracle
and the same known answers through the public ``integrate`` call.
"""
import numpy as np
import pytest

from oracle import build_oracle

import systems_for_tests as tsys
from systems_for_tests import full_params

CHAIN_ROOTS = [0.11299628093544488, 0.20674119231833346, 0.3541828705348678]  # test_cvode.py:165
_orcs = {}


def _orc(system):
    key = system._native.gen.digest
    if key not in _orcs:
        _orcs[key] = build_oracle(system._native)
    return _orcs[key]


# ---- CPU: oracle against the reference's known answers ------------------------------------------------------
@pytest.mark.parametrize('method,tol', [('dopri5', 1e-11), ('rosenbrock4', 1e-11), ('bdf', 1e-9)])
def test_oracle_growth_root(method, tol):
    orc = _orc(tsys.growth_root(compile=False))
    for ror in (False, True):
        sink = {}
        xy, nrows, st, cn = orc.integrate_adaptive([[1.0]], [0.0], [2.0], [[1.0]], 1e-12, 1e-12, dx0=1e-12,
                                                   mxsteps=5000, method=method, max_roots=4, roots_sink=sink,
                                                   return_on_root=ror)
        assert st[0] == 0 and sink['nroots_found'][0] == 1
        idx = sink['root_indices'][0, 0]
        x = xy[0, :nrows[0], 0]
        assert abs(x[idx] - 1) < tol and np.min(np.abs(x - 1)) < tol
        assert abs(xy[0, idx, 1] - np.e) < 100*tol
        assert np.all(np.diff(x) > 0)
        if ror:
            assert idx == nrows[0] - 1  # the root is the last row (_cvode_wrapper.pyx:208-209)
        else:
            assert x[-1] == 2.0


def test_oracle_predefined_root_out():
    orc = _orc(tsys.growth_root(compile=False))
    sink = {}
    y, nreached, st, cn = orc.integrate_predefined([[1.0]], [[0.0, 0.5, 1.5, 2.0]], [[1.0]], 1e-12, 1e-12,
                                                   mxsteps=5000, method='dopri5', max_roots=4, roots_sink=sink)
    assert st[0] == 0 and nreached[0] == 4 and sink['nroots_found'][0] == 1
    assert sink['root_indices'][0, 0] == 2  # the root lies in front of output point 2
    assert abs(sink['root_out'][0, 0, 0] - 1) < 1e-11 and abs(sink['root_out'][0, 0, 1] - np.e) < 1e-10
    assert np.allclose(y[0, :, 0], np.exp([0.0, 0.5, 1.5, 2.0]), rtol=1e-10)


@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4', 'bdf'])
def test_oracle_chain_crossings(method):
    y0, p = [[1.0, 0, 0]], [[7.0, 3.0]]
    for idx in range(3):
        orc = _orc(tsys.chain3_roots(idx, compile=False))
        sink = {}
        xy, nrows, st, cn = orc.integrate_adaptive(y0, [0.0], [0.7], full_params(y0, p), 1e-10, 1e-10, mxsteps=5000,
                                                   method=method, max_roots=4, roots_sink=sink, return_on_root=True)
        assert st[0] == 0 and abs(xy[0, nrows[0] - 1, 0] - CHAIN_ROOTS[idx]) < 1e-7
    # all three as one system: several events in one integration, in order
    orc = _orc(tsys.chain3_roots_all(compile=False))
    sink = {}
    xy, nrows, st, cn = orc.integrate_adaptive(y0, [0.0], [0.7], full_params(y0, p), 1e-10, 1e-10, mxsteps=5000,
                                               method=method, max_roots=8, roots_sink=sink)
    assert sink['nroots_found'][0] == 3
    assert np.allclose(sink['root_out'][0, :3, 0], CHAIN_ROOTS, atol=1e-7, rtol=0)
    assert np.allclose(xy[0, sink['root_indices'][0, :3], 0], CHAIN_ROOTS, atol=1e-7, rtol=0)
    assert xy[0, nrows[0] - 1, 0] == 0.7


def test_oracle_root_capacity_and_backward_direction():
    orc = _orc(tsys.chain3_roots_all(compile=False))
    y0, p = [[1.0, 0, 0]], [[7.0, 3.0]]
    sink = {}
    xy, nrows, st, cn = orc.integrate_adaptive(y0, [0.0], [0.7], full_params(y0, p), 1e-10, 1e-10, mxsteps=5000,
                                               method='dopri5', max_roots=2, roots_sink=sink)
    assert sink['nroots_found'][0] == 3 and np.allclose(sink['root_out'][0, :, 0], CHAIN_ROOTS[:2], atol=1e-7)
    # integrate forward, then backward from the end state (not all the way: y1 = y2 = 0 at x = 0 is a root
    # too): the same crossings in reverse order
    yend = xy[0, nrows[0] - 1, 1:][None]
    sink = {}
    orc.integrate_adaptive(yend, [0.7], [0.05], full_params(yend, p), 1e-11, 1e-11, mxsteps=5000,
                           method='dopri5', max_roots=8, roots_sink=sink)
    assert sink['nroots_found'][0] == 3
    assert np.allclose(sink['root_out'][0, :3, 0], CHAIN_ROOTS[::-1], atol=1e-5)


def test_oracle_invariant_guard():
    orc = _orc(tsys.leaky_invariant(compile=False))
    y0, p = [[1.0, 0.0]], [[2.0, 1.0, 0.05]]
    xo = np.linspace(0, 5, 11)[None]
    kw = dict(mxsteps=5000, method='rosenbrock4')
    y, nreached, st, cn = orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-10, 1e-10, **kw)
    drift = y[0].sum(axis=1) - 1
    assert st[0] == 0 and drift[-1] < -0.1
    for limit, mv in ((0.01, 0.01), (0.03, -0.015)):  # negative: |mv*I0| - mv (template:194)
        y, nreached, st, cn = orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-10, 1e-10,
                                                       max_invariant_violation=mv, **kw)
        n_ok = int(np.sum(np.abs(drift) <= limit))
        assert st[0] != 0 and nreached[0] == n_ok and np.isnan(y[0, n_ok:]).all()
        assert np.array_equal(y[0, :n_ok], orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-10, 1e-10,
                                                                    **kw)[0][0, :n_ok])
    # a limit that is never reached changes nothing
    y2 = orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-10, 1e-10, max_invariant_violation=0.5, **kw)[0]
    assert np.array_equal(y2, orc.integrate_predefined(y0, xo, full_params(y0, p), 1e-10, 1e-10, **kw)[0])


# ---- GPU: CUDA kernels against the oracle and the known answers ---------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize('method,tol', [('dopri5', 1e-11), ('rosenbrock4', 1e-11), ('bdf', 1e-9)])
def test_cuda_growth_root_like_reference(cuda_device, method, tol):
    native = tsys.growth_root()
    for ror in (False, True):
        res = native.integrate(2, [1], first_step=1e-12, atol=1e-12, rtol=1e-12, method=method, nsteps=5000,
                               return_on_root=ror)
        assert len(res.info['root_indices']) == 1 and res.info['success']
        assert np.min(np.abs(res.xout - 1)) < tol
        i = res.info['root_indices'][0]
        assert abs(res.xout[i] - 1) < tol and abs(res.yout[i, 0] - np.e) < 100*tol
        assert (i == len(res.xout) - 1) if ror else (res.xout[-1] == 2)


@pytest.mark.gpu
@pytest.mark.parametrize('method', ['dopri5', 'rosenbrock4', 'bdf'])
def test_cuda_roots_match_oracle_on_a_batch(cuda_device, method):
    native = tsys.chain3_roots_all()
    orc = _orc(native)
    rng = np.random.default_rng(11)
    n = 257
    y0 = np.column_stack((1 + rng.uniform(0, 0.5, n), rng.uniform(0, 0.1, n), np.zeros(n)))
    p = np.column_stack((rng.uniform(4, 9, n), rng.uniform(1, 3.5, n)))
    pf = full_params(y0, p)
    for ror in (False, True):
        sink_o, sink_c = {}, {}
        xy_o, nr_o, st_o, cn_o = orc.integrate_adaptive(y0, 0.0, 1.5, pf, 1e-10, 1e-10, mxsteps=5000, method=method,
                                                        max_roots=4, roots_sink=sink_o, return_on_root=ror)
        xy_c, nr_c, st_c, cn_c, _t = native._native.mod.integrate_adaptive_arrays(
            y0, np.zeros(n), np.full(n, 1.5), pf, 1e-10, 1e-10, mxsteps=5000, method=method, max_roots=4,
            roots_sink=sink_c, return_on_root=ror)
        assert np.array_equal(st_o, st_c) and (st_c == 0).all()
        assert np.array_equal(sink_o['nroots_found'], sink_c['nroots_found'])
        assert (sink_c['nroots_found'] == (1 if ror else 3)).all()
        k = 1 if ror else 3
        assert np.abs(sink_c['root_out'][:, :k] - sink_o['root_out'][:, :k]).max() < 1e-9
        # rows: the same count up to a step more or less (fp32 step-size control), root rows hold the roots
        assert np.abs(nr_c - nr_o).max() <= 3
        for i in range(0, n, 16):
            for j in range(k):
                r = sink_c['root_indices'][i, j]
                assert np.array_equal(xy_c[i, r], sink_c['root_out'][i, j])
            assert np.all(np.diff(xy_c[i, :nr_c[i], 0]) > 0)
            if ror:
                assert sink_c['root_indices'][i, 0] == nr_c[i] - 1
        # predefined mode: root_out / index of the next output point
        if not ror:
            xo = np.linspace(0, 1.5, 7)
            sink_o, sink_c = {}, {}
            y_o = orc.integrate_predefined(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method=method, max_roots=4,
                                           roots_sink=sink_o)[0]
            y_c = native._native.mod.integrate_predefined_arrays(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000,
                                                                 method=method, max_roots=4, roots_sink=sink_c)[0]
            assert np.abs(y_c - y_o).max() < 1e-8
            assert np.array_equal(sink_o['nroots_found'], sink_c['nroots_found'])
            assert np.array_equal(sink_o['root_indices'][:, :3], sink_c['root_indices'][:, :3])
            assert np.abs(sink_c['root_out'][:, :3] - sink_o['root_out'][:, :3]).max() < 1e-9
            xr = sink_c['root_out'][:, :3, 0]
            assert np.array_equal(sink_c['root_indices'][:, :3], np.searchsorted(xo, xr))


@pytest.mark.gpu
@pytest.mark.parametrize('idx', [0, 1, 2])
def test_cuda_chain_crossings_like_reference(cuda_device, idx):
    """ ``test_NativeSys__PartiallySolvedSystem__roots``: plain and partially solved, return_on_root """
    native = tsys.chain3_roots(idx)
    res = native.integrate(0.7, (1, 0, 0), (7, 3), return_on_root=True, method='bdf', nsteps=5000)
    assert abs(res.xout[-1] - CHAIN_ROOTS[idx]) < 1e-7 and res.info['success']
    res = native.integrate(0.7, (1, 0, 0), (7, 3), return_on_root=True, nsteps=5000)
    assert abs(res.xout[-1] - CHAIN_ROOTS[idx]) < 1e-7
    if idx == 1:
        pnative = tsys.chain3_roots_partsolved(idx)
        res = pnative.integrate(0.7, (1, 0, 0), (7, 3), return_on_root=True, nsteps=5000)
        assert abs(res.xout[-1] - CHAIN_ROOTS[idx]) < 1e-7
        assert res.yout.shape[1] == 3 and abs(res.yout[-1, 0] - res.yout[-1, 2]) < 1e-7


@pytest.mark.gpu
def test_cuda_invariant_guard_matches_oracle(cuda_device):
    native = tsys.leaky_invariant()
    orc = _orc(native)
    rng = np.random.default_rng(5)
    n = 300
    y0 = np.column_stack((1 + rng.uniform(0, 1, n), rng.uniform(0, 1, n)))
    p = np.column_stack((rng.uniform(1, 3, n), rng.uniform(0.5, 2, n), rng.uniform(0.0, 0.1, n)))
    p[::7, 2] = 0.0  # systems that do conserve A + B
    pf = full_params(y0, p)
    xo = np.linspace(0, 5, 11)
    for mv in (0.02, -0.01):
        y_o, nr_o, st_o, _c = orc.integrate_predefined(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4',
                                                       max_invariant_violation=mv)
        y_c, nr_c, st_c, _c, _t = native._native.mod.integrate_predefined_arrays(
            y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4', max_invariant_violation=mv)
        assert np.array_equal(nr_o, nr_c) and np.array_equal(st_o != 0, st_c != 0)
        assert (st_c[::7] == 0).all() and (st_c != 0).sum() > n//2
        ok = ~np.isnan(y_o)
        assert np.array_equal(ok, ~np.isnan(y_c)) and np.abs(y_c[ok] - y_o[ok]).max() < 1e-8
        I0 = y0.sum(axis=1)
        lim = mv if mv > 0 else np.abs(mv*I0) - mv
        drift = np.abs(np.nansum(y_c, axis=2) - I0[:, None])
        assert np.all(np.where(ok[..., 0], drift, 0) <= np.broadcast_to(np.atleast_1d(lim), (n,))[:, None]*(1 + 1e-9))
    with pytest.raises(RuntimeError):
        native.integrate(xo, y0[1], p[1], max_invariant_violation=1e-3, nsteps=5000)
    res = native.integrate(xo, y0[1], p[1], max_invariant_violation=1e-3, nsteps=5000, return_on_error=True)
    assert not res.info['success'] and res.info['nreached'] < 11


@pytest.mark.gpu
def test_cuda_invariant_guard_lane_group_kernel(cuda_device):
    """ ny = 6 chain with a leak: the one-system-per-thread and the lane-group Rosenbrock kernels give the same
    verdicts as the oracle; a limit that is never reached changes nothing """
    native = tsys.leaky_chain6()
    assert native._native.mod.ninvar == 1
    orc = _orc(native)
    rng = np.random.default_rng(6)
    n = 200
    y0 = np.zeros((n, 6))
    y0[:, 0] = 1 + rng.uniform(0, 1, n)
    p = rng.uniform(0.5, 3, (n, 6))
    p[:, 5] = rng.uniform(0, 0.3, n)
    p[::5, 5] = 0
    pf = full_params(y0, p)
    xo = np.linspace(0, 6, 13)
    y_o, nr_o, st_o, _c = orc.integrate_predefined(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4',
                                                   max_invariant_violation=0.01)
    assert (st_o[::5] == 0).all() and (st_o != 0).sum() > n//2
    for kernel in ('thread', 'group'):
        y_c, nr_c, st_c, _c, _t = native._native.mod.integrate_predefined_arrays(
            y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4', max_invariant_violation=0.01, kernel=kernel)
        assert np.array_equal(nr_o, nr_c) and np.array_equal(st_o != 0, st_c != 0)
        ok = ~np.isnan(y_o)
        assert np.array_equal(ok, ~np.isnan(y_c)) and np.abs(y_c[ok] - y_o[ok]).max() < 1e-8
        base = native._native.mod.integrate_predefined_arrays(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, kernel=kernel)
        loose = native._native.mod.integrate_predefined_arrays(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, kernel=kernel,
                                                               max_invariant_violation=10.0)
        assert (base[2] == 0).all() and np.array_equal(base[0], loose[0])


@pytest.mark.gpu
@pytest.mark.parametrize('kernel', ['thread', 'group'])
def test_cuda_roots_lane_group_kernel(cuda_device, kernel):
    """ ny = 6 chain, two event functions: the lane-group kernel (state distributed over the lanes, g evaluated
    from the state staged in shared memory) locates the same roots as the thread kernel and the oracle, in
    adaptive (rows inserted, return_on_root) and predefined mode """
    native = tsys.chain6_roots()
    orc = _orc(native)
    rng = np.random.default_rng(21)
    n = 130
    y0 = np.zeros((n, 6))
    y0[:, 0] = 1 + rng.uniform(0, 1, n)
    p = rng.uniform(1.0, 4.0, (n, 5))
    pf = full_params(y0, p)
    mod = native._native.mod
    for ror in (False, True):
        so, sc = {}, {}
        xy_o, nr_o, st_o, _c = orc.integrate_adaptive(y0, 0.0, 3.0, pf, 1e-10, 1e-10, mxsteps=5000,
                                                      method='rosenbrock4', max_roots=4, roots_sink=so,
                                                      return_on_root=ror)
        xy_c, nr_c, st_c, _c, _t = mod.integrate_adaptive_arrays(
            y0, np.zeros(n), np.full(n, 3.0), pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4', max_roots=4,
            roots_sink=sc, return_on_root=ror, kernel=kernel)
        assert (st_c == 0).all() and np.array_equal(so['nroots_found'], sc['nroots_found'])
        assert (sc['nroots_found'] >= 1).sum() > n//2
        for i in range(n):
            k = min(int(sc['nroots_found'][i]), 4)
            if k == 0:  # no crossing inside the span for these rates
                assert xy_c[i, nr_c[i] - 1, 0] == 3.0
                continue
            assert np.abs(sc['root_out'][i, :k] - so['root_out'][i, :k]).max() < 1e-9
            for j in range(k):
                assert np.array_equal(xy_c[i, sc['root_indices'][i, j]], sc['root_out'][i, j])
            if ror:
                assert sc['root_indices'][i, 0] == nr_c[i] - 1
            else:
                assert xy_c[i, nr_c[i] - 1, 0] == 3.0
            # the event functions vanish at the reported states
            yr = sc['root_out'][i, :k, 1:]
            assert np.min(np.abs(np.stack((yr[:, 0] - yr[:, 1], yr[:, 1] - yr[:, 2]))), axis=0).max() < 1e-9
    xo = np.linspace(0, 3, 7)
    so, sc = {}, {}
    y_o = orc.integrate_predefined(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4', max_roots=4,
                                   roots_sink=so)[0]
    y_c = mod.integrate_predefined_arrays(y0, xo, pf, 1e-10, 1e-10, mxsteps=5000, method='rosenbrock4', max_roots=4,
                                          roots_sink=sc, kernel=kernel)[0]
    assert np.abs(y_c - y_o).max() < 1e-8 and np.array_equal(so['nroots_found'], sc['nroots_found'])
    for i in range(n):
        k = min(int(sc['nroots_found'][i]), 4)
        assert np.array_equal(so['root_indices'][i, :k], sc['root_indices'][i, :k])
        assert k == 0 or np.abs(sc['root_out'][i, :k] - so['root_out'][i, :k]).max() < 1e-9
