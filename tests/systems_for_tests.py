# -*- coding: utf-8 -*-
"""
ODE systems that only the tests use.  ``__graft_entry__.build()`` imports :func:`all_systems` so
that every library is compiled in the build container and merely loaded on the GPU box.
"""
import numpy as np
import sympy as sp

from pyodesys_b200 import systems
from pyodesys_b200.native.cuda import NativeCudaSys
from pyodesys_b200.symbolic import SymbolicSys
from pyodesys_b200.transforms import TransformedSys, symmetricsys, PartiallySolvedSystem, get_logexp

_cache = {}


def _get(key, mk):
    if key not in _cache:
        _cache[key] = mk()
    return _cache[key]


def decay3_rhs(x, y, p):
    """ ``_decay3`` of pyodesys/tests/test_symbolic.py:40-49: every species decays """
    return [-p[0]*y[0], p[0]*y[0] - p[1]*y[1], p[1]*y[1] - p[2]*y[2]]


def expy(compile=True):
    """ dy/dx = e^y (``pyodesys/tests/test_symbolic.py:274-286``, G9) """
    return _get(('expy', compile), lambda: NativeCudaSys.from_callback(
        lambda x, y, p, be: [be.exp(y[0])], 1, 0, compile=compile))


def decay_err(compile=True):
    """ G8: rhs that reports a recoverable error for x > 0.5 - the reference's own override snippet
    (``pyodesys/native/tests/_tests.py:324-333``) compiled verbatim as device code. """
    return _get(('decay_err', compile), lambda: NativeCudaSys.from_callback(
        systems.decay_chain_rhs, 3, 2, compile=compile, namespace_override={
            'p_rhs': """
        f[0] = -m_p[0]*y[0];
        f[1] = m_p[0]*y[0] - m_p[1]*y[1];
        f[2] = m_p[1]*y[1];
        if (x > 0.5) return AnyODE::Status::recoverable_error;
        this->nfev++;
        return AnyODE::Status::success;
"""}))


def decay3_dxmax(compile=True):
    """ ``_test_NativeSys__get_dx_max_source_code`` (``_tests.py:275-292``) """
    return _get(('decay3_dxmax', compile), lambda: NativeCudaSys.from_callback(
        decay3_rhs, 3, 3, compile=compile, namespace_override={
            'p_get_dx_max': "AnyODE::ignore(y); return (1.0e-4 * x + 1.0e-3);"}))


def decay3_first_step(compile=True):
    """ ``_test_NativeSys__first_step_cb`` (``_tests.py:230-240``): first_step_expr = y0*1e-30 """
    return _get(('decay3_fs', compile), lambda: NativeCudaSys.from_callback(
        decay3_rhs, 3, 3, compile=compile,
        first_step_factory=lambda x, y, p, be: y[0]*1e-30))


def decay3(compile=True):
    return _get(('decay3', compile), lambda: NativeCudaSys.from_callback(
        decay3_rhs, 3, 3, compile=compile))


def decay_k43(compile=True):
    """ ``_test_NativeSys_two`` second system (``_tests.py:32-50``): hard-coded k=[4,3], no params """
    return _get(('decay_k43', compile), lambda: NativeCudaSys.from_callback(
        systems.decay_dydt_factory([4, 3]), 3, 0, compile=compile))


def decay_named(compile=True):
    """ ``_test_NativeSys__dep_by_name__single_varied`` (``_tests.py:305-318``) """
    def mk():
        y = sp.symarray('y', 3)
        f = systems.decay_dydt_factory([4, 3])(0, y)
        return NativeCudaSys(zip(y, f), names='a b c'.split(), dep_by_name=True, compile=compile)
    return _get(('decay_named', compile), mk)


def forced(compile=True):
    """ non-autonomous system: exercises df/dx in the Rosenbrock stepper """
    return _get(('forced', compile), lambda: NativeCudaSys.from_callback(
        lambda x, y, p, be: [-p[0]*(y[0] - be.sin(p[1]*x)), y[0] - y[1]], 2, 2, compile=compile))


def lu6(compile=True):
    """ any 6-variable system: gives the NY=6 instance of the register LU (AnyODE test_decomp.cpp:19-45) """
    return _get(('lu6', compile), lambda: NativeCudaSys.from_callback(
        systems.decay_chain_rhs, 6, 5, compile=compile))


def nojac(compile=True):
    return _get(('nojac', compile), lambda: NativeCudaSys.from_callback(
        systems.decay_f, 2, 1, jac=False, compile=compile))


def bounded(compile=True):
    """ decay with lower bounds -> recoverable error branch of rhs (template:163-190) """
    return _get(('bounded', compile), lambda: NativeCudaSys.from_callback(
        decay3_rhs, 3, 3, lower_bounds=[0, 0, 0], compile=compile))


def decay_rhs(t, y, k):
    """ linear decay chain, last species stable (``decay_rhs`` of pyodesys/native/tests/_tests.py) """
    ny = len(y)
    dydt = [0]*ny
    for idx in range(ny):
        if idx < ny - 1:
            dydt[idx] -= y[idx]*k[idx]
        if idx > 0:
            dydt[idx] += y[idx - 1]*k[idx - 1]
    return dydt


class TransformedCudaSys(TransformedSys, NativeCudaSys):
    """ the mixin the reference's tests build (``class TransformedNativeSys(TransformedSys, NativeSys)``) """


_logexp = (sp.log, sp.exp)


def loglog_decay4(compile=True):
    """ log-log transformed 4-species chain (``_test_symmetricsys_nativesys`` / ``_test_multiple_adaptive_chained``) """
    SS = symmetricsys(_logexp, _logexp, SuperClass=TransformedCudaSys)
    return _get(('loglog_decay4', compile), lambda: SS.from_callback(decay_rhs, 4, 3, compile=compile))


def lin_decay4(compile=True):
    return _get(('lin_decay4', compile), lambda: NativeCudaSys.from_callback(decay_rhs, 4, 3, compile=compile))


def loglog_robertson(compile=True):
    """ Robertson in log(c), log(t) (``_test_chained_multi_native`` with logc=logt=True) """
    SS = symmetricsys(_logexp, _logexp, SuperClass=TransformedCudaSys)
    return _get(('loglog_robertson', compile), lambda: SS.from_callback(systems.robertson_f, 3, 3, compile=compile))


def partsolved_loglog_decay3(multiple=False, compile=True):
    """ ``_get_transformed_partially_solved_system`` (pyodesys/native/tests/_tests.py:136-153): decay3 with y0 (and
    optionally y2) replaced by analytic expressions, then log-log transformed, then compiled """
    def mk():
        odesys = SymbolicSys.from_callback(decay3_rhs, 3, 3)
        if multiple:
            def substitutions(x0, y0, p0, be):
                analytic0 = y0[0]*be.exp(-p0[0]*(odesys.indep - x0))
                analytic2 = y0[0] + y0[1] + y0[2] - analytic0 - odesys.dep[1]
                return {odesys.dep[0]: analytic0, odesys.dep[2]: analytic2}
        else:
            def substitutions(x0, y0, p0, be):
                return {odesys.dep[0]: y0[0]*sp.exp(-p0[0]*(odesys.indep - x0))}
        partsys = PartiallySolvedSystem(odesys, substitutions)
        LogLogSys = symmetricsys(get_logexp(), get_logexp(), SuperClass=TransformedCudaSys)
        return LogLogSys.from_other(partsys, compile=compile)
    return _get(('partsolved_loglog_decay3', multiple, compile), mk)


def partsolved_loglog_decay3_multi(compile=True):
    return partsolved_loglog_decay3(True, compile)


def growth_root(compile=True):
    """ ``_test_NativeSys__roots`` (``pyodesys/native/tests/_tests.py:262-273``): y' = y, root of y - e at x = 1 """
    return _get(('growth_root', compile), lambda: NativeCudaSys.from_callback(
        lambda t, y: [y[0]], 1, 0, roots_cb=lambda t, y, p, backend: [y[0] - backend.exp(1)], compile=compile))


def chain3_rhs(x, y, p):
    """ A -> B -> C of ``test_NativeSys__PartiallySolvedSystem__roots`` (``native/tests/test_cvode.py:153-176``) """
    return [-p[0]*y[0], p[0]*y[0] - p[1]*y[1], p[1]*y[1]]


def chain3_roots(idx, compile=True):
    """ crossings y0 = y1, y0 = y2, y1 = y2 of the chain (reference x: 0.11299628, 0.20674119, 0.35418287) """
    return _get(('chain3_roots', idx, compile), lambda: NativeCudaSys.from_callback(
        chain3_rhs, 3, 2, compile=compile,
        roots_cb=lambda x, y: ([y[0] - y[1]], [y[0] - y[2]], [y[1] - y[2]])[idx]))


def chain3_roots_all(compile=True):
    """ the three crossings as one system with nroots = 3 (several events per integration) """
    return _get(('chain3_roots_all', compile), lambda: NativeCudaSys.from_callback(
        chain3_rhs, 3, 2, compile=compile, roots_cb=lambda x, y: [y[0] - y[1], y[0] - y[2], y[1] - y[2]]))


def chain3_roots_partsolved(idx, compile=True):
    """ same crossings with y0 solved analytically (``PartiallySolvedSystem``), natively compiled """
    def mk():
        ori = SymbolicSys.from_callback(chain3_rhs, 3, 2,
                                        roots_cb=lambda x, y: ([y[0] - y[1]], [y[0] - y[2]], [y[1] - y[2]])[idx])
        psys = PartiallySolvedSystem(ori, lambda t0, xyz, par0, be: {
            ori.dep[0]: xyz[0]*be.exp(-par0[0]*(ori.indep - t0))})
        return NativeCudaSys.from_other(psys, compile=compile)
    return _get(('chain3_roots_partsolved', idx, compile), mk)


def chain3_roots_0(compile=True):
    return chain3_roots(0, compile)


def chain3_roots_1(compile=True):
    return chain3_roots(1, compile)


def chain3_roots_2(compile=True):
    return chain3_roots(2, compile)


def chain3_roots_partsolved_1(compile=True):
    return chain3_roots_partsolved(1, compile)


def leaky_invariant(compile=True):
    """ A <-> B with a declared invariant A + B that the rhs does NOT conserve (a leak of rate p[2]):
    ``max_invariant_violation`` turns the drift into recoverable errors -> the integration fails where
    |A + B - (A0 + B0)| passes the limit (template:112-123,191-198). """
    return _get(('leaky_invariant', compile), lambda: NativeCudaSys.from_callback(
        lambda x, y, p: [-p[0]*y[0] + p[1]*y[1], p[0]*y[0] - p[1]*y[1] - p[2]*y[1]], 2, 3,
        linear_invariants=[[1, 1]], compile=compile))


def leaky_chain6(compile=True):
    """ six species decay chain whose last species leaks away (rate p[5]) although sum(y) is declared invariant;
    ny = 6 so that the lane-group kernel can run it too """
    def rhs(x, y, p):
        out = systems.decay_chain_rhs(x, y, p[:5])
        out[5] = out[5] - p[5]*y[5]
        return out
    return _get(('leaky_chain6', compile), lambda: NativeCudaSys.from_callback(
        rhs, 6, 6, linear_invariants=[[1]*6], compile=compile))


def scaled_chain4(compile=True):
    """ ``_test_ScaledSys_NativeSys`` (``pyodesys/native/tests/_tests.py:53-72``): dep_scaling = 1e8 """
    def mk():
        from pyodesys_b200.transforms import ScaledSys

        class ScaledCudaSys(ScaledSys, NativeCudaSys):
            pass
        y0, y1, y2, y3 = sp.symbols('y0 y1 y2 y3', real=True, positive=True)
        eqs = [(y0, -7*y0), (y1, 7*y0 - 3*y1), (y2, 3*y1 - 2*y2), (y3, 2*y2)]
        return ScaledCudaSys(eqs, dep_scaling=1e8, compile=compile)
    return _get(('scaled_chain4', compile), mk)


def _decay3_symbolic(**kw):
    return SymbolicSys.from_callback(decay3_rhs, 3, 3, **kw)


def decay3_nonneg(compile=True):
    """ ``_test_Decay_nonnegative`` (``_tests.py:98-104``) """
    return _get(('decay3_nonneg', compile), lambda: NativeCudaSys.from_other(
        _decay3_symbolic(lower_bounds=[0]*3), compile=compile))


def partsolved_family(compile=True):
    """ the native members of ``_test_PartiallySolvedSystem_Native`` (``_tests.py:107-134``): plain, scaled (42),
    reduced by the linear invariant, and the log-log transforms of the last two """
    def mk():
        from pyodesys_b200.transforms import ScaledSys
        logexp = get_logexp(1, 1e-24)
        NativeLogLog = symmetricsys(logexp, logexp, SuperClass=TransformedCudaSys)
        odesys = _decay3_symbolic(lower_bounds=[0, 0, 0], linear_invariants=[[1, 1, 1]])
        scaledsys = ScaledSys.from_other(odesys, dep_scaling=42)
        partsys = PartiallySolvedSystem.from_linear_invariants(scaledsys)
        return dict(odesys=odesys, scaledsys=scaledsys, partsys=partsys,
                    n_sys=NativeCudaSys.from_other(odesys, compile=compile),
                    ns_sys=NativeCudaSys.from_other(scaledsys, compile=compile),
                    np_sys=NativeCudaSys.from_other(partsys, compile=compile),
                    nll_scaledsys=NativeLogLog.from_other(scaledsys, compile=compile),
                    nll_partsys=NativeLogLog.from_other(partsys, compile=compile))
    return _get(('partsolved_family', compile), mk)


def band_chain3(compile=True):
    """ ``_test_NativeSys__band`` (``_tests.py:294-302``; xfail in the reference, ``native/_base.py:128-129``
    raises NotImplementedError): the same error here """
    def mk():
        y = sp.symarray('y', 3)
        f = systems.decay_dydt_factory([4, 3])(0, y)
        return NativeCudaSys(zip(y, f), band=(1, 0), compile=compile)
    return _get(('band_chain3', compile), mk)


def power_decay(compile=True):
    """ ``_test_chained_parameter_variation`` (``pyodesys/tests/test_symbolic.py:1602-1617``): a -> b with rate
    t**e; named variables and parameters, one linear invariant given by name """
    def f(t, y, p):
        k = t**p['e']
        return {'a': -k*y['a'], 'b': +k*y['a']}
    return _get(('power_decay', compile), lambda: NativeCudaSys.from_callback(
        f, names='ab', param_names='e', dep_by_name=True, par_by_name=True,
        linear_invariants=[{'a': 1, 'b': 1}], compile=compile))


def chain6_roots(compile=True):
    """ six species decay chain (ny = 6: thread AND lane-group kernels) with two event functions: y0 = y1, y1 = y2 """
    return _get(('chain6_roots', compile), lambda: NativeCudaSys.from_callback(
        systems.decay_chain_rhs, 6, 5, compile=compile, roots_cb=lambda x, y, p, be: [y[0] - y[1], y[1] - y[2]]))


def _damped_osc(x, y, p):
    """ damped harmonic oscillator of ``pyodesys/native/tests/test_cvode.py:198-204`` (nu=0.01, k=m=1) """
    return [y[1], -1.0 * y[0] - 0.01 * y[1]]


def osc_jtimes(compile=True):
    """ ``test_jtimes_native_cvode`` (``native/tests/test_cvode.py:198-217``): jac=True, jtimes=True """
    return _get(('osc_jtimes', compile), lambda: NativeCudaSys.from_callback(
        _damped_osc, 2, 0, jac=True, jtimes=True, compile=compile))


def osc_sparse(compile=True):
    """ ``test_sparse_jac_native_cvode`` (``native/tests/test_cvode.py:221-238``): jac=True, sparse=True """
    return _get(('osc_sparse', compile), lambda: NativeCudaSys.from_callback(
        _damped_osc, 2, 0, jac=True, sparse=True, compile=compile))


def decay_k43_sparse(compile=True):
    """ ``test_SymbolicSys_jacobian_sparse`` (``pyodesys/tests/test_symbolic.py:145-153``) """
    return _get(('decay_k43_sparse', compile), lambda: NativeCudaSys.from_callback(
        systems.decay_dydt_factory([4, 3]), 3, 0, sparse=True, compile=compile))


def robertson_jtimes_sparse(compile=True):
    """ Robertson (``pyodesys/tests/_robertson.py:98-101``) with both alternative Jacobian forms: non-linear
    entries, parameters in the products """
    return _get(('robertson_jts', compile), lambda: NativeCudaSys.from_callback(
        systems.robertson_f, 3, 3, jtimes=True, sparse=True, compile=compile))


def all_systems(compile=True):
    fns = [expy, decay_err, decay3_dxmax, decay3_first_step, decay3, decay_k43, decay_named, forced, lu6,
           nojac, bounded, loglog_decay4, lin_decay4, loglog_robertson, partsolved_loglog_decay3,
           partsolved_loglog_decay3_multi, growth_root, chain3_roots_0, chain3_roots_1, chain3_roots_2,
           chain3_roots_all, chain3_roots_partsolved_1, leaky_invariant, leaky_chain6, scaled_chain4, decay3_nonneg,
           partsolved_family, power_decay, chain6_roots, osc_jtimes, osc_sparse, decay_k43_sparse,
           robertson_jtimes_sparse]
    out = {}
    for f in fns:
        obj = f(compile=compile)
        if isinstance(obj, dict):  # a family: only its natively compiled members
            out.update({'%s.%s' % (f.__name__, k): v for k, v in obj.items() if hasattr(v, '_native')})
        else:
            out[f.__name__] = obj
    return out


def full_params(y0, p=None):
    """ params as the native layer receives them: [p..., y0...] (append_iv, core.py:410-411) """
    y0 = np.atleast_2d(np.asarray(y0, dtype=float))
    if p is None:
        return y0.copy()
    p = np.atleast_2d(np.asarray(p, dtype=float))
    if p.shape[0] == 1 and y0.shape[0] > 1:
        p = np.tile(p, (y0.shape[0], 1))
    return np.concatenate((p, y0), axis=1)
