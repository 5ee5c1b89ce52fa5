# -*- coding: utf-8 -*-
"""
The model systems of the five BASELINE.json configurations and their synthetic input batches.

Definitions follow the reference; generators follow SURVEY.md §8(d)
(``numpy.random.default_rng(seed)``, float64).  Used by ``bench.py``, the tests and
``__graft_entry__.build()`` (which pre-compiles them so that the GPU box only loads binaries).
"""
from __future__ import annotations

import json
import os

import numpy as np

from .native.cuda import NativeCudaSys

_data = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'data')


# --- right-hand sides -----------------------------------------------------------------------------
def vdp_f(t, y, p):
    """ van der Pol, ``pyodesys/tests/test_core.py:24-25`` """
    return [y[1], -y[0] + p[0]*y[1]*(1 - y[0]**2)]


def lorenz_f(t, y, p):
    """ Lorenz 1963 with p = (sigma, rho, beta), ``examples/lorenz.ipynb`` cell 2 """
    return [p[0]*(y[1] - y[0]), y[0]*(p[1] - y[2]) - y[1], y[0]*y[1] - p[2]*y[2]]


def robertson_f(t, y, p):
    """ Robertson, ``pyodesys/tests/_robertson.py:98-101``: r1=k1 A, r2=k2 B C, r3=k3 B^2 """
    r1 = p[0]*y[0]
    r2 = p[1]*y[1]*y[2]
    r3 = p[2]*y[1]**2
    return [r2 - r1, r1 - r2 - r3, r3]


def decay_f(t, y, p):
    """ ``pyodesys/tests/test_core.py:169-170`` """
    return [-y[0]*p[0], y[0]*p[0]]


def sine_f(t, y, p):
    """ harmonic oscillator y'' = -k^2 y, ``pyodesys/tests/test_core.py:262-265`` """
    k = p[0]
    return [y[1], -k**2 * y[0]]


def decay_chain_rhs(t, y, p):
    """ linear decay chain with rates in params, ``pyodesys/tests/test_symbolic.py`` ``decay_rhs`` """
    ny = len(y)
    out = [0]*ny
    for i in range(ny):
        if i < ny - 1:
            out[i] -= p[i]*y[i]
        if i > 0:
            out[i] += p[i - 1]*y[i - 1]
    return out


def decay_dydt_factory(k):
    """ decay chain with hard-coded rates, ``pyodesys/tests/test_symbolic.py`` ``decay_dydt_factory`` """
    def dydt(t, y, p=None):
        ny = len(k) + 1
        out = [0]*ny
        for i in range(ny):
            if i < ny - 1:
                out[i] -= k[i]*y[i]
            if i > 0:
                out[i] += k[i - 1]*y[i - 1]
        return out
    return dydt


def load_radiolysis():
    with open(os.path.join(_data, 'radiolysis1.json')) as fh:
        return json.load(fh)


def mass_action_rhs(reac, prod, inact_reac=None):
    """ ``examples/_chem_kinet.py:14-43``: rate_i = k_i prod_j C_j^reac_ij; dC = sum_i rate_i (prod-reac-inact) """
    nr, ns = len(reac), len(reac[0])
    if inact_reac is None:
        inact_reac = [[0]*ns]*nr

    def dCdt(t, C, params):
        rates = []
        for i in range(nr):
            r = params[i]
            for j, n in enumerate(reac[i]):
                if n:
                    r = r*C[j]**n
            rates.append(r)
        out = [0]*ns
        for r, sr, sp, si in zip(rates, reac, prod, inact_reac):
            for i in range(ns):
                coef = sp[i] - sr[i] - si[i]
                if coef:
                    out[i] += coef*r
        return out
    return dCdt


# --- system factories (cached: one compiled library per definition) -----------------------------------
_cache = {}


def _get(key, factory):
    if key not in _cache:
        _cache[key] = factory()
    return _cache[key]


def vdp(**kw):
    return _get(('vdp', tuple(sorted(kw.items()))), lambda: NativeCudaSys.from_callback(vdp_f, 2, 1, **kw))


def lorenz(**kw):
    return _get(('lorenz', tuple(sorted(kw.items()))), lambda: NativeCudaSys.from_callback(lorenz_f, 3, 3, **kw))


def robertson(**kw):
    return _get(('robertson', tuple(sorted(kw.items()))),
                lambda: NativeCudaSys.from_callback(robertson_f, 3, 3, names=('A', 'B', 'C'),
                                                    linear_invariants=[[1, 1, 1]], **kw))


def decay(**kw):
    return _get(('decay', tuple(sorted(kw.items()))), lambda: NativeCudaSys.from_callback(decay_f, 2, 1, **kw))


def sine(**kw):
    return _get(('sine', tuple(sorted(kw.items()))), lambda: NativeCudaSys.from_callback(sine_f, 2, 1, **kw))


def decay_chain(ny, **kw):
    return _get(('decay_chain', ny, tuple(sorted(kw.items()))),
                lambda: NativeCudaSys.from_callback(decay_chain_rhs, ny, ny - 1, **kw))


def radiolysis(**kw):
    def mk():
        d = load_radiolysis()
        return NativeCudaSys.from_callback(mass_action_rhs(d['reac'], d['prod'], d['inact_reac']),
                                           len(d['names']), len(d['reac']), names=tuple(d['names']), **kw)
    return _get(('radiolysis', tuple(sorted(kw.items()))), mk)


def radiolysis2(**kw):
    """ C4(ii), SURVEY.md 8d: 28 species = two copies of the reference network coupled by four first-order
    exchange reactions (H2O2, O2, H2, OH- of copy one <-> copy two); 112 rate parameters.  Exercises ny > 16. """
    def mk():
        d = load_radiolysis()
        ns, nr = len(d['names']), len(d['reac'])
        z = [0]*ns
        reac = [r + z for r in d['reac']] + [z + r for r in d['reac']]
        prod = [r + z for r in d['prod']] + [z + r for r in d['prod']]
        inact = [r + z for r in d['inact_reac']] + [z + r for r in d['inact_reac']]
        for name in ('H2O2', 'O2', 'H2', 'OH-'):
            i = d['names'].index(name)
            for a, b in ((i, ns + i), (ns + i, i)):
                rr, pp = [0]*(2*ns), [0]*(2*ns)
                rr[a], pp[b] = 1, 1
                reac.append(rr)
                prod.append(pp)
                inact.append([0]*(2*ns))
        names = tuple(d['names']) + tuple(n + "'" for n in d['names'])
        return NativeCudaSys.from_callback(mass_action_rhs(reac, prod, inact), 2*ns, len(reac), names=names, **kw)
    return _get(('radiolysis2', tuple(sorted(kw.items()))), mk)


def batch_radiolysis2(n=10**4, seed=5, tend=1e3):
    b1 = batch_radiolysis(n=n, seed=seed, tend=tend)
    b2 = batch_radiolysis(n=n, seed=seed + 1000, tend=tend)
    rng = np.random.default_rng(seed + 7)
    kx = 1e-2*10.0**rng.uniform(-1, 1, size=(n, 8))
    out = dict(b1)
    out['y0'] = np.concatenate((b1['y0'], b2['y0']), axis=1)
    out['params'] = np.concatenate((b1['params'], b2['params'], kx), axis=1)
    return out


# --- synthetic batches -----------------------------------------------------------------------
def batch_vdp(n=10**4, seed=1):
    """ C1: mu=1, y0 ~ U([-3,3]^2), t in [0,10] """
    rng = np.random.default_rng(seed)
    y0 = rng.uniform(-3, 3, size=(n, 2))
    params = np.ones((n, 1))
    return dict(x=(0.0, 10.0), y0=y0, params=params, atol=1e-8, rtol=1e-8, nsteps=1000, method='dopri5')


def batch_lorenz(n=10**7, nt=100, seed=2):
    """ C2: sigma=10, rho=28, beta=8/3, y0=(-10,-15,15)(1+1e-3 N(0,1)), xout=linspace(0,10,nt) """
    rng = np.random.default_rng(seed)
    y0 = np.array([-10.0, -15.0, 15.0])*(1 + 1e-3*rng.standard_normal((n, 3)))
    params = np.tile(np.array([10.0, 28.0, 8.0/3.0]), (n, 1))
    return dict(x=np.linspace(0, 10, nt), y0=y0, params=params, atol=1e-8, rtol=1e-8, nsteps=5000,
                method='dopri5')


def batch_robertson(n=10**6, seed=3, tend=1e5):
    """ C3/C5: k = (0.04,1e4,3e7) 10^U(-.5,.5), y0 = (1,0,0)+(U(0,.1),0,0) renormalised, x=(0,1e5) """
    rng = np.random.default_rng(seed)
    k = np.array([0.04, 1e4, 3e7])*10.0**rng.uniform(-0.5, 0.5, size=(n, 3))
    y0 = np.zeros((n, 3))
    y0[:, 0] = 1.0 + rng.uniform(0, 0.1, size=n)
    y0 /= y0.sum(axis=1, keepdims=True)
    return dict(x=(0.0, tend), y0=y0, params=k, atol=1e-12, rtol=1e-8, nsteps=50000,
                method='rosenbrock4', force_predefined=True)


def batch_radiolysis(n=10**5, seed=4, tend=1e3):
    """ C4: reference network, rates x 10^U(-.3,.3), y0 of ``_radiolysis.ipynb`` cell 2 """
    d = load_radiolysis()
    rng = np.random.default_rng(seed)
    names = d['names']
    k = np.array(d['params'])*10.0**rng.uniform(-0.3, 0.3, size=(n, len(d['params'])))
    y0 = np.zeros((n, len(names)))
    y0[:, names.index('H2O')] = 55.4e3
    y0[:, names.index('H+')] = 1e-4
    y0[:, names.index('OH-')] = 1e-4
    y0[:, names.index('O2')] = 1e-6
    y0[:, names.index('H2O2')] = rng.uniform(0, 1e-3, size=n)
    return dict(x=(0.0, tend), y0=y0, params=k, atol=1e-10, rtol=1e-8, nsteps=50000,
                method='rosenbrock4', force_predefined=True)


STOCK = {'vdp': vdp, 'lorenz': lorenz, 'robertson': robertson, 'decay': decay, 'sine': sine}
