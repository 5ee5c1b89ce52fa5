# -*- coding: utf-8 -*-
"""
Generates tests/golden/golden_r2.npz by IMPORTING THE PYTHON REFERENCE (read-only, /root/reference) in this
container: tight-tolerance solutions of seeded subsamples of every BASELINE.json configuration, computed by the
reference's own ``pyodesys.core.ODESys`` + scipy LSODA (``ODESys._integrate_scipy``), i.e. by code that shares
nothing with this repository's steppers.  The reference cannot travel to the GPU box, this fixture can.

    python tests/golden/make_golden_r2.py        (about 20 s on 8 cores)

Sections (SURVEY.md 8c "unpinned" list -> pinned by these):
  c3_*    Robertson, 1024 systems of the C3 generator (seed 3), xout = (0, 1e-3, 1, 1e2, 1e5)
  c1_*    van der Pol, 256 systems of the C1 generator (seed 1), xout = linspace(0, 10, 11)
  c2_*    Lorenz, 64 perturbed starts of the C2 generator (seed 2), xout = linspace(0, 2, 21) (chaotic beyond)
  c4_*    radiolysis ny = 14, 128 parameter sets of the C4 generator (seed 4), xout = (0, 1e-3, 1, 40, 1e3); the
          right-hand side and Jacobian handed to the reference are numpy restatements of the mass-action generator
          ``examples/_chem_kinet.py:14-43`` (NOT the generated code under test)
  c4x_*   the 28-species coupled network (two copies + 4 exchange reactions), 32 parameter sets, x = (0, 1e3)
Every section stores the inputs it was generated from, so the tests also check that the seeded generators of
``pyodesys_b200.systems`` still produce the same batches.
"""
import os
import sys
from concurrent.futures import ProcessPoolExecutor

import numpy as np

sys.path.insert(0, os.environ.get('PYODESYS_REFERENCE_ROOT', '/root/reference'))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))

from pyodesys import ODESys  # noqa: E402  (the reference)

RTOL = 1e-12


def rob_f(t, y, p):
    r1, r2, r3 = p[0]*y[0], p[1]*y[1]*y[2], p[2]*y[1]*y[1]
    return [r2 - r1, r1 - r2 - r3, r3]


def rob_j(t, y, p):
    return [[-p[0], p[1]*y[2], p[1]*y[1]],
            [p[0], -p[1]*y[2] - 2*p[2]*y[1], -p[1]*y[1]],
            [0, 2*p[2]*y[1], 0]]


def vdp_f(t, y, p):
    return [y[1], -y[0] + p[0]*y[1]*(1 - y[0]**2)]


def vdp_j(t, y, p):
    return [[0, 1], [-1 - p[0]*2*y[1]*y[0], p[0]*(1 - y[0]**2)]]


def lorenz_f(t, y, p):
    return [p[0]*(y[1] - y[0]), y[0]*(p[1] - y[2]) - y[1], y[0]*y[1] - p[2]*y[2]]


def lorenz_j(t, y, p):
    return [[-p[0], p[0], 0], [p[1] - y[2], -1, -y[0]], [y[1], y[0], -p[2]]]


class MassAction:
    """ numpy restatement of ``examples/_chem_kinet.py:14-43``: rate_i = k_i prod_j C_j^reac_ij,
    dC/dt = sum_i rate_i (prod_i - reac_i - inact_i); analytic Jacobian by the product rule """

    def __init__(self, reac, prod, inact):
        self.reac = np.asarray(reac, dtype=int)
        self.net = (np.asarray(prod, dtype=int) - self.reac - np.asarray(inact, dtype=int)).astype(float)
        self.terms = [[(j, int(n)) for j, n in enumerate(row) if n] for row in self.reac]

    def rates(self, C, k):
        r = np.array(k, dtype=float)
        for i, tl in enumerate(self.terms):
            for j, n in tl:
                r[i] *= C[j]**n
        return r

    def f(self, t, C, k):
        return self.rates(np.asarray(C), k) @ self.net

    def jac(self, t, C, k):
        C = np.asarray(C)
        ns = self.net.shape[1]
        drdC = np.zeros((len(self.terms), ns))
        for i, tl in enumerate(self.terms):
            for b, nb in tl:
                d = k[i]*nb*C[b]**(nb - 1)
                for j, n in tl:
                    if j != b:
                        d *= C[j]**n
                drdC[i, b] = d
        return self.net.T @ drdC


def network(copies):
    from pyodesys_b200.systems import load_radiolysis
    d = load_radiolysis()
    if copies == 1:
        return MassAction(d['reac'], d['prod'], d['inact_reac'])
    ns = len(d['names'])
    z = [0]*ns
    reac = [r + z for r in d['reac']] + [z + r for r in d['reac']]
    prod = [r + z for r in d['prod']] + [z + r for r in d['prod']]
    inact = [r + z for r in d['inact_reac']] + [z + r for r in d['inact_reac']]
    for name in ('H2O2', 'O2', 'H2', 'OH-'):  # same coupling as pyodesys_b200.systems.radiolysis2
        i = d['names'].index(name)
        for a, b in ((i, ns + i), (ns + i, i)):
            rr, pp = [0]*(2*ns), [0]*(2*ns)
            rr[a], pp[b] = 1, 1
            reac.append(rr)
            prod.append(pp)
            inact.append([0]*(2*ns))
    return MassAction(reac, prod, inact)


_SYS = {}


def _odesys(kind):
    if kind not in _SYS:
        if kind == 'rob':
            _SYS[kind] = ODESys(rob_f, rob_j)
        elif kind == 'vdp':
            _SYS[kind] = ODESys(vdp_f, vdp_j)
        elif kind == 'lorenz':
            _SYS[kind] = ODESys(lorenz_f, lorenz_j)
        else:
            net = network(1 if kind == 'c4' else 2)
            _SYS[kind] = ODESys(net.f, net.jac)
    return _SYS[kind]


def _solve(job):
    kind, xout, y0, p, atol = job
    r = _odesys(kind).integrate(xout, y0, p, integrator='scipy', name='lsoda', atol=atol, rtol=RTOL,
                                nsteps=10**6, force_predefined=True)
    assert r.info['success'], (kind, r.info)
    return r.yout


def solve_all(kind, xout, y0, p, atol):
    jobs = [(kind, xout, y0[i], p[i], atol) for i in range(len(y0))]
    with ProcessPoolExecutor(max_workers=os.cpu_count()) as ex:
        return np.array(list(ex.map(_solve, jobs, chunksize=4)))


def main():
    from pyodesys_b200.systems import batch_robertson, batch_vdp, batch_lorenz, batch_radiolysis, batch_radiolysis2
    out = {}
    b = batch_robertson(n=1024, seed=3)
    xo = np.array([0, 1e-3, 1.0, 1e2, 1e5])
    out.update(c3_xout=xo, c3_y0=b['y0'], c3_k=b['params'], c3_yout=solve_all('rob', xo, b['y0'], b['params'], 1e-16))
    print('c3 done', flush=True)
    b = batch_vdp(n=256, seed=1)
    xo = np.linspace(0, 10, 11)
    out.update(c1_xout=xo, c1_y0=b['y0'], c1_p=b['params'], c1_yout=solve_all('vdp', xo, b['y0'], b['params'], 1e-13))
    b = batch_lorenz(n=64, nt=100, seed=2)
    xo = np.linspace(0, 2, 21)
    out.update(c2_xout=xo, c2_y0=b['y0'], c2_p=b['params'],
               c2_yout=solve_all('lorenz', xo, b['y0'], b['params'], 1e-13))
    print('c1, c2 done', flush=True)
    b = batch_radiolysis(n=128, seed=4)
    xo = np.array([0.0, 1e-3, 1.0, 40.0, 1e3])
    out.update(c4_xout=xo, c4_y0=b['y0'], c4_k=b['params'], c4_yout=solve_all('c4', xo, b['y0'], b['params'], 1e-16))
    print('c4 done', flush=True)
    b = batch_radiolysis2(n=32)
    xo = np.array([0.0, 1e3])
    y0, k = b['y0'][:32], b['params'][:32]
    out.update(c4x_xout=xo, c4x_y0=y0, c4x_k=k, c4x_yout=solve_all('c4x', xo, y0, k, 1e-16))
    out['how'] = np.array('reference pyodesys ODESys._integrate_scipy lsoda rtol=%g, scipy %s' %
                          (RTOL, __import__('scipy').__version__))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden_r2.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
