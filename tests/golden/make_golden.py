# -*- coding: utf-8 -*-
"""
Generates tests/golden/golden.json by IMPORTING THE PYTHON REFERENCE (read-only, /root/reference)
in this container.  The reference cannot travel to the GPU box, the JSON fixture can.

    PYTHONPATH=/root/reference python tests/golden/make_golden.py

What of the reference runs here: ``pyodesys.core.ODESys`` + scipy (LSODA / predefined
dopri5), ``pyodesys.integrators.RK4_example_integrator`` and ``pyodesys/tests/bateman.py``.  The native
(cvode/odeint/gsl) back ends cannot be built here (no SUNDIALS/Boost/GSL, no pycodeexport).

Sections written:
  known_answers   literal "blessed" values copied as DATA from the reference's tests (G1, G6)
  rk4_predefined  G2: RK4_example_integrator.integrate_predefined on van der Pol, 150 points
  rk4_adaptive    G2: the first_step=1e-2 variant through ODESys.integrate
  bateman         G5: analytic decay chains via pyodesys/tests/bateman.py
  robertson_1e5   tight-tolerance LSODA runs (reference ODESys) of a seeded subsample of the C3 batch
  vdp_lsoda       tight-tolerance LSODA runs of a subsample of the C1 batch (t=10)
  lorenz_lsoda    tight-tolerance LSODA of the C2 start point up to t=2 (chaotic beyond)
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.environ.get('PYODESYS_REFERENCE_ROOT', '/root/reference'))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))

from pyodesys import ODESys  # noqa: E402  (the reference)
from pyodesys.integrators import RK4_example_integrator  # noqa: E402
from pyodesys.tests.bateman import bateman_full  # noqa: E402
from pyodesys.tests.test_core import vdp_f, vdp_j  # noqa: E402


def rob_f(t, y, p):
    r1, r2, r3 = p[0]*y[0], p[1]*y[1]*y[2], p[2]*y[1]*y[1]
    return [r2 - r1, r1 - r2 - r3, r3]


def rob_j(t, y, p):
    return [[-p[0], p[1]*y[2], p[1]*y[1]],
            [p[0], -p[1]*y[2] - 2*p[2]*y[1], -p[1]*y[1]],
            [0, 2*p[2]*y[1], 0]]


def lorenz_f(t, y, p):
    return [p[0]*(y[1] - y[0]), y[0]*(p[1] - y[2]) - y[1], y[0]*y[1] - p[2]*y[2]]


def lorenz_j(t, y, p):
    return [[-p[0], p[0], 0], [p[1] - y[2], -1, -y[0]], [y[1], y[0], -p[2]]]


def main():
    from pyodesys_b200.systems import batch_robertson, batch_vdp  # input generators only (numpy)
    out = {'generated_with': 'reference pyodesys @ /root/reference (ODESys + scipy %s)' %
           __import__('scipy').__version__}
    out['known_answers'] = {
        'vdp_mu2_t012': [[1, 0], [0.44449086, -1.32847148], [-1.89021896, -0.71633577]],
        'robertson_1e11': [0.2083340149701255e-7, 0.8333360770334713e-13, 0.9999999791665050],
    }
    # G2 -----------------------------------------------------------------------------------------
    odes = ODESys(vdp_f, vdp_j)
    xout = np.linspace(0, 2, 150)
    res = odes.integrate(xout, [1, 0], params=[2.0], integrator=RK4_example_integrator)
    out['rk4_predefined'] = {'xout': xout.tolist(), 'yout': res.yout.tolist(), 'nfev': int(res.info['nfev'])}
    res = odes.integrate([0, 2], [1, 0], params=[2.0], integrator=RK4_example_integrator, first_step=1e-2)
    out['rk4_adaptive'] = {'xout': res.xout.tolist(), 'yout': res.yout.tolist(), 'nfev': int(res.info['nfev'])}
    # G5 -----------------------------------------------------------------------------------------
    t = np.linspace(0, 2, 9)
    out['bateman'] = {
        'k43': {'k': [4, 3], 'y0': [5, 4, 2], 't': t.tolist(),
                'y': np.array(bateman_full([5, 4, 2], [4, 3, 0], t, exp=np.exp)).T.tolist()},
        'k732': {'k': [7., 3, 2], 'y0': [1, 0, 0, 0], 't': t[:5].tolist(),
                 'y': np.array(bateman_full([1, 0, 0, 0], [7., 3, 2, 0], t[:5], exp=np.exp)).T.tolist()},
    }
    # Robertson C3 subsample ------------------------------------------------------------------------
    b = batch_robertson(n=48, seed=3)
    rob = ODESys(rob_f, rob_j)
    ys, nfev = [], []
    for i in range(48):
        r = rob.integrate([0, 1e-3, 1.0, 1e2, 1e5], b['y0'][i], b['params'][i], integrator='scipy', name='lsoda',
                          atol=1e-15, rtol=1e-12, nsteps=100000)
        ys.append(r.yout.tolist())
        nfev.append(int(r.info['nfev']))
    out['robertson_1e5'] = {'seed': 3, 'n': 48, 'xout': [0, 1e-3, 1.0, 1e2, 1e5], 'y0': b['y0'].tolist(),
                            'k': b['params'].tolist(), 'yout': ys, 'nfev': nfev,
                            'how': 'reference ODESys._integrate_scipy lsoda atol=1e-15 rtol=1e-12'}
    # standard Robertson at 1e11 through the reference (cross-checks G6)
    r = rob.integrate([0, 1e11], [1, 0, 0], [0.04, 1e4, 3e7], integrator='scipy', name='lsoda',
                      atol=1e-18, rtol=1e-12, nsteps=100000, force_predefined=True)
    out['robertson_std_1e11_lsoda'] = r.yout[-1].tolist()
    # van der Pol C1 subsample ----------------------------------------------------------------------
    b = batch_vdp(n=32, seed=1)
    vd = ODESys(vdp_f, vdp_j)
    ys = []
    xo = np.linspace(0, 10, 11)
    for i in range(32):
        r = vd.integrate(xo, b['y0'][i], b['params'][i], integrator='scipy', name='lsoda',
                         atol=1e-13, rtol=1e-13, nsteps=100000)
        ys.append(r.yout.tolist())
    out['vdp_lsoda'] = {'seed': 1, 'n': 32, 'xout': xo.tolist(), 'y0': b['y0'].tolist(), 'mu': 1.0, 'yout': ys}
    # Lorenz short horizon ---------------------------------------------------------------------------
    lo = ODESys(lorenz_f, lorenz_j)
    xo = np.linspace(0, 2, 21)
    r = lo.integrate(xo, [-10.0, -15.0, 15.0], [10.0, 28.0, 8.0/3.0], integrator='scipy', name='lsoda',
                     atol=1e-13, rtol=1e-13, nsteps=100000)
    out['lorenz_lsoda'] = {'xout': xo.tolist(), 'y0': [-10.0, -15.0, 15.0], 'p': [10.0, 28.0, 8.0/3.0],
                           'yout': r.yout.tolist()}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden.json')
    with open(path, 'w') as fh:
        json.dump(out, fh)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
