# -*- coding: utf-8 -*-
"""
Small batches through every kernel family in one process (thread RK4 / DOPRI5 / Rosenbrock / BDF, matrix-free BDF,
lane-group and shared-memory Rosenbrock, event location in the group kernel, adaptive ragged output).  Written to be
run under ``compute-sanitizer --tool memcheck|racecheck``; that tool is closed on the round-1 GPU pool, so it serves as
a plain all-families smoke there (the parity tests remain the correctness evidence).  One line per case.
"""
import os
import sys

import numpy as np

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, _root)
sys.path.insert(0, os.path.join(_root, 'tests'))

from pyodesys_b200 import systems  # noqa: E402
import systems_for_tests as tsys  # noqa: E402


def main():
    os.environ.setdefault('PYODESYS_CUDA_STREAMING', '0')  # a kernel that waits for copies is not for a serialising tool
    rng = np.random.default_rng(0)
    n = 300
    k = np.array([0.04, 1e4, 3e7])*10**rng.uniform(-0.5, 0.5, size=(n, 3))
    y0 = np.tile([1.0, 0.0, 0.0], (n, 1))
    rob = systems.robertson()
    cases = []
    for method in ('rosenbrock4', 'bdf', 'dopri5'):
        tend = 1e3 if method in ('rosenbrock4', 'bdf') else 1e-2
        res = rob.integrate(np.linspace(0, tend, 5), y0, k, atol=1e-10, rtol=1e-8, method=method, nsteps=200000)
        cases.append(('robertson ' + method, all(r.info['success'] for r in res)))
    res = systems.vdp().integrate(np.linspace(0, 2, 150), rng.uniform(-2, 2, (n, 2)), np.full((n, 1), 2.0),
                                  method='rk4')  # fixed step (explicit: not for the stiff system)
    cases.append(('van der Pol rk4', all(r.info['success'] for r in res)))
    res = rob.integrate(1e2, y0, k, atol=1e-10, rtol=1e-8, nsteps=5000)  # adaptive: ragged rows
    cases.append(('robertson adaptive', all(r.info['success'] for r in res)))
    jts = tsys.robertson_jtimes_sparse()
    res = jts.integrate(np.linspace(0, 1e3, 5), y0, k, atol=1e-10, rtol=1e-8, method='bdf', nsteps=50000,
                        linear_solver='gmres', with_jtimes=True)
    cases.append(('robertson bdf gmres', all(r.info['success'] for r in res)))
    rad = systems.radiolysis()
    b = systems.batch_radiolysis(n=96, tend=1e1)
    for kernel in ('group', 'thread_smem'):
        res = rad.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'],
                            kernel=kernel, force_predefined=True)
        cases.append(('radiolysis ' + kernel, all(r.info['success'] for r in res)))
    ch = tsys.chain6_roots()
    y6 = np.zeros((64, 6))
    y6[:, 0] = 1 + rng.uniform(0, 1, 64)
    res = ch.integrate(3.0, y6, rng.uniform(1.0, 4.0, (64, 5)), atol=1e-10, rtol=1e-10, nsteps=5000, kernel='group')
    cases.append(('chain6 roots (group kernel events)', all(r.info['success'] for r in res)))
    ok = True
    for name, good in cases:
        print('%-40s %s' % (name, 'ok' if good else 'FAILED'))
        ok &= bool(good)
    sys.exit(0 if ok else 1)


if __name__ == '__main__':
    main()
