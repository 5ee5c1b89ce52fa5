# -*- coding: utf-8 -*-
"""
Parity at north_star's bar for EVERY BASELINE.json configuration: every state at every output time within
``max(10 rtol |y|, 10 atol)`` (tolerance written here: factor 10, no wider band anywhere in this file) of

* the fixtures of ``tests/golden/golden_r2.npz`` - tight-tolerance solutions computed by the reference's own
  ``pyodesys.core.ODESys`` + scipy LSODA on >= 1e3 Robertson / 256 van der Pol / 64 Lorenz / 128 radiolysis (ny = 14) /
  32 coupled-network (ny = 28) systems of the configurations' seeded generators (``tests/golden/make_golden_r2.py``);
* the CPU oracle on the same inputs.

CPU part (``-m "not gpu"``): the oracle against the fixtures (this is what pins the oracle for C1-C5).
GPU part (``-m gpu``): the CUDA path through the public API against both; achieved ratios are printed and, by
``scripts/parity_report.py``, written to ``profiles/r2_parity.json``.
"""
import os

import numpy as np
import pytest

from pyodesys_b200 import systems
from oracle import build_oracle
from systems_for_tests import full_params

_golden = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'golden_r2.npz')

# key: system, fixture prefix, params key, rtol, atol, nsteps, methods
CASES = {
    'c1': ('vdp', 'c1', 'p', 1e-8, 1e-8, 1000, ('dopri5',)),
    'c2': ('lorenz', 'c2', 'p', 1e-8, 1e-8, 5000, ('dopri5',)),
    'c3': ('robertson', 'c3', 'k', 1e-8, 1e-12, 50000, ('rosenbrock4', 'bdf')),
    'c4': ('radiolysis', 'c4', 'k', 1e-8, 1e-10, 50000, ('rosenbrock4', 'bdf')),
    'c4x': ('radiolysis2', 'c4x', 'k', 1e-8, 1e-10, 50000, ('rosenbrock4', 'bdf')),
}
PARAMS = [(k, m) for k, v in CASES.items() for m in v[6]]


@pytest.fixture(scope='module')
def g2():
    return np.load(_golden)


def band_ratio(y, ref, rtol, atol):
    """ |y - ref| in units of the band max(10 rtol |ref|, 10 atol) """
    return np.abs(y - ref)/np.maximum(10*rtol*np.abs(ref), 10*atol)


def _inputs(g2, key):
    sysname, pre, pk, rtol, atol, nsteps, _m = CASES[key]
    return sysname, g2[pre + '_xout'], g2[pre + '_y0'], g2[pre + '_' + pk], g2[pre + '_yout'], rtol, atol, nsteps


def test_fixture_inputs_are_the_configs_generators(g2):
    """ the fixture batches are the seeded generators of SURVEY.md 8(d) """
    b = systems.batch_robertson(n=1024, seed=3)
    assert np.array_equal(g2['c3_y0'], b['y0']) and np.array_equal(g2['c3_k'], b['params'])
    b = systems.batch_vdp(n=256, seed=1)
    assert np.array_equal(g2['c1_y0'], b['y0'])
    b = systems.batch_lorenz(n=64, nt=100, seed=2)
    assert np.array_equal(g2['c2_y0'], b['y0'])
    b = systems.batch_radiolysis(n=128, seed=4)
    assert np.array_equal(g2['c4_y0'], b['y0']) and np.array_equal(g2['c4_k'], b['params'])
    b = systems.batch_radiolysis2(n=32)
    assert np.array_equal(g2['c4x_k'], b['params'])
    assert g2['c3_yout'].shape == (1024, 5, 3) and g2['c4_yout'].shape == (128, 5, 14)
    # the references conserve what the systems conserve
    assert np.abs(g2['c3_yout'].sum(axis=2) - 1).max() < 1e-11


@pytest.mark.parametrize('key,method', PARAMS)
def test_oracle_within_band_of_reference_fixture(g2, key, method):
    sysname, xout, y0, p, ref, rtol, atol, nsteps = _inputs(g2, key)
    odesys = getattr(systems, sysname)(compile=False)
    orc = build_oracle(odesys._native)
    y, nr, st, cn = orc.integrate_predefined(y0, xout, full_params(y0, p), atol, rtol, mxsteps=nsteps, method=method)
    assert (st == 0).all()
    r = band_ratio(y, ref, rtol, atol)
    print('%s/%s oracle vs reference fixture: max err/band %.3g' % (key, method, r.max()))
    assert r.max() <= 1.0


@pytest.mark.gpu
@pytest.mark.parametrize('key,method', PARAMS)
def test_cuda_within_band_of_reference_fixture_and_oracle(cuda_device, g2, key, method):
    sysname, xout, y0, p, ref, rtol, atol, nsteps = _inputs(g2, key)
    odesys = getattr(systems, sysname)()
    res = odesys.integrate(xout, y0, p, atol=atol, rtol=rtol, nsteps=nsteps, method=method,
                           force_predefined=True)
    assert (res.info.status == 0).all()
    yg = np.array(res.yout)
    r_fix = band_ratio(yg, ref, rtol, atol)
    orc = build_oracle(odesys._native)
    yo, nr, st, cn = orc.integrate_predefined(y0, xout, full_params(y0, p), atol, rtol, mxsteps=nsteps, method=method)
    assert (st == 0).all()
    r_orc = band_ratio(yg, yo, rtol, atol)
    print('%s/%s cuda vs reference fixture: max err/band %.3g (per output time %s); vs oracle: %.3g' % (
        key, method, r_fix.max(), np.array2string(r_fix.max(axis=(0, 2)), precision=2), r_orc.max()))
    assert r_fix.max() <= 1.0
    assert r_orc.max() <= 1.0
    # same algorithm on both sides: about the same number of steps
    assert abs(float(res.info.counters[0].sum()) - float(cn[0].sum())) <= 0.05*float(cn[0].sum())
