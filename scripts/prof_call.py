"""cProfile of ONE public call (C3, 1e6 systems, page-locked inputs): where does the host time outside the library go?"""
import sys, os, cProfile, pstats, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyodesys_b200 import systems
n = 1000000
rob = systems.robertson()
b = systems.batch_robertson(n=n, seed=3)
y0 = torch.from_numpy(b['y0']).pin_memory().numpy(); k = torch.from_numpy(b['params']).pin_memory().numpy()
kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='rosenbrock4', force_predefined=True)
for _ in range(3):
    res = None
    res = rob.integrate(b['x'], y0, k, **kw)
ts = []
for _ in range(5):
    res = None
    t0 = time.perf_counter(); res = rob.integrate(b['x'], y0, k, **kw); ts.append(1e3*(time.perf_counter() - t0))
print('ms per call', ' '.join('%.3f' % t for t in ts), '| library total %.3f ms' % (1e3*res.info.timings['total']))
pr = cProfile.Profile()
res = None
pr.enable(); res = rob.integrate(b['x'], y0, k, **kw); pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(12)
# the failure screen: a batch with some failing systems must report them exactly as before
b2 = systems.batch_robertson(n=300000, seed=5)
b2['params'][4321, 0] = np.inf; b2['params'][123456, 1] = np.nan
r2 = rob.integrate(b2['x'], b2['y0'], b2['params'], atol=b2['atol'], rtol=b2['rtol'], nsteps=b2['nsteps'], method='rosenbrock4',
                   force_predefined=True, return_on_error=True)
st = np.array(r2.info.status)
print('failed systems', int((st != 0).sum()), 'library count', int(rob._native.mod.lib.pyodesys_cuda_query(b'last_failed_systems')),
      'indices', np.flatnonzero(st != 0)[:5])
try:
    rob.integrate(b2['x'], b2['y0'], b2['params'], atol=b2['atol'], rtol=b2['rtol'], nsteps=b2['nsteps'], method='rosenbrock4', force_predefined=True)
    print('no exception?!')
except RuntimeError as e:
    print('raises:', str(e)[:80])
