"""Where does the host time of one public-API call go?  (run on the GPU box)"""
import cProfile, pstats, sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from pyodesys_b200 import systems
rob = systems.robertson()
n = 1000000
b = systems.batch_robertson(n=n, seed=3)
y0 = torch.from_numpy(b['y0']).pin_memory().numpy()
k = torch.from_numpy(b['params']).pin_memory().numpy()
kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='rosenbrock4', force_predefined=True)
for _ in range(3):
    res = rob.integrate(b['x'], y0, k, **kw)
for _ in range(3):
    t0 = time.perf_counter(); res = rob.integrate(b['x'], y0, k, **kw); t1 = time.perf_counter()
    print('call %.2f ms' % ((t1 - t0)*1e3), {k_: round(v*1e3, 2) for k_, v in res.info.timings.items()})
pr = cProfile.Profile(); pr.enable()
res = rob.integrate(b['x'], y0, k, **kw)
pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(18)
