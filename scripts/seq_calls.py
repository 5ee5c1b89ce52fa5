import sys, os, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from pyodesys_b200 import systems
rob = systems.robertson()
for n in (250880, 125952, 500736):
    b = systems.batch_robertson(n=n, seed=77)
    y0 = torch.from_numpy(b['y0']).pin_memory().numpy(); k = torch.from_numpy(b['params']).pin_memory().numpy()
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='rosenbrock4', force_predefined=True)
    res = None
    for _ in range(3):
        res = None; res = rob.integrate(b['x'], y0, k, **kw)
    ts = []
    for _ in range(6):
        res = None
        t0 = time.perf_counter(); res = rob.integrate(b['x'], y0, k, **kw); ts.append(1e3*(time.perf_counter() - t0))
    print('n', n, 'ms per call', ' '.join('%.2f' % t for t in ts), 'kernel %.2f total %.2f' % (1e3*res.info.timings['kernel'], 1e3*res.info.timings['total']))
