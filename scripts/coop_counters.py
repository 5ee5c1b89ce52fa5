import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from pyodesys_b200 import systems
n = 4096
rad = systems.radiolysis()
b = systems.batch_radiolysis(n=n)
for method in ('bdf', 'rosenbrock4'):
    r = rad.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method=method, force_predefined=True)
    c = r.info.counters
    print(os.environ.get('PYODESYS_CUDA_NVCC_FLAGS', ''), method, 'ok', int((r.info.status == 0).sum()), 'steps %.1f nfev %.1f njev %.1f nlu %.1f nsolve %.1f nrej %.1f' % tuple(c[i].mean() for i in range(6)), 'ylast', repr(float(r.yout[7, -1, 3])))
