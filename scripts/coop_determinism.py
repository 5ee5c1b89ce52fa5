"""run the block-cooperative BDF twice on the same batch: a data race shows up as run-to-run differences"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pyodesys_b200 import systems
n = 2048
rad = systems.radiolysis()
b = systems.batch_radiolysis(n=n)
out = []
for rep in range(3):
    r = rad.integrate(b['x'], b['y0'], b['params'], atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='bdf', force_predefined=True)
    out.append((np.array(r.yout), np.array(r.info.counters)))
for rep in (1, 2):
    print('rep', rep, 'yout equal', np.array_equal(out[0][0], out[rep][0]), 'counters equal', np.array_equal(out[0][1], out[rep][1]),
          'systems with different step counts', int((out[0][1][0] != out[rep][1][0]).sum()))
c = out[0][1]
print('nrej histogram', np.bincount(c[5])[:40])
print('nrej mean %.2f max %d; systems with nrej > 30: %d; their mean steps %.1f (all: %.1f); top 10 nrej %s' % (
    c[5].mean(), c[5].max(), int((c[5] > 30).sum()), c[0][c[5] > 30].mean() if (c[5] > 30).any() else 0, c[0].mean(),
    np.sort(c[5])[-10:]))
