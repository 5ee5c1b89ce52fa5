"""does a system with many BDF rejections behave the same when it is integrated alone? (lock-step coupling check)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pyodesys_b200 import systems
n = 2048
rad = systems.radiolysis()
b = systems.batch_radiolysis(n=n)
kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='bdf', force_predefined=True)
r = rad.integrate(b['x'], b['y0'], b['params'], **kw)
c = np.array(r.info.counters)
worst = np.argsort(c[5])[-4:]
print('worst systems', worst, 'nrej', c[5][worst], 'steps', c[0][worst])
for i in worst:
    r1 = rad.integrate(b['x'], b['y0'][i:i + 1], b['params'][i:i + 1], **kw)
    c1 = np.array(r1.info.counters)[:, 0]
    lo = (i // 16)*16
    r16 = rad.integrate(b['x'], b['y0'][lo:lo + 16], b['params'][lo:lo + 16], **kw)
    c16 = np.array(r16.info.counters)[:, i - lo]
    print('system', i, 'in batch: steps %d nrej %d | alone: steps %d nrej %d | with its 16 block mates: steps %d nrej %d | y equal alone/batch %s' % (
        c[0][i], c[5][i], c1[0], c1[5], c16[0], c16[5], np.array_equal(np.array(r.yout)[i], np.array(r1.yout)[0])))
