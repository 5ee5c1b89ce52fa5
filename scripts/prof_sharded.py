import sys, os, time, cProfile, pstats
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from pyodesys_b200 import systems
n = 4000000
rob = systems.robertson()
b = systems.batch_robertson(n=n, seed=77)
y0 = torch.from_numpy(b['y0']).pin_memory().numpy(); k = torch.from_numpy(b['params']).pin_memory().numpy()
kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='rosenbrock4', force_predefined=True)
for _ in range(2):
    res = None
    res = rob.integrate(b['x'], y0, k, devices=[0, 0], **kw)
pr = cProfile.Profile()
res = None
pr.enable()
res = rob.integrate(b['x'], y0, k, devices=[0, 0], **kw)
pr.disable()
st = pstats.Stats(pr); st.sort_stats('cumulative').print_stats(18)
