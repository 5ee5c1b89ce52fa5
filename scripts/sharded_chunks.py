"""chunk-size sweep of the single-process multi-GPU path: python scripts/sharded_chunks.py <n> [chunk ...]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyodesys_b200 import systems
n = int(float(sys.argv[1])); chunks = [int(c) for c in sys.argv[2:]] or [0]
g = torch.cuda.device_count()
rob = systems.robertson()
b = systems.batch_robertson(n=n, seed=77)
y0 = torch.from_numpy(b['y0']).pin_memory().numpy(); k = torch.from_numpy(b['params']).pin_memory().numpy()
kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='rosenbrock4', force_predefined=True)
for c in chunks:
    res = None
    for _ in range(2):
        res = None
        res = rob.integrate(b['x'], y0, k, devices=list(range(g)), shard_chunk=c, **kw)
    t0 = time.perf_counter()
    for _ in range(3):
        res = None
        res = rob.integrate(b['x'], y0, k, devices=list(range(g)), shard_chunk=c, **kw)
        float(res.yout[-1, -1, -1])
    dt = (time.perf_counter() - t0)/3
    print('n %d gpus %d chunk %d: %.2f ms per call, %.1f M systems/s, schedule %s' % (n, g, c, 1e3*dt, n/dt/1e6, res.info.shard_stats))
