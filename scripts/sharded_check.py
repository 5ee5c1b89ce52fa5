#!/usr/bin/env python
"""One process, one host thread per GPU (pyodesys_b200.multi.integrate_sharded): timing and bit-equality of a
2e6-system Robertson batch on 1 and on all visible GPUs, from pageable and from page-locked inputs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyodesys_b200 import systems

rob = systems.robertson()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
b = systems.batch_robertson(n=n)
kw = dict(atol=b["atol"], rtol=b["rtol"], nsteps=b["nsteps"], method="rosenbrock4", force_predefined=True)
py0 = torch.from_numpy(b["y0"]).pin_memory().numpy()
pk = torch.from_numpy(b["params"]).pin_memory().numpy()
all_devs = list(range(torch.cuda.device_count()))
ref = refc = None
for name, (Y, K) in (("pageable", (b["y0"], b["params"])), ("pinned", (py0, pk))):
    for devs in ([0], all_devs) if len(all_devs) > 1 else ([0],):
        for _ in range(4):
            t0 = time.perf_counter()
            r = rob.integrate(b["x"], Y, K, devices=devs, **kw)
            t1 = time.perf_counter()
        print(name, devs, "call %.1f ms" % ((t1 - t0)*1e3), "ok", int((r.info.status == 0).sum()),
              {k: round(v*1e3, 2) for k, v in r.info.timings.items()})
        if ref is None:
            ref, refc = np.array(r.yout), np.array(r.info.counters)
        else:
            print("   equal to the first run:", np.array_equal(ref, np.array(r.yout)),
                  np.array_equal(refc, np.array(r.info.counters)))
