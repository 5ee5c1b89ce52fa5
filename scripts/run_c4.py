#!/usr/bin/env python
"""one device-resident launch of a C4-type batch (for ncu): python scripts/run_c4.py [system] [n] [kernel] [method]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyodesys_b200 import systems
name = sys.argv[1] if len(sys.argv) > 1 else 'radiolysis'
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
kernel = sys.argv[3] if len(sys.argv) > 3 else 'auto'
method = sys.argv[4] if len(sys.argv) > 4 else 'rosenbrock4'
odesys = getattr(systems, name)()
b = getattr(systems, 'batch_' + name)(n=n)
mod = odesys._native.mod
dev = torch.device('cuda', 0)
pf = np.concatenate((b['params'], b['y0']), axis=1)
d_y0, d_p = torch.from_numpy(b['y0']).to(dev), torch.from_numpy(pf).to(dev)
x = np.asarray(b['x'], dtype=float)
d_x = torch.from_numpy(x).to(dev)
nt, ny = x.size, odesys.ny
d_out = torch.empty((n, nt, ny), dtype=torch.float64, device=dev)
d_nr = torch.empty(n, dtype=torch.int32, device=dev)
d_st = torch.empty(n, dtype=torch.int32, device=dev)
d_cn = torch.zeros((6, n), dtype=torch.int32, device=dev)
reps = int(os.environ.get('REPS', '2'))
d_dbg = torch.zeros(16, dtype=torch.int64, device=dev)
profile = 'COOP_PROFILE' in os.environ.get('PYODESYS_CUDA_NVCC_FLAGS', '')
for _ in range(reps):
    tm = mod.integrate_predefined_device(n, nt, d_y0.data_ptr(), d_x.data_ptr(), d_p.data_ptr(), b['atol'], b['rtol'],
                                         d_out.data_ptr(), d_nr.data_ptr(), d_st.data_ptr(), d_cn.data_ptr(),
                                         mxsteps=b['nsteps'], method=method, xout_shared=True, kernel=kernel,
                                         debug_buffer=d_dbg.data_ptr() if profile else None)
    print('%s n=%d kernel=%s method=%s: %.3f ms, ok %d, steps %.1f' % (name, n, kernel, method, tm[0]*1e3,
          int((d_st == 0).sum()), float(d_cn[0].float().mean())))
if profile:
    v = d_dbg.cpu().numpy().astype(float)
    v[:5] /= reps
    v[6:] /= reps
    nw = v[5]/reps
    print('per warp (cycles): total %.4g  evaluations incl. barriers %.4g (%.1f %%)  evaluator body %.4g  inverse %.4g (%.1f %%)  '
          'solves %.4g (%.1f %%)' % (v[4]/nw, v[0]/nw, 100*v[0]/v[4], v[1]/nw, v[2]/nw, 100*v[2]/v[4], v[3]/nw, 100*v[3]/v[4]))
    print('loop top (retire/fetch) %.1f %%, all-done barrier %.1f %%, attempt %.1f %%, control/output %.1f %%' % tuple(100*v[k]/v[4] for k in (6, 7, 8, 9)))
