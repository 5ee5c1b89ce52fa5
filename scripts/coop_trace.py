"""per-attempt trace of ONE system on the block-cooperative BDF kernel (build with -DPYODESYS_COOP_TRACE):
   python scripts/coop_trace.py <index in batch_radiolysis(n=2048)> [method]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyodesys_b200 import systems
i = int(sys.argv[1]); method = sys.argv[2] if len(sys.argv) > 2 else 'bdf'
rad = systems.radiolysis(); mod = rad._native.mod
b = systems.batch_radiolysis(n=2048)
dev = torch.device('cuda', 0)
y0 = b['y0'][i:i + 1]; pf = np.concatenate((b['params'][i:i + 1], y0), axis=1)
d_y0, d_p = torch.from_numpy(y0).to(dev), torch.from_numpy(pf).to(dev)
x = np.asarray(b['x'], dtype=float); d_x = torch.from_numpy(x).to(dev)
nt, ny = x.size, rad.ny
d_out = torch.empty((1, nt, ny), dtype=torch.float64, device=dev)
d_nr = torch.empty(1, dtype=torch.int32, device=dev); d_st = torch.empty(1, dtype=torch.int32, device=dev)
d_cn = torch.zeros((6, 1), dtype=torch.int32, device=dev)
d_tr = torch.zeros(12*4000, dtype=torch.float64, device=dev)
mod.integrate_predefined_device(1, nt, d_y0.data_ptr(), d_x.data_ptr(), d_p.data_ptr(), b['atol'], b['rtol'], d_out.data_ptr(),
                                d_nr.data_ptr(), d_st.data_ptr(), d_cn.data_ptr(), mxsteps=b['nsteps'], method=method,
                                xout_shared=True, debug_buffer=d_tr.data_ptr())
n = int(d_cn[0, 0]); tr = d_tr.cpu().numpy().reshape(-1, 12)[:min(n, 4000)]
print('steps', n, 'nrej', int(d_cn[5, 0]))
for k, r in enumerate(tr):
    it = (int(r[7]) // 100) % 100
    print('%4d t %.6e h %.6e err %.3e ok %d bad %d order %d hnext %.6e code %d iters %d neq %d nrm %s' % (
        k + 1, r[0], r[1], r[2], r[3], r[4], r[5], r[6], int(r[7]) % 100, it, int(r[7]) // 10000,
        ' '.join('%.3e' % v for v in r[8:8 + max(it, 1)])))
