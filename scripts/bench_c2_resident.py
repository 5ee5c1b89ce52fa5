#!/usr/bin/env python
"""C2 at full size, device resident: Lorenz, DOPRI5, N trajectories x 100 dense-output samples (BASELINE.json
configs[1]: N = 1e7, 24 GB of trajectory output).  Inputs and outputs stay in HBM (ptr_kind 1); prints one JSON
line per output layout (row-major [N, nt, ny] as the reference returns it, and struct-of-arrays [nt, ny, N]):

    python scripts/bench_c2_resident.py [--n 10000000] [--reps 3]

Both candidate bounds of SURVEY.md 8(d) are reported: achieved output GB/s against the measured HBM peak and
algorithmic FP64 TFLOP/s (from the per-system counters) against the measured FP64 peak of bench.py.
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def main():
    from pyodesys_b200 import systems
    from bench_configs import flops
    args = sys.argv[1:]
    n = int(args[args.index('--n') + 1]) if '--n' in args else 10**7
    reps = int(args[args.index('--reps') + 1]) if '--reps' in args else 3
    nt = 100
    lor = systems.lorenz()
    mod = lor._native.mod
    gen = lor._native.gen
    b = systems.batch_lorenz(n=n, nt=nt)
    dev = torch.device('cuda', 0)
    d_y0 = torch.from_numpy(b['y0']).to(dev)
    d_p = torch.from_numpy(np.ascontiguousarray(b['params'])).to(dev)
    d_x = torch.from_numpy(np.ascontiguousarray(b['x'], dtype=np.float64)).to(dev)
    d_yout = torch.empty(n*nt*3, dtype=torch.float64, device=dev)
    d_nreached = torch.empty(n, dtype=torch.int32, device=dev)
    d_status = torch.empty(n, dtype=torch.int32, device=dev)
    d_counters = torch.empty((6, n), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    peaks = {}
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')) as fh:
            peaks = json.load(fh)
    except (OSError, ValueError):
        pass
    hbm_peak = float(peaks.get('hbm_gbs', 6547.5))
    for soa in (False, True):
        ts = []
        for r in range(reps + 1):
            tm = mod.integrate_predefined_device(
                n, nt, d_y0.data_ptr(), d_x.data_ptr(), d_p.data_ptr(), b['atol'], b['rtol'], d_yout.data_ptr(),
                d_nreached.data_ptr(), d_status.data_ptr(), d_counters.data_ptr(), mxsteps=b['nsteps'],
                method='dopri5', device=0, stream=stream.cuda_stream, xout_shared=True, yout_soa=soa,
                iv_from_y0=True)
            ts.append(float(tm[0]))
        torch.cuda.synchronize(dev)
        ker = min(ts[1:])
        counters = d_counters.cpu().numpy()
        status = d_status.cpu().numpy()
        fl = flops(counters, 3, gen.flops['rhs'], gen.flops['jac'], 'dopri5', nt)
        out_bytes = n*nt*3*8
        # spot check: trajectory 0 and n-1 against the other layout's meaning
        y = d_yout.view(nt, 3, n) if soa else d_yout.view(n, nt, 3)
        first = (y[:, :, 0] if soa else y[0]).cpu().numpy()
        print(json.dumps({
            "config": "C2 Lorenz dopri5 %d dense samples, device resident, yout %s" % (nt, "[nt,ny,N] (SoA)" if soa else "[N,nt,ny]"),
            "n": n, "ok": int((status == 0).sum()), "kernel_ms": ker*1e3, "systems_per_s": n/ker,
            "yout_GB": out_bytes*1e-9, "yout_GBps": out_bytes/ker*1e-9, "hbm_peak_GBps": hbm_peak,
            "hbm_frac": out_bytes/ker*1e-9/hbm_peak, "fp64_tflops_algorithmic": fl/ker*1e-12,
            "mean_steps": float(counters[0].mean()), "y_first_last_sample": first[-1].tolist()}), flush=True)


if __name__ == '__main__':
    main()
