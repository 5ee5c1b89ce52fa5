#!/usr/bin/env python
"""Secondary measurements for the BASELINE.json configs that are not bench.py's headline (C1 van der Pol,
C2 Lorenz dense output, C4 radiolysis; c3 = the headline workload with each stiff method).  Prints one JSON line per config; run on the GPU box:

    python scripts/bench_configs.py [c1] [c2] [c4] [--n N] [--reps K]
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def flops(counters, ny, f_rhs, f_jac, method, nt_out):
    nsteps, nfev, njev, nlu, nsolve = (counters[i].astype(np.float64).sum() for i in range(5))
    f_lu = ny*(ny - 1)/2 + (ny - 1)*ny*(2*ny - 1)/3
    f_solve = ny*(2*ny - 1)
    if method == 'rosenbrock4':
        f_step = 2*ny*10 + 2*ny*15 + ny*ny + ny + 4*ny
    elif method == 'dopri5':
        f_step = 2*ny*(1 + 2 + 3 + 4 + 5 + 6) + 2*ny*7 + 4*ny
    elif method == 'bdf':
        f_step = 2*ny*6 + 4*ny*3 + 4*ny  # predictor + psi, difference-table update, error norm (Newton work: nfev/nsolve)
    else:
        f_step = 16*ny
    return nsteps*f_step + nfev*f_rhs + njev*f_jac + nlu*f_lu + nsolve*f_solve + counters.shape[1]*nt_out*10*ny


def run(name, sysobj, b, n, reps, x, check=None, **extra):
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method=b['method'], force_predefined=True)
    kw.update(extra)
    res = None
    ts = []
    for r in range(reps + 2):
        t0 = time.perf_counter()
        res = sysobj.integrate(x, b['y0'], b['params'], **kw)
        ts.append(time.perf_counter() - t0)
    info = res.info
    gen = sysobj._native.gen
    nt = np.atleast_1d(x).shape[-1]
    fl = flops(info.counters, sysobj.ny, gen.flops['rhs'], gen.flops['jac'], b['method'], nt)
    ker = info.timings['kernel']
    out = {"config": name, "n": n, "method": b['method'], "ok": int((info.status == 0).sum()),
           "kernel_ms": ker*1e3, "systems_per_s_kernel": n/ker, "e2e_ms": min(ts[2:] or ts)*1e3,
           "systems_per_s_e2e": n/min(ts[2:] or ts), "fp64_tflops_algorithmic": fl/ker*1e-12,
           "mean_steps": float(info.counters[0].mean()), "max_steps": int(info.counters[0].max()),
           "yout_GBps": n*nt*sysobj.ny*8/ker*1e-9, "phases_ms": {k: v*1e3 for k, v in info.timings.items()}}
    if check is not None:
        out["check"] = check(res)
    print(json.dumps(out), flush=True)
    return res


def main():
    from pyodesys_b200 import systems
    args = sys.argv[1:]
    which = [a for a in args if a in ('c1', 'c2', 'c3', 'c3k', 'c4')] or ['c1', 'c2', 'c4']
    n_override = int(args[args.index('--n') + 1]) if '--n' in args else None
    reps = int(args[args.index('--reps') + 1]) if '--reps' in args else 3
    if 'c1' in which:
        n = n_override or 10**4
        b = systems.batch_vdp(n=n)
        run('C1 van der Pol dopri5 final state', systems.vdp(), b, n, reps, (0.0, 10.0))
    if 'c2' in which:
        n = n_override or 10**6
        b = systems.batch_lorenz(n=n, nt=100)
        run('C2 Lorenz dopri5 100 dense samples', systems.lorenz(), b, n, reps, b['x'])
    if 'c3' in which:
        n = n_override or 10**6
        b = systems.batch_robertson(n=n)
        from oracle import build_oracle
        rob = systems.robertson()
        for m in os.environ.get('C3_METHODS', 'rosenbrock4,bdf').split(','):
            b['method'] = m

            def check(res, m=m):
                k = min(n, 4096)
                orc = build_oracle(rob._native)
                p = np.concatenate((b['params'][:k], b['y0'][:k]), axis=1)
                y, nr, st, cn = orc.integrate_predefined(b['y0'][:k], np.array([0.0, b['x'][1]]), p, b['atol'], b['rtol'],
                                                         mxsteps=b['nsteps'], method=m)
                err = np.abs(res.yout[:k, -1] - y[:, -1])
                tol = np.maximum(10*b['rtol']*np.abs(y[:, -1]), 10*b['atol'])
                return {"vs_oracle_max_err_over_tol": float((err/tol).max()),
                        "max_mass_error": float(np.abs(res.yout[:, -1].sum(axis=1) - 1).max()),
                        "flags": os.environ.get('PYODESYS_CUDA_NVCC_FLAGS', '')}
            run('C3 Robertson %s final state' % m, rob, b, n, reps, b['x'], check=check)
    if 'c3k' in which:
        # C3 inputs through the matrix-free Newton-Krylov BDF (linear_solver='gmres' on the generated jtimes) and, for
        # comparison, the dense-LU BDF of the same library
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
        import systems_for_tests as tsys
        n = n_override or 10**6
        b = systems.batch_robertson(n=n)
        b['method'] = 'bdf'
        jts = tsys.robertson_jtimes_sparse()
        run('C3 Robertson bdf gmres (matrix-free, jtimes)', jts, b, n, reps, b['x'], linear_solver='gmres',
            with_jtimes=True)
        run('C3 Robertson bdf dense LU (sparse=True system)', jts, b, n, reps, b['x'])
    if 'c4' in which:
        n = n_override or 10**5
        b = systems.batch_radiolysis(n=n)
        from oracle import build_oracle
        rad = systems.radiolysis()

        def check(res):
            m = min(n, 256)
            orc = build_oracle(rad._native)
            p = np.concatenate((b['params'][:m], b['y0'][:m]), axis=1)
            y, nr, st, cn = orc.integrate_predefined(b['y0'][:m], np.array([0.0, b['x'][1]]), p, b['atol'], b['rtol'],
                                                     mxsteps=b['nsteps'], method='rosenbrock4')
            err = np.abs(res.yout[:m, -1] - y[:, -1])
            tol = np.maximum(10*b['rtol']*np.abs(y[:, -1]), 10*b['atol'])
            return {"vs_oracle_max_err_over_tol": float((err/tol).max()), "oracle_steps": float(cn[0].mean())}
        run('C4 radiolysis ny=14 rosenbrock4', rad, b, n, reps, b['x'], check=check,
            **({'kernel': os.environ['C4_KERNEL']} if 'C4_KERNEL' in os.environ else {}))


if __name__ == '__main__':
    main()
