#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""
bench.py - systems integrated per second on the batched native path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c3] [--systems-per-gpu n]

Workloads (``--config``; BASELINE.json ``configs``, generators of SURVEY.md 8d = pyodesys_b200.systems.batch_*):
  c3     (default, the one the metric is quoted on) stiff Robertson, 1e6 (k1,k2,k3,y0) sets, Rosenbrock-4, rtol 1e-8,
         atol 1e-12, t in [0, 1e5], final state only
  c3bdf  the same inputs with the BDF stepper (the cvode-default role)
  c1     van der Pol, 1e4 initial conditions, DOPRI5 in ADAPTIVE mode (every accepted step is a row), t in [0, 10]
  c2     Lorenz, 1e7 perturbed trajectories, DOPRI5, 100 dense-output samples each (24 GB of yout on the device;
         through the API in chunks of 1e6)
  c4     radiolysis network ny = 14, 1e5 rate-constant sets, Rosenbrock-4 (lane-group kernel); c4bdf: BDF
  c4x28  the 28-species coupled network, 1e5 sets, Rosenbrock-4 (32-lane groups)
  c5     Robertson sweep, 1e8 sets in total (1e8 / N per GPU), through the API in chunks of 1e7
One "step" = one pass of the hot path over one batch of n systems per GPU.  Multi-GPU: one process per GPU
(torchrun), contiguous index ranges, no collective on the data path; ``torch.distributed`` is used for the barrier
and the max-over-ranks of the step time only ("scaling": "weak").  Under torchrun rank 0 additionally times the
product's own single-process multi-GPU path (``integrate(..., devices=[0..N-1])``) -> ``sharded``.

value   - device resident: inputs already in HBM, one persistent kernel per step, CUDA events on the launching
          stream, L2 flushed between steps (not timed).
e2e     - the public API call ``NativeCudaSys.integrate(x, y0, params, ...)`` on pinned HOST arrays: H2D + kernel +
          D2H + host-side shaping inside the timed region.
roofline- FP64: algorithmic flops (per-system counters x codegen-time op counts) / kernel time, against
          the FP64 FMA peak measured live on the same GPU (csrc/fp64_peak.cu; MEASURED_PEAKS.json holds HBM and bf16
          figures only) with the theoretical SMs x 64 x 2 x clock beside it.  The HBM view is reported too.
cpu_baseline - the CPU oracle ("port": OpenMP restatement of the reference's native path; the reference's own
          cvode/odeint back ends cannot be built in this image) compiled with the reference's 'fast' flags
          (-O3 -march=native -ffast-math -funroll-loops, pyodesys/native/_base.py:41-42) on all host cores, on a
          bounded sample of the same workload.
check   - outside the timed region: a fixed random subsample of the timed results against the oracle
          (max err / band, band = max(10 rtol |y|, 10 atol)), failure counts, conserved quantities.
``--impl reference`` times the CPU path alone and prints the same line with "impl": "reference".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

_root = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, _root)

UNIT = "systems/s"

CONFIGS = {
    'c1': dict(workload="vdp_1e4_dopri5_adaptive_rtol1e-8", system='vdp', batch='batch_vdp', n=10000,
               method='dopri5', mode='adaptive',
               metric="systems integrated/s (van der Pol, 1e4 batch, DOPRI5 adaptive, rtol 1e-8)"),
    'c2': dict(workload="lorenz_1e7_dopri5_100samples_rtol1e-8", system='lorenz', batch='batch_lorenz', n=10000000,
               method='dopri5', mode='predefined', chunk=1000000, check_xmax=2.0,
               metric="systems integrated/s (Lorenz, 1e7 trajectories x 100 samples, DOPRI5, rtol 1e-8)"),
    'c3': dict(workload="robertson_1e6_rtol1e-8", system='robertson', batch='batch_robertson', n=1000000,
               method='rosenbrock4', mode='predefined',
               metric="systems integrated/s (Robertson, 1e6 batch, rtol 1e-8)"),
    'c3bdf': dict(workload="robertson_1e6_bdf_rtol1e-8", system='robertson', batch='batch_robertson', n=1000000,
                  method='bdf', mode='predefined',
                  metric="systems integrated/s (Robertson, 1e6 batch, BDF, rtol 1e-8)"),
    'c4': dict(workload="radiolysis14_1e5_rtol1e-8", system='radiolysis', batch='batch_radiolysis', n=100000,
               method='rosenbrock4', mode='predefined',
               metric="systems integrated/s (radiolysis ny=14, 1e5 batch, rtol 1e-8)"),
    'c4bdf': dict(workload="radiolysis14_1e5_bdf_rtol1e-8", system='radiolysis', batch='batch_radiolysis', n=100000,
                  method='bdf', mode='predefined',
                  metric="systems integrated/s (radiolysis ny=14, 1e5 batch, BDF, rtol 1e-8)"),
    'c4x28': dict(workload="radiolysis28_1e5_rtol1e-8", system='radiolysis2', batch='batch_radiolysis2', n=100000,
                  method='rosenbrock4', mode='predefined',
                  metric="systems integrated/s (coupled radiolysis network ny=28, 1e5 batch, rtol 1e-8)"),
    'c5': dict(workload="robertson_1e8_sweep_rtol1e-8", system='robertson', batch='batch_robertson', n=100000000,
               method='rosenbrock4', mode='predefined', chunk=10000000, total=True,
               metric="systems integrated/s (Robertson, 1e8 sweep, rtol 1e-8)"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--config', default='c3', choices=sorted(CONFIGS))
    ap.add_argument('--n', '--systems-per-gpu', dest='n', type=int, default=0,
                    help='systems per GPU and step (default: the configuration\'s size; use the long form under '
                         'torchrun: its parser trips over --n)')
    ap.add_argument('--cpu-seconds', type=float, default=15.0, help='target CPU time of the cpu_baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-sharded', action='store_true', help='skip the single-process multi-GPU arm under torchrun')
    ap.add_argument('--schedule', default='dynamic', choices=['static', 'dynamic'])
    ap.add_argument('--blocks-per-sm', type=int, default=0)
    ap.add_argument('--check-systems', type=int, default=4096)
    return ap.parse_args()


class Workload:
    """ One BASELINE configuration: system, seeded inputs, integrate arguments, flop and byte models. """

    def __init__(self, args, world, compile=True):
        from pyodesys_b200 import systems
        self.key = args.config
        self.cfg = cfg = CONFIGS[args.config]
        n = args.n or cfg['n']
        if cfg.get('total') and not args.n:
            n = cfg['n']//world  # a sweep of fixed total size
        self.n = n
        self.world = world
        self.method = cfg['method']
        self.adaptive = cfg['mode'] == 'adaptive'
        self.odesys = getattr(systems, cfg['system'])(compile=compile)
        self.batch_fn = getattr(systems, cfg['batch'])
        self.ny = self.odesys.ny
        self.nparams = len(self.odesys.params)
        self.chunk = min(n, cfg.get('chunk', n))
        b = self.batch_fn(n=2)
        self.atol, self.rtol, self.nsteps = b['atol'], b['rtol'], b['nsteps']
        x = np.asarray(b['x'], dtype=np.float64)
        self.x = x
        self.nt = 2 if self.adaptive else x.size
        self.tend = float(x[-1])
        self.max_rows = self.nsteps + 1

    def batch(self, n, seed):
        return self.batch_fn(n=n, seed=seed)

    def config(self, args):
        c = {"workload": self.cfg['workload'], "systems_per_gpu": self.n, "global_batch": self.n*self.world,
             "ny": self.ny, "nparams": self.nparams, "x": [float(self.x[0]), self.tend], "nt": int(self.nt),
             "mode": self.cfg['mode'], "rtol": self.rtol, "atol": self.atol, "method": self.method,
             "nsteps": self.nsteps, "inputs": "pyodesys_b200.systems.%s(seed = default + rank)" % self.cfg['batch'],
             "l2": "flushed between timed steps (256 MiB write, untimed)", "schedule": args.schedule,
             "parallelism": "systems sharded by contiguous range, %d GPU(s), no collective" % self.world}
        if self.chunk != self.n:
            c["e2e_chunk"] = self.chunk
        return c

    # ---- FP64 flop model (SURVEY.md 8d / DESIGN.md section 4) --------------------------------------------------------
    def flops(self, counters, nout_total):
        """ Sum over systems of the algorithmic FP64 operations, from the kernel's own per-system counters. """
        ny = self.ny
        gen = self.odesys._native.gen
        nsteps, nfev, njev, nlu, nsolve = (counters[i].astype(np.float64).sum() for i in range(5))
        f_lu = ny*(ny - 1)/2 + (ny - 1)*ny*(2*ny - 1)/3
        f_solve = ny*(2*ny - 1)
        if self.method == 'rosenbrock4':
            # stage sums y + sum a_ij g_j (1+2+3+4 terms), rhs sums sum c_ij g_j / h (1..5 terms), W = I/(gamma h) - J
            # (ny^2 + ny), error norm + new solution (4 ny)
            f_step = 2*ny*(1 + 2 + 3 + 4) + 2*ny*(1 + 2 + 3 + 4 + 5) + ny*ny + ny + 4*ny
            f_persolve = 0
        elif self.method == 'dopri5':
            f_step = 2*ny*(1 + 2 + 3 + 4 + 5 + 6) + 2*ny*7 + 4*ny
            f_persolve = 0
            nlu = nsolve = 0
        else:  # bdf: predictor + psi (2 * 5 ny), error estimate and table update (12 ny) per step; per Newton
            # iteration residual, update and norm (8 ny) besides the rhs and the solve; I - c J when refactored
            f_step = 22*ny
            f_persolve = 8*ny
            f_lu += 2*ny*ny
        f_dense = 2*ny*5
        return (nsteps*f_step + nfev*gen.flops['rhs'] + njev*gen.flops['jac'] + nlu*f_lu +
                nsolve*(f_solve + f_persolve) + nout_total*f_dense)

    def bytes_alg(self, n, nrows_total=None):
        gen = self.odesys._native.gen
        used = max(0, gen.np_used)
        inp = n*(self.ny + min(used, self.nparams))*8
        if self.adaptive:
            return inp + n*2*8 + int(nrows_total)*(1 + self.ny)*8 + n*(2 + 6)*4
        return inp + n*self.nt*self.ny*8 + n*(2 + 6)*4


class ClockSampler:
    """ nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe). """

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q,
                                          '--format=csv,noheader,nounits', '-lms', '20'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def mark(self):
        """ the timed region starts now: nvidia-smi was started before the warm-up (its start-up takes longer than a
        short timed region), only the samples from here on are reported """
        self.t_mark = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t_end = time.perf_counter()
        time.sleep(0.03)  # let the sample taken at the end of the region arrive
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        t0 = getattr(self, 't_mark', 0.0)
        window = [ln for t, ln in self.lines if t0 <= t <= t_end + 0.025]
        where = 'timed region'
        if not window:  # region shorter than the sampling period: the last samples of the warm-up (same load)
            window = [ln for _t, ln in self.lines[-3:]]
            where = 'end of warm-up (timed region shorter than the 20 ms sampling period)'
        for ln in window:
            parts = [p.strip() for p in ln.split(',')]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[4:8]):
                if val.lower().startswith('active'):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "window": where,
                "reasons": sorted(reasons)}


def fp64_peak(device, seconds=1.5):
    import ctypes
    from __graft_entry__ import build_fp64_peak
    lib = ctypes.CDLL(build_fp64_peak())
    best, sus = ctypes.c_double(), ctypes.c_double()
    rc = lib.pyodesys_cuda_fp64_peak(int(device), ctypes.c_double(seconds), ctypes.byref(best), ctypes.byref(sus))
    if rc != 0:
        return None, None
    return best.value, sus.value


def pin_to_gpu_numa_node(device, lib=None):
    """ bind this process to the CPUs of the NUMA node its GPU hangs off (round 1: all eight ranks ran on node 0 and
    the end-to-end arm lost 5 % at 8 GPUs to cross-socket pinned-memory traffic) """
    try:
        from pyodesys_b200.multi import gpu_cpu_affinity
        cpus = gpu_cpu_affinity(device, lib)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return "cpus %d-%d (%d)" % (min(cpus), max(cpus), len(cpus))
    except Exception as e:  # affinity is an optimisation, never a reason to fail
        return "unchanged (%s)" % type(e).__name__
    return "unchanged"


# ---- CPU arm ---------------------------------------------------------------------------------------------------------
def cpu_oracle_runner(wl, seconds, nthreads=None):
    """ Returns (sample size, cores, flags, run(n, seed) -> seconds) for the oracle on all host cores. """
    from oracle import build_oracle
    orc = build_oracle(wl.odesys._native, timing=True, method=wl.method)
    cores = nthreads or os.cpu_count() or 1

    def run(n, seed):
        b = wl.batch(n, seed)
        p = np.concatenate((b['params'], b['y0']), axis=1)
        t0 = time.perf_counter()
        if wl.adaptive:
            xy, nrows, st, cn = orc.integrate_adaptive(b['y0'], wl.x[0], wl.x[-1], p, wl.atol, wl.rtol,
                                                       mxsteps=wl.nsteps, method=wl.method, nthreads=cores)
        else:
            y, nr, st, cn = orc.integrate_predefined(b['y0'], wl.x, p, wl.atol, wl.rtol, mxsteps=wl.nsteps,
                                                     method=wl.method, nthreads=cores)
        dt = time.perf_counter() - t0
        assert (st == 0).all()
        return dt

    n0 = 16*cores
    dt0 = run(n0, 1003)
    n = int(max(n0, min(2000000, n0*seconds/max(dt0, 1e-6))))
    return n, cores, orc.flags, run


def main_reference(args, rank, world):
    if rank != 0:
        return None
    wl = Workload(args, world, compile=False)
    n, cores, flags, run = cpu_oracle_runner(wl, max(2.0, args.cpu_seconds/3))
    for w in range(args.warmup):
        run(n, 2000 + w)
    t = 0.0
    for k in range(args.steps):
        t += run(n, 3000 + k)
    value = n*args.steps/t
    sample = ("%d %s systems per step (same generator as the GPU arm), oracle/oracle.cpp method %s, g++ %s, OpenMP "
              "dynamic" % (n, wl.cfg['system'], wl.method, flags))
    cfg = wl.config(args)
    cfg.update(systems_per_step_measured=n, systems_per_gpu=n, global_batch=n,
               note="bounded sample of the workload named in 'workload'; throughput is per system")
    return {"metric": wl.cfg['metric'], "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3*t/args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference", "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference native back ends (cvode/odeint/gsl) cannot be built in this image (no SUNDIALS/"
                    "Boost/GSL/pycodeexport); this arm is the CPU port of that path on all host cores, compiled with "
                    "the reference's own 'fast' flags"}


class _StdoutToStderr:
    """ Everything written to file descriptor 1 while the bench runs goes to stderr (NCCL prints its version banner
    on stdout); the JSON line is the only thing this program puts on stdout. """

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def main():
    with _StdoutToStderr():
        line = _main()
    if line is not None:
        print(json.dumps(line), flush=True)


def check_against_oracle(wl, b, y_sample_fn, idx):
    """ fixed random subsample of the timed batch: CUDA results against the CPU oracle on the same inputs """
    from oracle import build_oracle
    orc = build_oracle(wl.odesys._native)
    y0, p = b['y0'][idx], b['params'][idx]
    pf = np.concatenate((p, y0), axis=1)
    if wl.adaptive:
        xy, nrows, st, cn = orc.integrate_adaptive(y0, wl.x[0], wl.x[-1], pf, wl.atol, wl.rtol, mxsteps=wl.nsteps,
                                                   method=wl.method)
        y_cpu = np.array([xy[i, nrows[i] - 1, 1:] for i in range(len(idx))])
        y_gpu = y_sample_fn(idx)
    else:
        y_cpu, nr, st, cn = orc.integrate_predefined(y0, wl.x, pf, wl.atol, wl.rtol, mxsteps=wl.nsteps,
                                                     method=wl.method)
        y_gpu = y_sample_fn(idx)
        xmax = wl.cfg.get('check_xmax')
        if xmax is not None:  # chaotic system: only the short horizon is comparable between two integrations
            sel = wl.x <= xmax
            y_cpu, y_gpu = y_cpu[:, sel], y_gpu[:, sel]
    ratio = np.abs(y_gpu - y_cpu)/np.maximum(10*wl.rtol*np.abs(y_cpu), 10*wl.atol)
    return {"subsample": int(len(idx)), "max_err_over_band": float(ratio.max()),
            "p999_err_over_band": float(np.quantile(ratio, 0.999)), "oracle_failures": int((st != 0).sum()),
            "band": "max(10 rtol |y|, 10 atol) vs oracle/oracle.cpp on the same inputs" +
                    (", x <= %g" % wl.cfg['check_xmax'] if wl.cfg.get('check_xmax') else "")}


def _main():
    args = parse()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        return main_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    wl = Workload(args, world)
    odesys, mod = wl.odesys, wl.odesys._native.mod
    affinity = pin_to_gpu_numa_node(local_rank, mod.lib) if world > 1 else "unchanged (single process)"
    n, ny, nt = wl.n, wl.ny, wl.nt
    b = wl.batch(n, {'c1': 1, 'c2': 2}.get(wl.key, 3) + rank)
    params_full = np.concatenate((b['params'], b['y0']), axis=1)  # append_iv layout [p..., y0...]
    launches0 = mod.launch_count()

    # ---- device resident arm ---------------------------------------------------------------------------
    d_y0 = torch.from_numpy(b['y0']).to(dev)
    d_p = torch.from_numpy(params_full).to(dev)
    d_status = torch.empty(n, dtype=torch.int32, device=dev)
    d_counters = torch.zeros((6, n), dtype=torch.int32, device=dev)
    d_nreached = torch.empty(n, dtype=torch.int32, device=dev)
    flush = torch.empty(256*1024*1024, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    common = dict(mxsteps=wl.nsteps, method=wl.method, device=local_rank, stream=stream.cuda_stream,
                  schedule=args.schedule, blocks_per_sm=args.blocks_per_sm)
    if wl.adaptive:
        d_x0 = torch.full((n,), float(wl.x[0]), dtype=torch.float64, device=dev)
        d_x1 = torch.full((n,), float(wl.x[-1]), dtype=torch.float64, device=dev)
        d_out = torch.empty((n, wl.max_rows, 1 + ny), dtype=torch.float64, device=dev)

        def step_resident():
            return mod.integrate_adaptive_device(
                n, d_y0.data_ptr(), d_x0.data_ptr(), d_x1.data_ptr(), d_p.data_ptr(), wl.atol, wl.rtol, wl.max_rows,
                d_out.data_ptr(), d_nreached.data_ptr(), d_status.data_ptr(), d_counters.data_ptr(), **common)
    else:
        d_x = torch.from_numpy(np.ascontiguousarray(wl.x)).to(dev)
        d_out = torch.empty((n, nt, ny), dtype=torch.float64, device=dev)

        def step_resident():
            return mod.integrate_predefined_device(
                n, nt, d_y0.data_ptr(), d_x.data_ptr(), d_p.data_ptr(), wl.atol, wl.rtol, d_out.data_ptr(),
                d_nreached.data_ptr(), d_status.data_ptr(), d_counters.data_ptr(), xout_shared=True, **common)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()  # before the peak measurement: nvidia-smi needs a few hundred ms to deliver its first line
    peak_best = peak_sus = None
    if rank == 0:
        peak_best, peak_sus = fp64_peak(local_rank)

    for _ in range(args.warmup):
        step_resident()
        flush.fill_(1)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler.mark()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_s = 0.0
    launches_t0 = mod.launch_count()
    for k in range(args.steps):
        flush.fill_(k)  # L2 flush, outside the event bracket
        ev[k][0].record(stream)
        tm = step_resident()
        ev[k][1].record(stream)
        kernel_s += tm[0]
    torch.cuda.synchronize(dev)
    launches_timed = mod.launch_count() - launches_t0  # this library's kernels inside the timed region
    if world > 1:
        dist.barrier()
    total_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms, kernel_s*1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, kernel_ms = float(t[0]), float(t[1])
    status = d_status.cpu().numpy()
    counters = d_counters.cpu().numpy()
    n_ok = int((status == 0).sum())
    value = n*world*args.steps/(total_ms*1e-3)

    rng = np.random.default_rng(12345)
    idx = np.sort(rng.choice(n, size=min(n, args.check_systems), replace=False))
    d_idx = torch.from_numpy(idx).to(dev)
    check = {"systems_ok": n_ok, "systems": n}
    if wl.adaptive:
        nrows = d_nreached.cpu().numpy()
        nrows_total = int(nrows.sum())
        nout_total = 0
        check["rows_written"] = nrows_total
        check["mean_rows_per_system"] = nrows_total/n

        def sample(ix):
            rows = torch.from_numpy(nrows[ix].astype(np.int64) - 1).to(dev)
            return d_out[d_idx, rows, 1:].cpu().numpy()
    else:
        nrows_total = None
        nout_total = n*(nt - 1)

        def sample(ix):
            return d_out[d_idx].cpu().numpy()
    if rank == 0:
        check.update(check_against_oracle(wl, b, sample, idx))
    if wl.cfg['system'] == 'robertson':
        yfin = d_out[:, -1, :].cpu().numpy()
        check["max_mass_error"] = float(np.abs(yfin.sum(axis=1) - 1.0).max())

    # ---- roofline ---------------------------------------------------------------------------------------
    flops = wl.flops(counters, nout_total)
    kern_avg_s = kernel_ms*1e-3/args.steps
    achieved_tf = flops/kern_avg_s*1e-12
    bytes_alg = wl.bytes_alg(n, nrows_total)
    peaks = {}
    try:
        with open(os.path.join(_root, 'MEASURED_PEAKS.json')) as fh:
            peaks = json.load(fh)
    except OSError:
        pass
    hbm_peak = peaks.get('hbm_gbs', 6650.0)
    # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the committed
    # ``ncu --set full`` capture (profiles/ncu_traffic.json names it); only quoted for the configuration it was taken on
    traffic = None
    try:
        with open(os.path.join(_root, 'profiles', 'ncu_traffic.json')) as fh:
            tr = json.load(fh)
        tr = tr.get(wl.key, tr if wl.key == 'c3' else {})
        if tr.get('systems_per_launch') == n and tr.get('method') == wl.method:
            traffic = tr['dram_bytes_per_launch']
    except (OSError, ValueError, KeyError):
        pass
    props = torch.cuda.get_device_properties(dev)
    sm_mhz = (clocks or {}).get('sm_mhz') or peaks.get('sm_max_mhz') or 1965.0
    peak_theory = props.multi_processor_count*64*2*sm_mhz*1e6*1e-12
    roofline = {"bound": "fp64", "achieved": achieved_tf, "peak": peak_sus, "unit": "TFLOP/s",
                "frac": (achieved_tf/peak_sus) if peak_sus else None, "traffic": traffic,
                "traffic_unit": "bytes per launch (ncu dram read+write; algorithmic bytes per launch: %d)" % bytes_alg,
                "peak_source": "measured live: csrc/fp64_peak.cu DFMA chains, sustained %.1f s (burst %.2f)" %
                               (1.5, peak_best or 0.0),
                "peak_theoretical": peak_theory, "frac_of_theoretical": achieved_tf/peak_theory,
                "peak_theoretical_source": "%d SMs x 64 FP64 FMA/clk x 2 x %.0f MHz (median SM clock of the timed "
                                           "region)" % (props.multi_processor_count, sm_mhz),
                "flops_per_launch": flops, "kernel_ms": kern_avg_s*1e3,
                "mean_steps_per_system": float(counters[0].mean()), "max_steps_per_system": int(counters[0].max()),
                "reject_rate": float(counters[5].sum()/max(1, counters[0].sum())),
                "hbm": {"achieved": bytes_alg/kern_avg_s*1e-9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": bytes_alg/kern_avg_s*1e-9/hbm_peak, "bytes_per_launch": bytes_alg,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}
    del d_out
    torch.cuda.empty_cache()

    # ---- end to end through the public API, pinned host buffers ---------------------------------------------
    ch = wl.chunk
    h_y0 = torch.from_numpy(b['y0']).pin_memory()
    h_k = torch.from_numpy(b['params']).pin_memory()
    e2e_steps = max(1, min(args.steps, 3))
    kw = dict(atol=wl.atol, rtol=wl.rtol, nsteps=wl.nsteps, method=wl.method, device=local_rank,
              schedule=args.schedule)
    if wl.adaptive:
        xarg = np.tile(np.array([wl.x[0], wl.x[-1]]), (n, 1))
    else:
        xarg = wl.x
        kw['force_predefined'] = True

    def e2e_pass():
        last, d2h, phases = 0.0, 0, {}
        for lo in range(0, n, ch):
            hi = min(n, lo + ch)
            res = odesys.integrate(xarg[lo:hi] if wl.adaptive else xarg, h_y0.numpy()[lo:hi], h_k.numpy()[lo:hi], **kw)
            if wl.adaptive:
                last = float(res[hi - lo - 1].yout[-1, -1])  # the result is on the host
                d2h += int(getattr(res.info, 'd2h_bytes', 0)) or \
                    (hi - lo)*(wl.max_rows*(1 + ny)*8 + 4 + 4 + 6*4)
            else:
                last = float(res.yout[-1, -1, -1])
                d2h += (hi - lo)*(nt*ny*8 + 4 + 4 + 6*4)
            for k_, v_ in res.info.timings.items():
                phases[k_] = phases.get(k_, 0.0) + 1e3*v_
        return last, d2h, phases

    for _ in range(max(3, args.warmup) if n <= 2000000 else 1):  # device buffers and the pinned pool reach steady state
        e2e_pass()
    if world > 1:
        dist.barrier()
    launches_e0 = mod.launch_count()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        last, d2h_bytes, phases = e2e_pass()
    e2e_s = time.perf_counter() - t0
    launches_e2e = mod.launch_count() - launches_e0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t[0])
    h2d = n*(ny + wl.nparams)*8 + (n*2*8 if wl.adaptive else nt*8*(-(-n//ch)))
    e2e = {"value": n*world*e2e_steps/e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
           "d2h_bytes_per_step": int(d2h_bytes), "steps": e2e_steps, "ms_per_step": 1e3*e2e_s/e2e_steps,
           "phases_ms": phases, "calls_per_step": -(-n//ch), "gpu_launches": int(launches_e2e),
           "api": "NativeCudaSys.integrate(x, y0[N,%d], params[N,%d], atol, rtol, nsteps%s)" % (
               ny, wl.nparams, "" if wl.adaptive else ", force_predefined=True")}
    check["last"] = last
    launches = mod.launch_count() - launches0

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ncpu, cores, flags, run = cpu_oracle_runner(wl, args.cpu_seconds)
        dt = run(ncpu, 3)
        cpu = {"value": ncpu/dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d systems of the same generator (seed 3), oracle/oracle.cpp method %s, g++ %s, OpenMP "
                         "schedule(dynamic), %.1f s" % (ncpu, wl.method, flags, dt)}

    sharded = None
    if world > 1 and not args.no_sharded and not wl.adaptive:
        sharded = sharded_arm(args, wl, dist, rank, world)

    line = None
    if rank == 0:
        line = {"metric": wl.cfg['metric'], "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms/args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": wl.config(args), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": int(launches_timed), "gpu_launches_whole_run": int(launches), "clocks": clocks, "check": check, "cpu_affinity": affinity}
        if sharded is not None:
            line["sharded"] = sharded
    if world > 1:
        dist.destroy_process_group()
    return line


def sharded_arm(args, wl, dist, rank, world):
    """ The product's own multi-GPU path, timed by rank 0 while the other ranks wait at a barrier: ONE process, one
    pinned host batch, ``integrate(..., devices=[0..N-1])`` (host threads + dynamic chunk counter, results gathered
    in one pinned array).  strong: the configuration's batch (n systems in total) over N GPUs; weak: N * n systems. """
    import torch
    out = None
    # The waiting ranks must not hold their GPU: an NCCL barrier is a spinning kernel, and two processes on one GPU
    # are time-sliced - rank 0's kernels on device r would get half of it (measured: 2.2x slower chunks on device 1).
    # Hence a CPU-side (gloo) barrier, after every rank has drained its own device.
    torch.cuda.synchronize()
    os.environ.setdefault('GLOO_SOCKET_IFNAME', 'lo')  # single node: the container's hostname may not resolve
    try:
        cpu_group = dist.new_group(backend='gloo')
    except Exception as e:  # no gloo transport: fall back to the NCCL barrier (the waiting GPUs are then time-sliced)
        print("bench.py: gloo group unavailable (%s): the sharded arm runs against NCCL barrier kernels" % e,
              file=sys.stderr)
        cpu_group = None
    dist.barrier(group=cpu_group)
    if rank == 0:
        os.sched_setaffinity(0, range(os.cpu_count() or 1))  # this arm drives every GPU: undo the per-rank binding
        devices = list(range(world))
        kw = dict(atol=wl.atol, rtol=wl.rtol, nsteps=wl.nsteps, method=wl.method, force_predefined=True,
                  schedule=args.schedule)
        out = {"api": "NativeCudaSys.integrate(x, y0, params, ..., devices=%r)" % devices, "unit": UNIT}
        for label, ntot in (("strong", wl.n), ("weak", wl.n*world)):
            bb = wl.batch(ntot, 77)
            h_y0 = torch.from_numpy(bb['y0']).pin_memory()
            h_k = torch.from_numpy(bb['params']).pin_memory()
            res = None
            for _ in range(2):
                res = None
                res = wl.odesys.integrate(wl.x, h_y0.numpy(), h_k.numpy(), devices=devices, **kw)
            reps = 3
            t0 = time.perf_counter()
            for _ in range(reps):
                res = None  # hand the page-locked result block back to the pool before the next call
                res = wl.odesys.integrate(wl.x, h_y0.numpy(), h_k.numpy(), devices=devices, **kw)
                float(res.yout[-1, -1, -1])
            dt = (time.perf_counter() - t0)/reps
            out[label] = {"systems": int(ntot), "ms_per_call": 1e3*dt, "value": ntot/dt,
                          "ok": int((res.info.status == 0).sum())}
            if label == "strong":
                for _ in range(2):
                    r1 = wl.odesys.integrate(wl.x, h_y0.numpy(), h_k.numpy(), device=0, **kw)
                t0 = time.perf_counter()
                for _ in range(reps):
                    r1 = None
                    r1 = wl.odesys.integrate(wl.x, h_y0.numpy(), h_k.numpy(), device=0, **kw)
                    float(r1.yout[-1, -1, -1])
                one = (time.perf_counter() - t0)/reps
                same = bool(np.array_equal(np.array(r1.yout), np.array(res.yout), equal_nan=True))
                out[label].update(ms_one_gpu=1e3*one, speedup=one/dt, bit_equal_to_one_gpu=same)
            sched = getattr(res.info, 'shard_stats', None)
            if sched:
                out[label]["schedule"] = sched
            res = None
    dist.barrier(group=cpu_group)
    return out


if __name__ == '__main__':
    main()
