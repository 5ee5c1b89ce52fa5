#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""
bench.py - systems integrated per second on the batched native path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--n SYSTEMS_PER_GPU]

Workload (BASELINE.json configs[2], the one the metric is quoted on): stiff Robertson, per-system
(k1,k2,k3,y0) sets, Rosenbrock-4, rtol 1e-8, atol 1e-12, t in [0, 1e5], final state only.  One
"step" = one pass of the hot path over one batch of ``--n`` systems per GPU (default 1e6).
Multi-GPU: one process per GPU (torchrun), contiguous index ranges, no collective on the data path;
``torch.distributed`` is used for the barrier and the max-over-ranks of the step time only
("scaling": "weak": per-GPU work is fixed).

value   - device resident: inputs already in HBM, one persistent kernel per step, CUDA events on the
          launching stream, L2 flushed between steps (not timed).
e2e     - the public API call ``NativeCudaSys.integrate(x, y0, params, ...)`` on pinned HOST arrays:
          H2D + kernel + D2H + host-side shaping inside the timed region.
roofline- FP64: algorithmic flops (per-system counters x codegen-time op counts) / kernel
          time, against the FP64 FMA peak measured live on the same GPU (csrc/fp64_peak.cu), because
          MEASURED_PEAKS.json holds HBM and bf16 figures only.  The HBM view is reported beside it.
cpu_baseline - the CPU oracle (oracle/, "port": OpenMP restatement of the reference's native path; the
          reference's own cvode/odeint back ends cannot be built in this image) on all host cores, on
          a bounded sample of the same workload.
``--impl reference`` times that CPU path alone and prints the same line with "impl": "reference".
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

METRIC = "systems integrated/s (Robertson, 1e6 batch, rtol 1e-8)"
UNIT = "systems/s"
TEND = 1e5


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--n', '--systems-per-gpu', dest='n', type=int, default=1000000,
                    help='systems per GPU and step (use the long form under torchrun: its parser trips over --n)')
    ap.add_argument('--cpu-seconds', type=float, default=15.0, help='target CPU time of the cpu_baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--schedule', default='dynamic', choices=['static', 'dynamic'])
    ap.add_argument('--blocks-per-sm', type=int, default=0)
    return ap.parse_args()


def workload_config(args, world):
    return {"workload": "robertson_1e6_rtol1e-8", "systems_per_gpu": args.n, "global_batch": args.n*world,
            "ny": 3, "nparams": 3, "x": [0.0, TEND], "rtol": 1e-8, "atol": 1e-12, "method": "rosenbrock4",
            "nsteps": 50000, "inputs": "k=(0.04,1e4,3e7)*10^U(-.5,.5), y0=(1+U(0,.1),0,0)/sum, seed 3+rank",
            "l2": "flushed between timed steps (256 MiB write, untimed)", "schedule": args.schedule,
            "parallelism": "systems sharded by contiguous range, %d GPU(s), no collective" % world}


# ---- FP64 flop model (SURVEY.md 8d / DESIGN.md "Flop model") ---------------------------------------------
def flops_per_system_totals(counters, ny, f_rhs, f_jac, nt_out=1):
    """ Sum over systems of the algorithmic FP64 operations, from the kernel's own counters. """
    nsteps, nfev, njev, nlu, nsolve = (counters[i].astype(np.float64).sum() for i in range(5))
    f_lu = ny*(ny - 1)/2 + (ny - 1)*ny*(2*ny - 1)/3
    f_solve = ny*(2*ny - 1)
    # Rosenbrock-4 step: stage sums y + sum a_ij g_j (1+2+3+4 terms), rhs sums sum c_ij g_j / h (1..5 terms),
    # W = I/(gamma h) - J (ny^2 + ny), error norm + new solution (4 ny)
    f_step = 2*ny*(1 + 2 + 3 + 4) + 2*ny*(1 + 2 + 3 + 4 + 5) + ny*ny + ny + 4*ny
    f_dense = 2*ny*5
    n = counters.shape[1]
    return nsteps*f_step + nfev*f_rhs + njev*f_jac + nlu*f_lu + nsolve*f_solve + n*nt_out*f_dense


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


def cpu_oracle_rate(rob, batch_fn, seconds, nthreads=None):
    """ Times the CPU oracle (all host cores) on a bounded sample; returns (systems/s, sample size, cores). """
    from oracle import build_oracle
    orc = build_oracle(rob._native)
    cores = nthreads or os.cpu_count() or 1

    def run(n, seed):
        b = batch_fn(n=n, seed=seed, tend=TEND)
        p = np.concatenate((b['params'], b['y0']), axis=1)
        xo = np.array([0.0, TEND])
        t0 = time.perf_counter()
        y, nr, st, cn = orc.integrate_predefined(b['y0'], xo, p, b['atol'], b['rtol'], mxsteps=b['nsteps'],
                                                 method='rosenbrock4', nthreads=cores)
        dt = time.perf_counter() - t0
        assert (st == 0).all()
        return dt

    n0 = 64*cores
    dt0 = run(n0, 1003)
    n = int(max(n0, min(2000000, n0*seconds/max(dt0, 1e-6))))
    return n, cores, run


def main_reference(args, rank, world):
    if rank != 0:
        return None
    from pyodesys_b200 import systems
    rob = systems.robertson(compile=False)
    n, cores, run = cpu_oracle_rate(rob, systems.batch_robertson, max(2.0, args.cpu_seconds/3))
    for w in range(args.warmup):
        run(n, 2000 + w)
    t = 0.0
    for k in range(args.steps):
        t += run(n, 3000 + k)
    value = n*args.steps/t
    sample = "%d Robertson systems per step (same generator as the GPU arm), oracle/oracle.cpp, OpenMP dynamic" % n
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3*t/args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": workload_config(args, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference native back ends (cvode/odeint/gsl) cannot be built in this image (no SUNDIALS/"
                    "Boost/GSL/pycodeexport); this arm is the CPU port of that path on all host cores"}
    return line


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

    from pyodesys_b200 import systems
    rob = systems.robertson()
    mod = rob._native.mod
    n = args.n
    b = systems.batch_robertson(n=n, seed=3 + rank, tend=TEND)
    params_full = np.concatenate((b['params'], b['y0']), axis=1)  # append_iv layout [k1,k2,k3,y0...]
    launches0 = mod.launch_count()

    # ---- device resident arm ---------------------------------------------------------------------------
    d_y0 = torch.from_numpy(b['y0']).to(dev)
    d_p = torch.from_numpy(params_full).to(dev)
    d_x = torch.tensor([0.0, TEND], dtype=torch.float64, device=dev)
    d_yout = torch.empty((n, 2, 3), dtype=torch.float64, device=dev)
    d_nreached = torch.empty(n, dtype=torch.int32, device=dev)
    d_status = torch.empty(n, dtype=torch.int32, device=dev)
    d_counters = torch.zeros((6, n), dtype=torch.int32, device=dev)
    flush = torch.empty(256*1024*1024, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    def step_resident():
        return mod.integrate_predefined_device(
            n, 2, d_y0.data_ptr(), d_x.data_ptr(), d_p.data_ptr(), b['atol'], b['rtol'], d_yout.data_ptr(),
            d_nreached.data_ptr(), d_status.data_ptr(), d_counters.data_ptr(), mxsteps=b['nsteps'],
            method='rosenbrock4', device=local_rank, stream=stream.cuda_stream, xout_shared=True,
            schedule=args.schedule, blocks_per_sm=args.blocks_per_sm)

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
    for k in range(args.steps):
        flush.fill_(k)  # L2 flush, outside the event bracket
        ev[k][0].record(stream)
        tm = step_resident()
        ev[k][1].record(stream)
        kernel_s += tm[0]
    torch.cuda.synchronize(dev)
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
    yfin = d_yout[:, 1, :].cpu().numpy()
    n_ok = int((status == 0).sum())
    mass_err = float(np.abs(yfin.sum(axis=1) - 1.0).max())
    value = n*world*args.steps/(total_ms*1e-3)

    # ---- roofline ---------------------------------------------------------------------------------------
    gen = rob._native.gen
    flops = flops_per_system_totals(counters, 3, gen.flops['rhs'], gen.flops['jac'])
    kern_avg_s = kernel_ms*1e-3/args.steps
    achieved_tf = flops/kern_avg_s*1e-12
    bytes_alg = n*(3 + 3)*8 + n*2*3*8 + n*(2 + 6)*4  # y0 + used params in, yout + status/nreached/counters out
    peaks = {}
    try:
        with open(os.path.join(_root, 'MEASURED_PEAKS.json')) as fh:
            peaks = json.load(fh)
    except OSError:
        pass
    hbm_peak = peaks.get('hbm_gbs', 6650.0)
    # dram__bytes_read.sum + dram__bytes_write.sum of this kernel, per launch of 1e6 systems, from the committed
    # ``ncu --set full`` capture (the summary profiles/ncu_traffic.json names); only quoted for the configuration it was taken on
    traffic = None
    try:
        with open(os.path.join(_root, 'profiles', 'ncu_traffic.json')) as fh:
            tr = json.load(fh)
        if tr.get('systems_per_launch') == n and tr.get('method') == 'rosenbrock4':
            traffic = tr['dram_bytes_per_launch']
    except (OSError, ValueError, KeyError):
        pass
    roofline = {"bound": "fp64", "achieved": achieved_tf, "peak": peak_sus, "unit": "TFLOP/s",
                "frac": (achieved_tf/peak_sus) if peak_sus else None, "traffic": traffic,
                "traffic_unit": "bytes per launch (ncu dram read+write; algorithmic bytes per launch: %d)" % bytes_alg,
                "peak_source": "measured live: csrc/fp64_peak.cu DFMA chains, sustained %.1f s (burst %.2f)" %
                               (1.5, peak_best or 0.0),
                "flops_per_launch": flops, "kernel_ms": kern_avg_s*1e3,
                "mean_steps_per_system": float(counters[0].mean()), "max_steps_per_system": int(counters[0].max()),
                "reject_rate": float(counters[5].sum()/max(1, counters[0].sum())),
                "hbm": {"achieved": bytes_alg/kern_avg_s*1e-9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": bytes_alg/kern_avg_s*1e-9/hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}

    # ---- end to end through the public API, pinned host buffers ---------------------------------------------
    h_y0 = torch.from_numpy(b['y0']).pin_memory()
    h_k = torch.from_numpy(b['params']).pin_memory()
    e2e_steps = max(1, min(args.steps, 3))
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], method='rosenbrock4', force_predefined=True,
              device=local_rank, schedule=args.schedule)
    for _ in range(max(3, args.warmup)):  # warm-up: device buffers and the pinned output pool reach steady state
        res = rob.integrate(b['x'], h_y0.numpy(), h_k.numpy(), **kw)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        res = rob.integrate(b['x'], h_y0.numpy(), h_k.numpy(), **kw)
        last = float(res.yout[-1, -1, -1])  # the result is on the host
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t[0])
    e2e = {"value": n*world*e2e_steps/e2e_s, "unit": UNIT,
           "h2d_bytes_per_step": int(n*(3 + 3)*8 + 2*8), "d2h_bytes_per_step": int(n*(2*3*8 + 4 + 4 + 6*4)),
           "steps": e2e_steps, "ms_per_step": 1e3*e2e_s/e2e_steps,
           "phases_ms": {k: 1e3*v for k, v in res.info.timings.items()},
           "api": "NativeCudaSys.integrate(x, y0[N,3], params[N,3], atol, rtol, nsteps, force_predefined=True)"}
    launches = mod.launch_count() - launches0

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ncpu, cores, run = cpu_oracle_rate(rob, systems.batch_robertson, args.cpu_seconds)
        dt = run(ncpu, 3)
        cpu = {"value": ncpu/dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d systems of the same generator (seed 3), oracle/oracle.cpp Rosenbrock-4, OpenMP "
                         "schedule(dynamic), %.1f s" % (ncpu, dt)}

    line = None
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms/args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(args, world), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": int(launches), "clocks": clocks,
                "check": {"systems_ok": n_ok, "systems": n, "max_mass_error": mass_err, "last": last}}
    if world > 1:
        dist.destroy_process_group()
    return line


if __name__ == '__main__':
    main()
