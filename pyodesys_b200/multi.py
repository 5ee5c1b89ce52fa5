# -*- coding: utf-8 -*-
"""
Sharding of a batch over GPUs / ranks.

The batch is embarrassingly parallel (the reference's only parallelism is an OpenMP "bag of
tasks" over systems, ``README.rst:199-201``), so the partition is a contiguous index range per
shard and there is NO collective on the data path:

* :func:`shard_range` - the partition rule (used by ``bench.py`` for one-process-per-GPU runs under
  ``torchrun`` and by :func:`integrate_sharded`);
* :func:`integrate_sharded` - one host thread per device inside one process; every thread calls the
  C ABI (ctypes releases the GIL) on its slice and writes into disjoint slices of the shared
  output arrays, so the "gather" is free;
* :func:`gather_to_rank0` - optional host-side gather for multi-process runs (``torch.distributed``,
  gloo or nccl) when a caller wants all results in one place.
"""
from __future__ import annotations

import threading

import numpy as np

from .core import BatchInfo


def shard_range(n, nshards, k):
    """ Half-open range ``[lo, hi)`` of shard ``k``: ``ceil(n/nshards)`` systems each, last one short. """
    if nshards < 1 or not 0 <= k < nshards:
        raise ValueError("invalid shard %d of %d" % (k, nshards))
    per = -(-n // nshards)
    lo = min(n, k*per)
    hi = min(n, lo + per)
    return lo, hi


def integrate_sharded(mod, devices, intern_x, y0, params, force_predefined, atol, rtol, nsteps,
                      first_step, kwargs):
    """ Runs ``mod`` (a :class:`~pyodesys_b200.native.cuda.CudaModule`) on several devices at once. """
    from .native.cuda import _Ragged, _tdict
    n = y0.shape[0]
    devices = list(devices)
    g = len(devices)
    adaptive = intern_x.shape[-1] == 2 and not force_predefined
    ny = mod.ny
    # one page-locked array per output for the whole batch: every device thread hands row (column) ranges of them
    # to the C ABI, the copies back from the devices ARE the gather
    status = mod.pinned.empty((n,), np.int32)
    counters = mod.pinned.empty((6, n), np.int32)
    timings = np.zeros((g, 4))
    errors = [None]*g
    kw = dict(kwargs)
    kw.pop('device', None)
    if adaptive:
        steps = int(nsteps) if nsteps else 500
        max_rows = kw.pop('max_rows', None) or steps*(1 + int(kw.get('autorestart', 0))) + 1
        xyout = mod.pinned.empty((n, max_rows, 1 + ny), np.float64)
        nrows = mod.pinned.empty((n,), np.int32)
    else:
        shared_x = intern_x.shape[0] > 1 and intern_x.strides[0] == 0  # one broadcast row (core.to_arrays)
        xout = np.ascontiguousarray(intern_x[0] if shared_x else intern_x, dtype=np.float64)
        nt = xout.shape[-1]
        yout = mod.pinned.empty((n, nt, ny), np.float64)
        nreached = mod.pinned.empty((n,), np.int32)

    def work(k):
        lo, hi = shard_range(n, g, k)
        if hi <= lo:
            return
        try:
            if adaptive:
                r = mod.integrate_adaptive_arrays(
                    y0[lo:hi], intern_x[lo:hi, 0], intern_x[lo:hi, 1], params[lo:hi], atol, rtol,
                    dx0=first_step, mxsteps=nsteps, device=devices[k], max_rows=max_rows, out=xyout[lo:hi],
                    out_aux=(nrows[lo:hi], status[lo:hi], counters[:, lo:hi], n), **kw)
            else:
                r = mod.integrate_predefined_arrays(
                    y0[lo:hi], xout if shared_x else xout[lo:hi], params[lo:hi], atol, rtol, dx0=first_step,
                    mxsteps=nsteps,
                    device=devices[k], out=yout[lo:hi],
                    out_aux=(nreached[lo:hi], status[lo:hi], counters[:, lo:hi], n), **kw)
            timings[k] = r[4]
        except Exception as e:  # re-raised on the calling thread (anyode_parallel.hpp:6-27 semantics)
            errors[k] = e

    threads = [threading.Thread(target=work, args=(k,)) for k in range(g)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for e in errors:
        if e is not None:
            raise e
    tmax = timings.max(axis=0)
    roe = kwargs.get('return_on_error', False)
    method = kwargs.get('method') or mod.default_method()
    iv = y0 if kwargs.get('iv_from_y0', False) else None
    if adaptive:
        return BatchInfo('adaptive', status, None, counters, 0.0, 0.0, _Ragged(xyout, nrows, 0),
                         _Ragged(xyout, nrows, 1), params, roe, timings=_tdict(tmax), method=method, iv=iv)
    return BatchInfo('predefined', status, nreached, counters, 0.0, 0.0, intern_x, yout, params, roe,
                     nt=nt, timings=_tdict(tmax), method=method, iv=iv)


def gather_to_rank0(local, n_total, dist=None):
    """ Host-side gather of per-rank row blocks (contiguous ranges of :func:`shard_range`) on rank 0.

    ``local``: this rank's array whose first axis is its shard.  Returns the full array on rank 0 and
    ``None`` elsewhere.  Works with the gloo backend on CPU tensors (used by the tests) and needs no
    collective on the integration path itself.
    """
    import torch
    if dist is None:
        import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    per = -(-n_total // world)
    shape = (per,) + tuple(local.shape[1:])
    buf = torch.zeros(shape, dtype=torch.from_numpy(np.asarray(local)).dtype)
    buf[:local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local))
    out = [torch.zeros_like(buf) for _ in range(world)] if rank == 0 else None
    dist.gather(buf, out, dst=0)
    if rank != 0:
        return None
    parts = []
    for k in range(world):
        lo, hi = shard_range(n_total, world, k)
        parts.append(out[k][:hi - lo].numpy())
    return np.concatenate(parts, axis=0)
