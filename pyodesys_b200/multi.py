# -*- coding: utf-8 -*-
"""
Sharding of a batch over GPUs / ranks.

The batch is embarrassingly parallel (the reference's only parallelism is an OpenMP "bag of
tasks" over systems, ``README.rst:199-201``), so the partition is a contiguous index range per
shard and there is NO collective on the data path:

* :func:`shard_range` - the partition rule (used by ``bench.py`` for one-process-per-GPU runs under
  ``torchrun`` and by :func:`integrate_sharded`);
* :func:`integrate_sharded` - one process, two host threads per device (the two call contexts of a device,
  ``options.ctx_lane``); the batch is cut into chunks that the threads draw from a shared counter (dynamic
  scheduling: stiffness varies with the parameters), every thread is bound to the CPUs of its GPU's
  NUMA node, calls the C ABI (ctypes releases the GIL) on its chunk and writes into disjoint slices of batch-wide
  page-locked output arrays, so the "gather" is free;
* :func:`gather_to_rank0` - optional host-side gather for multi-process runs (``torch.distributed``,
  gloo or nccl) when a caller wants all results in one place.
"""
from __future__ import annotations

import itertools
import os
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


_affinity_cache = {}


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(','):
        if not part:
            continue
        lo, _, hi = part.partition('-')
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_cpu_affinity(device, lib=None):
    """ CPUs local to CUDA device ``device`` (``/sys/bus/pci/devices/<bus id>/local_cpulist``), or ``None``.
    ``lib``: a loaded system library (``CudaModule.lib``) - it resolves the bus id of the CUDA device index, which
    honours ``CUDA_VISIBLE_DEVICES``; without one ``nvidia-smi`` order is used. """
    if device in _affinity_cache:
        return _affinity_cache[device]
    cpus = None
    try:
        bus = None
        if lib is not None and hasattr(lib, 'pyodesys_cuda_device_pci_bus_id'):
            import ctypes
            buf = ctypes.create_string_buffer(32)
            if lib.pyodesys_cuda_device_pci_bus_id(int(device), buf, 32) == 0:
                bus = buf.value.decode().lower()
        if bus is None:
            import pynvml
            pynvml.nvmlInit()
            info = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(int(device)))
            bus = info.busId.decode() if isinstance(info.busId, bytes) else info.busId
            bus = bus.lower()[-12:]  # nvml prints an 8-digit domain
        with open('/sys/bus/pci/devices/%s/local_cpulist' % bus) as fh:
            cpus = _parse_cpulist(fh.read())
        allowed = os.sched_getaffinity(0)
        cpus = (cpus & allowed) or None
    except Exception:
        cpus = None
    _affinity_cache[device] = cpus
    return cpus


def chunk_size(n, ndevices, target=1000000, minimum=32768):
    """ Chunk length of the multi-device schedule.  A chunk is one call of the single-device path, which overlaps its
    own copies with its kernel (streaming launch) and ends in the ~1 ms makespan tail of one-system-per-thread
    scheduling, so chunks are made LARGE: measured on two GPUs (``scripts/sharded_chunks.py``) 2e6 systems take 16.9
    ms in chunks of 1e6 against 18.8 ms in chunks of 2.5e5, 2e7 systems 160 against 178 ms.  Rule: a batch of fewer
    than two ``target``-sized chunks per device is dealt in equal shares (one call per device); a larger one is cut
    into chunks of ``target`` systems that the device lanes draw dynamically (a slow chunk - stiffness varies with the
    parameters - then does not leave the other GPUs idle at the end).  Never below ``minimum`` unless the batch is
    smaller, a multiple of 1024. """
    g = max(1, ndevices)
    if n < 2*target*g:
        c = max(min(minimum, n), -(-n // g))
    else:
        c = target
    if c >= n:
        return n
    return -(-c // 1024)*1024


def integrate_sharded(mod, devices, intern_x, y0, params, force_predefined, atol, rtol, nsteps,
                      first_step, kwargs):
    """ Runs ``mod`` (a :class:`~pyodesys_b200.native.cuda.CudaModule`) on several devices at once. """
    from .native.cuda import _RaggedPieces, _tdict
    n = y0.shape[0]
    devices = list(devices)
    g = len(devices)
    adaptive = intern_x.shape[-1] == 2 and not force_predefined
    ny = mod.ny
    # one page-locked array per output for the whole batch: every device thread hands row (column) ranges of them
    # to the C ABI, the copies back from the devices ARE the gather
    status = mod.pinned.empty((n,), np.int32)
    counters = mod.pinned.empty((6, n), np.int32)
    kw = dict(kwargs)
    kw.pop('device', None)
    csize = int(kw.pop('shard_chunk', 0)) or chunk_size(n, g)
    nchunks = -(-n // csize) if n else 0
    lanes = 2 if nchunks > g else 1  # more than one chunk per device: two call contexts keep a kernel running
    roots = kw.pop('roots_sink', None)
    return_on_root, max_roots = kw.pop('return_on_root', False), int(kw.pop('max_roots', 8))
    root_arrays = None
    if roots is not None and mod.nroots and max_roots > 0:
        root_arrays = (mod.pinned.empty((n, max_roots, 1 + ny), np.float64),
                       mod.pinned.empty((n, max_roots), np.int32), mod.pinned.empty((n,), np.int32))
        roots.update(root_out=root_arrays[0], root_indices=root_arrays[1], nroots_found=root_arrays[2],
                     max_roots=max_roots)
    if adaptive:
        steps = int(nsteps) if nsteps else 500
        max_rows = kw.pop('max_rows', None) or steps*(1 + int(kw.get('autorestart', 0))) + 1
        if root_arrays is not None:
            max_rows += max_roots
        pieces = []  # compacted rows of every chunk: (first system, rows, offsets)
    else:
        shared_x = intern_x.shape[0] > 1 and intern_x.strides[0] == 0  # one broadcast row (core.to_arrays)
        xout = np.ascontiguousarray(intern_x[0] if shared_x else intern_x, dtype=np.float64)
        nt = xout.shape[-1]
        yout = mod.pinned.empty((n, nt, ny), np.float64)
        nreached = mod.pinned.empty((n,), np.int32)

    if not adaptive and root_arrays is None and os.environ.get('PYODESYS_SHARDED_NATIVE', '1') == '1':
        # predefined output, no event functions: ONE call; chunk scheduling, device threads and NUMA binding are
        # native (pyodesys_cuda_integrate_predefined_multi) - per-chunk Python work under the GIL cost ~4 ms per call
        kwn = dict(kw)
        kwn.pop('ctx_lane', None)
        yout, nreached, status, counters, busy_d, taken_d = mod.integrate_predefined_multi(
            devices, lanes, csize, y0, xout, params, atol, rtol, dx0=first_step, mxsteps=nsteps,
            out=(yout, nreached, status, counters), **kwn)
        binfo = BatchInfo('predefined', status, nreached, counters, 0.0, 0.0, intern_x, yout, params,
                          kwargs.get('return_on_error', False), nt=nt,
                          timings=_tdict(np.array([float(busy_d.max())/max(lanes, 1), 0.0, 0.0, float(busy_d.max())/max(lanes, 1)])),
                          method=kwargs.get('method') or mod.default_method(),
                          iv=y0 if kwargs.get('iv_from_y0', False) else None)
        binfo.shard_stats = {"chunk": int(csize), "chunks": int(nchunks), "lanes_per_device": lanes, "driver": "native",
                             "chunks_per_device": [int(v) for v in taken_d],
                             "busy_s_per_device": [float(v) for v in busy_d]}
        return binfo

    draw = itertools.count()  # the shared chunk counter: next() is atomic under the GIL
    nworkers = g*lanes
    timings = np.zeros((nworkers, 4))
    busy = np.zeros(nworkers)
    taken = [0]*nworkers
    errors = [None]*nworkers
    stop = threading.Event()

    def work(w):
        dev, lane = devices[w // lanes], w % lanes
        try:
            cpus = gpu_cpu_affinity(dev, mod.lib)
            if cpus:
                os.sched_setaffinity(0, cpus)  # pid 0 = the calling thread
        except OSError:
            pass
        try:
            while not stop.is_set():
                k = next(draw)
                if k >= nchunks:
                    break
                lo, hi = k*csize, min(n, (k + 1)*csize)
                extra = {}
                if root_arrays is not None:
                    extra = dict(roots_sink={}, return_on_root=return_on_root, max_roots=max_roots,
                                 roots_out=tuple(a[lo:hi] for a in root_arrays))
                if adaptive:
                    pcs, _st, _cn, tm = mod.integrate_adaptive_compact(
                        y0[lo:hi], intern_x[lo:hi, 0], intern_x[lo:hi, 1], params[lo:hi], atol, rtol,
                        dx0=_rows(first_step, lo, hi, n), mxsteps=nsteps, device=dev, max_rows=max_rows,
                        out_aux=(status[lo:hi], counters[:, lo:hi], n), ctx_lane=lane, **extra,
                        **_row_kw(kw, lo, hi, n))
                    pieces.extend((lo + plo, rows, off) for plo, rows, off in pcs)  # list.extend is atomic enough
                    r = (None, None, None, None, tm)
                else:
                    r = mod.integrate_predefined_arrays(
                        y0[lo:hi], xout if shared_x else xout[lo:hi], params[lo:hi], atol, rtol,
                        dx0=_rows(first_step, lo, hi, n), mxsteps=nsteps, device=dev, out=yout[lo:hi],
                        out_aux=(nreached[lo:hi], status[lo:hi], counters[:, lo:hi], n), ctx_lane=lane,
                        **extra, **_row_kw(kw, lo, hi, n))
                timings[w] += r[4]
                busy[w] += r[4][3]
                taken[w] += 1
        except Exception as e:  # re-raised on the calling thread (anyode_parallel.hpp:6-27 semantics)
            errors[w] = e
            stop.set()

    threads = [threading.Thread(target=work, args=(w,)) for w in range(nworkers)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for e in errors:
        if e is not None:
            raise e
    tmax = timings.max(axis=0) if nworkers else np.zeros(4)
    roe = kwargs.get('return_on_error', False)
    method = kwargs.get('method') or mod.default_method()
    iv = y0 if kwargs.get('iv_from_y0', False) else None
    if adaptive:
        pieces.sort(key=lambda pc: pc[0])
        binfo = BatchInfo('adaptive', status, None, counters, 0.0, 0.0, _RaggedPieces(pieces, n, 0),
                          _RaggedPieces(pieces, n, 1), params, roe, timings=_tdict(tmax), method=method, iv=iv)
        binfo.d2h_bytes = sum(r_.nbytes + o_.nbytes for _lo, r_, o_ in pieces) + status.nbytes + counters.nbytes
    else:
        binfo = BatchInfo('predefined', status, nreached, counters, 0.0, 0.0, intern_x, yout, params, roe,
                          nt=nt, timings=_tdict(tmax), method=method, iv=iv)
    binfo.shard_stats = {"chunk": int(csize), "chunks": int(nchunks), "lanes_per_device": lanes,
                         "chunks_per_device": [int(sum(taken[d*lanes:(d + 1)*lanes])) for d in range(g)],
                         "busy_s_per_device": [float(busy[d*lanes:(d + 1)*lanes].sum()) for d in range(g)]}
    return binfo


def _rows(v, lo, hi, n):
    """ per-system step-size arrays follow their rows into a chunk; scalars pass through """
    if v is None or np.ndim(v) == 0 or np.size(v) == 1:
        return v
    v = np.asarray(v).reshape(-1)
    if v.size < n:
        raise ValueError('dx0 too short')
    return v[lo:hi]


def _row_kw(kw, lo, hi, n):
    out = dict(kw)
    for key in ('dx_min', 'dx_max'):
        if key in out and out[key] is not None and np.size(out[key]) > 1:
            v = np.asarray(out[key]).reshape(-1)
            if v.size < n:
                raise ValueError('%s too short' % key)
            out[key] = v[lo:hi]
    return out


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
