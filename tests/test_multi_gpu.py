# -*- coding: utf-8 -*-
"""
The product's multi-GPU path: ``integrate(..., devices=[...])`` -> :func:`pyodesys_b200.multi.integrate_sharded`
(contiguous chunks drawn from a shared host-side counter, one pinned output array for the whole batch,
no collective).  Systems are independent, so every array must equal the one-device call bit for bit - whatever the
chunking, the number of devices or which device integrated which chunk.

With one GPU visible the same driver is exercised with ``devices=[0, 0]`` (both shards land on device 0: two host
threads per shard entry, the two call contexts of the device); with two or more the real thing runs.
"""
import numpy as np
import pytest

from pyodesys_b200 import systems, multi

import systems_for_tests as tsys


def test_chunk_rule_and_cpulist_parser():
    # small batches: equal shares, one call per device; large ones: chunks of ~1e6 drawn dynamically
    assert multi.chunk_size(1000000, 1) == 1000000 and multi.chunk_size(1000000, 8) == 125952
    assert multi.chunk_size(100000000, 8) == 1000448 and multi.chunk_size(5000, 4) == 5000
    assert multi.chunk_size(2000000, 2) == 1000448 and multi.chunk_size(50000000, 2) == 1000448
    for n, g in ((1000000, 8), (123457, 3), (40000, 2)):
        c = multi.chunk_size(n, g)
        assert c == n or c % 1024 == 0
    assert multi._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert multi.gpu_cpu_affinity(12345) is None  # no such device: no binding, no exception


def _devices():
    import torch
    return [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]


@pytest.mark.gpu
@pytest.mark.parametrize('chunk', [0, 37000])
def test_sharded_predefined_bit_equal_to_one_device(cuda_device, chunk):
    rob = systems.robertson()
    n = 300000
    b = systems.batch_robertson(n=n, seed=5, tend=1e3)
    b['params'][4321, 0] = np.inf  # one system that fails (non-finite right-hand side)
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=2000, method='rosenbrock4', force_predefined=True,
              return_on_error=True, first_step=np.linspace(1e-6, 1e-5, n))
    xo = np.array([0.0, 1.0, 1e2, 1e3])
    one = rob.integrate(xo, b['y0'], b['params'], **kw)
    two = rob.integrate(xo, b['y0'], b['params'], devices=_devices(), shard_chunk=chunk, **kw)
    assert np.array_equal(np.array(one.yout), np.array(two.yout), equal_nan=True)
    for name in ('status', 'nreached', 'counters'):
        assert np.array_equal(np.array(getattr(one.info, name)), np.array(getattr(two.info, name))), name
    st = two.info.shard_stats
    assert sum(st['chunks_per_device']) == st['chunks'] == -(-n // st['chunk'])
    assert st['chunks'] >= (4 if chunk else 2) and two.info.status[4321] != 0  # default rule: one share per device
    r = two[12345]
    assert r.info['success'] and r.yout.shape == (4, 3) and abs(r.yout[-1].sum() - 1) < 1e-9


@pytest.mark.gpu
def test_sharded_adaptive_and_roots(cuda_device):
    """ ragged adaptive output and event functions through the sharded driver (round 1 dropped the root arrays
    silently when several devices were given) """
    native = tsys.chain3_roots_all()
    rng = np.random.default_rng(11)
    n = 70000
    y0 = np.column_stack((1 + rng.uniform(0, 0.5, n), rng.uniform(0, 0.1, n), np.zeros(n)))
    p = np.column_stack((rng.uniform(4, 9, n), rng.uniform(1, 3.5, n)))
    x = np.tile([0.0, 1.5], (n, 1))
    kw = dict(atol=1e-8, rtol=1e-8, nsteps=300, method='dopri5', max_roots=4)
    one = native.integrate(x, y0, p, **kw)
    two = native.integrate(x, y0, p, devices=_devices(), shard_chunk=33000, **kw)
    assert two.info.shard_stats['chunks'] == 3
    for i in (0, 32999, 33000, 69999):
        a, c = one[i], two[i]
        assert np.array_equal(a.xout, c.xout) and np.array_equal(a.yout, c.yout)
        assert a.info['root_indices'] == c.info['root_indices'] and len(c.info['root_indices']) == 3
    assert np.array_equal(one.info.roots['root_out'], two.info.roots['root_out'])


@pytest.mark.gpu
def test_two_real_devices(cuda_device):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU visible")
    rob = systems.robertson()
    b = systems.batch_robertson(n=200000, seed=8, tend=1e2)
    kw = dict(atol=b['atol'], rtol=b['rtol'], nsteps=b['nsteps'], force_predefined=True)
    one = rob.integrate(b['x'], b['y0'], b['params'], **kw)
    two = rob.integrate(b['x'], b['y0'], b['params'], devices=[0, 1], **kw)
    assert np.array_equal(np.array(one.yout), np.array(two.yout))
    assert all(c > 0 for c in two.info.shard_stats['chunks_per_device'])
