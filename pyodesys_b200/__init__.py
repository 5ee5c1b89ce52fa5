# -*- coding: utf-8 -*-
"""
pyodesys_b200 - a B200-native (sm_100a) back end for the batched native integration path of
pyodesys: ``NativeSys.from_callback(...).integrate(x, y0[N,ny], params[N,np], atol, rtol, nsteps)``.

Only that path is provided: symbolic definition -> CUDA code generation ->
hand-written integrator kernels behind a C ABI -> ``Result`` / ``info`` in the reference's layout.
"""
from .core import (ODESys, Result, BatchResult, BatchInfo, integrate_auto_switch, integrate_chained,
                   chained_parameter_variation)
from .symbolic import SymbolicSys
from .transforms import TransformedSys, ScaledSys, symmetricsys, PartiallySolvedSystem, get_logexp

__version__ = '0.1.0'
__all__ = ['ODESys', 'Result', 'BatchResult', 'BatchInfo', 'SymbolicSys', 'TransformedSys', 'ScaledSys',
           'symmetricsys', 'PartiallySolvedSystem', 'get_logexp', 'integrate_auto_switch', 'integrate_chained',
           'chained_parameter_variation', 'native_sys']


def __getattr__(name):
    if name == 'native_sys':
        from .native import native_sys
        return native_sys
    raise AttributeError(name)
