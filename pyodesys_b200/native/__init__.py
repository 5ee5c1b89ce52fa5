# -*- coding: utf-8 -*-
"""
``native_sys`` registry, same role as ``pyodesys/native/__init__.py:18-22`` where the
reference lists ``cvode``, ``gsl`` and ``odeint``; this package contributes ``cuda``.
"""
from .cuda import NativeCudaSys, NativeCudaCode, CudaModule

native_sys = {'cuda': NativeCudaSys}
__all__ = ['native_sys', 'NativeCudaSys', 'NativeCudaCode', 'CudaModule']
