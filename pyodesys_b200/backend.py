# -*- coding: utf-8 -*-
"""
SymPy-only symbolic back end.

The reference obtains its CAS through ``sym.Backend`` (``pyodesys/symbolic.py:23``), a
package which is not installed in this image.  Only the small surface the batched native
path touches is provided here: symbol factories, ``Matrix``,
``cse``, ``ccode``, ``numbered_symbols``, ``Lambdify`` and a handful of functions.
"""
from __future__ import annotations

import numpy as np
import sympy


class _Lambdify:
    """Vectorised numeric callback: ``cb(arr)`` where ``arr[..., i]`` is the value of ``args[i]``.

    Same calling convention the reference's ``_Callback`` relies on
    (``pyodesys/util.py:246-261``): leading axes broadcast, last axis enumerates arguments,
    the result has the expressions on its last axis.
    """

    def __init__(self, args, exprs):
        self.args = tuple(args)
        exprs = list(exprs)
        self.nout = len(exprs)
        self._fn = sympy.lambdify(self.args, exprs, modules='numpy', cse=True)

    def __call__(self, inp):
        inp = np.asarray(inp, dtype=np.float64)
        if inp.shape[-1] != len(self.args):
            raise ValueError("Expected last axis of length %d" % len(self.args))
        cols = [inp[..., i] for i in range(len(self.args))]
        out = self._fn(*cols)
        lead = inp.shape[:-1]
        res = np.empty(lead + (self.nout,), dtype=np.float64)
        for i, o in enumerate(out):
            res[..., i] = o
        return res


class Backend:
    """ Minimal stand-in for ``sym.Backend('sympy')``. """

    name = 'sympy'

    def __init__(self, *_ignored):
        pass

    # symbol factories -----------------------------------------------------------------
    @staticmethod
    def Symbol(name, real=True, **kw):
        return sympy.Symbol(name, real=real, **kw)

    @staticmethod
    def Dummy(name=None, **kw):
        return sympy.Dummy(name, **kw)

    @staticmethod
    def real_symarray(prefix, n):
        return tuple(sympy.Symbol('%s_%d' % (prefix, i), real=True) for i in range(n))

    symarray = real_symarray

    # containers / algorithms ----------------------------------------------------------
    Matrix = staticmethod(sympy.Matrix)
    numbered_symbols = staticmethod(sympy.numbered_symbols)
    flatten = staticmethod(sympy.flatten)
    Lambdify = _Lambdify

    @staticmethod
    def cse(exprs, symbols=None, ignore=()):
        return sympy.cse(list(exprs), symbols=symbols, ignore=tuple(ignore))

    @staticmethod
    def ccode(expr):
        return sympy.ccode(expr)

    @staticmethod
    def sparse_jacobian_csc(exprs, dep):
        """ CSC triple (data expressions, colptrs, rowvals) — layout of ``symbolic.py:641``. """
        data, colptrs, rowvals = [], [0], []
        for dj in dep:
            for ri, e in enumerate(exprs):
                d = sympy.diff(e, dj)
                if d != 0:
                    data.append(d)
                    rowvals.append(ri)
            colptrs.append(len(data))
        return sympy.Matrix(1, len(data), data), colptrs, rowvals

    # functions ------------------------------------------------------------------------
    exp = staticmethod(sympy.exp)
    log = staticmethod(sympy.log)
    sqrt = staticmethod(sympy.sqrt)
    sin = staticmethod(sympy.sin)
    cos = staticmethod(sympy.cos)
    tanh = staticmethod(sympy.tanh)
    Abs = staticmethod(sympy.Abs)
    Min = staticmethod(sympy.Min)
    Max = staticmethod(sympy.Max)
    pi = sympy.pi

    def __getattr__(self, key):  # anything else sympy has (be.tan, be.Piecewise, ...)
        return getattr(sympy, key)


def get_backend(be=None):
    if isinstance(be, Backend):
        return be
    return Backend()
