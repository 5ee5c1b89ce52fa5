# -*- coding: utf-8 -*-
"""
Symbolic system definition feeding the CUDA code generator.

Carries the part of ``pyodesys.symbolic.SymbolicSys`` that the native path consumes:
construction from ``(dep, expr)`` pairs or from a callback (``pyodesys/symbolic.py:231-318``,
``:384-463``), analytic Jacobian / ``df/dx`` derivation (``:637-650``, ``:674-684``),
initial-value symbols appended to the parameter vector (``append_iv``, ``:268-273``) and
the invariants list (``all_invariants``).  System *transformations* (``TransformedSys``,
``PartiallySolvedSystem`` …) are CAS work outside the data-parallel path.
"""
from __future__ import annotations

import inspect

import numpy as np

from .backend import Backend, get_backend
from .core import ODESys


def _ensure_4args(func):
    """ ``pyodesys/util.py:191-221``: accept ``f(x, y)``, ``f(x, y, p)`` or ``f(x, y, p, be)``. """
    if func is None:
        return None
    try:
        nargs = len(inspect.signature(func).parameters)
    except (TypeError, ValueError):
        nargs = 4
    if nargs >= 4:
        return func
    if nargs == 3:
        return lambda x, y, p, be: func(x, y, p)
    if nargs == 2:
        return lambda x, y, p, be: func(x, y)
    raise ValueError("Incorrect number of arguments")


def _is_autonomous(indep, exprs):
    if indep is None:
        return True
    return all(indep not in e.free_symbols for e in exprs if hasattr(e, 'free_symbols'))


def _get_indep_name(names):
    if 'x' not in names:
        return 'x'
    i = 0
    while 'indep%d' % i in names:
        i += 1
    return 'indep%d' % i


class SymbolicSys(ODESys):
    """ ODE system from symbolic expressions (see module docstring for the covered surface). """

    _attrs_to_copy = ('first_step_expr', 'names', 'param_names', 'dep_by_name', 'par_by_name',
                      'description', 'linear_invariants', 'nonlinear_invariants',
                      'to_arrays_callbacks')
    nnz = -1
    band = None
    sparse = False

    def __init__(self, dep_exprs, indep=None, params=None, jac=True, dfdx=True,
                 jtimes=False, first_step_expr=None, roots=None, backend=None, lower_bounds=None,
                 upper_bounds=None, linear_invariants=None, nonlinear_invariants=None,
                 init_indep=None, init_dep=None, sparse=False, **kwargs):
        self.be = get_backend(backend)
        pairs = list(dep_exprs.items()) if isinstance(dep_exprs, dict) else list(dep_exprs)
        self.dep, exprs = zip(*pairs)
        self.dep = tuple(self.dep)
        self.exprs = tuple(self.be.sympify(e) for e in exprs)
        self.indep = indep
        if params is True or params is None:
            free = set()
            for e in self.exprs:
                free |= e.free_symbols
            all_free = tuple(s for s in sorted(free, key=str) if s not in self.dep + (self.indep,))
            if params is None and all_free:
                raise ValueError("Pass params explicitly or pass True to have them deduced.")
            params = all_free
        self.params = tuple(params)
        self._jac = jac
        self._dense_jac = jac
        self._jtimes = tuple(jtimes) if hasattr(jtimes, '__iter__') else jtimes  # symbolic.py:247
        self._dfdx = dfdx
        self.first_step_expr = first_step_expr
        self.roots = roots
        self.nroots = 0 if roots is None else len(roots)

        names = kwargs.get('names', None)
        if names is True:
            kwargs['names'] = [y.name for y in self.dep]
        if kwargs.get('param_names', None) is True:
            kwargs['param_names'] = [p.name for p in self.params]

        if init_indep is True:
            init_indep = self.be.Symbol('i_' + (str(self.indep) if self.indep is not None else 'indep'))
        if init_dep is True:
            init_dep = tuple(self.be.Symbol('i_' + str(d)) for d in self.dep)
        self.init_indep, self.init_dep = init_indep, init_dep
        if (init_indep is None) != (init_dep is None):
            raise ValueError("Need both or neither of init_indep & init_dep.")
        if init_dep is not None:
            kwargs['append_iv'] = True

        ny = len(self.exprs)
        self.lower_bounds = None if lower_bounds is None else np.array(lower_bounds)*np.ones(ny)
        self.upper_bounds = None if upper_bounds is None else np.array(upper_bounds)*np.ones(ny)
        self._nonnegative = kwargs.pop('nonnegative', False)
        # banded / sparse Jacobian storage is a choice of cvode's linear solver (pyodesys/symbolic.py:160-170); the
        # CUDA kernels factorise dense ny x ny matrices, so the hint is recorded and otherwise ignored
        self.band = kwargs.pop('band', None)
        self.autonomous_exprs = _is_autonomous(self.indep, self.exprs)
        super().__init__(**kwargs)
        self.linear_invariants = linear_invariants
        self.nonlinear_invariants = nonlinear_invariants
        self.sparse = sparse
        if self.sparse:  # symbolic.py:302-303
            if jac is not True:
                raise ValueError("sparse=True derives the CSC Jacobian itself: pass jac=True")
            self.get_jac()
            self.nnz = len(self._rowvals)

    @property
    def ny(self):
        return len(self.exprs)

    @property
    def linear_invariants(self):
        return self._linear_invariants

    @linear_invariants.setter
    def linear_invariants(self, lin_invar):
        if lin_invar is None or len(lin_invar) == 0:
            self._linear_invariants = None
        else:
            if isinstance(lin_invar[0], dict) and self.names:  # rows keyed by name (symbolic.py:73-74)
                lin_invar = [[d.get(n, 0) for n in self.names] for d in lin_invar]
            m = self.be.Matrix(len(lin_invar), self.ny, lambda i, j: lin_invar[i][j])
            self._linear_invariants = m

    def all_invariants(self, linear_invariants=None, nonlinear_invariants=None, dep=None):
        """ ``pyodesys/symbolic.py:720-727``: expressions constant along a trajectory. """
        linear_invariants = linear_invariants or self.linear_invariants
        nonlinear_invariants = nonlinear_invariants or self.nonlinear_invariants
        dep = dep or self.dep
        out = []
        if linear_invariants is not None:
            for i in range(linear_invariants.shape[0]):
                out.append(sum(linear_invariants[i, j]*dep[j] for j in range(len(dep))))
        return out + list(nonlinear_invariants or [])

    def get_invariants_callback(self):
        """ ``pyodesys/symbolic.py:755-759``: ``cb(x[nt], y[nt, ny], p[np_tot]) -> [nt, ninvar]`` """
        invar = self.all_invariants()
        if len(invar) == 0:
            return None
        fn = self._callback(invar)

        def cb(x, y, p):
            y = np.asarray(y, dtype=np.float64)
            xs = np.broadcast_to(np.asarray(x, dtype=np.float64), y.shape[:-1])[..., None]
            p = np.asarray(p, dtype=np.float64)
            npar = len(self.params) + (len(self.init_dep) if self.append_iv and self.init_dep else 0)
            p = np.broadcast_to(p[..., :npar], y.shape[:-1] + (npar,))
            return fn(np.concatenate((xs, y, p), axis=-1))
        return cb

    # -- derivation ---------------------------------------------------------------------
    def get_jac(self):
        """ ``pyodesys/symbolic.py:637-650``: the dense ny x ny matrix, or with ``sparse=True`` the row of the nnz
        structurally non-zero entries in compressed-sparse-column order (``_colptrs``, ``_rowvals`` beside it). """
        if self._jac is True:
            if self.sparse is True:
                self._jac, self._colptrs, self._rowvals = self.be.sparse_jacobian_csc(self.exprs, self.dep)
            else:
                self._jac = self.get_dense_jac()
        elif self._jac is False:
            return False
        elif not hasattr(self._jac, 'tolist'):
            self._jac = self.be.Matrix(self._jac)
        return self._jac

    def get_dense_jac(self):
        """ The ny x ny Jacobian whatever the storage hint: the CUDA kernels factorise dense matrices in registers
        or shared memory.  For a sparse system it is rebuilt from the CSC triple, so the kernels consume exactly
        the entries ``sparse_jac_csc`` produces. """
        if self._dense_jac is True:
            if self.sparse is True:
                data = self.get_jac()
                rows = [[0]*self.ny for _ in range(self.ny)]
                for col in range(self.ny):
                    for k in range(self._colptrs[col], self._colptrs[col + 1]):
                        rows[self._rowvals[k]][col] = data[k]
                self._dense_jac = self.be.Matrix(rows)
            else:
                f = self.be.Matrix(1, self.ny, self.exprs)
                self._dense_jac = f.jacobian(self.be.Matrix(1, self.ny, self.dep))
        elif self._dense_jac is False:
            return False
        elif not hasattr(self._dense_jac, 'tolist'):
            self._dense_jac = self.be.Matrix(self._dense_jac)
        return self._dense_jac

    def get_dfdx(self):
        """ ``pyodesys/symbolic.py:674-684`` """
        if self._dfdx is True:
            if self.indep is None:
                self._dfdx = self.be.Matrix(1, self.ny, [0]*self.ny)
            else:
                self._dfdx = self.be.Matrix(1, self.ny, [e.diff(self.indep) for e in self.exprs])
        elif self._dfdx is False:
            return False
        elif not hasattr(self._dfdx, 'tolist'):
            self._dfdx = self.be.Matrix(1, self.ny, list(self._dfdx))
        return self._dfdx

    def get_jtimes(self):
        """ ``pyodesys/symbolic.py:652-665``: ``(v, J(y) v)`` as the directional derivative of ``exprs`` along v. """
        if self._jtimes is False:
            return False
        if self._jtimes is True:
            r = self.be.Dummy('r')
            v = tuple(self.be.Dummy('v_{0}'.format(i)) for i in range(self.ny))
            f = self.be.Matrix(1, self.ny, self.exprs)
            f = f.subs([(x_i, x_i + r*v_i) for x_i, v_i in zip(self.dep, v)])
            self._jtimes = tuple(zip(v, f.diff(r).subs(r, 0)))
        return tuple(zip(*self._jtimes))

    # -- numeric callbacks (used by tests to cross-check generated device code) -----------
    def _callback(self, exprs):
        args = ((self.indep,) if self.indep is not None else (self.be.Dummy('x'),)) \
            + self.dep + self.params + (tuple(self.init_dep) if self.append_iv and self.init_dep else ())
        return self.be.Lambdify(args, exprs)

    def f_numeric(self, x, y, p):
        """ Evaluate the right-hand side with numpy; ``p`` must include appended initial values. """
        y = np.asarray(y, dtype=np.float64)
        xs = np.broadcast_to(np.asarray(x, dtype=np.float64), y.shape[:-1])[..., None]
        p = np.broadcast_to(np.asarray(p, dtype=np.float64), y.shape[:-1] + (np.shape(p)[-1],))
        return self._callback(self.exprs)(np.concatenate((xs, y, p), axis=-1))

    def j_numeric(self, x, y, p):
        y = np.asarray(y, dtype=np.float64)
        xs = np.broadcast_to(np.asarray(x, dtype=np.float64), y.shape[:-1])[..., None]
        p = np.broadcast_to(np.asarray(p, dtype=np.float64), y.shape[:-1] + (np.shape(p)[-1],))
        jac = self.get_dense_jac()
        flat = [jac[i, j] for i in range(self.ny) for j in range(self.ny)]
        out = self._callback(flat)(np.concatenate((xs, y, p), axis=-1))
        return out.reshape(y.shape[:-1] + (self.ny, self.ny))

    # -- alternate constructors ---------------------------------------------------------
    @classmethod
    def from_callback(cls, rhs, ny=None, nparams=None, first_step_factory=None,
                      roots_cb=None, indep_name=None, **kwargs):
        """ ``pyodesys/symbolic.py:384-463`` — call ``rhs(x, y, p, backend)`` on symbols. """
        names = tuple(kwargs.pop('names', None) or ())
        if ny is None:
            if not names:
                raise ValueError("Need ny or names")
            ny = len(names)
        if nparams is None:
            nparams = len(kwargs.get('param_names', ()) or ())
        be = get_backend(kwargs.pop('backend', None))
        indep_name = indep_name or _get_indep_name(names)
        x = be.Symbol(indep_name, real=True)
        y = be.real_symarray('y', ny)
        p = be.real_symarray('p', nparams)
        _y = dict(zip(names, y)) if kwargs.get('dep_by_name', False) else y
        _p = dict(zip(kwargs['param_names'], p)) if kwargs.get('par_by_name', False) else p
        exprs = _ensure_4args(rhs)(x, _y, _p, be)
        try:
            if len(exprs) != ny:
                raise ValueError("Callback returned unexpected (%d) number of expressions: %d"
                                 % (ny, len(exprs)))
        except TypeError:
            raise ValueError("Callback did not return an array_like of expressions: %s" % str(exprs))
        if roots_cb is not None:
            kwargs['roots'] = _ensure_4args(roots_cb)(x, _y, _p, be)
        if first_step_factory is not None:
            kwargs['first_step_expr'] = _ensure_4args(first_step_factory)(x, _y, _p, be)
        if kwargs.get('dep_by_name', False):
            exprs = [exprs[k] for k in names]
        return cls(zip(y, exprs), x, kwargs.pop('params', None) if len(p) == 0 else p,
                   backend=be, names=names, **kwargs)

    @classmethod
    def from_other(cls, ori, **kwargs):
        """ ``pyodesys/symbolic.py:465-508`` — same expressions, another class / options. """
        for k in cls._attrs_to_copy + ('params', 'roots', 'lower_bounds', 'upper_bounds'):
            if k not in kwargs:
                val = getattr(ori, k, None)
                if k == 'linear_invariants' and val is not None:
                    val = val.tolist()
                kwargs[k] = val
        if 'jac' not in kwargs and getattr(ori, '_dense_jac', True) is not True and not getattr(ori, 'sparse', False):
            kwargs['jac'] = ori._dense_jac
        if 'dfdx' not in kwargs and getattr(ori, '_dfdx', True) is not True:
            kwargs['dfdx'] = ori._dfdx
        kwargs.setdefault('backend', ori.be)
        # processors of the template stay innermost (``symbolic.py:489-497``): its pre-processors run first, its
        # post-processors last
        if ori.pre_processors:
            kwargs['pre_processors'] = list(kwargs.get('pre_processors') or []) + list(ori.pre_processors)
        if ori.post_processors:
            kwargs['post_processors'] = list(ori.post_processors) + list(kwargs.get('post_processors') or [])
        return cls(zip(ori.dep, ori.exprs), ori.indep, **kwargs)
