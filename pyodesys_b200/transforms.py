# -*- coding: utf-8 -*-
"""
Variable transformations in front of the batched native path.

Covers what the reference's native tests reach through ``class T(TransformedSys, NativeSys)``
(``pyodesys/native/tests/_tests.py:54-55,81-82,198-227``): :class:`TransformedSys`
(``pyodesys/symbolic.py:798-958``), :func:`symmetricsys` (``:961-1027``) and :class:`ScaledSys`
(``:1030-1122``); the expression rewriting follows ``pyodesys/util.py:126-187``
(``transform_exprs_dep`` / ``transform_exprs_indep``).

Difference in kind: the reference's pre/post processors recurse over the systems of a batch
(``zip(*[... for ...])``, ``symbolic.py:937-958``); here both directions are ONE vectorised numpy
evaluation over the whole batch (``[N, nt, ny]`` in one call), so a log-transformed batch of 1e6 systems
does not spend its time in a Python loop.  The transformed right-hand side itself is compiled to CUDA
like any other system, because the mixin order puts :class:`NativeCudaSys` behind this class.
"""
from __future__ import annotations

from itertools import repeat

import numpy as np

from .backend import get_backend
from .symbolic import SymbolicSys, _ensure_4args, _get_indep_name


def _check_pair(fw, bw, sym, be):
    """ bw(fw(sym)) == sym (``pyodesys/util.py:104-123``) """
    back = bw.subs(sym, fw)
    diff = be.simplify(back - sym) if hasattr(be, 'simplify') else back - sym
    if diff != 0:
        import sympy
        if sympy.simplify(sympy.expand_log(sympy.powsimp(back - sym, force=True), force=True)) != 0:
            raise ValueError("Incorrect (did you set real=True?) fw: %s, bw: %s" % (fw, bw))


class TransformedSys(SymbolicSys):
    """ ``SymbolicSys`` integrated in transformed variables ``z_i = fw_i(y_i)``, ``xi = fw(x)``.

    Parameters as in the reference: ``dep_transf`` = iterable of (forward, backward) expression pairs in
    the dependent-variable symbols, ``indep_transf`` = one such pair in ``indep``, ``exprs_process_cb``
    post-processes the transformed expressions.
    """

    def __init__(self, dep_exprs, indep=None, dep_transf=None, indep_transf=None, params=(),
                 exprs_process_cb=None, check_transforms=True, **kwargs):
        pairs = list(dep_exprs.items()) if isinstance(dep_exprs, dict) else list(dep_exprs)
        dep, exprs = zip(*pairs)
        be = get_backend(kwargs.get('backend'))
        exprs = [be.sympify(e) for e in exprs]
        lin = kwargs.pop('linear_invariants', None)
        nonlin = list(kwargs.pop('nonlinear_invariants', None) or [])
        if lin is not None and len(lin) > 0:
            # linear in y is not linear in z: carried as general invariants
            nonlin = [sum(row[j]*dep[j] for j in range(len(dep))) for row in lin] + nonlin
        roots = kwargs.pop('roots', None)

        if dep_transf is not None:
            dep_transf = list(dep_transf)
            if len(dep_transf) != len(dep):
                raise ValueError("Incompatible lengths")
            self.dep_fw, self.dep_bw = map(tuple, zip(*dep_transf))
            if check_transforms:
                for fw, bw, y in zip(self.dep_fw, self.dep_bw, dep):
                    _check_pair(fw, bw, y, be)
            back = dict(zip(dep, self.dep_bw))
            # dz/dx = fw'(y) f(y) with y = bw(z); the symbols of y are reused for z
            exprs = [(e*fw.diff(y)).subs(back, simultaneous=True) for fw, y, e in zip(self.dep_fw, dep, exprs)]
            nonlin = [iv.subs(back, simultaneous=True) for iv in nonlin]
            if roots is not None:
                roots = [r.subs(back, simultaneous=True) for r in roots]
        else:
            self.dep_fw = self.dep_bw = None

        if indep_transf is not None:
            if indep is None:
                raise ValueError("indep_transf needs the independent variable")
            self.indep_fw, self.indep_bw = indep_transf
            if check_transforms:
                _check_pair(self.indep_fw, self.indep_bw, indep, be)
            dfw = self.indep_fw.diff(indep)
            # dy/dxi = f / fw'(x) with x = bw(xi)
            exprs = [(e/dfw).subs(indep, self.indep_bw) for e in exprs]
            nonlin = [iv.subs(indep, self.indep_bw) for iv in nonlin]
            if roots is not None:
                roots = [r.subs(indep, self.indep_bw) for r in roots]
        else:
            self.indep_fw = self.indep_bw = None

        if exprs_process_cb is not None:
            exprs = exprs_process_cb(exprs)
        lower_b = kwargs.pop('lower_bounds', None)
        upper_b = kwargs.pop('upper_bounds', None)
        pre = list(kwargs.pop('pre_processors', None) or [])
        post = list(kwargs.pop('post_processors', None) or [])
        # the numeric transformation callbacks take (x, y..., p...) on the last axis
        tr_args = ((indep,) if indep is not None else (be.Dummy('x'),)) + tuple(dep) + tuple(params or ())
        mk = lambda ex: None if ex is None else be.Lambdify(tr_args, ex)  # noqa: E731
        self._f_dep, self._b_dep = mk(self.dep_fw), mk(self.dep_bw)
        self._f_indep = mk(None if self.indep_fw is None else [self.indep_fw])
        self._b_indep = mk(None if self.indep_bw is None else [self.indep_bw])
        self._n_tr_params = len(tuple(params or ()))
        # invariants that are still linear after the transformation (ScaledSys) stay linear rows, the rest become
        # general expressions (``_group_invariants``, pyodesys/symbolic.py:777-795)
        lin_rows, nonlin_kept = [], []
        for iv in nonlin:
            derivs = [be.sympify(iv).diff(y) for y in dep]
            if all(d.is_Number for d in derivs):
                lin_rows.append(derivs)
            else:
                nonlin_kept.append(iv)
        nonlin = nonlin_kept
        super().__init__(zip(dep, exprs), indep, params, roots=roots,
                         linear_invariants=lin_rows or None, nonlinear_invariants=nonlin or None,
                         pre_processors=pre + [self._forward_transform_xy],
                         post_processors=[self._back_transform_out] + post, **kwargs)
        nan_p = [float('nan')]*self._n_tr_params
        for name, b in (('lower_bounds', lower_b), ('upper_bounds', upper_b)):
            if b is not None:
                b = np.asarray(b, dtype=np.float64)*np.ones(len(dep))
                if self._f_dep is not None:
                    b = self._f_dep(np.concatenate(([float('nan')], b, nan_p)))
                setattr(self, name, b)

    # ---- vectorised processors: any number of leading batch axes --------------------------------------
    def _stack(self, x, y, p):
        """ argument rows [x, y..., p...] with x[...], y[..., ny], p[..., np] broadcast together """
        y = np.asarray(y, dtype=np.float64)
        lead = np.broadcast_shapes(np.shape(x), y.shape[:-1])
        p = np.asarray(p, dtype=np.float64)[..., :self._n_tr_params]
        p = np.broadcast_to(p.reshape(p.shape[:-1] + (1,)*(len(lead) - (p.ndim - 1)) + p.shape[-1:]),
                            lead + (self._n_tr_params,))
        return np.concatenate((np.broadcast_to(x, lead)[..., None], np.broadcast_to(y, lead + y.shape[-1:]), p),
                              axis=-1)

    def _forward_transform_xy(self, x, y, p):
        """ (x[..., nt], y0[..., ny], p[..., np]) -> internal; ``symbolic.py:949-958`` without the per-system loop """
        x, y, p = (np.asarray(a, dtype=np.float64) for a in (x, y, p))
        _y = y if self._f_dep is None else self._f_dep(self._stack(x[..., 0], y, p))
        if self._f_indep is None:
            _x = x
        else:
            _x = self._f_indep(self._stack(x, y[..., None, :], p))[..., 0]
        return _x, _y, p

    def _back_transform_out(self, xout, yout, params):
        """ (xout[..., nt], yout[..., nt, ny], p[..., np]) -> user variables; ragged input (adaptive batches)
        is handled row by row """
        if isinstance(xout, (list, tuple)) or (hasattr(xout, '__getitem__') and not hasattr(xout, 'ndim')):
            out = [self._back_transform_out(xout[i], yout[i], params[i]) for i in range(len(xout))]
            return [o[0] for o in out], [o[1] for o in out], params
        if self._b_indep is None and self._b_dep is None:
            return xout, yout, params
        rows = self._stack(xout, yout, np.asarray(params))
        x = xout if self._b_indep is None else self._b_indep(rows)[..., 0]
        y = yout if self._b_dep is None else self._b_dep(rows)
        return x, y, params

    # ---- construction from callbacks -------------------------------------------------------------------------
    @classmethod
    def from_callback(cls, cb, ny=None, nparams=None, dep_transf_cbs=None, indep_transf_cbs=None,
                      roots_cb=None, **kwargs):
        """ ``pyodesys/symbolic.py:885-935``: ``dep_transf_cbs`` = iterable of (fw, bw) callables in ``y_i``,
        ``indep_transf_cbs`` = one such pair in ``x``. """
        names = tuple(kwargs.get('names', None) or ())
        if ny is None:
            ny = len(names)
        if nparams is None:
            nparams = len(kwargs.get('param_names', ()) or ())
        be = get_backend(kwargs.pop('backend', None))
        x = be.Symbol(kwargs.pop('indep_name', None) or _get_indep_name(names), real=True)
        y = be.real_symarray('y', ny)
        p = be.real_symarray('p', nparams)
        _y = dict(zip(names, y)) if kwargs.get('dep_by_name', False) else y
        _p = dict(zip(kwargs['param_names'], p)) if kwargs.get('par_by_name', False) else p
        exprs = _ensure_4args(cb)(x, _y, _p, be)
        if kwargs.get('dep_by_name', False):
            exprs = [exprs[k] for k in names]
        dep_transf = None if dep_transf_cbs is None else [(fw(yi), bw(yi)) for (fw, bw), yi in zip(dep_transf_cbs, y)]
        indep_transf = None if indep_transf_cbs is None else (indep_transf_cbs[0](x), indep_transf_cbs[1](x))
        if roots_cb is not None:
            kwargs['roots'] = _ensure_4args(roots_cb)(x, _y, _p, be)
        return cls(list(zip(y, exprs)), x, dep_transf=dep_transf, indep_transf=indep_transf, params=p,
                   backend=be, **kwargs)


def symmetricsys(dep_tr=None, indep_tr=None, SuperClass=TransformedSys, **kwargs):
    """ Class factory: the same (forward, backward) callables for every dependent variable
    (``pyodesys/symbolic.py:961-1027``). """
    for tr, what in ((dep_tr, 'dep_tr'), (indep_tr, 'indep_tr')):
        if tr is not None and not (callable(tr[0]) and callable(tr[1])):
            raise ValueError("Expected %s to be a pair of callables" % what)

    class _SymmetricSys(SuperClass):
        def __init__(self, dep_exprs, indep=None, **inner_kwargs):
            kw = dict(kwargs)
            kw.update(inner_kwargs)
            pairs = list(dep_exprs)
            if 'dep_transf' not in kw and dep_tr is not None:
                kw['dep_transf'] = [(dep_tr[0](d), dep_tr[1](d)) for d, _ in pairs]
            if 'indep_transf' not in kw and indep_tr is not None:
                kw['indep_transf'] = (indep_tr[0](indep), indep_tr[1](indep))
            super().__init__(pairs, indep, **kw)

        @classmethod
        def from_callback(cls, cb, ny=None, nparams=None, **inner_kwargs):
            kw = dict(kwargs)
            kw.update(inner_kwargs)
            return SuperClass.from_callback.__func__(
                cls, cb, ny, nparams, dep_transf_cbs=repeat(dep_tr) if dep_tr is not None else None,
                indep_transf_cbs=indep_tr, **kw)

    return _SymmetricSys


class ScaledSys(TransformedSys):
    """ Linear scaling of dependent / independent variables (``pyodesys/symbolic.py:1030-1122``). """

    @staticmethod
    def _scale_fw_bw(scaling):
        return (lambda v: scaling*v, lambda v: v/scaling)

    def __init__(self, dep_exprs, indep=None, dep_scaling=1, indep_scaling=1, params=(), **kwargs):
        pairs = list(dep_exprs)
        try:
            scal = list(dep_scaling)
        except TypeError:
            scal = [dep_scaling]*len(pairs)
        kwargs.setdefault('dep_transf', [(s*d, d/s) for s, (d, _) in zip(scal, pairs)])
        if indep is not None:
            kwargs.setdefault('indep_transf', (indep_scaling*indep, indep/indep_scaling))
        super().__init__(pairs, indep, params=params, **kwargs)

    @classmethod
    def from_callback(cls, cb, ny=None, nparams=None, dep_scaling=1, indep_scaling=1, **kwargs):
        return TransformedSys.from_callback.__func__(
            cls, cb, ny, nparams, dep_transf_cbs=repeat(cls._scale_fw_bw(dep_scaling)),
            indep_transf_cbs=cls._scale_fw_bw(indep_scaling), **kwargs)


def get_logexp(a=1, b=0, a2=None, b2=None, backend=None):
    """ (forward, backward) callables for ``u = ln(a*x + b)`` (``pyodesys/symbolic.py:1361-1390``); a small shift
    ``b`` (forward only, ``b2 = 0``) keeps the transformation defined at x = 0. """
    a2 = a if a2 is None else a2
    b2 = b if b2 is None else b2
    be = get_backend(backend)
    return (lambda x: be.log(a*x + b), lambda x: (be.exp(x) - b2)/a2)


class PartiallySolvedSystem(SymbolicSys):
    """ An ODE system in which some dependent variables are replaced by analytic expressions
    (``pyodesys/symbolic.py:1134-1358``).

    ``analytic_factory(x0, y0, p0[, backend])`` returns ``{dep_symbol: expression}`` in terms of the initial values;
    those variables leave the system, the rest is integrated with the expressions substituted, and the full state
    is rebuilt afterwards.  The initial values reach the integrated system as ordinary extra parameters
    ``[..., x0, y0_0 .. y0_{ny-1}]`` (the reference does the same, ``symbolic.py:1240-1249``).  Both processors
    are single vectorised evaluations over the whole batch.
    """

    def __init__(self, original_system, analytic_factory, **kwargs):
        ori = original_system
        be = ori.be
        if 'pre_processors' in kwargs or 'post_processors' in kwargs:
            raise NotImplementedError("Cannot override pre-/postprocessors")
        self._ori_sys = ori
        self.original_dep = tuple(ori.dep)
        # own symbols for the initial values: a native subclass adds its ``i_<dep>`` symbols for the REDUCED system
        x0 = be.Symbol('ps0_' + (str(ori.indep) if ori.indep is not None else 'indep'))
        y0 = tuple(be.Symbol('ps0_' + str(d)) for d in ori.dep)
        pars = dict(zip(ori.param_names, ori.params)) if ori.par_by_name else ori.params
        dep0 = dict(zip(ori.names, y0)) if ori.dep_by_name else y0
        self.analytic_exprs = dict(_ensure_4args(analytic_factory)(x0, dep0, pars, be))
        if len(self.analytic_exprs) == 0:
            raise ValueError("Failed to produce any analytic expressions.")
        for d in self.analytic_exprs:
            if d not in self.original_dep:
                raise ValueError("%s is not a dependent variable of the original system" % d)
        self._analytic_idx = [self.original_dep.index(d) for d in self.analytic_exprs]   # in factory order
        self._keep_idx = [i for i in range(len(self.original_dep)) if i not in self._analytic_idx]
        new_dep = [self.original_dep[i] for i in self._keep_idx]
        new_exprs = [ori.exprs[i].subs(self.analytic_exprs) for i in self._keep_idx]
        self._n_ori_params = len(ori.params)
        new_pars = tuple(ori.params) + (x0,) + y0
        kw = dict(kwargs)
        for attr in ('dep_by_name', 'par_by_name', 'description', 'to_arrays_callbacks'):
            if attr not in kw and getattr(ori, attr, None) is not None:
                kw[attr] = getattr(ori, attr)
        if ori.names:
            self.free_names = [ori.names[i] for i in self._keep_idx]
            kw.setdefault('names', self.free_names)
        else:
            self.free_names = None
        if ori.param_names:
            kw.setdefault('param_names', list(ori.param_names) + [str(x0)] + [str(s) for s in y0])
        for b in ('lower_bounds', 'upper_bounds'):
            if b not in kw and getattr(ori, b, None) is not None:
                kw[b] = np.asarray(getattr(ori, b))[self._keep_idx]
        if kw.get('linear_invariants') is None and ori.linear_invariants is not None:
            # only meaningful when they do not involve the eliminated variables
            rows = [r for r in ori.linear_invariants.tolist() if all(r[i] == 0 for i in self._analytic_idx)]
            kw['linear_invariants'] = [[r[i] for i in self._keep_idx] for r in rows] or None
        roots = kw.pop('roots', ori.roots)
        if roots is not None:
            roots = [r.subs(self.analytic_exprs) for r in roots]
        args = ((ori.indep,) if ori.indep is not None else (be.Dummy('x'),)) + tuple(new_dep) + new_pars
        self._analytic_cb = be.Lambdify(args, list(self.analytic_exprs.values()))
        self._n_new_params = len(new_pars)
        super().__init__(zip(new_dep, new_exprs), ori.indep, new_pars, backend=be, roots=roots,
                         pre_processors=list(ori.pre_processors) + [self._reduce_xyp],
                         post_processors=[self._rebuild_xyp] + list(ori.post_processors), **kw)

    def _reduce_xyp(self, x, y, p):
        """ (x[..., nt], y0[..., ny_ori], p[..., np]) -> (x, y0 of the kept variables, [p, x0, y0]) """
        x, y, p = (np.asarray(a, dtype=np.float64) for a in (x, y, p))
        lead = y.shape[:-1]
        x0 = np.broadcast_to(x[..., :1], lead + (1,))
        p = np.broadcast_to(p, lead + p.shape[-1:])
        return x, y[..., self._keep_idx], np.concatenate((p, x0, y), axis=-1)

    def _rebuild_xyp(self, x, y, p):
        """ (x[..., nt], y[..., nt, ny_kept], p[..., >= np + 1 + ny_ori]) -> full state, original parameters """
        if isinstance(x, (list, tuple)) or (hasattr(x, '__getitem__') and not hasattr(x, 'ndim')):
            out = [self._rebuild_xyp(x[i], y[i], p[i]) for i in range(len(x))]
            return [o[0] for o in out], [o[1] for o in out], np.asarray(p)[..., :self._n_ori_params]
        x, y, p = (np.asarray(a, dtype=np.float64) for a in (x, y, p))
        lead = y.shape[:-1]
        pp = p[..., :self._n_new_params]
        pp = np.broadcast_to(pp.reshape(pp.shape[:-1] + (1,)*(len(lead) - (pp.ndim - 1)) + pp.shape[-1:]),
                             lead + (self._n_new_params,))
        rows = np.concatenate((np.broadcast_to(x, lead)[..., None], y, pp), axis=-1)
        analytic = self._analytic_cb(rows)
        full = np.empty(lead + (len(self.original_dep),))
        full[..., self._keep_idx] = y
        full[..., self._analytic_idx] = analytic
        return x, full, p[..., :self._n_ori_params]

    @classmethod
    def from_linear_invariants(cls, ori_sys, preferred=None, **kwargs):
        """ Eliminate one variable per linear invariant (``pyodesys/symbolic.py:1275-1336``): variable ``tgt`` of
        invariant row ``r`` becomes ``y0_tgt - sum_{c != tgt} A[r,c] (y_c - y0_c) / A[r,tgt]``. """
        A = ori_sys.be.Matrix(ori_sys.linear_invariants)
        if len(A.rref()[1]) < A.rows:
            raise NotImplementedError("Linear invariants contain linear dependencies.")
        cols = [(r, [c for c in range(A.cols) if A[r, c] != 0]) for r in range(A.rows)]
        if preferred is None:
            preferred = list(ori_sys.names[:A.rows]) if ori_sys.dep_by_name else list(range(A.rows))
        targets = [ori_sys.names.index(d) if ori_sys.dep_by_name else (d if isinstance(d, int) else ori_sys.dep.index(d))
                   for d in preferred]
        chosen = []
        for r, cs in sorted(cols, key=lambda rc: len(rc[1])):
            hit = next((t for t in targets if t in cs), None)
            if hit is not None:
                chosen.append((r, hit))
                targets.remove(hit)
            if not targets:
                break
        if targets:
            raise ValueError("Could not find a solutions for: %s" % targets)

        def analytic_factory(x0, y0, p0, be):
            iv = (lambda c: y0[ori_sys.names[c]]) if ori_sys.dep_by_name else (lambda c: y0[c])
            return {ori_sys.dep[t]: iv(t) - sum(A[r, c]*(ori_sys.dep[c] - iv(c)) for c in range(A.cols) if c != t)/A[r, t]
                    for r, t in chosen}
        used_rows, used_cols = zip(*chosen)
        rest = [[cell for c, cell in enumerate(row) if c not in used_cols]
                for r, row in enumerate(A.tolist()) if r not in used_rows]
        return cls(ori_sys, analytic_factory, linear_invariants=rest or None, **kwargs)

    def __getitem__(self, key):
        d = self.original_dep[self._ori_sys.names.index(key)]
        return self.analytic_exprs.get(d, d)

    def integrate(self, *args, **kwargs):
        """ per-component ``atol`` is given for the ORIGINAL variables (``symbolic.py:1345-1358``) """
        atol = kwargs.get('atol', None)
        if isinstance(atol, dict):
            kwargs['atol'] = [atol[k] for k in self.free_names]
        elif atol is not None and np.ndim(atol) == 1 and len(atol) == len(self.original_dep):
            kwargs['atol'] = [atol[i] for i in self._keep_idx]
        return super().integrate(*args, **kwargs)
