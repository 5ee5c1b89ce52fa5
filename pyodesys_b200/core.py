# -*- coding: utf-8 -*-
"""
Host-side driver for the batched native path: argument shaping, dispatch and results.

Mirrors the slice of ``pyodesys.core.ODESys`` that the batched native ``integrate`` call
runs through (``pyodesys/core.py:240-290`` ``to_arrays``/``pre_process``,
``core.py:400-449`` ``integrate``) and ``pyodesys.results.Result``
(``pyodesys/results.py:13-57``).  Everything below the ``_integrate_native`` dispatch is
CUDA (see :mod:`pyodesys_b200.native.cuda`).

The only deliberate extension is :class:`BatchResult`: the reference builds ``N`` ``Result``
objects and ``N`` info dicts in a Python loop (``core.py:436-441``, ``native/_base.py:331-340``)
which at N=1e6 costs orders of magnitude more than the device work; ``BatchResult`` keeps
the arrays and materialises a ``Result`` on ``__getitem__``.
"""
from __future__ import annotations

import numpy as np


class Result:
    """ Same attribute and tuple-unpacking layout as ``pyodesys/results.py:13-40``. """

    def __init__(self, xout, yout, params, info, odesys):
        self.xout = xout
        self.yout = yout
        self.params = params
        self.info = info
        self.odesys = odesys
        self.names = getattr(odesys, 'names', ())

    def copy(self):
        return Result(self.xout.copy(), self.yout.copy(), self.params.copy(),
                      self.info.copy(), self.odesys)

    def __len__(self):
        return 3

    def __getitem__(self, key):
        if key == 0:
            return self.xout
        elif key == 1:
            return self.yout
        elif key == 2:
            return self.info
        elif key == 3:
            raise StopIteration
        raise KeyError("Invalid key: %s (for backward compatibility reasons)." % str(key))

    def named_param(self, param_name):
        return self.params[self.odesys.param_names.index(param_name)]

    def named_dep(self, name):
        return self.yout[..., self.odesys.names.index(name)]

    def between(self, lower, upper, xdata=None, ydata=None):
        xdata = self.xout if xdata is None else xdata
        ydata = self.yout if ydata is None else ydata
        sel = (xdata > lower) & (xdata < upper)
        return xdata[..., sel], ydata[..., sel, :]

    def calc_invariant_violations(self):
        """ invariants along the trajectory minus their value at the first point (``pyodesys/results.py:265-268``),
        evaluated on the internal variables the system was integrated in """
        cb = self.odesys.get_invariants_callback()
        x, y, p = (np.asarray(self.info['internal_' + k]) for k in ('xout', 'yout', 'params'))
        val = cb(x, y, p)
        return val - val[0, :]

    def extend_by_integration(self, xend, params=None, odesys=None, autonomous=None, npoints=1, **kwargs):
        """ continue from the last point to ``xend`` (``pyodesys/results.py:277-314``): rows are appended, counters
        summed, ``success`` and-ed, ``*_indices`` shifted, ``internal_*`` concatenated when the system is the same """
        odesys = odesys or self.odesys
        if autonomous is None:
            autonomous = bool(getattr(odesys, 'autonomous_interface', False) == True)  # noqa: E712
        x0 = self.xout[-1]
        nx0 = self.xout.size
        span = np.linspace(0.0, xend - x0, npoints + 1) if autonomous else np.linspace(x0, xend, npoints + 1)
        res = odesys.integrate(span, self.yout[..., -1, :], self.params if params is None else params, **kwargs)
        self._append_leg(res, x0 if autonomous else 0.0, nx0, odesys is self.odesys)
        return self

    def _append_leg(self, res, x_shift, nx0, same_system):
        self.xout = np.concatenate((self.xout, np.asarray(res.xout)[1:] + x_shift))
        self.yout = np.concatenate((self.yout, np.asarray(res.yout)[..., 1:, :]))
        info = {k: v for k, v in self.info.items() if not (k.startswith('internal') and not same_system)}
        for k, v in res.info.items():
            if k.startswith('internal'):
                if same_system and k in info and k != 'internal_params':
                    tail = np.asarray(v)[1:]
                    info[k] = np.concatenate((np.asarray(info[k]), tail + x_shift if k == 'internal_xout' else tail))
            elif k == 'success':
                info[k] = bool(info.get(k, True)) and bool(v)
            elif k.endswith('_indices'):
                info[k] = list(info.get(k, [])) + [itm + nx0 - 1 for itm in v]
            elif k in ('mode', 'method', 'status', 'nreached', 'root_out'):
                info[k] = v
            elif isinstance(v, str):
                info[k] = (info[k] if isinstance(info.get(k), list) else [info.get(k)]) + [v]
            elif k in info:
                info[k] = info[k] + v
            else:
                info[k] = v
        self.info = info


class BatchInfo:
    """ Array-backed per-system info: ``batch_info[i]`` builds the reference's dict.

    Keys per system follow ``_odeint_wrapper.pyx:25-34`` (``mode``, ``success``,
    ``nreached`` + the integer/double ``Info`` maps) and ``native/_base.py:331-340``
    (``internal_xout``, ``internal_yout``, ``internal_params``, ``nfev``/``njev`` aliases).
    """

    counter_names = ('n_steps', 'nfev', 'njev', 'nlu', 'nsolve', 'nreject')

    def __init__(self, mode, status, nreached, counters, time_wall, time_cpu, xout, yout, params,
                 return_on_error, nt=None, timings=None, method=None, iv=None):
        self.mode = mode
        self.status = status
        self.nreached = nreached
        self.counters = counters
        self.time_wall = time_wall
        self.time_cpu = time_cpu
        self.xout = xout
        self.yout = yout
        self._params = params
        self._iv = iv  # initial values still to be appended to the parameter rows (lazy append_iv)
        self.return_on_error = return_on_error
        self.nt = nt
        self.timings = timings or {}
        self.method = method
        self.roots = None  # event-function outputs (root_out / root_indices / nroots_found arrays), if any
        self.krylov = False  # matrix-free Newton iteration: the factorisation counter holds the J v products

    @property
    def params(self):
        """ internal parameter rows ``[p..., y0...]`` as the reference hands them to native code
        (``core.py:410-411``); concatenated on first access only """
        if self._iv is not None:
            self._params = np.concatenate((self._params, self._iv), axis=-1)
            self._iv = None
        return self._params

    def __len__(self):
        return len(self.status)

    def success(self):
        return self.status == 0

    def __getitem__(self, i):
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        c = self.counters[:, i]  # SoA [ncounters, N] as written by the kernel
        d = {k: int(c[j]) for j, k in enumerate(self.counter_names)}
        d['n_rhs_evals'] = d['nfev']
        d['dense_n_dls_jac_evals'] = d['njev']
        d['n_lu'] = d.pop('nlu')
        if self.krylov:  # cvode's name for the products (_cvode_wrapper.pyx:59-75; asserted by test_cvode.py:215)
            d['njvev'], d['n_lu'] = d['n_lu'], 0
        d['n_solve'] = d.pop('nsolve')
        d['n_err_test_fails'] = d.pop('nreject')
        d['time_wall'] = self.time_wall
        d['time_cpu'] = self.time_cpu
        d['mode'] = self.mode
        d['status'] = int(self.status[i])
        d['success'] = bool(self.status[i] == 0)
        if self.mode == 'predefined':
            d['nreached'] = int(self.nreached[i])
        d['internal_xout'] = self.xout[i]
        d['internal_yout'] = self.yout[i]
        d['internal_params'] = self.params[i, ...]
        if self.method is not None:
            d['method'] = self.method
        if self.roots:
            from .native.cuda import root_info
            d.update(root_info(self.roots, i, self.mode))
        return d

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]


class BatchResult:
    """ Lazy list of :class:`Result` (``len``, indexing, slicing and iteration like a list).

    ``xout``/``yout``/``params`` are the *internal* arrays; post-processors run when an
    element is materialised, exactly where ``core.py:439-441`` runs them.
    """

    def __init__(self, info, odesys):
        self.info = info
        self.odesys = odesys
        self._cache = {}

    @property
    def xout(self):
        return self.info.xout

    @property
    def yout(self):
        return self.info.yout

    @property
    def params(self):
        return self.info.params

    def __len__(self):
        return len(self.info)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        if i not in self._cache:
            nfo = self.info[i]
            x, y, p = self.odesys.post_process(nfo['internal_xout'], nfo['internal_yout'],
                                               nfo['internal_params'])
            r = Result(x, y, p, nfo, self.odesys)
            if len(self._cache) > 4096:
                self._cache.clear()
            self._cache[i] = r
        return self._cache[i]

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    def processed(self):
        """ ``(xout, yout, params)`` of the WHOLE batch after the system's post-processors (one vectorised call,
        e.g. the back-transformation of a :class:`~pyodesys_b200.transforms.TransformedSys`) instead of one
        ``Result`` at a time. """
        return self.odesys.post_process(self.info.xout, self.info.yout, self.info.params)


class ODESys:
    """ Argument shaping + dispatch of the reference's ``ODESys`` for the native path. """

    append_iv = False
    numpy = np

    def __init__(self, names=(), param_names=(), dep_by_name=False, par_by_name=False,
                 pre_processors=None, post_processors=None, append_iv=False,
                 to_arrays_callbacks=None, indep_name=None, description=None, **kwargs):
        if kwargs:
            raise TypeError("Unsupported keyword arguments: %s" % ', '.join(sorted(kwargs)))
        self.names = tuple(names or ())
        self.param_names = tuple(param_names or ())
        self.dep_by_name = dep_by_name
        self.par_by_name = par_by_name
        self.pre_processors = list(pre_processors or [])
        self.post_processors = list(post_processors or [])
        self.append_iv = append_iv
        self.to_arrays_callbacks = to_arrays_callbacks
        self.indep_name = indep_name
        self.description = description
        self.taken_names = ()

    # -- ``core.py:209-238`` ------------------------------------------------------------
    @staticmethod
    def _array_from_dict(d, keys):
        vals = [d[k] for k in keys]
        lens = [len(v) if hasattr(v, '__len__') else 0 for v in vals]
        n = max(lens) if lens else 0
        if n == 0:
            return np.array(vals, dtype=np.float64)
        if any(ln not in (0, n) for ln in lens):
            raise ValueError("Mixed lengths in dictionary.")
        out = np.empty((n, len(keys)), dtype=np.float64)
        for idx, v in enumerate(vals):
            out[:, idx] = v
        return out

    def _conditional_from_dict(self, cont, by_name, names):
        if isinstance(cont, dict):
            if not by_name:
                raise ValueError("not by name, yet a dictionary was passed.")
            cont = self._array_from_dict(cont, names)
        return cont

    # -- ``core.py:240-284`` ------------------------------------------------------------
    def to_arrays(self, x, y, p, callbacks=None, reshape=True):
        try:
            nx = len(x)
        except TypeError:
            _x = 0*x, x
        else:
            # a lone end point means (0, xend) (core.py:247-252); a single ROW of points does not
            _x = (0*x[0], x[0]) if nx == 1 and np.ndim(x[0]) == 0 else x
        _y = self._conditional_from_dict(y, self.dep_by_name, self.names)
        _p = self._conditional_from_dict(p, self.par_by_name, self.param_names)
        callbacks = callbacks or self.to_arrays_callbacks
        if callbacks is not None:
            if len(callbacks) != 3:
                raise ValueError("Need 3 callbacks/None values.")
            _x, _y, _p = [e if cb is None else cb(e) for cb, e in zip(callbacks, [_x, _y, _p])]
        arrs = [np.atleast_1d(np.asarray(a, dtype=np.float64)) for a in (_x, _y, _p)]
        if reshape:
            extra = None
            for a in arrs:
                if a.ndim == 1:
                    continue
                elif a.ndim == 2:
                    if extra is None:
                        extra = a.shape[0]
                    elif extra != a.shape[0]:
                        raise ValueError("Size mismatch!")
                else:
                    raise NotImplementedError("Only 2 dimensions currently supported.")
            if extra is not None:
                # broadcast views instead of the reference's numpy.tile (core.py:282-283):
                # no O(N) copy until the native shim needs contiguous memory.
                arrs = [a if a.ndim == 2 else np.broadcast_to(a, (extra, a.shape[0])) for a in arrs]
        return arrs

    def pre_process(self, xout, y0, params=()):
        for pre in self.pre_processors:
            xout, y0, params = pre(xout, y0, params)
        return [np.atleast_1d(a) for a in (xout, y0, params)]

    def post_process(self, xout, yout, params):
        for post in self.post_processors:
            xout, yout, params = post(xout, yout, params)
        return xout, yout, params

    def adaptive(self, y0, x0, xend, params=(), **kwargs):
        """ ``core.py:298-321`` """
        return self.integrate((x0, xend), y0, params=params, **kwargs)

    def predefined(self, y0, xout, params=(), **kwargs):
        """ ``core.py:323-345`` """
        xout, yout, info = self.integrate(xout, y0, params=params, force_predefined=True, **kwargs)
        return yout, info

    # -- ``core.py:347-449`` ------------------------------------------------------------
    def integrate(self, x, y0, params=(), atol=1e-8, rtol=1e-8, **kwargs):
        arrs = self.to_arrays(x, y0, params)
        _x, _y, _p = _arrs = self.pre_process(*arrs)
        ndims = [a.ndim for a in _arrs]
        if ndims == [1, 1, 1]:
            twodim = False
        elif ndims == [2, 2, 2]:
            twodim = True
        else:
            raise ValueError("Pre-processor made ndims inconsistent?")

        if self.append_iv:
            if twodim and getattr(self, '_append_iv_on_device', False):
                # the CUDA kernels read the appended initial values from y0 themselves (options.iv_from_y0)
                kwargs['iv_from_y0'] = True
            else:
                _p = np.concatenate((_p, _y), axis=-1)

        if hasattr(self, 'ny') and _y.shape[-1] != self.ny:
            raise ValueError("Incorrect shape of intern_y0")
        if isinstance(atol, dict):
            kwargs['atol'] = [atol[k] for k in self.names]
        else:
            kwargs['atol'] = atol
        kwargs['rtol'] = rtol

        integrator = kwargs.pop('integrator', None)
        if integrator is None:
            integrator = 'native'
        if not isinstance(integrator, str):
            raise ValueError("Only named integrators are supported by this back end")
        args = tuple(map(np.atleast_2d, (_x, _y, _p)))
        self._current_integration_kwargs = kwargs
        binfo = getattr(self, '_integrate_' + integrator)(*args, **kwargs)
        if twodim:
            return BatchResult(binfo, self)
        nfo = binfo[0]
        self._internal = (np.array(nfo['internal_xout']), np.array(nfo['internal_yout']), _p.copy())
        return Result(*(tuple(self.post_process(nfo['internal_xout'], nfo['internal_yout'], _p))
                        + (nfo, self)))


def integrate_auto_switch(odes, kw, x, y0, params=(), **kwargs):
    """ Switch between formulations of one ODE system (``pyodesys/core.py:824-915``).

    ``odes[0]`` integrates first with its entry of every iterable in ``kw`` (typically a small ``nsteps`` and
    ``return_on_error=True``); systems that did not reach the end are restarted from where they stopped with
    ``odes[1]`` and so on.  Adaptive mode only (``x`` = (x0, xend) per system), like the reference.  The
    restart is one batched call per formulation over the UNFINISHED systems only.  Returns one :class:`Result` (or a list for a batch) whose ``info`` sums ``nfev, njev, time_cpu,
    time_wall`` and carries ``nsys`` = number of formulations used.
    """
    x_arr = np.asarray(x, dtype=np.float64)
    if x_arr.shape[-1] > 2:
        raise NotImplementedError("Only adaptive support return_on_error for now")
    batch = x_arr.ndim == 2
    X = np.atleast_2d(x_arr).copy()
    n = X.shape[0]
    Y0 = np.array(np.broadcast_to(np.asarray(y0, dtype=np.float64), (n,) + np.shape(y0)[-1:]))
    sum_keys = ('nfev', 'njev', 'time_cpu', 'time_wall')
    autonomous = lambda o: bool(getattr(o, 'autonomous_interface', False) == True)  # noqa: E712,E731
    nxt_aut = autonomous(odes[0])
    offset = X[:, 0].copy() if nxt_aut else np.zeros(n)      # what has to be added to xout of an autonomous leg
    xs = [[np.array([0.0 if nxt_aut else X[i, 0]])] for i in range(n)]
    ys = [[Y0[i:i + 1].copy()] for i in range(n)]
    tot = [dict.fromkeys(sum_keys, 0) for _ in range(n)]
    last_params = None
    used = 0
    # only what has not finished yet is handed to the next formulation (``pyodesys/core.py:865-908`` restarts what
    # did not finish; a batch that mostly succeeds in its first formulation costs one small second call)
    P = params
    per_system_params = False
    try:
        parr = np.asarray(params, dtype=np.float64)
        if parr.ndim == 2 and parr.shape[0] == n and n > 1:
            P, per_system_params = parr, True
    except (TypeError, ValueError):
        pass
    todo = np.arange(n)
    last_params = [None]*n
    nsys_used = np.zeros(n, dtype=int)
    for oi, ode in enumerate(odes):
        used = oi + 1
        if oi < len(odes) - 1:
            nxt_aut = autonomous(odes[oi + 1])
        call_kw = dict(kwargs)
        call_kw.update({k: v[oi] for k, v in kw.items()})
        p_arg = P[todo] if per_system_params else P
        res = ode.integrate(X[todo] if batch else X[0], Y0[todo] if batch else Y0[0], p_arg, **call_kw)
        legs = res if batch else [res]
        unfinished = []
        for j, r in enumerate(legs):
            i = int(todo[j])
            xs[i].append(np.asarray(r.xout)[1:] + offset[i])
            ys[i].append(np.asarray(r.yout)[1:])
            for k in sum_keys:
                if k in r.info:
                    tot[i][k] += r.info[k]
            tot[i]['success'] = r.info['success']
            last_params[i] = r.params
            nsys_used[i] = used
            if not r.info['success']:
                unfinished.append(i)
                xlast = np.asarray(r.xout)[-1]
                if nxt_aut:
                    X[i] = (0.0, abs(X[i, -1] - xlast))
                    offset[i] += xlast
                else:
                    X[i, 0] = xlast
                Y0[i] = np.asarray(r.yout)[-1]
        if not unfinished or oi == len(odes) - 1:
            break
        todo = np.array(unfinished)
    out = [Result(np.concatenate(xs[i]), np.concatenate(ys[i], axis=0), last_params[i], dict(nsys=used, **tot[i]),
                  odes[0]) for i in range(n)]
    return out if batch else out[0]


integrate_chained = integrate_auto_switch  # the reference keeps the old name (core.py:918)


def chained_parameter_variation(subject, durations, y0, varied_params, default_params=None,
                                integrate_kwargs=None, x0=None, npoints=1, numpy=None):
    """ Integrate over consecutive durations with some parameters changed in between
    (``pyodesys/core.py:921-1000``).

    ``varied_params`` maps a parameter name (``par_by_name``) or index to one value per duration - a scalar, or
    for a batch (``y0`` two-dimensional) optionally one value per system.  The reference extends every
    ``Result`` of a batch separately (one native call per system and duration); here every duration is ONE
    batched call over all systems, started from the final states of the previous one.  ``result.params`` keeps
    the parameters of the first duration, as in the reference.
    """
    if len(durations) == 0:
        raise ValueError("need at least 1 duration (preferably many)")
    if npoints <= 0:
        raise ValueError("need at least 1 point per duration")
    for k, v in varied_params.items():
        if len(v) != len(durations):
            raise ValueError("Mismathced lengths of durations and varied_params")
    integrate = subject.integrate if isinstance(subject, ODESys) else subject
    default_params = {} if default_params is None else default_params
    integrate_kwargs = integrate_kwargs or {}
    ends = np.cumsum(durations)
    if x0 is None:
        x0 = ends[0]*0
    out = None
    for idx, xend in enumerate(ends):
        params = dict(default_params) if isinstance(default_params, dict) else list(default_params)
        for k, v in varied_params.items():
            params[k] = v[idx]
        if out is None:
            out = integrate(np.linspace(x0, xend, npoints + 1), y0, params, **integrate_kwargs)
            if not isinstance(out, Result):
                out = [out[i] for i in range(len(out))]  # materialise: the legs are appended in place
            continue
        if isinstance(out, Result):
            out.extend_by_integration(xend, params, npoints=npoints, **integrate_kwargs)
            continue
        # batch: one call for all systems
        odesys = out[0].odesys
        autonomous = bool(getattr(odesys, 'autonomous_interface', False) == True)  # noqa: E712
        xlast = np.array([r.xout[-1] for r in out])
        ylast = np.array([r.yout[-1] for r in out])
        if autonomous:
            span = np.array([np.linspace(0.0, xend - xl, npoints + 1) for xl in xlast])
        else:
            span = np.array([np.linspace(xl, xend, npoints + 1) for xl in xlast])
        legs = odesys.integrate(span, ylast, params, **integrate_kwargs)
        for i, r in enumerate(out):
            r._append_leg(legs[i], xlast[i] if autonomous else 0.0, r.xout.size, True)
    return out
