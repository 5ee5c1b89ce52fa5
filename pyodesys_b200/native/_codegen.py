# -*- coding: utf-8 -*-
"""
Code generation: SymPy expressions -> one self-contained C++/CUDA header.

This is the CUDA counterpart of ``_NativeCodeBase.variables`` (``pyodesys/native/_base.py:126-281``)
plus the mako template it feeds (``pyodesys/native/sources/odesys_anyode_template.cpp``):

* symbols are replaced by C identifiers ``y[i]``, ``x``, ``m_p[i]`` (``_base.py:157-164``);
* CSE level 1 ("common"): sub-expressions that only involve parameters are hoisted into
  ``m_p_cse[k]`` and evaluated once per system (``_base.py:181-195``, ``template:104-110``);
* CSE level 2 per function group: rhs, jac(+dfdt), first_step, roots
  (``_base.py:197-226``);
* ``namespace_override`` lets the user replace a generated body with raw source
  (``_base.py:275-279``) - the same member names (``m_p``, ``this->nfev``,
  ``AnyODE::Status::recoverable_error``) are valid inside the generated ``struct OdeSys`` so
  the reference's snippets compile unchanged.

The emitted text is plain C++ guarded by ``ODESYS_FN`` (``__host__ __device__ __forceinline__``
under nvcc, ``inline`` under g++) so that the very same functions are used by the sm_100a
kernels and by the CPU oracle (``oracle/``).  Operation counts of the CSE'd forms are
emitted as ``ODESYS_FLOPS_*`` and feed the FP64 roofline.
"""
from __future__ import annotations

import hashlib
import os

import sympy
from sympy.printing.c import C99CodePrinter


class _Printer(C99CodePrinter):
    """ C printer that never emits ``pow`` for small integer exponents. """

    def _print_Pow(self, expr):
        base, exp = expr.base, expr.exp
        if exp.is_Integer and 1 < abs(int(exp)) <= 16:
            b = self._print(base)
            n = abs(int(exp))
            if n == 2:
                s = '((%s)*(%s))' % (b, b)
            else:
                s = 'odesys_ipow<%d>(%s)' % (n, b)
            return s if int(exp) > 0 else '(1.0/%s)' % s
        if exp == -1:
            return '(1.0/(%s))' % self._print(base)
        return super()._print_Pow(expr)

    def _print_Symbol(self, expr):
        return expr.name  # names such as ``y[0]``/``m_p[2]`` must survive verbatim

    def _print_Float(self, expr):
        return repr(float(expr))

    def _print_Integer(self, expr):
        return '%d.0' % int(expr) if abs(int(expr)) < 2**53 else repr(float(int(expr)))

    def _print_Rational(self, expr):
        return '(%d.0/%d.0)' % (expr.p, expr.q)


_printer = _Printer({'contract': False})


def _ccode(expr):
    return _printer.doprint(sympy.sympify(expr))


def _count(exprs_and_cses):
    tot = 0
    for e in exprs_and_cses:
        tot += int(sympy.count_ops(e, visual=False))
    return tot


class GeneratedCode:
    """ Result of :func:`generate`: header text + the compile-time facts about the system. """

    def __init__(self, text, ny, nparams, np_tot, np_used, npcse, flops, has_jac, autonomous,
                 has_first_step, has_dx_max, nroots, jac_nnz, nnz=-1, has_jtimes=False):
        self.text = text
        self.ny = ny
        self.nparams = nparams
        self.np_tot = np_tot
        self.np_used = np_used
        self.npcse = npcse
        self.flops = flops
        self.has_jac = has_jac
        self.autonomous = autonomous
        self.has_first_step = has_first_step
        self.has_dx_max = has_dx_max
        self.nroots = nroots
        self.jac_nnz = jac_nnz
        self.nnz = nnz                # stored CSC entries (sparse=True) or -1
        self.has_jtimes = has_jtimes

    @property
    def digest(self):
        return hashlib.sha256(self.text.encode('utf-8')).hexdigest()[:20]


def _body(cses, assigns, ret='return AnyODE::Status::success;', indent='        '):
    lines = []
    for tok, expr in cses:
        lines.append('%sconst double %s = %s;' % (indent, tok, expr))
    for lhs, expr in assigns:
        lines.append('%s%s = %s;' % (indent, lhs, expr))
    if ret:
        lines.append(indent + ret)
    return '\n'.join(lines)


def _lazy_lines(cses, outs, cc, indent='        '):
    """ Statements for ``outs`` = [(format-with-%s, expr)] where every CSE temporary is defined right before
    the first output that needs it: short live ranges instead of sympy's "all temporaries first" order
    (101 temporaries up front cost ~200 registers for the radiolysis Jacobian). """
    defs = {sy: e for sy, e in cses}
    order = [sy for sy, _e in cses]
    pos = {sy: k for k, sy in enumerate(order)}
    done = set()
    lines = []

    def need(expr):
        todo = sorted((sy for sy in expr.free_symbols if sy in defs and sy not in done), key=pos.get)
        for sy in todo:
            if sy in done:
                continue
            need(defs[sy])
            done.add(sy)
            lines.append('%sconst double %s = %s;' % (indent, sy.name, cc(defs[sy])))

    for fmt, expr in outs:
        expr = sympy.sympify(expr)
        need(expr)
        lines.append(indent + fmt % cc(expr))
    return lines


def _accumulate_lines(cses, exprs, cc, prefix='acc', indent='        '):
    """ Sum-of-rates form: every output is an accumulator; a temporary that enters outputs as a plain
    additive term (numeric coefficient times the temporary - a reaction rate in mass-action kinetics) is
    added to them right after its definition and can die immediately.  Returns (lines, accumulator names). """
    temps = [sy for sy, _e in cses]
    tempset = set(temps)
    uses = {sy: [] for sy in temps}
    leftovers = []
    for i, expr in enumerate(exprs):
        expr = sympy.sympify(expr)
        rest_terms = []
        for term in sympy.Add.make_args(expr):
            c, rest = term.as_coeff_Mul()
            if rest in tempset and c.is_number:
                uses[rest].append((i, c))
            else:
                rest_terms.append(term)
        leftovers.append(rest_terms)
    lines = []
    started = [False]*len(exprs)

    def add(i, text, negative=False):
        # the first contribution declares the accumulator (no ``0.0 +`` left for the FP64 pipe)
        if not started[i]:
            started[i] = True
            lines.append('%sdouble %s%d = %s%s;' % (indent, prefix, i, '-' if negative else '', text))
        else:
            lines.append('%s%s%d %s= %s;' % (indent, prefix, i, '-' if negative else '+', text))

    for sy, e in cses:
        lines.append('%sconst double %s = %s;' % (indent, sy.name, cc(e)))
        for i, c in uses[sy]:
            if c == 1:
                add(i, sy.name)
            elif c == -1:
                add(i, sy.name, negative=True)
            else:
                add(i, '(%s)' % cc(c*sy))
    for i, terms in enumerate(leftovers):
        if terms:
            add(i, '(%s)' % cc(sympy.Add(*terms)))
        elif not started[i]:
            add(i, '0.0')
    return lines, ['%s%d' % (prefix, i) for i in range(len(exprs))]



EVAL_PARTS = 4       # the block-cooperative kernel has four warps: each evaluates one part
EVAL_PARTS_MIN_NY = 9


def _balanced_parts(nitems, part_cost, weights, nparts=EVAL_PARTS, sweeps=2):
    """ Split ``range(nitems)`` into ``nparts`` groups that minimise the largest ``part_cost(group)`` (the operation
    count of the group AFTER its own common-subexpression elimination: rows of a kinetics right-hand side share
    their reaction rates, so the cost of a group is far from additive).  Greedy start by weight, then single-item
    moves away from the most expensive group while that lowers the maximum. """
    order = sorted(range(nitems), key=lambda i: -weights[i])
    parts = [[] for _ in range(nparts)]
    load = [0]*nparts
    for i in order:
        q = min(range(nparts), key=lambda k: load[k])
        parts[q].append(i)
        load[q] += weights[i]
    cost = [part_cost(sorted(pt)) if pt else 0 for pt in parts]
    for _ in range(sweeps):
        improved = False
        worst = max(range(nparts), key=lambda k: cost[k])
        for i in list(parts[worst]):
            if len(parts[worst]) <= 1:
                break
            c_without = part_cost(sorted(j for j in parts[worst] if j != i))
            best = None
            for q in range(nparts):
                if q == worst:
                    continue
                c_with = part_cost(sorted(parts[q] + [i]))
                if max(c_with, c_without) < cost[worst] and (best is None or c_with < best[1]):
                    best = (q, c_with)
            if best is not None:
                parts[worst].remove(i)
                parts[best[0]].append(i)
                cost[worst], cost[best[0]] = c_without, best[1]
                improved = True
                break
        if not improved:
            break
    return [sorted(pt) for pt in parts], cost


def generate(odesys, namespace_override=None, namespace_extend=None):
    """ Build the header for ``odesys`` (a :class:`pyodesys_b200.symbolic.SymbolicSys`). """
    ns_over = dict(namespace_override or {})
    ns_ext = dict(namespace_extend or {})
    be = odesys.be
    ny = odesys.ny
    if odesys.band is not None:
        raise NotImplementedError("Banded jacobian not yet implemented.")  # _base.py:128-129
    nparams = len(odesys.params)
    np_tot = nparams + (ny if odesys.append_iv else 0)

    jac = odesys.get_dense_jac() if hasattr(odesys, 'get_dense_jac') else odesys.get_jac()
    all_exprs = list(odesys.exprs)
    if jac is not False:
        dfdx = odesys.get_dfdx()
        jac_flat = [jac[i, j] for i in range(ny) for j in range(ny)]
        dfdx_flat = [dfdx[i] for i in range(ny)] if dfdx is not False else [sympy.S.Zero]*ny
        all_exprs += jac_flat + dfdx_flat
        nj = len(jac_flat) + len(dfdx_flat)
    else:
        nj = 0
    first_step = odesys.first_step_expr
    nfs = 0
    if first_step is not None and 'p_first_step' not in ns_over:
        all_exprs.append(be.sympify(first_step))
        nfs = 1
    roots = odesys.roots
    nroots = 0 if roots is None else len(roots)
    if nroots:
        all_exprs += [be.sympify(r) for r in roots]
    # _base.py:134-155 - CSC Jacobian entries (``sparse=True``) and the Jacobian-vector product (``jtimes=True``);
    # appended last so that systems without them generate exactly the text they did before
    nnz = odesys.nnz if (jac is not False and getattr(odesys, 'sparse', False)) else -1
    if nnz >= 0:
        csc_data = odesys.get_jac()
        all_exprs += [csc_data[k] for k in range(nnz)]
    jtimes = odesys.get_jtimes() if hasattr(odesys, 'get_jtimes') else False
    njtimes = 0
    if jtimes is not False:
        jt_v, jt_exprs = jtimes
        all_exprs += [be.sympify(e) for e in jt_exprs]
        njtimes = len(jt_exprs)

    invar = [be.sympify(e) for e in odesys.all_invariants()] if hasattr(odesys, 'all_invariants') else []
    ninvar = len(invar)

    xsym = be.Symbol('x')
    subsd = {k: be.Symbol('y[%d]' % i) for i, k in enumerate(odesys.dep)}
    if odesys.indep is not None:
        subsd[odesys.indep] = xsym
    subsd.update({k: be.Symbol('m_p[%d]' % i) for i, k in enumerate(odesys.params)})
    if odesys.append_iv and odesys.init_dep is not None:
        subsd.update({k: be.Symbol('m_p[%d]' % (nparams + i)) for i, k in enumerate(odesys.init_dep)})
    if njtimes:
        subsd.update({k: be.Symbol('v[%d]' % i) for i, k in enumerate(jt_v)})  # _base.py:160-162

    use_cse = os.getenv('PYODESYS_NATIVE_CSE', '1') == '1'  # _base.py:175-179

    def cse_cb(exprs, symbols, ignore=()):
        if not use_cse or len(exprs) == 0:
            return [], list(exprs)
        return be.cse(exprs, symbols=symbols, ignore=ignore)

    # ---- level 1: parameter-only sub-expressions -> m_p_cse[] -------------------------------
    ignore = tuple(odesys.dep) + ((odesys.indep,) if odesys.indep is not None else ()) + \
        (tuple(jt_v) if njtimes else ())
    common_cses, common_exprs = cse_cb(all_exprs, be.numbered_symbols('cse_temporary'), ignore)
    common_subs, k = {}, 0
    for symb, _sub in common_cses:
        if any(symb in e.free_symbols for e in common_exprs):
            common_subs[symb] = be.Symbol('m_p_cse[%d]' % k)
            k += 1
    npcse = k
    common_cses = [(s.xreplace(common_subs), e.xreplace(common_subs)) for s, e in common_cses]
    common_exprs = [e.xreplace(common_subs) for e in common_exprs]

    def cc(e):
        return _ccode(sympy.sympify(e).xreplace(subsd))

    rhs_cses, rhs_exprs = cse_cb(common_exprs[:ny], be.numbered_symbols('cse'))
    off = ny
    if nj:
        jac_cses, jac_exprs = cse_cb(common_exprs[off:off + nj], be.numbered_symbols('cse'))
        off += nj
    if nfs:
        fs_cses, fs_exprs = cse_cb(common_exprs[off:off + 1], be.numbered_symbols('cse'))
        off += 1
    if nroots:
        roots_cses, roots_exprs = cse_cb(common_exprs[off:off + nroots], be.numbered_symbols('cse'))
        off += nroots
    if nnz >= 0:
        csc_cses, csc_exprs = cse_cb(common_exprs[off:off + nnz], be.numbered_symbols('cse'))
        off += nnz
    if njtimes:
        jt_cses, jt_exprs_c = cse_cb(common_exprs[off:off + njtimes], be.numbered_symbols('cse'))

    flops = {
        'pcse': _count([e for _s, e in common_cses]),
        'rhs': _count([e for _s, e in rhs_cses] + list(rhs_exprs)),
        'jac': _count([e for _s, e in jac_cses] + list(jac_exprs)) if nj else 0,
    }
    jac_nnz = sum(1 for e in jac_exprs[:ny*ny] if e != 0) if nj else 0

    # ---- which leading entries of m_p are referenced at all ---------------------------------
    used = -1
    # (the invariant expressions are emitted outside the CSE'd groups but read m_p like everything else)
    all_final = [e for _s, e in common_cses] + list(common_exprs) + list(invar)
    psyms = list(odesys.params) + (list(odesys.init_dep) if odesys.append_iv and odesys.init_dep else [])
    for i, ps in enumerate(psyms):
        if any(ps in sympy.sympify(e).free_symbols for e in all_final):
            used = i
    np_used = used + 1
    if any(key in ns_over for key in ('p_rhs', 'p_jac', 'p_first_step', 'p_get_dx_max', 'p_constructor')):
        np_used = np_tot  # raw user source may read any entry

    # ---- emit ---------------------------------------------------------------------------------
    setup_lines = []
    for symb, expr in common_cses:
        tok = symb.name
        if tok.startswith('m_p_cse'):
            setup_lines.append('        %s = %s;' % (tok, cc(expr)))
        else:
            setup_lines.append('        const double %s = %s;' % (tok, cc(expr)))
    for line in ns_over.get('p_constructor', []) + ns_ext.get('p_constructor', []):
        setup_lines.append('        ' + line)

    if 'p_rhs' in ns_over:
        rhs_body = ns_over['p_rhs']
    else:
        acc_lines, acc_names = _accumulate_lines(rhs_cses, rhs_exprs, cc)
        rhs_body = '\n'.join(acc_lines + ['        f[%d] = %s;' % (i, nm) for i, nm in enumerate(acc_names)] +
                             ['        this->nfev++;', '        return AnyODE::Status::success;'])

    if 'p_jac' in ns_over:
        jac_body = ns_over['p_jac']
    elif nj:
        dfdt_nz = [(i, e) for i, e in enumerate(jac_exprs[ny*ny:]) if e != 0]
        outs = [('jac[ldim*%d + %d] = %%s;' % (idx // ny, idx % ny), e) for idx, e in enumerate(jac_exprs[:ny*ny])]
        outs += [('if (dfdt) dfdt[%d] = %%s;' % i, e) for i, e in dfdt_nz]
        jac_body = '\n'.join(_lazy_lines(jac_cses, outs, cc))
        jac_body += '\n        if (dfdt) {\n%s\n        }\n' % '\n'.join(
            '            dfdt[%d] = 0.0;' % i for i, e in enumerate(jac_exprs[ny*ny:]) if e == 0)
        jac_body += '        this->njev++;\n        return AnyODE::Status::success;'
    else:
        jac_body = '        return AnyODE::Status::unrecoverable_error;'

    # "sink" variants for the group-per-system kernels: every lane of a group evaluates the expressions
    # and the sink keeps the rows that lane owns (``s.set(i, v)`` with literal i, resolved at compile time)
    if 'p_rhs' in ns_over:
        rhs_sink_body = '''        double f_all[ODESYS_NY];
        const AnyODE::Status st = rhs(x, y, f_all);
#pragma unroll
        for (int i = 0; i < ODESYS_NY; ++i) s.set(i, f_all[i]);
        return st;'''
    else:
        acc_lines, acc_names = _accumulate_lines(rhs_cses, rhs_exprs, cc)
        rhs_sink_body = '\n'.join(acc_lines + ['        s.set(%d, %s);' % (i, nm) for i, nm in enumerate(acc_names)] +
                                  ['        this->nfev++;', '        return AnyODE::Status::success;'])
    if 'p_jac' in ns_over or not nj:
        jac_sink_body = '''        double j_all[ODESYS_NY*ODESYS_NY], d_all[ODESYS_NY];
        const AnyODE::Status st = dense_jac_rmaj(x, y, nullptr, j_all, ODESYS_NY, d_all);
#pragma unroll
        for (int i = 0; i < ODESYS_NY; ++i) {
#pragma unroll
            for (int j = 0; j < ODESYS_NY; ++j) s.setj(i, j, j_all[i*ODESYS_NY + j]);
            s.setdfdt(i, d_all[i]);
        }
        return st;'''
    else:
        outs = [('s.setj(%d, %d, %%s);' % (idx // ny, idx % ny), e) for idx, e in enumerate(jac_exprs[:ny*ny])]
        outs += [('s.setdfdt(%d, %%s);' % i, e) for i, e in enumerate(jac_exprs[ny*ny:])]
        jac_sink_body = '\n'.join(_lazy_lines(jac_cses, outs, cc))
        jac_sink_body += '\n        this->njev++;\n        return AnyODE::Status::success;'

    # ---- evaluation in EVAL_PARTS independent pieces (block-cooperative kernel: one piece per warp) -------------
    # rows of f / columns of J are dealt to the parts so that the largest part - after its own CSE - is as small as
    # possible; each part is a self-contained sink function that sets only its rows / columns
    parts_rhs_body = parts_jac_body = None
    eval_parts_cost = None
    if (use_cse and ny >= EVAL_PARTS_MIN_NY and nj and 'p_rhs' not in ns_over and 'p_jac' not in ns_over and
            os.getenv('PYODESYS_EVAL_PARTS', '1') == '1'):
        f_common = common_exprs[:ny]
        j_common = common_exprs[ny:ny + ny*ny]
        d_common = common_exprs[ny + ny*ny:ny + nj]

        def cost_rows(rows):
            c, e2 = cse_cb([f_common[i] for i in rows], be.numbered_symbols('cse'))
            return _count([e for _s, e in c] + list(e2))

        def cost_cols(cols):
            c, e2 = cse_cb([j_common[i*ny + j] for j in cols for i in range(ny)], be.numbered_symbols('cse'))
            return _count([e for _s, e in c] + list(e2))

        rparts, rcost = _balanced_parts(ny, cost_rows, [sympy.count_ops(e) + 1 for e in f_common])
        cparts, ccost = _balanced_parts(ny, cost_cols,
                                        [sum(sympy.count_ops(j_common[i*ny + j]) for i in range(ny)) + 1
                                         for j in range(ny)])
        eval_parts_cost = {'rhs': rcost, 'jac': ccost}
        chunks_r, chunks_j = [], []
        for q in range(EVAL_PARTS):
            rows, cols = rparts[q], cparts[q]
            c, e2 = cse_cb([f_common[i] for i in rows], be.numbered_symbols('cse'))
            acc_lines, acc_names = _accumulate_lines(c, e2, cc, indent='            ')
            body = acc_lines + ['            s.set(%d, %s);' % (i, nm) for i, nm in zip(rows, acc_names)]
            chunks_r.append('        %sif constexpr (P == %d) {\n%s\n        }' % ('' if q == 0 else 'else ', q,
                                                                                 '\n'.join(body)))
            idx = [(i, j) for j in cols for i in range(ny)]
            c, e2 = cse_cb([j_common[i*ny + j] for i, j in idx], be.numbered_symbols('cse'))
            outs = [('s.setj(%d, %d, %%s);' % ij, e) for ij, e in zip(idx, e2)]
            if q == 0:  # the time derivative goes with the first part
                cd, ed = cse_cb(list(d_common), be.numbered_symbols('dcse'))
                c = list(c) + list(cd)
                outs += [('s.setdfdt(%d, %%s);' % i, e) for i, e in enumerate(ed)]
            chunks_j.append('        %sif constexpr (P == %d) {\n%s\n        }' % (
                '' if q == 0 else 'else ', q, '\n'.join(_lazy_lines(c, outs, cc, indent='            '))))
        parts_rhs_body = '\n'.join(chunks_r) + '\n        return AnyODE::Status::success;'
        parts_jac_body = '\n'.join(chunks_j) + '\n        return AnyODE::Status::success;'

    if 'p_first_step' in ns_over:
        fs_body = '        ' + ns_over['p_first_step']
        has_fs = True
    elif nfs:
        fs_body = _body([(s.name, cc(e)) for s, e in fs_cses], [], ret='return %s;' % cc(fs_exprs[0]))
        has_fs = True
    else:
        fs_body = '        return 0.0;  // invokes the default behaviour of the stepper'
        has_fs = False

    if ns_over.get('p_get_dx_max'):
        dxmax_body = '        ' + ns_over['p_get_dx_max']
        has_dx_max = True
    else:
        dxmax_body = '        return INFINITY;'
        has_dx_max = False

    if nroots:
        roots_body = _body([(s.name, cc(e)) for s, e in roots_cses],
                           [('out[%d]' % i, cc(e)) for i, e in enumerate(roots_exprs)])
    else:
        roots_body = '        return AnyODE::Status::success;'

    # template:311-344 - data[nnz] in CSC order plus the (constant) index arrays
    if nnz >= 0:
        csc_body = _body([(sy.name, cc(e)) for sy, e in csc_cses],
                         [('data[%d]' % k, cc(e)) for k, e in enumerate(csc_exprs)] +
                         [('rowvals[%d]' % k, '%d' % r) for k, r in enumerate(odesys._rowvals)] +
                         [('colptrs[%d]' % k, '%d' % c) for k, c in enumerate(odesys._colptrs)],
                         ret='this->njev++;\n        return AnyODE::Status::success;')
    else:
        csc_body = '        return AnyODE::Status::unrecoverable_error;'
    # template:214-243
    if 'p_jtimes' in ns_over:
        jtimes_body = '        ' + ns_over['p_jtimes'] + '\n        this->njvev++;\n        return AnyODE::Status::success;'
    elif njtimes:
        jtimes_body = _body([(sy.name, cc(e)) for sy, e in jt_cses],
                            [('Jv[%d]' % i, cc(e)) for i, e in enumerate(jt_exprs_c)],
                            ret='this->njvev++;\n        return AnyODE::Status::success;')
    else:
        jtimes_body = '        return AnyODE::Status::unrecoverable_error;'
    has_jtimes = bool(njtimes) or 'p_jtimes' in ns_over

    # template:112-123,191-198 - invariants: values at the initial state, and the violation test of ``rhs``
    if ninvar:
        inv_cses, inv_exprs = cse_cb(invar, be.numbered_symbols('cse'))
        invar_body = _body([(sy.name, cc(e)) for sy, e in inv_cses],
                           [('out[%d]' % i, cc(e)) for i, e in enumerate(inv_exprs)], ret='')
    else:
        invar_body = ''

    lb = odesys.lower_bounds
    ub = odesys.upper_bounds

    checks = []
    if lb is not None:
        checks += ['y[%d] < %r' % (i, float(v)) for i, v in enumerate(lb)]
    if ub is not None:
        checks += ['y[%d] > %r' % (i, float(v)) for i, v in enumerate(ub)]
    bounds_body = '        return %s;' % (' || '.join(checks) if checks else 'false')

    anon = ns_over.get('p_anon') or ''
    includes = ''.join('#include %s\n' % inc for inc in ns_ext.get('p_includes', []))
    macros = [
        ('ODESYS_NY', ny), ('ODESYS_NPARAMS', nparams), ('ODESYS_NP_TOT', np_tot),
        ('ODESYS_NP_USED', np_used), ('ODESYS_NPCSE', npcse), ('ODESYS_HAS_JAC', int(bool(nj) or 'p_jac' in ns_over)),
        ('ODESYS_AUTONOMOUS', int(odesys.autonomous_exprs)), ('ODESYS_HAS_FIRST_STEP', int(has_fs)),
        ('ODESYS_HAS_DX_MAX', int(has_dx_max)), ('ODESYS_NROOTS', nroots),
        ('ODESYS_FLOPS_RHS', flops['rhs']), ('ODESYS_FLOPS_JAC', flops['jac']),
        ('ODESYS_FLOPS_PCSE', flops['pcse']), ('ODESYS_JAC_NNZ', jac_nnz),
        ('ODESYS_HAS_BOUNDS', int(lb is not None or ub is not None)),
        ('ODESYS_NONNEGATIVE', int(bool(getattr(odesys, '_nonnegative', False)))),
        ('ODESYS_NINVAR', ninvar), ('ODESYS_NNZ', nnz), ('ODESYS_HAS_JTIMES', int(has_jtimes)),
        ('ODESYS_EVAL_PARTS', EVAL_PARTS if parts_rhs_body else 0),
    ]
    struct_body = '''
    int nfev, njev;
#if ODESYS_HAS_JTIMES
    int njvev;
#endif

    // template:163-190 - true when a lower/upper bound is violated
    ODESYS_FN bool outside_bounds(const double * const y) const {
        AnyODE::ignore(y);
%(bounds)s
    }
    // template:92-125 - parameter-only sub-expressions, once per system
    ODESYS_FN void setup() {
        nfev = 0; njev = 0;
#if ODESYS_HAS_JTIMES
        njvev = 0;
#endif
%(setup)s
    }
    // template:147-204
    ODESYS_FN AnyODE::Status rhs(double x, const double * const ODESYS_RESTRICT y,
                                 double * const ODESYS_RESTRICT f) {
        AnyODE::ignore(x); AnyODE::ignore(y);
%(rhs)s
    }
    // template:246-293 (row-major: jac[ldim*row + col]); all ny*ny entries are written
    ODESYS_FN AnyODE::Status dense_jac_rmaj(double x, const double * const ODESYS_RESTRICT y,
                                            const double * const ODESYS_RESTRICT fy,
                                            double * const ODESYS_RESTRICT jac, long int ldim,
                                            double * const ODESYS_RESTRICT dfdt) {
        AnyODE::ignore(x); AnyODE::ignore(y); AnyODE::ignore(fy); AnyODE::ignore(ldim);
%(jac)s
    }
    // same expressions, results handed to a sink: s.set(i, f_i) / s.setj(i, j, J_ij) / s.setdfdt(i, .)
    template <class Sink> ODESYS_FN AnyODE::Status rhs_sink(double x, const double * const y, Sink &s) {
        AnyODE::ignore(x); AnyODE::ignore(y);
%(rhs_sink)s
    }
    template <class Sink> ODESYS_FN AnyODE::Status dense_jac_sink(double x, const double * const y, Sink &s) {
        AnyODE::ignore(x); AnyODE::ignore(y);
%(jac_sink)s
    }
#if ODESYS_EVAL_PARTS > 0
    // the same expressions dealt to ODESYS_EVAL_PARTS self-contained pieces: part P sets its rows of f / its columns
    // of J only (the counters nfev / njev are left to the caller)
    template <int P, class Sink> ODESYS_FN AnyODE::Status rhs_sink_part(double x, const double * const y, Sink &s) {
        AnyODE::ignore(x); AnyODE::ignore(y); AnyODE::ignore(s);
%(parts_rhs)s
    }
    template <int P, class Sink> ODESYS_FN AnyODE::Status dense_jac_sink_part(double x, const double * const y, Sink &s) {
        AnyODE::ignore(x); AnyODE::ignore(y); AnyODE::ignore(s);
%(parts_jac)s
    }
#endif
    // template:127-133 - number of stored Jacobian entries, -1 for dense storage
    ODESYS_FN int get_nnz() const { return ODESYS_NNZ; }
#if ODESYS_NNZ >= 0
    // template:311-344 - compressed sparse column: data[nnz], colptrs[ny + 1], rowvals[nnz]
    template <class Index>
    ODESYS_FN AnyODE::Status sparse_jac_csc(double x, const double * const ODESYS_RESTRICT y,
                                            const double * const ODESYS_RESTRICT fy,
                                            double * const ODESYS_RESTRICT data, Index * const ODESYS_RESTRICT colptrs,
                                            Index * const ODESYS_RESTRICT rowvals) {
        AnyODE::ignore(x); AnyODE::ignore(y); AnyODE::ignore(fy);
%(csc)s
    }
#endif
#if ODESYS_HAS_JTIMES
    // template:214-243 - Jv = J(x, y) v without forming J
    ODESYS_FN AnyODE::Status jtimes(const double * const ODESYS_RESTRICT v, double * const ODESYS_RESTRICT Jv,
                                    double x, const double * const ODESYS_RESTRICT y,
                                    const double * const ODESYS_RESTRICT fy) {
        AnyODE::ignore(x); AnyODE::ignore(y); AnyODE::ignore(fy); AnyODE::ignore(v);
%(jtimes)s
    }
#endif
    // template:295-309
    ODESYS_FN double get_dx0(double x, const double * const y) {
        AnyODE::ignore(x); AnyODE::ignore(y);
%(fs)s
    }
    // template:346-375
    ODESYS_FN double get_dx_max(double x, const double * const y) {
        AnyODE::ignore(x); AnyODE::ignore(y);
%(dxmax)s
    }
    // template:377-396
    ODESYS_FN AnyODE::Status roots(double x, const double * const y, double * const out) {
        AnyODE::ignore(x); AnyODE::ignore(y); AnyODE::ignore(out);
%(roots)s
    }
#if ODESYS_NINVAR > 0
    // template:112-123 (m_invar0 from the initial state) and :191-198 (violation -> recoverable error)
    double m_invar0[ODESYS_NINVAR];
    double m_max_invariant_violation;  // 0 = guard off (kernels instantiated without the guard only ever store 0)
    ODESYS_FN void invariants(double x, const double * const y, double * const out) const {
        AnyODE::ignore(x); AnyODE::ignore(y);
%(invar)s
    }
    ODESYS_FN void init_invariants(double max_violation, double x, const double * const y) {
        m_max_invariant_violation = max_violation;
        if (max_violation != 0.0) invariants(x, y, m_invar0);
    }
    ODESYS_FN bool invariant_violated(double x, const double * const y) const {
        if (m_max_invariant_violation == 0.0) return false;
        double cur[ODESYS_NINVAR];
        invariants(x, y, cur);
        bool bad = false;
#pragma unroll
        for (int i = 0; i < ODESYS_NINVAR; ++i) {
            const double lim = (m_max_invariant_violation > 0) ? m_max_invariant_violation
                : fabs(m_max_invariant_violation*m_invar0[i]) - m_max_invariant_violation;
            if (fabs(cur[i] - m_invar0[i]) > lim) bad = true;
        }
        return bad;
    }
#endif
''' % dict(csc=csc_body, jtimes=jtimes_body, invar=invar_body, bounds=bounds_body, setup='\n'.join(setup_lines), rhs=rhs_body, jac=jac_body,
           rhs_sink=rhs_sink_body, jac_sink=jac_sink_body, parts_rhs=parts_rhs_body or '', parts_jac=parts_jac_body or '', fs=fs_body, dxmax=dxmax_body, roots=roots_body)

    text = '''// -*- mode: read-only -*-
// Generated by pyodesys_b200.native._codegen (CUDA counterpart of
// pyodesys/native/sources/odesys_anyode_template.cpp).  Do not edit.
#pragma once
%(macros)s
#include "odesys_prelude.h"
%(includes)s
%(anon)s

namespace odesys_anyode {
// parameters held by value (registers): one system per thread, CPU oracle
struct OdeSys {
    double m_p[ODESYS_NP_TOT > 0 ? ODESYS_NP_TOT : 1];
    double m_p_cse[ODESYS_NPCSE > 0 ? ODESYS_NPCSE : 1];
%(body)s};
// parameters behind an array-like handle (shared memory): ``double *`` for one system per lane group,
// a strided view (element i of thread t at base[i*BLOCK + t]) for the shared-memory thread kernels
template <class Arr> struct OdeSysT {
    Arr m_p;
    Arr m_p_cse;
%(body)s};
typedef OdeSysT<double *> OdeSysRef;
}  // namespace odesys_anyode
''' % dict(
        macros='\n'.join('#define %s %d' % kv for kv in macros),
        includes=includes, anon=anon, body=struct_body)

    return GeneratedCode(text, ny, nparams, np_tot, np_used, npcse, flops,
                         has_jac=bool(nj) or 'p_jac' in ns_over,
                         autonomous=bool(odesys.autonomous_exprs), has_first_step=has_fs,
                         has_dx_max=has_dx_max, nroots=nroots, jac_nnz=jac_nnz, nnz=nnz,
                         has_jtimes=has_jtimes)
