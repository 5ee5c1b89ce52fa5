// Variable-step, variable-order (1..5) BDF / NDF on the block-cooperative kernel (ny > 8).
//
// Role in the reference: cvode's default ``method='BDF'`` (pyodesys/native/sources/_cvode_wrapper.pyx:88,244), which is
// what the reference's radiolysis example runs through (examples/_radiolysis.ipynb cells 1-7).  Same method and same
// control logic as BdfStepper (bdf_integrator.cuh; oracle/oracle.cpp ``Bdf``) - quasi-constant step size NDF in backward
// differences, simplified Newton on the stored Jacobian, order/step selection after order + 1 equal steps - laid out
// for coop_kernel_body (coop_integrators.cuh):
//   * vectors are distributed over the lanes of a group (R rows per lane); the difference table D[8][ny] of a system
//     sits in shared memory (it is addressed by the current order), everything else in registers;
//   * (I - c J)^-1 is kept in registers between steps (c_invert / c_apply) and re-formed - from the Jacobian that stays
//     in the block's shared-memory staging area - only when c = h / alpha_k or the Jacobian changed;
//   * every Newton iteration is one evaluation slot of the block (right-hand side of all its systems by one warp);
//     the number of iterations of a step attempt is the maximum over the systems of the block (block votes);
//   * Jacobians are re-evaluated only for the systems that ask (request flags), after a convergence failure.
// Requires coop_integrators.cuh first.
#pragma once

namespace pyodesys_cuda {

#if ODESYS_HAS_JAC

#ifndef PYODESYS_COOP_BDF_G
#define PYODESYS_COOP_BDF_G (PYODESYS_COOP_FITS(8) ? 8 : PYODESYS_COOP_FITS(16) ? 16 : 32)
#endif

namespace cbdf {
constexpr int MAXORD = 5;
constexpr int NEWTON_MAXITER = 4;
constexpr int ND = MAXORD + 3;
// gamma_k = sum_{j<=k} 1/j, kappa of the NDF formulas (Shampine & Reichelt 1997)
__constant__ double kGamma[7] = {0.0, 1.0, 1.5, 1.0 + 0.5 + 1.0 / 3, 1.0 + 0.5 + 1.0 / 3 + 0.25,
                                 1.0 + 0.5 + 1.0 / 3 + 0.25 + 0.2, 1.0 + 0.5 + 1.0 / 3 + 0.25 + 0.2 + 1.0 / 6};
__constant__ double kKappa[7] = {0.0, -0.1850, -1.0 / 9, -0.0823, -0.0415, 0.0, 0.0};
}  // namespace cbdf

template <int G> struct CoopBdf {
    typedef CoopLayout<G, cbdf::ND * NY> L;
    static constexpr int R = L::R;
    static constexpr bool kAlwaysInit = true;  // init() evaluates f(t0, y0) and the first Jacobian
    static constexpr int kExtraDoubles = cbdf::ND * NY;
    // systems start on a grid of kEpoch block iterations: the order / step selections of all systems of the block -
    // rescaling, new Jacobian, new inverse - fall into the same iteration (odesys_prelude.h)
    static constexpr int kEpoch = ODESYS_BDF_EPOCH;
#ifdef PYODESYS_COOP_TRACE
    int trace_iter, trace_code;  // debugging aid: Newton iterations and outcome of the last attempt
    float trace_nrm[4];
    __device__ __forceinline__ double trace_order() const { return (double)order; }
    __device__ __forceinline__ double trace_flags() const { return trace_code + 100.0 * trace_iter + 10000.0 * n_equal; }
    __device__ __forceinline__ double trace_extra(int k) const { return (double)trace_nrm[k]; }
#endif

    double W[R][NY];  // (I - c J)^-1 (c_invert layout)
    int ord[R];
    bool identity;
    double d[R];
    float iscale[R];
    double h_cur, c_lu;
    float newton_tol, newton_floor, safety;
    int order, n_equal, n_attempt;
    bool have_lu, jac_current, fresh, newton_failed, refresh_jac;

    __device__ __forceinline__ double &Dx(const Coop<G> &c, const Grp<G> &g, int j, int r) const {
        extern __shared__ double c_smem[];
        // padding rows (r * G + gl >= NY) alias the slot of row 0 of the same lane group: they only ever hold zeros
        const int i = g.owns(r) ? r * G + g.gl : NY - 1;
        return c_smem[L::oExtra + (c.sysl * cbdf::ND + j) * NY + i];
    }
    __device__ __forceinline__ double alpha(int k) const { return (1.0 - cbdf::kKappa[k]) * cbdf::kGamma[k]; }
    __device__ __forceinline__ double errconst(int k) const { return fma(cbdf::kKappa[k], cbdf::kGamma[k], 1.0 / (k + 1)); }
    __device__ __forceinline__ void set_scale(const Grp<G> &g, const KernelArgs &a, const double *y) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double at = g.owns(r) ? a.atol[r * G + g.gl] : 1.0;
            iscale[r] = g.owns(r) ? rcp_ftz(fmaxf((float)fma(a.rtol, fabs(y[r]), at), 1e-37f)) : 0.0f;
        }
    }
    // weighted rms norm over the system (group reduction with full-mask xor shuffles), single precision
    __device__ __forceinline__ float rms(const double *v) const {
        float s = 0.0f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float q = (float)v[r] * iscale[r];
            s = fmaf(q, q, s);
        }
#pragma unroll
        for (int off = G / 2; off > 0; off /= 2) s += __shfl_xor_sync(FULL_MASK, s, off);
        float out;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(out) : "f"(s * (1.0f / NY)));
        return out;
    }

    __device__ __forceinline__ void begin() {
        order = 1;
        n_equal = 0;
        n_attempt = 0;
        refresh_jac = false;
        have_lu = false;
        fresh = true;
        newton_failed = false;
        jac_current = false;
        safety = 0.9f;
        h_cur = 0.0;
        c_lu = 0.0;
        newton_tol = 0.03f;
        newton_floor = 0.0f;
        identity = true;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            ord[r] = -1;
            d[r] = 0.0;
            iscale[r] = 0.0f;
#pragma unroll
            for (int j = 0; j < NY; ++j) W[r][j] = 0.0;
        }
    }

    // f(t0, y0) -> D[1] (scaled by h at the first attempt), the first Jacobian, and - where no first step was given -
    // Hairer's starting step for order 1.  Executed by the whole block; ``want`` marks the groups that use it.
    __device__ __forceinline__ double init(Coop<G> &c, const Grp<G> &g, const KernelArgs &a, double t0, const double *y,
                                           double span, bool want, bool have_h, int &nfev, int &njev) {
        newton_tol = (float)fmax(10.0 * DBL_EPSILON / a.rtol, ODESYS_BDF_NEWTON_TOL);  // odesys_prelude.h
        newton_floor = fmaxf(0.03f * newton_tol, (float)(3.0 * DBL_EPSILON / a.rtol));  // see BdfStepper
        double f0[R], y1[R], f1[R];
        c.template eval<2>(g, a, t0, y, f0, want);
        if (want) {
            nfev += 1;
            njev += 1;
            jac_current = true;
            have_lu = false;
#pragma unroll
            for (int j = 0; j < cbdf::ND; ++j)
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (g.owns(r)) Dx(c, g, j, r) = (j == 0) ? y[r] : (j == 1 ? f0[r] : 0.0);
        }
        double d0 = 0.0, d1 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!g.owns(r)) continue;
            const double isk = fast_rcp<1>(a.atol[r * G + g.gl] + a.rtol * fabs(y[r]) + 1e-300);
            d0 = fmax(d0, fabs(y[r]) * isk);
            d1 = fmax(d1, fabs(f0[r]) * isk);
        }
        d0 = gmax_d<G>(g, d0);
        d1 = gmax_d<G>(g, d1);
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, fabs(span));
        if (!(h0 > 0.0)) h0 = 1e-6;
        const double dir = span >= 0 ? 1.0 : -1.0;
#pragma unroll
        for (int r = 0; r < R; ++r) y1[r] = fma(dir * h0, f0[r], y[r]);
        const int st = c.template eval<0>(g, a, t0 + dir * h0, y1, f1);
        if (want && !have_h) nfev += 1;
        double d2 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!g.owns(r)) continue;
            const double isk = fast_rcp<1>(a.atol[r * G + g.gl] + a.rtol * fabs(y[r]) + 1e-300);
            d2 = fmax(d2, fabs(f1[r] - f0[r]) * isk);
        }
        d2 = gmax_d<G>(g, d2) / h0;
        if (st != 0 || !(d2 <= DBL_MAX)) d2 = 0.0;
        const double dmax = fmax(d1, d2);
        const double h1 = (dmax <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : 0.01 * fast_rcp<1>(dmax);  // order 1: (0.01/dmax)^(1/1)
        return fmin(fmin(100.0 * h0, h1), fabs(span)) * dir;
    }

    // D[0..order] <- U^T (R(factor)^T D[0..order]);  R_ij = prod_{m=1..i} (m - 1 - factor j)/m,  U = R(1)
    __device__ __forceinline__ void change_D(const Coop<G> &c, const Grp<G> &g, double factor) {
        double T[cbdf::MAXORD + 1][R];
#pragma unroll
        for (int j = 0; j <= cbdf::MAXORD; ++j) {
            double rr = 1.0;
#pragma unroll
            for (int r = 0; r < R; ++r) T[j][r] = Dx(c, g, 0, r);
            if (j > 0 && j <= order) {
#pragma unroll
                for (int m = 1; m <= cbdf::MAXORD; ++m) {
                    rr *= fma(-factor, (double)j, (double)(m - 1)) * (1.0 / m);
                    if (m <= order) {
#pragma unroll
                        for (int r = 0; r < R; ++r) T[j][r] = fma(rr, Dx(c, g, m, r), T[j][r]);
                    }
                }
            }
        }
#pragma unroll
        for (int j = 1; j <= cbdf::MAXORD; ++j) {
            if (j <= order) {
                double cb = 1.0, acc[R];
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = T[0][r];
#pragma unroll
                for (int m = 1; m <= j; ++m) {
                    cb = -cb * (double)(j - m + 1) / (double)m;
#pragma unroll
                    for (int r = 0; r < R; ++r) acc[r] = fma(cb, T[m][r], acc[r]);
                }
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (g.owns(r)) Dx(c, g, j, r) = acc[r];
            }
        }
    }

    __device__ __forceinline__ int attempt(Coop<G> &c, const Grp<G> &g, const KernelArgs &a, double t, const double *,
                                           double h, double *ynew, double &err, Counters &cnt, int &nfev, int &njev,
                                           bool live) {
        extern __shared__ double c_smem[];
        if (live) n_attempt += 1;
        if (fresh) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (g.owns(r)) Dx(c, g, 1, r) *= h;
            h_cur = h;
            fresh = false;
        } else if (h != h_cur) {
            change_D(c, g, h * fast_rcp<2>(h_cur));
            h_cur = h;
            n_equal = 0;
            have_lu = false;
        }
        const double tnew = t + h;
        double ypred[R], psi[R], yk[R], f[R], dy[R];
        const double ialpha = fast_rcp<2>(alpha(order));
#pragma unroll
        for (int r = 0; r < R; ++r) {
            double s = Dx(c, g, 0, r), ps = 0.0;
#pragma unroll
            for (int j = 1; j <= cbdf::MAXORD; ++j) {
                if (j <= order) {
                    const double dj = Dx(c, g, j, r);
                    s += dj;
                    ps = fma(cbdf::kGamma[j], dj, ps);
                }
            }
            ypred[r] = g.owns(r) ? s : 0.0;
            psi[r] = g.owns(r) ? ps * ialpha : 0.0;
            ynew[r] = g.owns(r) ? Dx(c, g, 0, r) : 0.0;
        }
        set_scale(g, a, ypred);
        const double cc = h * ialpha;
        bool converged = false;
        int n_iter = 0, bad = 0;
        newton_failed = false;
        err = 2.0;
        // ---- a new order / step was selected: new Jacobian at (tnew, ypred).  The slot also yields f(tnew, ypred),
        // which is what the first Newton iteration needs.
        bool have_f0 = false;
        int st0 = 0;
        double f0p[R];
        {
            const bool want_j = live && refresh_jac && !jac_current;
            if (live) refresh_jac = false;
            if (__syncthreads_or(want_j)) {
                st0 = c.template eval<2>(g, a, tnew, ypred, f0p, want_j);
                have_f0 = true;
                if (want_j) {
                    njev += 1;
                    if (st0 != 0) bad = 1;
                    jac_current = true;
                    have_lu = false;
                }
            }
        }
        for (int pass = 0; pass < 2; ++pass) {
            // ---- (I - c J)^-1 where c or J changed (warp vote: the inverse runs with full-warp shuffles) ----------
            const bool need_lu = live && !converged && !bad && (!have_lu || cc != c_lu);
            if (__any_sync(FULL_MASK, need_lu)) {
                if (need_lu) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const int i = g.owns(r) ? r * G + g.gl : 0;
                        const double *row = c_smem + L::oJ + (i * L::NYP) * L::NSYSP + c.sysl;
#pragma unroll
                        for (int j = 0; j < NY; ++j) {
                            const double v = row[j * L::NSYSP];
                            W[r][j] = g.owns(r) ? fma(-cc, v, (i == j) ? 1.0 : 0.0) : 0.0;
                        }
                    }
                }
                // groups that keep their inverse only take part in the collectives of the elimination
                const bool ok = c_invert<G>(g, W, ord, identity, need_lu);
                if (need_lu) {
                    cnt.nlu += 1;
                    c_lu = cc;
                    have_lu = ok;
                    if (!ok) bad = 1;
                }
            }
            // ---- simplified Newton iteration, one evaluation slot of the block per iteration -----------------------
            bool gdone = !live || converged || bad != 0;  // this group takes no further part in the iteration
            float nrm_old = -1.0f;
            if (!converged) {  // (a group that converged in the first pass keeps its correction through the second)
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    d[r] = 0.0;
                    yk[r] = ypred[r];
                }
            }
            for (int k = 0; k < cbdf::NEWTON_MAXITER; ++k) {
                if (!__syncthreads_or(!gdone)) break;
                int st;
                if (k == 0 && have_f0) {  // (block-uniform) f(tnew, ypred) came with the Jacobian
                    st = st0;
#pragma unroll
                    for (int r = 0; r < R; ++r) f[r] = f0p[r];
                } else {
                    st = c.template eval<0>(g, a, tnew, yk, f);
                }
                if (!gdone) {
                    nfev += 1;
                    n_iter = k + 1;
                }
                if (!gdone && st != 0) gdone = true;  // recoverable error of the right-hand side: not converged
#pragma unroll
                for (int r = 0; r < R; ++r) dy[r] = fma(cc, f[r], -psi[r]) - d[r];
                c_apply<G>(g, W, ord, identity, dy);  // warp collectives: every group, results used where !gdone
                const float nrm = rms(dy);
#ifdef PYODESYS_COOP_TRACE
                if (!gdone) trace_nrm[k] = nrm;
#endif
                if (!gdone) {
                    cnt.nsolve += 1;
                    bool stop = !(nrm <= 3.0e38f);
                    float rate = -1.0f;
                    const bool tiny = nrm <= newton_floor;
                    if (!stop && nrm_old >= 0.0f) rate = nrm * rcp_ftz(nrm_old);
                    if (!stop && !tiny && rate >= 0.0f) {
                        float rp = rate;
                        if (k <= 2) rp *= rate;
                        if (k <= 1) rp *= rate;
                        if (rate >= 1.0f || rp * nrm > newton_tol * (1.0f - rate)) stop = true;
                    }
                    if (stop) {
                        gdone = true;
                    } else {
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            yk[r] += dy[r];
                            d[r] += dy[r];
                        }
                        if (tiny || (rate >= 0.0f && rate * nrm < newton_tol * (1.0f - rate))) {
                            converged = true;
                            gdone = true;
                        }
                        nrm_old = nrm;
                    }
                }
            }
            if (converged) {
#pragma unroll
                for (int r = 0; r < R; ++r) ynew[r] = yk[r];
            }
            // ---- not converged with a Jacobian from an earlier step: refresh it at (tnew, ypred), second pass ----------
            const bool refresh = live && !converged && !bad && !jac_current && pass == 0;
            if (!__syncthreads_or(refresh)) break;
            st0 = c.template eval<2>(g, a, tnew, ypred, f0p, refresh);
            have_f0 = true;  // the second pass starts from ypred again
            if (refresh) {
                nfev += 1;
                njev += 1;
                if (st0 != 0) bad = 1;
                jac_current = true;
                have_lu = false;
            }
        }
#ifdef PYODESYS_COOP_TRACE
        trace_iter = n_iter;
        trace_code = bad ? 3 : (!converged ? (jac_current ? 2 : 1) : 0);
#endif
        if (bad) return 1;
        if (!converged) {
            newton_failed = true;
            return 0;
        }
        safety = 0.9f * (2 * cbdf::NEWTON_MAXITER + 1) / (float)(2 * cbdf::NEWTON_MAXITER + n_iter);
        set_scale(g, a, ynew);
        const float ef = (float)errconst(order) * rms_masked(g, d);
        bool fin = true;
#pragma unroll
        for (int r = 0; r < R; ++r) fin = fin && (fabs(ynew[r]) <= DBL_MAX);
        if (!(ef <= 3.0e38f) || __any_sync(g.mask, !fin)) return 1;
        err = (double)ef;
        return 0;
    }

    // accept / reject, difference-table update, order and step selection (BdfStepper::control).  Called by live
    // groups only, whole groups at a time: its reductions use the group mask.
    __device__ __forceinline__ bool control(const Coop<G> &c, const Grp<G> &g, double err, double &hnext, double h) {
        if (newton_failed) {
            newton_failed = false;
            hnext = 0.5 * h;
            return false;
        }
        if (err > 1.0) {
            hnext = h * (double)fmaxf(0.2f, safety * fast_powf(err, -1.0f / (float)(order + 1)));
            return false;
        }
        jac_current = false;
        n_equal += 1;
        // D[order+2] = d - D[order+1]; D[order+1] = d; D[j] += D[j+1] for j = order..0
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!g.owns(r)) continue;
            Dx(c, g, order + 2, r) = d[r] - Dx(c, g, order + 1, r);
            Dx(c, g, order + 1, r) = d[r];
            for (int j = order; j >= 0; --j) Dx(c, g, j, r) += Dx(c, g, j + 1, r);
        }
        hnext = h;
        if (n_equal < order + 1 || n_attempt % (order <= 2 ? ODESYS_BDF_EPOCH / 2 : ODESYS_BDF_EPOCH) != 0) return true;
        double vm[R], vp[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            vm[r] = (g.owns(r) && order > 1) ? Dx(c, g, order, r) : 0.0;
            vp[r] = (g.owns(r) && order < cbdf::MAXORD) ? Dx(c, g, order + 2, r) : 0.0;
        }
        const float inf = __int_as_float(0x7f800000);
        float fm = -1.0f, fp = -1.0f;
        const float em = (float)errconst(order > 1 ? order - 1 : 1) * rms_masked(g, vm);
        const float ep = (float)errconst(order < cbdf::MAXORD ? order + 1 : cbdf::MAXORD) * rms_masked(g, vp);
        if (order > 1) fm = em == 0.0f ? inf : fast_powf(em, -1.0f / (float)order);
        if (order < cbdf::MAXORD) fp = ep == 0.0f ? inf : fast_powf(ep, -1.0f / (float)(order + 2));
        const float fc = err == 0.0 ? inf : fast_powf(err, -1.0f / (float)(order + 1));
        float best = fm;
        int delta = -1;
        if (fc > best) {
            best = fc;
            delta = 0;
        }
        if (fp > best) {
            best = fp;
            delta = 1;
        }
        order += delta;
        hnext = h * (double)fminf(10.0f, safety * best);
        n_equal = 0;
        have_lu = false;
        refresh_jac = ODESYS_BDF_REFRESH_JAC != 0;
        return true;
    }
    // the same norm with the group's own mask (control runs under per-group divergence)
    __device__ __forceinline__ float rms_masked(const Grp<G> &g, const double *v) const {
        float s = 0.0f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float q = (float)v[r] * iscale[r];
            s = fmaf(q, q, s);
        }
#pragma unroll
        for (int off = G / 2; off > 0; off /= 2) s += __shfl_xor_sync(g.mask, s, off);
        float out;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(out) : "f"(s * (1.0f / NY)));
        return out;
    }

    // Newton backward interpolation through the last order + 1 solution values; theta in (0, 1] of the last step
    __device__ __forceinline__ void dense(const Coop<G> &c, const Grp<G> &g, double theta, double, const double *,
                                          const double *, double *out) {
#pragma unroll
        for (int r = 0; r < R; ++r) out[r] = g.owns(r) ? Dx(c, g, 0, r) : 0.0;
        double p = 1.0;
        for (int j = 0; j < cbdf::MAXORD; ++j) {
            p *= (theta - 1.0 + j) * (1.0 / (1 + j));
            if (j < order) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (g.owns(r)) out[r] = fma(Dx(c, g, j + 1, r), p, out[r]);
            }
        }
    }
};

template <int G, int MODE>
__global__ void __launch_bounds__(kCoopBlock, PYODESYS_COOP_MIN_BLOCKS) coop_bdf_kernel(const KernelArgs a) {
    coop_kernel_body<CoopBdf<G>, G, MODE>(a);
}

#endif  // ODESYS_HAS_JAC

}  // namespace pyodesys_cuda
