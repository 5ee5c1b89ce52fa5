// Variable-step, variable-order (1..5) BDF / NDF stepper for the one-system-per-thread kernel.
//
// Role in the reference: cvode's default ``method='BDF'`` (pyodesys/native/sources/_cvode_wrapper.pyx:88,244;
// batch loop = multi_predefined / multi_adaptive call sites :190-197,339-345).  SUNDIALS is not part of the
// reference tree; this is the published quasi-constant step size NDF formulation in backward differences
// (Shampine & Reichelt, "The MATLAB ODE suite", SIAM J. Sci. Comput. 18 (1997), section 2; order/step
// selection after Byrne & Hindmarsh 1975), i.e. a multistep method with a simplified Newton corrector on the
// stored Jacobian - the same class of method as CVODE's BDF, not its Nordsieck implementation.
//
// Per lane:
//   D[j], j = 0..7   backward differences of the solution on the current uniform grid (D[0] = y): SHARED memory,
//                    element (j, i) of thread t at smem[(j*NY + i)*BLOCK + t] (conflict free).  The table is
//                    addressed by the current order; in registers the compiler turned the order-predicated
//                    copies into dynamic indexing and moved the whole stepper (472 B) to local memory
//                    (153 ms per 1e6 Robertson systems against 20 ms for Rosenbrock-4)
//   Jm               the stored Jacobian (re-evaluated with every order / step selection and after a Newton
//                    convergence failure): SHARED memory behind D, it is read once per factorisation
//   W, piv           LU of I - c Jm, c = h / alpha_k, re-formed only when c or Jm changed, registers
// One step: predictor sum, <= 4 simplified Newton iterations (1 rhs + 1 triangular solve each), error
// estimate errconst_k * d; after k+1 equal steps the estimates for orders k-1, k, k+1 choose the next order
// and step, and D is rescaled to the new step by D <- U^T (R(factor)^T D) with U_ij = (-1)^i C(j, i).
// Requires integrators.cuh (KernelArgs, Tol, lu_factor/lu_solve, eval_rhs, initial_step) first.
#pragma once

namespace pyodesys_cuda {

#if ODESYS_HAS_JAC

namespace bdfc {
constexpr int MAXORD = 5;
constexpr int NEWTON_MAXITER = 4;
// The tolerances reach the stepper already calibrated (kappa = ODESYS_KAPPA_BDF of odesys_prelude.h, applied by the
// shim; oracle/oracle.cpp applies the same table).
}  // namespace bdfc
// gamma_k = sum_{j<=k} 1/j, alpha_k = (1 - kappa_k) gamma_k, errconst_k = kappa_k gamma_k + 1/(k+1),
// kappa = {0, -0.1850, -1/9, -0.0823, -0.0415, 0}
__constant__ double kBdfGamma[7] = {0.0, 1.0, 1.5, 1.0 + 0.5 + 1.0 / 3, 1.0 + 0.5 + 1.0 / 3 + 0.25,
                                    1.0 + 0.5 + 1.0 / 3 + 0.25 + 0.2, 1.0 + 0.5 + 1.0 / 3 + 0.25 + 0.2 + 1.0 / 6};
__constant__ double kBdfKappa[7] = {0.0, -0.1850, -1.0 / 9, -0.0823, -0.0415, 0.0, 0.0};

struct BdfSink {
    double *J;   // this thread's column of the stored Jacobian (shared memory)
    int stride;
    __device__ __forceinline__ void setj(int i, int j, double v) { J[(i * NY + j) * stride] = v; }
    __device__ __forceinline__ void setdfdt(int, double) {}
};

#if ODESYS_HAS_JTIMES
// Matrix-free linear solve of the Newton iteration (the role of cvode's SPGMR with the user's jtimes,
// ``linear_solver='gmres', with_jtimes=True``: pyodesys/native/tests/test_cvode.py:198-217): full GMRES on
// A = I - c J(t, yref), A v formed as v - c * OdeSys::jtimes(v).  Modified Gram-Schmidt Arnoldi with Givens
// rotations, at most NY iterations (exact in exact arithmetic for an NY-dimensional system), x0 = 0.
// ``b`` holds the right-hand side on entry and the solution on return; returns the number of J v products.
// The basis is indexed at run time and lives in local memory: this is the option for systems where forming and
// factorising J is the expensive part, not the throughput path of the headline configs.
#if ODESYS_NY <= 4
#define PYODESYS_GMRES_INLINE __forceinline__  // unrolled, all in registers; an out-of-line call would push sys to the stack
#else
#define PYODESYS_GMRES_INLINE __noinline__
#endif
template <class Sys>
__device__ PYODESYS_GMRES_INLINE int gmres_solve(Sys &sys, double t, const double *yref, double c, double *b) {
    double V[NY + 1][NY], H[NY + 1][NY], cs[NY], sn[NY], g[NY + 1];
    double beta = 0.0;
    for (int i = 0; i < NY; ++i) beta = fma(b[i], b[i], beta);
    beta = sqrt(beta);
    if (!(beta > 0.0)) return 0;  // b == 0 (x = 0), or NaN: the caller's norm test sees it
    const double ibeta = 1.0 / beta;
    for (int i = 0; i < NY; ++i) V[0][i] = b[i] * ibeta;
    for (int i = 0; i <= NY; ++i) g[i] = 0.0;
    g[0] = beta;
    int m = 0, nmv = 0;
    // ny <= 4: fully unrolled, every index a literal -> basis and Hessenberg matrix in registers
    constexpr int kUnroll = NY <= 4 ? NY : 1;
#pragma unroll kUnroll
    for (int j = 0; j < NY; ++j) {
        double w[NY], Jv[NY];
        sys.jtimes(V[j], Jv, t, yref, nullptr);
        nmv += 1;
        for (int i = 0; i < NY; ++i) w[i] = fma(-c, Jv[i], V[j][i]);
#pragma unroll kUnroll
        for (int k = 0; k <= j; ++k) {
            double d = 0.0;
            for (int i = 0; i < NY; ++i) d = fma(w[i], V[k][i], d);
            H[k][j] = d;
            for (int i = 0; i < NY; ++i) w[i] = fma(-d, V[k][i], w[i]);
        }
#pragma unroll kUnroll
        for (int k = 0; k <= j; ++k) {  // second Gram-Schmidt pass: keeps the basis orthogonal to rounding
            double d = 0.0;
            for (int i = 0; i < NY; ++i) d = fma(w[i], V[k][i], d);
            H[k][j] += d;
            for (int i = 0; i < NY; ++i) w[i] = fma(-d, V[k][i], w[i]);
        }
        double hn = 0.0;
        for (int i = 0; i < NY; ++i) hn = fma(w[i], w[i], hn);
        hn = sqrt(hn);
#pragma unroll kUnroll
        for (int k = 0; k < j; ++k) {
            const double hk = H[k][j], hk1 = H[k + 1][j];
            H[k][j] = fma(cs[k], hk, sn[k] * hk1);
            H[k + 1][j] = fma(-sn[k], hk, cs[k] * hk1);
        }
        const double hjj = H[j][j];
        const double den = sqrt(fma(hjj, hjj, hn * hn));
        if (!(den > 0.0)) break;  // A singular on the Krylov space (or NaN): solve with what is there
        cs[j] = hjj / den;
        sn[j] = hn / den;
        H[j][j] = den;
        g[j + 1] = -sn[j] * g[j];
        g[j] = cs[j] * g[j];
        m = j + 1;
        if (hn <= 1e-14 * den || fabs(g[j + 1]) <= 1e-16 * beta || j + 1 == NY) break;
        const double ihn = 1.0 / hn;
        for (int i = 0; i < NY; ++i) V[j + 1][i] = w[i] * ihn;
    }
    // back substitution on the leading m x m block; loop bounds are compile-time, m enters as a predicate
    double yk[NY];
#pragma unroll kUnroll
    for (int i = NY - 1; i >= 0; --i) {
        double sacc = g[i];
#pragma unroll kUnroll
        for (int k = i + 1; k < NY; ++k)
            if (k < m) sacc = fma(-H[i][k], yk[k], sacc);
        yk[i] = (i < m) ? sacc / H[i][i] : 0.0;
    }
#pragma unroll
    for (int i = 0; i < NY; ++i) b[i] = 0.0;
#pragma unroll kUnroll
    for (int k = 0; k < NY; ++k) {
        if (k < m) {
#pragma unroll
            for (int i = 0; i < NY; ++i) b[i] = fma(yk[k], V[k][i], b[i]);
        }
    }
    return nmv;
}
#endif

// KRYLOV = false: LU of I - c J in registers;  true: matrix-free (gmres_solve), a kernel of its own so that the
// dense path carries none of its registers
// gamma_j for a compile-time j (folds to an immediate in unrolled loops)
__device__ __forceinline__ double bdf_gamma_c(int j) {
    return j == 1 ? 1.0 : j == 2 ? 1.5 : j == 3 ? 11.0 / 6 : j == 4 ? 25.0 / 12 : j == 5 ? 137.0 / 60 : j == 6 ? 2.45 : 0.0;
}

template <bool KRYLOV> struct BdfStepper {
    typedef OdeSys Sys;
    static constexpr bool kAdaptive = true;
    static constexpr int kOrder = 1;
    // systems are started on a grid of kEpoch loop iterations, so that the step attempts of the lanes of a warp are
    // numbered alike and their order / step selections (odesys_prelude.h) coincide
    static constexpr int kEpoch = ODESYS_BDF_EPOCH;
    static constexpr int ND = bdfc::MAXORD + 3;
    static constexpr int kSharedDoublesPerThread = ND * NY + NY * NY;
    double *Dsm;   // this thread's column of the difference table in shared memory
    int dstride;   // = blockDim.x
    __device__ __forceinline__ void bind_shared(double *base, int tid, int nthreads) {
        Dsm = base + tid;
        dstride = nthreads;
    }
    __device__ __forceinline__ double &Dx(int j, int i) { return Dsm[(j * NY + i) * dstride]; }
    __device__ __forceinline__ double &Jx(int r, int q) { return Dsm[(ND * NY + r * NY + q) * dstride]; }
    RegArr<NY * NY> W;  // factors of I - c Jm (ny == 3: its explicit inverse, see lin_factor in integrators.cuh)
    int piv[NY];
    double d[NY];
    float iscale[NY];   // 1 / (atol_i + rtol |y_i|): the norms only steer Newton and the step size -> fp32
    double h_cur, c_lu;
    float newton_tol, newton_floor, safety;
    int order, n_equal, n_attempt;
    bool have_lu, jac_current, fresh, newton_failed, refresh_jac;
    static constexpr bool krylov = KRYLOV;  // Newton corrections by gmres_solve on OdeSys::jtimes: no Jacobian, no LU

    __device__ __forceinline__ double alpha(int k) const { return (1.0 - kBdfKappa[k]) * kBdfGamma[k]; }
    __device__ __forceinline__ double errconst(int k) const { return fma(kBdfKappa[k], kBdfGamma[k], 1.0 / (k + 1)); }
    __device__ __forceinline__ void set_scale(const KernelArgs &a, const double *y) {
#pragma unroll
        for (int i = 0; i < NY; ++i) iscale[i] = rcp_ftz(fmaxf((float)fma(a.rtol, fabs(y[i]), a.atol[i]), 1e-37f));
    }
    // weighted rms norm in single precision (sqrt.approx: one MUFU instead of the IEEE sqrt subroutine)
    __device__ __forceinline__ float rms(const double *v) const {
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            const float q = (float)v[i] * iscale[i];
            s = fmaf(q, q, s);
        }
        float r;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s * (1.0f / NY)));
        return r;
    }
    __device__ __forceinline__ int eval_jac(Sys &sys, double t, const double *y) {
        BdfSink sink{&Jx(0, 0), dstride};
        const int njev0 = sys.njev;
        const int bad = sys.dense_jac_sink(t, y, sink);
        sys.njev = njev0 + 1;
        return bad;
    }
    __device__ __forceinline__ int begin(Sys &sys, const KernelArgs &a, const Tol &tol, double t, const double *y) {
        newton_tol = (float)fmax(10.0 * DBL_EPSILON / tol.rtol, ODESYS_BDF_NEWTON_TOL);  // odesys_prelude.h
        // a correction this small is at the rounding level of the residual: converged, whatever the rate looks like
        newton_floor = fmaxf(0.03f * newton_tol, (float)(3.0 * DBL_EPSILON / tol.rtol));
        order = 1;
        n_equal = 0;
        n_attempt = 0;
        refresh_jac = false;
        have_lu = false;
        fresh = true;
        newton_failed = false;
        safety = 0.9f;
        h_cur = 0.0;
        c_lu = 0.0;
#pragma unroll
        for (int j = 0; j < ND; ++j)
#pragma unroll
            for (int i = 0; i < NY; ++i) Dx(j, i) = 0.0;
#pragma unroll
        for (int i = 0; i < NY; ++i) Dx(0, i) = y[i];
        double f0[NY];
        int bad = eval_rhs(sys, a, t, y, f0);  // D[1] = f0, scaled by h at the first attempt
#pragma unroll
        for (int i = 0; i < NY; ++i) Dx(1, i) = f0[i];
        if (!krylov) bad |= eval_jac(sys, t, y);
        jac_current = true;
        return bad;
    }
    __device__ __forceinline__ void auto_step(Sys &sys, const KernelArgs &a, const Tol &tl, double t, const double *y,
                                              double span, double &h) {
        double f0[NY];
#pragma unroll
        for (int i = 0; i < NY; ++i) f0[i] = Dx(1, i);
        h = initial_step(sys, a, tl, t, y, f0, span, 1) * (span >= 0 ? 1.0 : -1.0);
    }
    // D[0..order] <- U^T (R(factor)^T D[0..order]);  R_ij = prod_{m=1..i} (m - 1 - factor j)/m,  U = R(1)
    __device__ __forceinline__ void change_D(double factor) {
        double T[bdfc::MAXORD + 1][NY];
#pragma unroll
        for (int j = 0; j <= bdfc::MAXORD; ++j) {
            double r = 1.0;
#pragma unroll
            for (int i = 0; i < NY; ++i) T[j][i] = Dx(0, i);
            if (j > 0 && j <= order) {
#pragma unroll
                for (int m = 1; m <= bdfc::MAXORD; ++m) {
                    r *= fma(-factor, (double)j, (double)(m - 1)) * (1.0 / m);
                    if (m <= order) {
#pragma unroll
                        for (int i = 0; i < NY; ++i) T[j][i] = fma(r, Dx(m, i), T[j][i]);
                    }
                }
            }
        }
#pragma unroll
        for (int j = 1; j <= bdfc::MAXORD; ++j) {
            if (j <= order) {
                // (-1)^i C(j, i), i = 0..j
                double cb = 1.0;
#pragma unroll
                for (int i = 0; i < NY; ++i) Dx(j, i) = T[0][i];
#pragma unroll
                for (int m = 1; m <= j; ++m) {
                    cb = -cb * (double)(j - m + 1) / (double)m;
#pragma unroll
                    for (int i = 0; i < NY; ++i) Dx(j, i) = fma(cb, T[m][i], Dx(j, i));
                }
            }
        }
    }
    __device__ __forceinline__ int attempt(Sys &sys, const KernelArgs &a, const Tol &, double t, const double *,
                                           double h, double *ynew, double &err, Counters &cnt) {
        n_attempt += 1;
        if (fresh) {
#pragma unroll
            for (int i = 0; i < NY; ++i) Dx(1, i) *= h;
            h_cur = h;
            fresh = false;
        } else if (h != h_cur) {
            change_D(h * fast_rcp<2>(h_cur));
            h_cur = h;
            n_equal = 0;
            have_lu = false;
        }
        const double tnew = t + h;
        double ypred[NY], psi[NY];
        const double ialpha = fast_rcp<2>(alpha(order));
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            double s = Dx(0, i), ps = 0.0;
#pragma unroll
            for (int j = 1; j <= bdfc::MAXORD; ++j) {
                if (j <= order) {
                    const double dj = Dx(j, i);
                    s += dj;
                    ps = fma(bdf_gamma_c(j), dj, ps);
                }
            }
            ypred[i] = s;
            psi[i] = ps * ialpha;
        }
        set_scale(a, ypred);
        const double c = h * ialpha;
        bool converged = false;
        int n_iter = 0;
        newton_failed = false;
        err = 2.0;
#pragma unroll
        for (int i = 0; i < NY; ++i) ynew[i] = Dx(0, i);
        if (krylov) jac_current = true;  // J is applied at (tnew, ypred) by every product
        if (refresh_jac) {  // a new order / step was selected: the iteration matrix is re-formed, with a new Jacobian
            refresh_jac = false;
            if (!krylov && !jac_current) {
                if (eval_jac(sys, tnew, ypred)) return 1;
                jac_current = true;
                have_lu = false;
            }
        }
        for (int pass = 0; pass < 2 && !converged; ++pass) {
            if (!krylov && (!have_lu || c != c_lu)) {
#pragma unroll
                for (int r = 0; r < NY; ++r)
#pragma unroll
                    for (int q = 0; q < NY; ++q) W[r * NY + q] = fma(-c, Jx(r, q), r == q ? 1.0 : 0.0);
                const bool ok = lin_factor<false>(W, piv, a.rtol, cnt);
                cnt.nlu += 1;
                c_lu = c;
                have_lu = ok;
                if (!ok) return 1;
            }
            double yk[NY], f[NY], dy[NY];
            float nrm_old = -1.0f;
#pragma unroll
            for (int i = 0; i < NY; ++i) {
                d[i] = 0.0;
                yk[i] = ypred[i];
            }
            for (int k = 0; k < bdfc::NEWTON_MAXITER; ++k) {
                n_iter = k + 1;
                const int st = eval_rhs(sys, a, tnew, yk, f);
                if (st != 0) break;
#pragma unroll
                for (int i = 0; i < NY; ++i) dy[i] = fma(c, f[i], -psi[i]) - d[i];
#if ODESYS_HAS_JTIMES
                if (krylov) cnt.nlu += gmres_solve(sys, tnew, ypred, c, dy);  // the nlu slot counts J v products (njvev)
                else
#endif
                    lin_solve<false>(W, piv, dy);
                cnt.nsolve += 1;
                const float nrm = rms(dy);
                if (!(nrm <= 3.0e38f)) break;  // NaN / inf from the rhs
                // contraction rate of the simplified Newton iteration; the tests are the usual ones
                // (rate >= 1, or rate^(MAXITER-k)/(1-rate) |dy| > tol: will not converge in time) without divisions
                float rate = -1.0f;
                const bool tiny = nrm <= newton_floor;
                if (nrm_old >= 0.0f) rate = nrm * rcp_ftz(nrm_old);
                if (!tiny && rate >= 0.0f) {
                    float rp = rate;
                    if (k <= 2) rp *= rate;
                    if (k <= 1) rp *= rate;
                    if (rate >= 1.0f || rp * nrm > newton_tol * (1.0f - rate)) break;
                }
#pragma unroll
                for (int i = 0; i < NY; ++i) {
                    yk[i] += dy[i];
                    d[i] += dy[i];
                }
                if (tiny || (rate >= 0.0f && rate * nrm < newton_tol * (1.0f - rate))) {
                    converged = true;
                    break;
                }
                nrm_old = nrm;
            }
            if (converged) {
#pragma unroll
                for (int i = 0; i < NY; ++i) ynew[i] = yk[i];
            } else if (krylov || jac_current) {
                break;
            } else {
                double fp[NY];
                int bad = eval_rhs(sys, a, tnew, ypred, fp);
                bad |= eval_jac(sys, tnew, ypred);
                if (bad) return 1;
                jac_current = true;
                have_lu = false;
            }
        }
        if (!converged) {
            newton_failed = true;
            return 0;
        }
        safety = 0.9f * (2 * bdfc::NEWTON_MAXITER + 1) / (float)(2 * bdfc::NEWTON_MAXITER + n_iter);
        set_scale(a, ynew);
        const float ef = (float)errconst(order) * rms(d);
        if (!(ef <= 3.0e38f) || !finite_all(ynew)) return 1;
        err = (double)ef;
        return 0;
    }
    __device__ __forceinline__ bool control(double err, double, double &hnext, double h) {
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
        // (shared memory is addressed by the order directly; the running sums stay in registers)
        double carry[NY];
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            const double di = d[i];
            Dx(order + 2, i) = di - Dx(order + 1, i);
            Dx(order + 1, i) = di;
            carry[i] = di;
        }
#pragma unroll
        for (int j = bdfc::MAXORD; j >= 0; --j) {
            if (j <= order) {  // one (mostly warp-uniform) branch per row instead of selects per element
#pragma unroll
                for (int i = 0; i < NY; ++i) {
                    carry[i] += Dx(j, i);
                    Dx(j, i) = carry[i];
                }
            }
        }
        hnext = h;
        // selection only on the grid of attempts (odesys_prelude.h; half the period while the order is low)
        if (n_equal < order + 1 || n_attempt % (order <= 2 ? ODESYS_BDF_EPOCH / 2 : ODESYS_BDF_EPOCH) != 0) return true;
        // error estimates one order down / up: errconst_{k-1} D[k], errconst_{k+1} D[k+2]
        double vm[NY], vp[NY];
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            vm[i] = 0.0;
            vp[i] = 0.0;
        }
        if (order > 1) {
#pragma unroll
            for (int i = 0; i < NY; ++i) vm[i] = Dx(order, i);
        }
#pragma unroll
        for (int i = 0; i < NY; ++i) vp[i] = Dx(order + 2, i);
        const float inf = __int_as_float(0x7f800000);
        float fm = -1.0f, fp = -1.0f;
        if (order > 1) {
            const float e = (float)errconst(order - 1) * rms(vm);
            fm = e == 0.0f ? inf : fast_powf(e, -1.0f / (float)order);
        }
        if (order < bdfc::MAXORD) {
            const float e = (float)errconst(order + 1) * rms(vp);
            fp = e == 0.0f ? inf : fast_powf(e, -1.0f / (float)(order + 2));
        }
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
    __device__ __forceinline__ void accepted(const double *) {}
    // Newton backward interpolation through the last order+1 solution values; theta in (0, 1] of the last step
    __device__ __forceinline__ void dense(double theta, double, const double *, const double *, double *out) {
#pragma unroll
        for (int i = 0; i < NY; ++i) out[i] = Dx(0, i);
        double p = 1.0;
#pragma unroll
        for (int j = 0; j < bdfc::MAXORD; ++j) {
            p *= (theta - 1.0 + j) * (1.0 / (1 + j));
            if (j < order) {
#pragma unroll
                for (int i = 0; i < NY; ++i) out[i] = fma(Dx(j + 1, i), p, out[i]);
            }
        }
    }
};
template <> struct Stepper<M_BDF> : BdfStepper<false> {};
#if ODESYS_HAS_JTIMES
template <> struct Stepper<M_BDF_KRYLOV> : BdfStepper<true> {};
#endif

#endif  // ODESYS_HAS_JAC

}  // namespace pyodesys_cuda
