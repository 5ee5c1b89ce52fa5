// Batched ODE integrators for sm_100a: ONE SYSTEM PER THREAD, state in registers.
//
// This file replaces the hot loop of the reference's batched native path: the OpenMP loop over
// systems inside ``multi_predefined<OdeSys>`` / ``multi_adaptive<OdeSys>`` (call sites
// pyodesys/native/sources/_odeint_wrapper.pyx:89-92,163-166; _cvode_wrapper.pyx:190-197,339-345),
// where each iteration owns one integrator instance and calls ``OdeSys::rhs`` /
// ``dense_jac_rmaj`` through virtual functions.  Here every lane of a persistent grid owns one
// system at a time; when a lane finishes it fetches the next index (grid-stride or a
// warp-aggregated atomic counter = work stealing), so divergent step counts do not idle the SM.
//
// The steppers restate published algorithms (the reference delegates them to Boost.odeint via
// pyodeint, which is not vendored in the reference tree):
//   * RK4      - classical fixed step (pyodesys/integrators.py:17-61 is the in-tree oracle);
//   * DOPRI5   - Dormand & Prince 5(4) with FSAL, Hairer's 4th order dense output, odeint's
//                ``controlled_runge_kutta`` error measure and step-size rule;
//   * ROS4     - six-stage stiffly accurate Rosenbrock 4(3) (odeint ``rosenbrock4``, the default
//                ``method`` of _odeint_wrapper.pyx:43,116) with its predictive controller and
//                dense output; the per-system ny x ny LU lives in registers.
// Requires the generated header (struct odesys_anyode::OdeSys, ODESYS_* macros) to be included
// first.  float64 throughout; no tensor cores (per-system matrices are 2x2 .. 8x8 with pivoting).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <float.h>

namespace pyodesys_cuda {

using odesys_anyode::OdeSys;
using odesys_anyode::OdeSysT;
constexpr int NY = ODESYS_NY;
constexpr int NP_TOT = ODESYS_NP_TOT;
constexpr int NP_USED = ODESYS_NP_USED;
constexpr unsigned FULL_MASK = 0xffffffffu;

enum : int { M_RK4 = 0, M_DOPRI5 = 1, M_ROS4 = 2, M_BDF = 3, M_BDF_KRYLOV = 4 };  // 4: BDF, Newton corrections by GMRES on jtimes
enum : int { ST_OK = 0, ST_MXSTEPS = 1, ST_STEP_TOO_SMALL = 2, ST_RHS = 3, ST_ROWS = 4 };
enum : int { C_NSTEPS = 0, C_NFEV = 1, C_NJEV = 2, C_NLU = 3, C_NSOLVE = 4, C_NREJECT = 5, NCOUNT = 6 };

struct KernelArgs {
    int64_t n;
    int nt;                 // predefined: number of output points per system
    const double *y0;       // [n, NY]
    const double *params;   // [n, params_ld]  (params_ld = NP_TOT, or the user's np columns with iv_from_y0)
    int64_t params_ld;
    int iv_from_y0;         // m_p[np + i] = y0[i] instead of a host-side concatenation (core.py:410-411)
    const double *xout;     // predefined: [n, nt] or [nt] (xout_stride == 0)
    int64_t xout_stride;
    const double *x0;       // adaptive: [n]
    const double *xend;     // adaptive: [n]
    const double *dx0, *dx_min, *dx_max;  // [n] or nullptr
    double atol[NY];
    double rtol;
    int64_t mxsteps;
    int autorestart;
    double get_dx_max_factor;
    int error_outside_bounds;
    double max_invariant_violation;  // != 0: rhs reports a recoverable error when an invariant drifts (template:191-198)
    int linear_solver;      // BDF: 0 = dense LU of I - c J, 1 = matrix-free GMRES on OdeSys::jtimes (no Jacobian formed)
    // event functions (OdeSys::roots, template:377-396): located on the dense output of every accepted step
    int max_roots;          // capacity per system of root_out / root_indices (0 = no root finding)
    int return_on_root;     // adaptive mode: the first root ends the integration (its row is the last one)
    double *root_out;       // [n, max_roots, 1 + NY]: (x_root, y(x_root))
    int32_t *root_indices;  // [n, max_roots]: adaptive: row of xyout; predefined: index of the next output point
    int32_t *nroots_found;  // [n]
    // outputs
    double *yout;           // predefined: [n, nt, NY]  (or [nt, NY, n] when yout_soa)
    int yout_soa;
    double *xyout;          // adaptive: [n, max_rows, 1+NY]
    int64_t max_rows;
    int32_t *nrows;         // adaptive
    int32_t *nreached;      // predefined
    int32_t *status;
    int32_t *counters;      // [NCOUNT, counters_ld] or nullptr
    int64_t counters_ld;    // row length of ``counters`` (= n of the whole call; a chunk sees a column window)
    unsigned long long *work_counter;  // dynamic scheduling (nullptr = static grid-stride)
    // Streaming host-pointer calls (one persistent launch, copies run underneath on a second stream):
    const unsigned long long *ready;   // device word: inputs of systems [0, *ready) have arrived (nullptr: all of them)
    unsigned int *slice_done;          // device [ceil(n / slice_len)]: retired systems per output slice (nullptr: off)
    unsigned int *slice_flags;         // host-mapped [nslices]: set to 1 once a slice is complete and visible
    int64_t slice_len;
    unsigned int *abort_flag;          // host-mapped word: != 0 ends a streaming call early (1: host, 2: input time-out)
    unsigned long long wait_timeout_ns;
};

// leader lane: block until the inputs of systems [0, hi) are in device memory.  Returns false when the call was
// aborted: the host raised ``abort_flag`` (its copy thread failed) or nothing arrived for ``wait_timeout_ns`` of wall
// clock (%globaltimer) - the kernel then retires without drawing further systems and the call returns an error; the
// CUDA context stays usable (a trap would have killed every CUDA user of the process).
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ bool wait_inputs(const unsigned long long *ready, volatile unsigned int *abort_flag,
                                            unsigned long long timeout_ns, unsigned long long hi) {
    const volatile unsigned long long *r = ready;
    if (*r < hi) {
        const unsigned long long t0 = global_ns();
        unsigned spin = 0;
        while (*r < hi) {
            __nanosleep(200);
            if ((++spin & 255u) == 0u) {  // the flag lives in host-mapped memory: look at it every ~50 us only
                if (*abort_flag != 0u) return false;
                if (global_ns() - t0 > timeout_ns) {
                    *abort_flag = 2u;  // tells the host why, and every other waiting warp to stop as well
                    return false;
                }
            }
        }
    }
    __threadfence();  // loads of the inputs stay behind the flag read
    return true;
}

// a system of an output slice retired: the last one of the slice tells the host, which then copies the slice back
__device__ __forceinline__ void slice_retire(const KernelArgs &a, int64_t s) {
    __threadfence();  // this system's results before the count
    const int64_t j = s / a.slice_len;
    const int64_t lo = j * a.slice_len;
    const unsigned int len = (unsigned int)((lo + a.slice_len > a.n ? a.n : lo + a.slice_len) - lo);
    if (atomicAdd(a.slice_done + j, 1u) + 1u == len) {
        __threadfence_system();
        *(volatile unsigned int *)(a.slice_flags + j) = 1u;
    }
}

struct Counters {
    int nsteps, nlu, nsolve, nreject;
};

struct Tol {
    double rtol;
    double atol[NY];
};

// ----------------------------------------------------------------------------------------------
// small helpers
// ----------------------------------------------------------------------------------------------
// Integer view of |x|'s high word: exponent and top 20 mantissa bits.  Used where only the magnitude class
// matters (finite tests, pivot search) so that the work runs on the integer pipe, not on the FP64 pipe that
// bounds the kernel.
#ifndef PYODESYS_INT_CHECKS
#define PYODESYS_INT_CHECKS 0  /* measured on B200: 20.96 ms with, 20.05 ms without (LOP3 + ISETP replace one DSETP) */
#endif
__device__ __forceinline__ int hi_abs(double x) { return __double2hiint(x) & 0x7fffffff; }
// fixed-point log2 of |x|: biased exponent << 18 | top 18 mantissa bits (three of them can be added in 32 bits)
__device__ __forceinline__ int lg_fix(double x) { return (__double2hiint(x) >> 2) & 0x1fffffff; }
__device__ __forceinline__ bool finite_all(const double *v) {
    bool ok = true;
#if PYODESYS_INT_CHECKS
#pragma unroll
    for (int i = 0; i < NY; ++i) ok = ok && (hi_abs(v[i]) < 0x7ff00000);
#else
#pragma unroll
    for (int i = 0; i < NY; ++i) ok = ok && (fabs(v[i]) <= DBL_MAX);
#endif
    return ok;
}

// x^(1/4), x^(-1/5), x^(-1/3) for step-size factors.  The factor only steers the step size, it
// never enters the solution, so single precision SFU math (ex2/lg2) is ample.
// Single-instruction SFU forms (flush-to-zero variants: no denormal pre/post-scaling code around the MUFU).
__device__ __forceinline__ float lg2_ftz(float x) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float ex2_ftz(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rcp_ftz(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_powf(float x, float e) {
    return ex2_ftz(e * lg2_ftz(fminf(fmaxf(x, 1e-30f), 1e30f)));
}
__device__ __forceinline__ float fast_powf(double x, float e) {
    const float xf = fminf(fmaxf((float)x, 1e-30f), 1e30f);
    return ex2_ftz(e * lg2_ftz(xf));
}
__device__ __forceinline__ double fast_pow(double x, float e) { return (double)fast_powf(x, e); }

// Reciprocal without the IEEE division's special-case slow path: MUFU.RCP64H seed (>= 20 bits) and
// NEWTON Newton-Raphson steps (1: ~40 bits - step-size control; 2: full double to ~1 ulp - LU pivots,
// 1/h).  Inputs are finite, non-zero and normal wherever this is used (checked by the callers'
// ``ok`` / finite tests); a zero gives inf/NaN like the division would.
template <int NEWTON> __device__ __forceinline__ double fast_rcp(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
#pragma unroll
    for (int k = 0; k < NEWTON; ++k) {
        const double e = fma(-x, r, 1.0);
        r = fma(r, e, r);
    }
    return r;
}

template <class Sys>
__device__ __forceinline__ AnyODE::Status eval_rhs(Sys &sys, const KernelArgs &a, double x, const double *y,
                                                   double *f) {
    // Every call counts as one evaluation (the reference's template increments nfev before its
    // bound checks, odesys_anyode_template.cpp:161).  The counter is advanced HERE rather than by the
    // generated/user body: with a body that skips ``nfev++`` on its error path (the override snippet of
    // pyodesys/native/tests/_tests.py:324-333) nvcc 12.9 -O3 folded the whole step into "always fails".
    const int nfev0 = sys.nfev;
    AnyODE::Status st = sys.rhs(x, y, f);
    sys.nfev = nfev0 + 1;
#if ODESYS_HAS_BOUNDS
    if (a.error_outside_bounds && sys.outside_bounds(y)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NINVAR > 0
    if (sys.invariant_violated(x, y)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NONNEGATIVE
#pragma unroll
    for (int i = 0; i < NY; ++i)
        if (y[i] < 0) st = AnyODE::Status::recoverable_error;
#endif
    return st;
}

template <int N> struct RegArr {
    double v[N];
    __device__ __forceinline__ double &operator[](int i) { return v[i]; }
    __device__ __forceinline__ const double &operator[](int i) const { return v[i]; }
};

// In-register LU with partial (row) pivoting, row-major a[N*N]; the reciprocal pivots are kept
// on the diagonal so the six triangular solves per Rosenbrock step are division free.
// Same factorisation as AnyODE::getrf_callback (anyode_blasless.hpp:10-60): Doolittle, row
// pivot = largest magnitude in the column (first one wins on ties).
// Row interchanges are chains of 64-bit selects (2 FSEL each, ~14 % of the executed instructions for
// ny = 3).  Guarding them with a warp vote (skip when no lane swaps - the common case for stiff
// kinetics) was measured SLOWER on B200 (20.3 -> 21.4..22.5 ms per 1e6 Robertson systems): the
// branches cost more issue slots and scheduling freedom than the selects they remove.
template <int N, bool BRANCH_SWAP, class Mat> __device__ __forceinline__ bool lu_factor(Mat &a, int *piv) {
    bool ok = true;
#pragma unroll
    for (int k = 0; k < N; ++k) {
        int p = k;
#if PYODESYS_INT_CHECKS
        // magnitudes compared on the high words (sign stripped): candidates that differ only below the top 20
        // mantissa bits are equally good pivots, and the compare runs on the integer pipe
        int amax = hi_abs(a[k * N + k]);
#pragma unroll
        for (int i = k + 1; i < N; ++i) {
            const int v = hi_abs(a[i * N + k]);
            if (v > amax) {
                amax = v;
                p = i;
            }
        }
#else
        double amax = fabs(a[k * N + k]);
#pragma unroll
        for (int i = k + 1; i < N; ++i) {
            const double v = fabs(a[i * N + k]);
            if (v > amax) {
                amax = v;
                p = i;
            }
        }
#endif
        piv[k] = p;
        if (BRANCH_SWAP) {
            // shared-memory matrix: a real (rare) interchange instead of selects on every element
            if (p != k) {
#pragma unroll
                for (int i = k + 1; i < N; ++i) {
                    if (p == i) {
#pragma unroll
                        for (int j = 0; j < N; ++j) {
                            const double u = a[k * N + j];
                            a[k * N + j] = a[i * N + j];
                            a[i * N + j] = u;
                        }
                    }
                }
            }
        } else {
#pragma unroll
            for (int i = k + 1; i < N; ++i) {
                const bool sw = (p == i);
#pragma unroll
                for (int j = 0; j < N; ++j) {
                    const double u = a[k * N + j], w = a[i * N + j];
                    a[k * N + j] = sw ? w : u;
                    a[i * N + j] = sw ? u : w;
                }
            }
        }
#if PYODESYS_INT_CHECKS
        ok = ok && (amax >= 0x00100000) && (amax < 0x7ff00000);  // normal, finite pivot
#else
        ok = ok && (amax > 0.0) && (amax <= DBL_MAX);
#endif
        const double inv = fast_rcp<2>(a[k * N + k]);
        a[k * N + k] = inv;
        // The pivot row is read once into registers.  Every row below it is staged: all its loads, the
        // arithmetic, all its stores - and the loads of row i+1 are issued before the stores of row i (the
        // shared-memory variant runs one warp per scheduler, so the ILP has to come from the code itself;
        // for the register variant the staging arrays vanish).
        double prow[N], cur[N], nxt[N];
#pragma unroll
        for (int j = k + 1; j < N; ++j) prow[j] = a[k * N + j];
        if (k + 1 < N) {
#pragma unroll
            for (int j = k; j < N; ++j) cur[j] = a[(k + 1) * N + j];
        }
#pragma unroll
        for (int i = k + 1; i < N; ++i) {
            if (i + 1 < N) {
#pragma unroll
                for (int j = k; j < N; ++j) nxt[j] = a[(i + 1) * N + j];
            }
            const double l = cur[k] * inv;
            a[i * N + k] = l;
#pragma unroll
            for (int j = k + 1; j < N; ++j) a[i * N + j] = fma(-l, prow[j], cur[j]);
#pragma unroll
            for (int j = k; j < N; ++j) cur[j] = nxt[j];
        }
    }
    return ok;
}

// Solves L U x = P b in place (anyode_blasless.hpp:63-95).
template <int N, bool BRANCH_SWAP, class Mat>
__device__ __forceinline__ void lu_solve(const Mat &a, const int *piv, double *b) {
    // the multipliers were swapped together with their rows (LAPACK convention), hence all row
    // interchanges are applied to b before the forward substitution starts
#pragma unroll
    for (int k = 0; k < N; ++k) {
#pragma unroll
        for (int i = k + 1; i < N; ++i) {
            const bool sw = (piv[k] == i);
            const double u = b[k], w = b[i];
            b[k] = sw ? w : u;
            b[i] = sw ? u : w;
        }
    }
#pragma unroll
    for (int k = 0; k < N; ++k) {
#pragma unroll
        for (int i = k + 1; i < N; ++i) b[i] = fma(-a[i * N + k], b[k], b[i]);
    }
#pragma unroll
    for (int i = N - 1; i >= 0; --i) {
        double s = b[i];
#pragma unroll
        for (int j = i + 1; j < N; ++j) s = fma(-a[i * N + j], b[j], s);
        b[i] = s * a[i * N + i];
    }
}

// Hairer's starting step-size heuristic (dopri5.f / rodas.f ``HINIT``) with the max norm that the
// controllers below use.  ``f0`` = f(t0, y0).  Costs one rhs evaluation.
template <class Sys>
__device__ __forceinline__ double initial_step(Sys &sys, const KernelArgs &a, const Tol &tol, double t0,
                                               const double *y, const double *f0, double span, int order) {
    // Division free: this runs once per system on the refill path, where a single lane of the warp is active
    // (ncu: IEEE divisions with their zero-operand slow path made the refill path 8 % of all issued instructions).
    // It is a heuristic - ~40-bit reciprocals are ample.
    double d0 = 0.0, d1 = 0.0, isk[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) {
        isk[i] = fast_rcp<1>(tol.atol[i] + tol.rtol * fabs(y[i]) + 1e-300);
        d0 = fmax(d0, fabs(y[i]) * isk[i]);
        d1 = fmax(d1, fabs(f0[i]) * isk[i]);
    }
    double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 * fast_rcp<1>(d1);
    h0 = fmin(h0, fabs(span));
    const double dir = span >= 0 ? 1.0 : -1.0;
    double y1[NY], f1[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) y1[i] = fma(dir * h0, f0[i], y[i]);
    AnyODE::Status st = eval_rhs(sys, a, t0 + dir * h0, y1, f1);
    double d2 = 0.0;
#pragma unroll
    for (int i = 0; i < NY; ++i) d2 = fmax(d2, fabs(f1[i] - f0[i]) * isk[i]);
    d2 *= fast_rcp<1>(h0);
    if (st != AnyODE::Status::success || !(d2 <= DBL_MAX)) d2 = 0.0;
    const double dmax = fmax(d1, d2);
    // the starting step is a heuristic: single precision exp2/log2 instead of the double precision pow
    const double h1 = (dmax <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : fast_pow(0.01 * fast_rcp<1>(dmax), 1.0f / (float)order);
    return fmin(fmin(100.0 * h0, h1), fabs(span));
}

// ----------------------------------------------------------------------------------------------
// Steppers.  Common protocol (all inlined into the persistent kernel):
//   begin(...)      once per system (and after an autorestart)
//   attempt(...)    one trial step t -> t+h producing ynew and the scaled error norm
//   control(...)    accept/reject + next step size
//   dense(...)      interpolate inside the last accepted step
// ----------------------------------------------------------------------------------------------
template <int METHOD> struct Stepper;

// ---- classical RK4, fixed step -----------------------------------------------------------------
template <> struct Stepper<M_RK4> {
    typedef OdeSys Sys;
    static constexpr bool kAdaptive = false;
    static constexpr int kOrder = 4;
    static constexpr int kSharedDoublesPerThread = 0;
    static constexpr int kEpoch = 0;  // no start grid (see BdfStepper)
    __device__ __forceinline__ void bind_shared(double *, int, int) {}
    __device__ __forceinline__ int begin(OdeSys &, const KernelArgs &, const Tol &, double, const double *) { return 0; }
    __device__ __forceinline__ void auto_step(OdeSys &, const KernelArgs &, const Tol &, double, const double *,
                                              double, double &) {}  // the nominal step never changes
    // Same evaluation order as RK4_example_integrator.integrate_predefined (integrators.py:44-61).
    __device__ __forceinline__ int attempt(OdeSys &sys, const KernelArgs &a, const Tol &, double t, const double *y,
                                           double h, double *ynew, double &err, Counters &) {
        double k1[NY], k2[NY], k3[NY], k4[NY], tmp[NY];
        int bad = 0;
        bad |= eval_rhs(sys, a, t, y, k1);
#pragma unroll
        for (int i = 0; i < NY; ++i) tmp[i] = y[i] + h / 2 * k1[i];
        bad |= eval_rhs(sys, a, t + h / 2, tmp, k2);
#pragma unroll
        for (int i = 0; i < NY; ++i) tmp[i] = y[i] + h / 2 * k2[i];
        bad |= eval_rhs(sys, a, t + h / 2, tmp, k3);
#pragma unroll
        for (int i = 0; i < NY; ++i) tmp[i] = y[i] + h * k3[i];
        bad |= eval_rhs(sys, a, t + h, tmp, k4);
#pragma unroll
        for (int i = 0; i < NY; ++i) ynew[i] = y[i] + h / 6 * (k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i]);
        err = 0.0;
        return bad;
    }
    __device__ __forceinline__ bool control(double, double, double &hnext, double h) {
        hnext = h;
        return true;
    }
    __device__ __forceinline__ void accepted(const double *) {}
    __device__ __forceinline__ void dense(double theta, double, const double *yold, const double *ynew, double *out) {
#pragma unroll
        for (int i = 0; i < NY; ++i) out[i] = yold[i] + theta * (ynew[i] - yold[i]);  // only hit at theta == 1
    }
};

// ---- Dormand-Prince 5(4) -----------------------------------------------------------------------
// Coefficients in constant memory (an FP64 instruction takes a constant-bank operand for free; as 64-bit
// literals they cost two UMOV each - 14 % of the executed instructions in profiles/r1g_c2_dopri5_ncu_summary.md).
namespace dp5c {
enum { A21, A31, A32, A41, A42, A43, A51, A52, A53, A54, A61, A62, A63, A64, A65, B1, B3, B4, B5, B6,
       E1, E3, E4, E5, E6, E7, D1, D3, D4, D5, D6, D7, C2, C3, C4, C5, NCOEF };
}
__constant__ double kDp5[dp5c::NCOEF] = {
    0.2, 3.0 / 40.0, 9.0 / 40.0, 44.0 / 45.0, -56.0 / 15.0, 32.0 / 9.0,
    19372.0 / 6561.0, -25360.0 / 2187.0, 64448.0 / 6561.0, -212.0 / 729.0,
    9017.0 / 3168.0, -355.0 / 33.0, 46732.0 / 5247.0, 49.0 / 176.0, -5103.0 / 18656.0,
    35.0 / 384.0, 500.0 / 1113.0, 125.0 / 192.0, -2187.0 / 6784.0, 11.0 / 84.0,
    71.0 / 57600.0, -71.0 / 16695.0, 71.0 / 1920.0, -17253.0 / 339200.0, 22.0 / 525.0, -1.0 / 40.0,
    -12715105075.0 / 11282082432.0, 87487479700.0 / 32700410799.0, -10690763975.0 / 1880347072.0,
    701980252875.0 / 199316789632.0, -1453857185.0 / 822651844.0, 69997945.0 / 29380423.0,
    0.2, 0.3, 0.8, 8.0 / 9.0};
#define DC(name) kDp5[dp5c::name]
template <> struct Stepper<M_DOPRI5> {
    typedef OdeSys Sys;
    static constexpr bool kAdaptive = true;
    static constexpr int kOrder = 5;
    static constexpr int kSharedDoublesPerThread = 0;
    static constexpr int kEpoch = 0;  // no start grid (see BdfStepper)
    __device__ __forceinline__ void bind_shared(double *, int, int) {}
    double k1[NY];                                   // f(t, y), FSAL
    double k3[NY], k4[NY], k5[NY], k6[NY];           // stages of the last attempt (dense output forms from them)
    double k7[NY];                                   // f(t+h, ynew) of the last attempt

    __device__ __forceinline__ int begin(OdeSys &sys, const KernelArgs &a, const Tol &, double t, const double *y) {
        return eval_rhs(sys, a, t, y, k1);
    }
    __device__ __forceinline__ void auto_step(OdeSys &sys, const KernelArgs &a, const Tol &tol, double t,
                                              const double *y, double span, double &h) {
        h = initial_step(sys, a, tol, t, y, k1, span, 5) * (span >= 0 ? 1.0 : -1.0);
    }
    __device__ __forceinline__ int attempt(OdeSys &sys, const KernelArgs &a, const Tol &tol, double t, const double *y,
                                           double h, double *ynew, double &err, Counters &) {
        double k2[NY], tmp[NY];
        int bad = 0;
#pragma unroll
        for (int i = 0; i < NY; ++i) tmp[i] = fma(h, DC(A21) * k1[i], y[i]);
        bad |= eval_rhs(sys, a, fma(DC(C2), h, t), tmp, k2);
#pragma unroll
        for (int i = 0; i < NY; ++i) tmp[i] = fma(h, fma(DC(A32), k2[i], DC(A31) * k1[i]), y[i]);
        bad |= eval_rhs(sys, a, fma(DC(C3), h, t), tmp, k3);
#pragma unroll
        for (int i = 0; i < NY; ++i)
            tmp[i] = fma(h, fma(DC(A43), k3[i], fma(DC(A42), k2[i], DC(A41) * k1[i])), y[i]);
        bad |= eval_rhs(sys, a, fma(DC(C4), h, t), tmp, k4);
#pragma unroll
        for (int i = 0; i < NY; ++i)
            tmp[i] = fma(h, fma(DC(A54), k4[i], fma(DC(A53), k3[i], fma(DC(A52), k2[i], DC(A51) * k1[i]))), y[i]);
        bad |= eval_rhs(sys, a, fma(DC(C5), h, t), tmp, k5);
#pragma unroll
        for (int i = 0; i < NY; ++i)
            tmp[i] = fma(h, fma(DC(A65), k5[i], fma(DC(A64), k4[i], fma(DC(A63), k3[i],
                            fma(DC(A62), k2[i], DC(A61) * k1[i])))), y[i]);
        bad |= eval_rhs(sys, a, t + h, tmp, k6);
#pragma unroll
        for (int i = 0; i < NY; ++i)
            ynew[i] = fma(h, fma(DC(B6), k6[i], fma(DC(B5), k5[i], fma(DC(B4), k4[i],
                             fma(DC(B3), k3[i], DC(B1) * k1[i])))), y[i]);
        bad |= eval_rhs(sys, a, t + h, ynew, k7);
        // error estimate and odeint's measure: |err_i| / (atol + rtol (|y_i| + |h| |y'_i|))
        // the error measure only steers the step size: formed in single precision (see Ros4Stepper::attempt)
        float ef = 0.0f;
        const double ah = fabs(h);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            const double e = h * fma(DC(E7), k7[i], fma(DC(E6), k6[i], fma(DC(E5), k5[i], fma(DC(E4), k4[i],
                                     fma(DC(E3), k3[i], DC(E1) * k1[i])))));
            const double sc = fma(tol.rtol, fma(ah, fabs(k1[i]), fabs(y[i])), tol.atol[i]);
            ef = fmaxf(ef, fabsf((float)e) * rcp_ftz(fmaxf((float)sc, 1e-37f)));
        }
        if (!(ef <= 3.0e38f)) bad |= 1;
        err = (double)ef;
        return bad;
    }
    // odeint default_step_adjuster: shrink by max(0.9 err^-1/3, 1/5); grow (only if err < 0.5) by
    // 0.9 err^-1/5 capped at 5.
    __device__ __forceinline__ bool control(double err, double, double &hnext, double h) {
        const float ef = (float)err;
        if (ef > 1.0f) {
            hnext = h * (double)fmaxf(0.9f * ex2_ftz((-1.0f / 3.0f) * lg2_ftz(fminf(ef, 1e30f))), 0.2f);
            return false;
        }
        hnext = h;
        if (ef < 0.5f) {
            const float e = fmaxf(3.2e-4f /* 5^-5 */, ef);
            hnext = h * (double)fminf(0.9f * ex2_ftz(-0.2f * lg2_ftz(e)), 5.0f);
        }
        return true;
    }
    __device__ __forceinline__ void accepted(const double *) {
#pragma unroll
        for (int i = 0; i < NY; ++i) k1[i] = k7[i];
    }
    // continuous extension (Hairer, Norsett, Wanner: Solving ODEs I, II.6; dopri5.f CONTD5), formed from the stages
    // only when an output point (or an event) falls inside the step - not in every attempt
    __device__ __forceinline__ void dense(double theta, double h, const double *yold, const double *ynew, double *out) {
        const double th1 = 1.0 - theta;
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            const double ydiff = ynew[i] - yold[i];
            const double r3 = fma(h, k1[i], -ydiff);
            const double r4 = ydiff - fma(h, k7[i], r3);
            const double r5 = h * fma(DC(D7), k7[i], fma(DC(D6), k6[i], fma(DC(D5), k5[i], fma(DC(D4), k4[i],
                                      fma(DC(D3), k3[i], DC(D1) * k1[i])))));
            out[i] = yold[i] + theta * (ydiff + th1 * (r3 + theta * (r4 + th1 * r5)));
        }
    }
};
#undef DC

#if ODESYS_HAS_JAC
// ---- Rosenbrock 4(3), six stages (RODAS-type; coefficients proven by oracle/check_ros4_coeffs.py) ---
// The coefficients live in constant memory: FP64 instructions take a constant-bank operand for free,
// whereas 64-bit literals cost two UMOV each (14 % of all executed instructions in the first ncu capture).
namespace ros4c {
enum { A21, A31, A32, A41, A42, A43, A51, A52, A53, A54, C21, C31, C32, C41, C42, C43, C51, C52, C53, C54,
       C61, C62, C63, C64, C65, D1, D2, D3, D4, C2X, C3X, C4X, D21, D22, D23, D24, D25, D31, D32, D33, D34, D35,
       NCOEF };
}
__constant__ double kRos4[ros4c::NCOEF] = {
    0.1544000000000000e+01, 0.9466785280815826e+00, 0.2557011698983284e+00, 0.3314825187068521e+01,
    0.2896124015972201e+01, 0.9986419139977817e+00, 0.1221224509226641e+01, 0.6019134481288629e+01,
    0.1253708332932087e+02, -0.6878860361058950e+00,
    -0.5668800000000000e+01, -0.2430093356833875e+01, -0.2063599157091915e+00, -0.1073529058151375e+00,
    -0.9594562251023355e+01, -0.2047028614809616e+02, 0.7496443313967647e+01, -0.1024680431464352e+02,
    -0.3399990352819905e+02, 0.1170890893206160e+02, 0.8083246795921522e+01, -0.7981132988064893e+01,
    -0.3152159432874371e+02, 0.1631930543123136e+02, -0.6058818238834054e+01,
    0.2500000000000000e+00, -0.1043000000000000e+00, 0.1035000000000000e+00, -0.3620000000000023e-01,
    0.386, 0.21, 0.63,
    0.1012623508344586e+02, -0.7487995877610167e+01, -0.3480091861555747e+02, -0.7992771707568823e+01,
    0.1025137723295662e+01, -0.6762803392801253e+00, 0.6087714651680015e+01, 0.1643084320892478e+02,
    0.2476722511418386e+02, -0.6594389125716872e+01};
#define RC(name) kRos4[ros4c::name]

// The linear algebra of one implicit step: one factorisation of W, several solves (six for Rosenbrock-4, one per
// Newton iteration for BDF).
//
// ny == 3 (register kernel): W^-1 is formed explicitly and the six solves are 3x3 matrix-vector products
// (9 independent FMAs, depth 3) instead of six permuted forward/back substitutions (9 FP64 + 12 selects each,
// depth ~10): the serial substitution chains and the select-based row interchanges were the dominant
// latency and issue-slot cost of the step (profiles/r1b_ros4_ncu_summary.md).  The inverse comes from the
// cofactor formula when that is safe: the determinant's cancellation is measured (|det| against the sum of
// the magnitudes of its three terms) and has to keep the relative error of W^-1 a factor 100 below rtol;
// otherwise this lane falls back to the pivoted LU (AnyODE::getrf semantics) and builds W^-1 from three
// unit-vector solves.
#ifndef PYODESYS_ROS4_INV3
#define PYODESYS_ROS4_INV3 1
#endif
#ifndef PYODESYS_INV3_INTCHECK
#define PYODESYS_INV3_INTCHECK 1  /* conditioning test of the cofactor inverse on the integer pipe (0: in double) */
#endif
template <bool BRANCH_SWAP, class Mat>
__device__ __forceinline__ bool lin_factor(Mat &W, int *piv, double rtol, Counters &cnt) {
    if constexpr (PYODESYS_ROS4_INV3 && NY == 3 && !BRANCH_SWAP) {
        const double a00 = W[0], a01 = W[1], a02 = W[2], a10 = W[3], a11 = W[4], a12 = W[5], a20 = W[6],
                     a21 = W[7], a22 = W[8];
        const double c00 = fma(a11, a22, -a12 * a21), c01 = fma(a12, a20, -a10 * a22), c02 = fma(a10, a21, -a11 * a20);
#if PYODESYS_INV3_INTCHECK
        // Is the cofactor formula accurate enough?  Its rounding error relative to |W^-1| is ~ eps * mag / |det| with
        // mag = the sum of the magnitudes of the six triple products of the determinant (this includes the
        // cancellation inside the 2x2 cofactors, not only that of the final three-term sum - a cofactor that cancels
        // makes W^-1 inaccurate while |t0| + |t1| + |t2| still looks harmless; round-1 advisor finding).  The test
        // only needs magnitude classes, so it runs on the integer pipe: lg(x) = (exponent | top 18 mantissa bits) is a
        // fixed-point log2 (piecewise linear, error < 0.09 bit), products are sums, the six-term sum is replaced by
        // its largest term (at most 2.6 bits too small); 4 bits of margin cover both (checked against the double
        // formula on 4e5 random and nearly singular matrices: never "safe" where the double test says "unsafe").
        // The FP64 pipe, which bounds this kernel, sees none of it.
        const double det = fma(a00, c00, fma(a01, c01, a02 * c02));
        const int v00 = lg_fix(a00), v01 = lg_fix(a01), v02 = lg_fix(a02), v10 = lg_fix(a10), v11 = lg_fix(a11),
                  v12 = lg_fix(a12), v20 = lg_fix(a20), v21 = lg_fix(a21), v22 = lg_fix(a22);
        const int m0 = v00 + max(v11 + v22, v12 + v21), m1 = v01 + max(v12 + v20, v10 + v22),
                  m2 = v02 + max(v10 + v21, v11 + v20);
        const int vmag = max(m0, max(m1, m2)), vdet = lg_fix(det);
        // log2(|det| / mag) >= log2(100 eps / rtol) + 4, all in units of 2^-18 bit; 2 * (1023 << 18) puts the single
        // factor det on the scale of a triple product.  Also false for det = 0, denormal, inf, NaN.
        const int need = (int)((-45.36f + 4.0f - lg2_ftz((float)rtol)) * 262144.0f);
        const bool safe = (vdet + 2 * (1023 << 18) - vmag >= need) && vdet >= (1 << 18) && vdet < (0x7ff00000 >> 2);
#else
        const double t0 = a00 * c00, t1 = a01 * c01, t2 = a02 * c02;
        const double det = t0 + t1 + t2;
        // the magnitudes of all six triple products: includes the cancellation inside the 2x2 cofactors
        const double mag = fabs(a00) * (fabs(a11 * a22) + fabs(a12 * a21)) + fabs(a01) * (fabs(a12 * a20) + fabs(a10 * a22)) +
                           fabs(a02) * (fabs(a10 * a21) + fabs(a11 * a20));
        // eps * mag / |det| <= rtol / 100  (also false for det = 0, NaN, inf)
        const bool safe = fabs(det) * rtol >= (100.0 * DBL_EPSILON) * mag && fabs(det) <= DBL_MAX && mag > 0.0;
#endif
        bool ok = true;
        if (__builtin_expect(!safe, 0)) {
            ok = lu_factor<NY, false>(W, piv);
            double e[NY], inv[NY * NY];
#pragma unroll
            for (int j = 0; j < NY; ++j) {
#pragma unroll
                for (int i = 0; i < NY; ++i) e[i] = (i == j) ? 1.0 : 0.0;
                lu_solve<NY, false>(W, piv, e);
#pragma unroll
                for (int i = 0; i < NY; ++i) inv[i * NY + j] = e[i];
            }
#pragma unroll
            for (int i = 0; i < NY * NY; ++i) W[i] = inv[i];
            cnt.nsolve += NY;
        } else {
            const double r = fast_rcp<2>(det);
            W[0] = c00 * r;
            W[3] = c01 * r;
            W[6] = c02 * r;
            W[1] = fma(a02, a21, -a01 * a22) * r;
            W[4] = fma(a00, a22, -a02 * a20) * r;
            W[7] = fma(a01, a20, -a00 * a21) * r;
            W[2] = fma(a01, a12, -a02 * a11) * r;
            W[5] = fma(a02, a10, -a00 * a12) * r;
            W[8] = fma(a00, a11, -a01 * a10) * r;
        }
        return ok;
    } else {
        return lu_factor<NY, BRANCH_SWAP>(W, piv);
    }
}
template <bool BRANCH_SWAP, class Mat>
__device__ __forceinline__ void lin_solve(const Mat &W, const int *piv, double *b) {
    if constexpr (PYODESYS_ROS4_INV3 && NY == 3 && !BRANCH_SWAP) {
        const double b0 = b[0], b1 = b[1], b2 = b[2];
        b[0] = fma(W[0], b0, fma(W[1], b1, W[2] * b2));
        b[1] = fma(W[3], b0, fma(W[4], b1, W[5] * b2));
        b[2] = fma(W[6], b0, fma(W[7], b1, W[8] * b2));
    } else {
        lu_solve<NY, BRANCH_SWAP>(W, piv, b);
    }
}

// ---- storage: everything in registers (ny <= ~8); larger systems use smem_integrators.cuh / group_integrators.cuh
struct RegStore {
    typedef OdeSys Sys;
    typedef RegArr<NY> Vec;
    typedef RegArr<NY * NY> Mat;
    static constexpr bool kBranchSwap = false;
    Vec g1, g2, g3, g4, g5;
    __device__ __forceinline__ void attach(double *, Sys &) {}
    __device__ __forceinline__ Mat matrix() const { return Mat(); }
};
// receives J_ij from the generated code and stores W_ij = delta_ij/(gamma h) - J_ij right away
template <class Mat> struct JacSink {
    Mat &w;
    double *dfdt;
    double diag;
    __device__ __forceinline__ void setj(int i, int j, double v) { w[i * NY + j] = (i == j ? diag : 0.0) - v; }
    __device__ __forceinline__ void setdfdt(int i, double v) { dfdt[i] = v; }
};

template <class Store> struct Ros4Stepper : Store {
    typedef typename Store::Sys Sys;
    static constexpr bool kAdaptive = true;
    static constexpr int kOrder = 4;
    static constexpr int kSharedDoublesPerThread = 0;
    static constexpr int kEpoch = 0;  // no start grid (see BdfStepper)
    __device__ __forceinline__ void bind_shared(double *, int, int) {}
    using Store::g1;
    using Store::g2;
    using Store::g3;
    using Store::g4;
    using Store::g5;
    float h_old, lg_err_old;                        // predictive (Gustafsson) controller memory (single precision)
    bool first, last_rejected;

    __device__ __forceinline__ int begin(Sys &, const KernelArgs &, const Tol &, double, const double *) {
        first = true;
        last_rejected = false;
        h_old = 0.0f;
        lg_err_old = 0.0f;
        return 0;
    }
    template <class Mat> __device__ __forceinline__ bool lin_factor(Mat &W, int *piv, double rtol, Counters &cnt) {
        return pyodesys_cuda::lin_factor<Store::kBranchSwap>(W, piv, rtol, cnt);
    }
    template <class Mat> __device__ __forceinline__ void lin_solve(const Mat &W, const int *piv, double *b) {
        pyodesys_cuda::lin_solve<Store::kBranchSwap>(W, piv, b);
    }
    __device__ __forceinline__ void auto_step(Sys &sys, const KernelArgs &a, const Tol &tol, double t,
                                              const double *y, double span, double &h) {
        double f0[NY];
        eval_rhs(sys, a, t, y, f0);
        h = initial_step(sys, a, tol, t, y, f0, span, 4) * (span >= 0 ? 1.0 : -1.0);
    }
    __device__ __forceinline__ int attempt(Sys &sys, const KernelArgs &a, const Tol &tol, double t, const double *y,
                                           double h, double *ynew, double &err, Counters &cnt) {
        typename Store::Mat W = this->matrix();
        double dfdt[NY], f[NY], tmp[NY];
        int piv[NY];
        int bad = 0;
        bad |= eval_rhs(sys, a, t, y, f);
        // W = I/(gamma h) - J, formed entry by entry as the generated code produces J
        const double ih = fast_rcp<2>(h);
        {
            JacSink<typename Store::Mat> sink{W, dfdt, 4.0 * ih /* 1/gamma = 4 */};
            const int njev0 = sys.njev;
            bad |= sys.dense_jac_sink(t, y, sink);
            sys.njev = njev0 + 1;
        }
        if (!lin_factor(W, piv, tol.rtol, cnt)) bad |= 1;
        cnt.nlu += 1;
        cnt.nsolve += 6;
        double b[NY];
#pragma unroll
        for (int i = 0; i < NY; ++i) b[i] = ODESYS_AUTONOMOUS ? f[i] : fma(h * RC(D1), dfdt[i], f[i]);
        lin_solve(W, piv, b);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            g1[i] = b[i];
            tmp[i] = fma(RC(A21), b[i], y[i]);
        }
        bad |= eval_rhs(sys, a, fma(RC(C2X), h, t), tmp, f);
        const double c21 = RC(C21) * ih;
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            b[i] = fma(c21, g1[i], f[i]);
            if (!ODESYS_AUTONOMOUS) b[i] = fma(h * RC(D2), dfdt[i], b[i]);
        }
        lin_solve(W, piv, b);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            g2[i] = b[i];
            tmp[i] = fma(RC(A32), b[i], fma(RC(A31), g1[i], y[i]));
        }
        bad |= eval_rhs(sys, a, fma(RC(C3X), h, t), tmp, f);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            b[i] = fma(ih, fma(RC(C32), g2[i], RC(C31) * g1[i]), f[i]);
            if (!ODESYS_AUTONOMOUS) b[i] = fma(h * RC(D3), dfdt[i], b[i]);
        }
        lin_solve(W, piv, b);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            g3[i] = b[i];
            tmp[i] = fma(RC(A43), b[i], fma(RC(A42), g2[i], fma(RC(A41), g1[i], y[i])));
        }
        bad |= eval_rhs(sys, a, fma(RC(C4X), h, t), tmp, f);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            b[i] = fma(ih, fma(RC(C43), g3[i], fma(RC(C42), g2[i], RC(C41) * g1[i])), f[i]);
            if (!ODESYS_AUTONOMOUS) b[i] = fma(h * RC(D4), dfdt[i], b[i]);
        }
        lin_solve(W, piv, b);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            g4[i] = b[i];
            tmp[i] = fma(RC(A54), b[i], fma(RC(A53), g3[i], fma(RC(A52), g2[i], fma(RC(A51), g1[i], y[i]))));
        }
        bad |= eval_rhs(sys, a, t + h, tmp, f);
#pragma unroll
        for (int i = 0; i < NY; ++i)
            b[i] = fma(ih, fma(RC(C54), g4[i], fma(RC(C53), g3[i], fma(RC(C52), g2[i], RC(C51) * g1[i]))), f[i]);
        lin_solve(W, piv, b);
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            g5[i] = b[i];
            tmp[i] += b[i];
        }
        bad |= eval_rhs(sys, a, t + h, tmp, f);
#pragma unroll
        for (int i = 0; i < NY; ++i)
            b[i] = fma(ih, fma(RC(C65), g5[i], fma(RC(C64), g4[i], fma(RC(C63), g3[i],
                               fma(RC(C62), g2[i], RC(C61) * g1[i])))), f[i]);
        lin_solve(W, piv, b);
        // The error measure only steers the step size: max_i |b_i| / (atol_i + rtol max(|y_i|, |ynew_i|)) is formed
        // in single precision (a double fmax costs ~7 instructions, a double reciprocal 4 FP64-pipe slots; the
        // FP64 pipe bounds this kernel).  max(|y|, |ynew|) is decided on the high words.
        float ef = 0.0f;
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            ynew[i] = tmp[i] + b[i];
            const double ymax = (hi_abs(ynew[i]) > hi_abs(y[i])) ? fabs(ynew[i]) : fabs(y[i]);
            const float sk = fmaxf((float)fma(tol.rtol, ymax, tol.atol[i]), 1e-37f);
            ef = fmaxf(ef, fabsf((float)b[i]) * rcp_ftz(sk));
        }
        if (!(ef <= 3.0e38f)) bad |= 1;  // NaN or overflow: handled like a failed evaluation
        err = (double)ef;
        return bad;
    }
    // odeint rosenbrock4_controller::try_step: fac in [1/6, 5], predictive correction after the
    // first accepted step, no growth directly after a rejection.  Single precision: the factor only
    // chooses the next step size.
    __device__ __forceinline__ bool control(double err, double, double &hnext, double h) {
        constexpr float isafe = 1.0f / 0.9f, fac1 = 5.0f, fac2 = 1.0f / 6.0f;
        const float ef = fminf(fmaxf((float)err, 1e-30f), 1e30f);
        const float lg = lg2_ftz(ef);
        float fac = fmaxf(fac2, fminf(fac1, ex2_ftz(0.25f * lg) * isafe));
        if (err <= 1.0) {
            const float hf = (float)h;
            if (!first) {
                // (h_old/h) (err^2/err_old)^(1/4) / safe
                float fp = h_old * rcp_ftz(hf) * ex2_ftz(0.25f * (2.0f * lg - lg_err_old)) * isafe;
                fp = fmaxf(fac2, fminf(fac1, fp));
                fac = fmaxf(fac, fp);
            }
            float rfac = rcp_ftz(fac);
            if (last_rejected) rfac = fminf(rfac, 1.0f);  // no growth directly after a rejection
            first = false;
            h_old = hf;
            lg_err_old = fmaxf(-6.643856f /* log2(0.01) */, lg);  // log2(max(0.01, err))
            last_rejected = false;
            hnext = h * (double)rfac;
            return true;
        }
        last_rejected = true;
        hnext = h * (double)rcp_ftz(fac);
        return false;
    }
    __device__ __forceinline__ void accepted(const double *) {}
    // rosenbrock4_dense_output: the two continuation vectors are linear in the stage increments; they
    // are formed here (only when an output point falls inside the step) instead of in every attempt.
    __device__ __forceinline__ void dense(double theta, double, const double *yold, const double *ynew, double *out) {
        const double s1 = 1.0 - theta;
#pragma unroll
        for (int i = 0; i < NY; ++i) {
            const double c3 = fma(RC(D25), g5[i], fma(RC(D24), g4[i], fma(RC(D23), g3[i], fma(RC(D22), g2[i], RC(D21) * g1[i]))));
            const double c4 = fma(RC(D35), g5[i], fma(RC(D34), g4[i], fma(RC(D33), g3[i], fma(RC(D32), g2[i], RC(D31) * g1[i]))));
            out[i] = yold[i] * s1 + theta * (ynew[i] + s1 * (c3 + theta * c4));
        }
    }
};
template <> struct Stepper<M_ROS4> : Ros4Stepper<RegStore> {};  // instantiated for ny <= 16 only (shim)
#undef RC
#endif  // ODESYS_HAS_JAC

// ----------------------------------------------------------------------------------------------
// Event functions (OdeSys::roots, odesys_anyode_template.cpp:377-396; consumers _cvode_wrapper.pyx:89,206-213,
// 354-355).  After every accepted step the signs of g(t, y) at both ends are compared; a change is located on the
// step's dense output with a bracketing secant iteration (Illinois weights, the scheme CVODE's rootfinding
// documents: tolerance 100 eps (|t| + |h|), the component whose secant root is earliest drives the iteration, the
// reported root is the upper end of the final bracket).  Several roots inside one step are found in order.
// ----------------------------------------------------------------------------------------------
#if ODESYS_NROOTS > 0
constexpr int NROOTS = ODESYS_NROOTS;
__device__ __forceinline__ bool root_crossed(double g_from, double g_to) {
    return g_from != 0.0 && (g_to == 0.0 || (g_from < 0.0) != (g_to < 0.0));
}

// Returns true when the integration has to stop at a root (return_on_root): tnew / ynew are then the root.
template <int MODE, class St, class Sys>
__device__ __noinline__ bool locate_roots(St &st, Sys &sys, const KernelArgs &a, int64_t s, double t, double hh,
                                          double &tnew, const double *y, double *ynew, double *g_old, int &n_roots,
                                          int &k_out, double dir, int &status) {
    double g_lo[NROOTS], g_hi[NROOTS];
    sys.roots(tnew, ynew, g_hi);
#pragma unroll
    for (int r = 0; r < NROOTS; ++r) g_lo[r] = g_old[r];
    double lo = 0.0;
    bool stop = false;
    const double ttol = 100.0 * DBL_EPSILON * (fabs(t) + fabs(hh)) / fabs(hh);  // in units of the step
    for (int found = 0; found < 16 && !stop; ++found) {
        bool any = false;
#pragma unroll
        for (int r = 0; r < NROOTS; ++r) any = any || root_crossed(g_lo[r], g_hi[r]);
        if (!any) break;
        double ta = lo, tb = 1.0, ga[NROOTS], gb[NROOTS], yb[NY];
#pragma unroll
        for (int r = 0; r < NROOTS; ++r) {
            ga[r] = g_lo[r];
            gb[r] = g_hi[r];
        }
#pragma unroll
        for (int i = 0; i < NY; ++i) yb[i] = ynew[i];
        double alpha = 1.0;
        int side = 0, sideprev = -1;
        for (int it = 0; it < 200 && (tb - ta) > ttol; ++it) {
            double best = -1.0, gai = 0.0, gbi = 0.0;
#pragma unroll
            for (int r = 0; r < NROOTS; ++r)
                if (root_crossed(ga[r], gb[r])) {
                    const double frac = fabs(gb[r] / (gb[r] - ga[r]));
                    if (frac > best) {
                        best = frac;
                        gai = ga[r];
                        gbi = gb[r];
                    }
                }
            if (sideprev == side) alpha = (side == 2) ? alpha * 2.0 : alpha * 0.5;
            else alpha = 1.0;
            double tm = tb - (tb - ta) * gbi / (gbi - alpha * gai);
            const double width = tb - ta;
            if (!(fabs(tm - ta) >= 0.5 * ttol)) {
                const double fracint = width / ttol;
                tm = ta + (fracint > 5.0 ? 0.1 : 0.5 / fracint) * width;
            }
            if (!(fabs(tb - tm) >= 0.5 * ttol)) {
                const double fracint = width / ttol;
                tm = tb - (fracint > 5.0 ? 0.1 : 0.5 / fracint) * width;
            }
            double ym[NY], gm[NROOTS];
            st.dense(tm, hh, y, ynew, ym);
            sys.roots(fma(tm, hh, t), ym, gm);
            bool chg = false;
#pragma unroll
            for (int r = 0; r < NROOTS; ++r) chg = chg || root_crossed(ga[r], gm[r]);
            sideprev = side;
            if (chg) {
                tb = tm;
#pragma unroll
                for (int r = 0; r < NROOTS; ++r) gb[r] = gm[r];
#pragma unroll
                for (int i = 0; i < NY; ++i) yb[i] = ym[i];
                side = 1;
            } else {
                ta = tm;
#pragma unroll
                for (int r = 0; r < NROOTS; ++r) ga[r] = gm[r];
                side = 2;
            }
        }
        const bool at_end = (tb == 1.0);
        const double x_root = at_end ? tnew : fma(tb, hh, t);
        int idx = k_out;
        if (MODE == 0) {
            const double *xo = a.xout + s * a.xout_stride;
            while (idx < a.nt && (xo[idx] - x_root) * dir < 0.0) idx += 1;
        }
        const bool stored = n_roots < a.max_roots;  // roots beyond the capacity are counted only
        if (stored) {
            double *ro = a.root_out + (s * a.max_roots + n_roots) * (1 + NY);
            ro[0] = x_root;
#pragma unroll
            for (int i = 0; i < NY; ++i) ro[1 + i] = yb[i];
            a.root_indices[s * a.max_roots + n_roots] = idx;
        }
        n_roots += 1;
        if (MODE == 1 && a.return_on_root) {
            // the root becomes the end of this (shortened) step; the driver writes its row and retires the system
            tnew = x_root;
#pragma unroll
            for (int i = 0; i < NY; ++i) ynew[i] = yb[i];
            stop = true;
        } else if (MODE == 1 && !at_end && stored) {
            if (k_out >= a.max_rows) {
                status = ST_ROWS;
                stop = true;
            } else {
                double *row = a.xyout + (s * a.max_rows + k_out) * (1 + NY);
                row[0] = x_root;
#pragma unroll
                for (int i = 0; i < NY; ++i) row[1 + i] = yb[i];
                k_out += 1;
            }
        }
        // continue behind the root: components that crossed have their new sign (or are exactly zero and stay
        // disarmed until they leave zero)
        lo = tb;
#pragma unroll
        for (int r = 0; r < NROOTS; ++r) g_lo[r] = gb[r];
        if (at_end) break;
    }
#pragma unroll
    for (int r = 0; r < NROOTS; ++r) g_old[r] = g_hi[r];
    return stop && status < 0;
}
#endif  // ODESYS_NROOTS

// ----------------------------------------------------------------------------------------------
// The persistent kernel.  MODE 0: predefined output points, MODE 1: every accepted step.
// ----------------------------------------------------------------------------------------------
#ifndef PYODESYS_MIN_BLOCKS
#define PYODESYS_MIN_BLOCKS (ODESYS_NY <= 4 ? 4 : 1)
#endif
// BDF keeps its difference table and the Jacobian in shared memory; ny <= 4: 4 blocks of 128 per SM (128 registers,
// ~20 spilled values outside the Newton loop) beat 3 blocks without spills: 36.0 against 40.4 ms per 1e6 Robertson
#ifndef PYODESYS_BDF_MIN_BLOCKS
#define PYODESYS_BDF_MIN_BLOCKS (ODESYS_NY <= 4 ? 4 : 1)
#endif
#define PYODESYS_MIN_BLOCKS_FOR(METHOD) (((METHOD) == M_BDF || (METHOD) == M_BDF_KRYLOV) ? PYODESYS_BDF_MIN_BLOCKS : PYODESYS_MIN_BLOCKS)
// GUARD: compiled with the invariant test of rhs (max_invariant_violation != 0); the unguarded instantiation is the
// one every other call runs (the test costs 3.6 % on C3 even when it is switched off at run time).
template <int METHOD, int MODE, int BLOCK, bool GUARD = false>
__global__ void __launch_bounds__(BLOCK, PYODESYS_MIN_BLOCKS_FOR(METHOD)) integrate_kernel(const KernelArgs a) {
    using St = Stepper<METHOD>;
    const unsigned lane = threadIdx.x & 31u;
    const int64_t nthreads = (int64_t)gridDim.x * BLOCK;
    int64_t s = (int64_t)blockIdx.x * BLOCK + threadIdx.x - nthreads;  // static schedule cursor

    typename St::Sys sys;
    St st;
    if (St::kSharedDoublesPerThread > 0) {
        extern __shared__ double stepper_smem[];
        st.bind_shared(stepper_smem, threadIdx.x, BLOCK);
    }
    Tol tol;
    tol.rtol = a.rtol;
#pragma unroll
    for (int i = 0; i < NY; ++i) tol.atol[i] = a.atol[i];

    double y[NY], t = 0, h = 0, tend = 0, dir = 1, hmin = 0, hmax = INFINITY;
    int64_t steps_left = 0;
    int k_out = 0;        // MODE 0: next output index; MODE 1: rows written
    double x_next = 0;    // MODE 0: xout[k_out] (dir * inf once all points are written), kept in a register
    int restarts = 0, nfail = 0;
    Counters cnt = {0, 0, 0, 0};
    bool active = false, exhausted = false;
#if ODESYS_NROOTS > 0
    double g_old[NROOTS];
    int n_roots = 0;
#endif
    // St::kEpoch > 0: idle lanes draw their next system only every kEpoch-th iteration of this loop (an active lane
    // makes exactly one step attempt per iteration, so the attempts of all lanes of the warp are numbered alike
    // modulo kEpoch; a lane waits 2.5 iterations on average, of the hundreds its next system takes)
    int epoch_phase = 0;

    for (;;) {
        bool may_fetch = true;
        if (St::kEpoch > 0) {
            may_fetch = epoch_phase == 0;
            epoch_phase = epoch_phase + 1 == St::kEpoch ? 0 : epoch_phase + 1;
        }
        // ---- (A) fetch a new system for every idle lane ---------------------------------------
        if (!may_fetch) {
        } else if (a.work_counter != nullptr) {
            const unsigned need = __ballot_sync(FULL_MASK, !active && !exhausted);
            if (need) {
                const int leader = __ffs(need) - 1;
                unsigned long long base = 0;
                if ((int)lane == leader) {
                    base = atomicAdd(a.work_counter, (unsigned long long)__popc(need));
                    if (a.ready != nullptr && base < (unsigned long long)a.n) {
                        const unsigned long long hi = base + (unsigned long long)__popc(need);
                        if (!wait_inputs(a.ready, a.abort_flag, a.wait_timeout_ns,
                                         hi < (unsigned long long)a.n ? hi : (unsigned long long)a.n))
                            base = ~0ull;  // aborted: the drawn indices are dropped, the call returns an error
                    }
                }
                base = __shfl_sync(FULL_MASK, base, leader);
                if (!active && !exhausted) s = (base == ~0ull) ? a.n : (int64_t)base + __popc(need & ((1u << lane) - 1u));
            }
        } else if (!active && !exhausted) {
            s += nthreads;
        }
        if (may_fetch && !active && !exhausted) {
            if (s >= a.n) {
                exhausted = true;
            } else {
                // ---- load one system (row-major inputs; a warp reads one contiguous span) -------
#pragma unroll
                for (int i = 0; i < NY; ++i) y[i] = a.y0[s * NY + i];
#pragma unroll
                for (int i = 0; i < NP_USED; ++i) {
                    if (i >= ODESYS_NPARAMS && a.iv_from_y0) sys.m_p[i] = y[i - ODESYS_NPARAMS];
                    else sys.m_p[i] = a.params[s * a.params_ld + i];
                }
                sys.setup();
                cnt = {0, 0, 0, 0};
                restarts = a.autorestart;
                nfail = 0;
                if (MODE == 0) {
                    const double *xo = a.xout + s * a.xout_stride;
                    t = xo[0];
                    tend = xo[a.nt - 1];
                    if (a.yout_soa) {
#pragma unroll
                        for (int i = 0; i < NY; ++i) a.yout[(int64_t)i * a.n + s] = y[i];
                    } else {
#pragma unroll
                        for (int i = 0; i < NY; ++i) a.yout[(s * a.nt) * NY + i] = y[i];
                    }
                    k_out = 1;
                    x_next = a.nt > 1 ? xo[1] : INFINITY;
                } else {
                    t = a.x0[s];
                    tend = a.xend[s];
                    double *row = a.xyout + (s * a.max_rows) * (1 + NY);
                    row[0] = t;
#pragma unroll
                    for (int i = 0; i < NY; ++i) row[1 + i] = y[i];
                    k_out = 1;
                }
                dir = (tend >= t) ? 1.0 : -1.0;
                if (MODE == 0 && a.nt <= 1) x_next = dir * INFINITY;
                hmin = a.dx_min ? a.dx_min[s] : 0.0;
                hmax = a.dx_max ? a.dx_max[s] : INFINITY;
                if (!(hmax > 0.0)) hmax = INFINITY;
                steps_left = a.mxsteps > 0 ? a.mxsteps : INT64_MAX;
#if ODESYS_NINVAR > 0
                // GUARD = false: the limit is the literal 0 and the tests inside eval_rhs fold away
                sys.init_invariants(GUARD ? a.max_invariant_violation : 0.0, t, y);
#endif
#if ODESYS_NROOTS > 0
                n_roots = 0;
                if (a.max_roots > 0) sys.roots(t, y, g_old);
#endif
                int bad = st.begin(sys, a, tol, t, y);
                h = a.dx0 ? a.dx0[s] : 0.0;
#if ODESYS_HAS_FIRST_STEP
                if (h == 0.0) h = sys.get_dx0(t, y);
#endif
                if (h != 0.0) {
                    h = fabs(h) * dir;
                } else if (METHOD == M_RK4) {
                    // no step given: one step per requested interval / 100 steps over the span
                    h = (MODE == 0) ? dir * INFINITY : (tend - t) / 100.0;
                } else {
                    st.auto_step(sys, a, tol, t, y, tend - t, h);
                }
                active = true;
                if (bad || !finite_all(y)) {
                    nfail = 1 << 20;  // fail on the first pass through (D)
                }
                if (tend == t) {
                    // zero-length span: every requested point equals the start
                    if (MODE == 0) {
                        for (; k_out < a.nt; ++k_out) {
#pragma unroll
                            for (int i = 0; i < NY; ++i) {
                                if (a.yout_soa) a.yout[((int64_t)k_out * NY + i) * a.n + s] = y[i];
                                else a.yout[(s * a.nt + k_out) * NY + i] = y[i];
                            }
                        }
                        x_next = dir * INFINITY;
                    }
                }
            }
        }
        if (__all_sync(FULL_MASK, !active && exhausted)) break;
        if (!active) continue;

        // ---- (B) termination / failure bookkeeping for this lane --------------------------------
        int status = -1;  // -1: keep stepping
        if (nfail >= (1 << 20)) status = ST_RHS;
        else if (t == tend) status = ST_OK;
        else if (steps_left <= 0) {
            if (restarts > 0) {
                // autorestart (_cvode_wrapper.pyx:90): new step budget, re-initialised stepper
                restarts -= 1;
                steps_left = a.mxsteps;
                st.begin(sys, a, tol, t, y);
                st.auto_step(sys, a, tol, t, y, tend - t, h);
            } else {
                status = ST_MXSTEPS;
            }
        }

        if (status < 0) {
            // ---- (C) one trial step --------------------------------------------------------------
            double hh = h;
            if (fabs(hh) > hmax) hh = dir * hmax;
#if ODESYS_HAS_DX_MAX
            if (a.get_dx_max_factor > 0.0) {
                const double m = a.get_dx_max_factor * sys.get_dx_max(t, y);
                if (fabs(hh) > m) hh = dir * m;
            }
#endif
            if (St::kAdaptive && fabs(hh) < hmin) hh = dir * hmin;
            // Mandatory landing point: the end of the span, or (fixed-step RK4 with requested output
            // points, integrators.py:44-61) the next requested point.  Never step over it and snap onto
            // it when the remainder is a rounding artefact.
            double target = tend;
            if (METHOD == M_RK4 && MODE == 0) target = x_next;
            const double rem = target - t;
            const bool lands = fabs(hh) >= fabs(rem) * (1.0 - 1e-13);
            double tnew = t + hh;
            if (lands) {
                hh = rem;
                tnew = target;
            }
            const bool hits_end = lands && (target == tend);
            double ynew[NY], err = 0.0, hnext = hh;
            int bad = st.attempt(sys, a, tol, t, y, hh, ynew, err, cnt);
            cnt.nsteps += 1;
            steps_left -= 1;
            bool ok;
            if (bad || !finite_all(ynew)) {
                // recoverable error (AnyODE::Status::recoverable_error) or NaN: retry with h/4
                ok = false;
                hnext = 0.25 * hh;
                nfail += 1;
                if (nfail >= 12 || !St::kAdaptive) nfail = 1 << 20;
                double ignored;
                st.control(2.0, 0.0, ignored, hh);  // let the controller know a rejection happened
            } else {
                ok = st.control(err, 0.0, hnext, hh);
                if (ok) nfail = 0;
            }
            if (!ok) {
                cnt.nreject += 1;
                if (St::kAdaptive && (fabs(hnext) <= 16.0 * DBL_EPSILON * fabs(t) || fabs(hnext) == 0.0 ||
                                      (hmin > 0.0 && fabs(hh) <= hmin)))
                    status = ST_STEP_TOO_SMALL;
                h = hnext;
            } else {
                // ---- (D) accepted: emit outputs ---------------------------------------------------
                // x_next is tested from a register; everything an output needs (the address of the xout row, the
                // dense-output vectors) stays inside this branch - the compiler used to hoist the loop invariant
                // dense-output arithmetic (30 FP64 instructions for ny = 3) in front of the test, i.e. into every step.
                bool stop_at_root = false;
#if ODESYS_NROOTS > 0
                if (a.max_roots > 0)
                    stop_at_root = locate_roots<MODE>(st, sys, a, s, t, hh, tnew, y, ynew, g_old, n_roots, k_out, dir, status);
#endif
                if (MODE == 0 && (x_next - tnew) * dir <= 0.0) {
                    const double *xo = a.xout + s * a.xout_stride;
                    while (k_out < a.nt) {
                        const double xk = x_next;
                        if ((xk - tnew) * dir > 0.0) break;
                        double yo[NY];
                        if (xk == tnew) {
#pragma unroll
                            for (int i = 0; i < NY; ++i) yo[i] = ynew[i];
                        } else {
                            st.dense((xk - t) / hh, hh, y, ynew, yo);
                        }
                        if (a.yout_soa) {
#pragma unroll
                            for (int i = 0; i < NY; ++i) a.yout[((int64_t)k_out * NY + i) * a.n + s] = yo[i];
                        } else {
#pragma unroll
                            for (int i = 0; i < NY; ++i) a.yout[(s * a.nt + k_out) * NY + i] = yo[i];
                        }
                        k_out += 1;
                        x_next = k_out < a.nt ? xo[k_out] : dir * INFINITY;
                        steps_left = a.mxsteps > 0 ? a.mxsteps : INT64_MAX;  // budget is per output interval
                    }
                } else if (MODE == 1) {
                    if (k_out >= a.max_rows) {
                        status = ST_ROWS;
                    } else {
                        double *row = a.xyout + (s * a.max_rows + k_out) * (1 + NY);
                        row[0] = tnew;
#pragma unroll
                        for (int i = 0; i < NY; ++i) row[1 + i] = ynew[i];
                        k_out += 1;
                    }
                }
                st.accepted(ynew);
#pragma unroll
                for (int i = 0; i < NY; ++i) y[i] = ynew[i];
                t = tnew;
                if (hits_end && !stop_at_root) t = tend;
                if (St::kAdaptive) h = hnext;  // a fixed-step method keeps its nominal step
                if (stop_at_root && status < 0) status = ST_OK;
            }
        }

        if (status >= 0) {
            // ---- (E) retire the system --------------------------------------------------------------
            a.status[s] = status;
            if (MODE == 0) {
                a.nreached[s] = k_out;
                if (status != ST_OK) {
                    const double qnan = nan("");
                    for (int k = k_out; k < a.nt; ++k)
#pragma unroll
                        for (int i = 0; i < NY; ++i) {
                            if (a.yout_soa) a.yout[((int64_t)k * NY + i) * a.n + s] = qnan;
                            else a.yout[(s * a.nt + k) * NY + i] = qnan;
                        }
                }
            } else {
                a.nrows[s] = k_out;
            }
#if ODESYS_NROOTS > 0
            if (a.max_roots > 0) a.nroots_found[s] = n_roots;
#endif
            if (a.counters) {
                a.counters[C_NSTEPS * a.counters_ld + s] = cnt.nsteps;
                a.counters[C_NFEV * a.counters_ld + s] = sys.nfev;
                a.counters[C_NJEV * a.counters_ld + s] = sys.njev;
                a.counters[C_NLU * a.counters_ld + s] = cnt.nlu;
                a.counters[C_NSOLVE * a.counters_ld + s] = cnt.nsolve;
                a.counters[C_NREJECT * a.counters_ld + s] = cnt.nreject;
            }
            if (a.slice_done != nullptr) slice_retire(a, s);
            active = false;
        }
    }
}

// ----------------------------------------------------------------------------------------------
// Test kernels: evaluate the generated functions / the register LU for a batch (parity tests).
// ----------------------------------------------------------------------------------------------
__global__ void eval_kernel(int64_t n, const double *x, const double *y, const double *params, double *f,
                            double *jac, double *dfdt) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    OdeSys sys;
    double yy[NY], ff[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) yy[i] = y[s * NY + i];
#pragma unroll
    for (int i = 0; i < NP_TOT; ++i) sys.m_p[i] = params[s * NP_TOT + i];
    sys.setup();
    sys.rhs(x[s], yy, ff);
#pragma unroll
    for (int i = 0; i < NY; ++i) f[s * NY + i] = ff[i];
#if ODESYS_HAS_JAC
    if (jac) {
        double J[NY * NY], dd[NY];
        sys.dense_jac_rmaj(x[s], yy, ff, J, NY, dd);
#pragma unroll
        for (int i = 0; i < NY * NY; ++i) jac[s * NY * NY + i] = J[i];
        if (dfdt) {
#pragma unroll
            for (int i = 0; i < NY; ++i) dfdt[s * NY + i] = dd[i];
        }
    }
#endif
}

// Jv = J(x, y) v by the generated matrix-free product (template:214-243), one system per thread
__global__ void eval_jtimes_kernel(int64_t n, const double *x, const double *y, const double *params, const double *v,
                                   double *Jv) {
#if ODESYS_HAS_JTIMES
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    OdeSys sys;
    double yy[NY], vv[NY], out[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) {
        yy[i] = y[s * NY + i];
        vv[i] = v[s * NY + i];
    }
#pragma unroll
    for (int i = 0; i < NP_TOT; ++i) sys.m_p[i] = params[s * NP_TOT + i];
    sys.setup();
    sys.jtimes(vv, out, x[s], yy, nullptr);
#pragma unroll
    for (int i = 0; i < NY; ++i) Jv[s * NY + i] = out[i];
#endif
}

// data[n, nnz] of the CSC Jacobian (template:311-344); the index arrays are the same for every system and are
// written once by thread 0
__global__ void eval_sparse_jac_kernel(int64_t n, const double *x, const double *y, const double *params, double *data,
                                       int32_t *colptrs, int32_t *rowvals) {
#if ODESYS_NNZ >= 0
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    OdeSys sys;
    double yy[NY], dat[ODESYS_NNZ > 0 ? ODESYS_NNZ : 1];
    int32_t cp[NY + 1], rv[ODESYS_NNZ > 0 ? ODESYS_NNZ : 1];
#pragma unroll
    for (int i = 0; i < NY; ++i) yy[i] = y[s * NY + i];
#pragma unroll
    for (int i = 0; i < NP_TOT; ++i) sys.m_p[i] = params[s * NP_TOT + i];
    sys.setup();
    sys.sparse_jac_csc(x[s], yy, nullptr, dat, cp, rv);
#pragma unroll
    for (int k = 0; k < ODESYS_NNZ; ++k) data[s * ODESYS_NNZ + k] = dat[k];
    if (s == 0) {
#pragma unroll
        for (int k = 0; k <= NY; ++k) colptrs[k] = cp[k];
#pragma unroll
        for (int k = 0; k < ODESYS_NNZ; ++k) rowvals[k] = rv[k];
    }
#endif
}

#if ODESYS_NY <= 16
__global__ void lu_kernel(int64_t n, const double *A, const double *b, double *xsol, int32_t *ok) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    RegArr<NY * NY> a;
    double v[NY];
    int piv[NY];
#pragma unroll
    for (int i = 0; i < NY * NY; ++i) a[i] = A[s * NY * NY + i];
#pragma unroll
    for (int i = 0; i < NY; ++i) v[i] = b[s * NY + i];
    const bool good = lu_factor<NY, false>(a, piv);
    lu_solve<NY, false>(a, piv, v);
#pragma unroll
    for (int i = 0; i < NY; ++i) xsol[s * NY + i] = v[i];
    ok[s] = good ? 1 : 0;
}
#endif

}  // namespace pyodesys_cuda
