// Batched stiff integrator for sm_100a: ONE SYSTEM PER THREAD, ALL PER-SYSTEM ARRAYS IN SHARED MEMORY,
// ROLLED LOOPS.  For ny ~ 9..20.
//
// Why this file exists (measured on B200, radiolysis network ny = 14, np = 54; profiles/r1_c4_*.md):
//   * the register kernel of integrators.cuh needs W = I/(gamma h) - J (ny^2 doubles) + stage vectors +
//     parameters per thread: 255 registers + 4.4 KB of local-memory spills, L1 hit rate 22 %, spills reach
//     DRAM, FP64 pipe 6 % busy;
//   * keeping the same fully unrolled code but with W / stage vectors / parameters in shared memory removed
//     the spills, yet ran no faster: one Rosenbrock step is ~15 000 straight-line instructions (250 KB of
//     SASS against a 32 KB L1.5 instruction cache); ncu's dominant stall became "no instruction";
//   * one system per lane group (group_integrators.cuh) pays for every generated expression G times.
// So here the matrix AND the vectors live in shared memory (element e of thread t at col[e*32], conflict
// free), indices are dynamic, the linear algebra is written as small rolled loops, and the big pieces that
// are used six times per step (generated rhs, triangular solves) are __noinline__ functions: the whole step
// is a few thousand instructions that stay in the instruction cache.  One warp per block; the per-thread
// footprint decides how many warps share an SM (two for ny = 14).
//
// Same algorithm as Stepper<M_ROS4> (odeint rosenbrock4 + controller + dense output; LU with partial
// pivoting as AnyODE::getrf_callback/getrs_callback, anyode_blasless.hpp:10-95, pivots applied LAPACK style).
// Requires integrators.cuh first.
#pragma once

namespace pyodesys_cuda {

#if ODESYS_HAS_JAC

constexpr int kSB = 32;  // threads per block = stride between consecutive elements of one thread
// per-thread column layout, in doubles
constexpr int S_W = 0;                     // W[NY][NY] row-major, becomes its LU factors
constexpr int S_Y = NY * NY;               // vectors of length NY:
constexpr int S_YNEW = S_Y + NY;
constexpr int S_F = S_YNEW + NY;
constexpr int S_G1 = S_F + NY;             // g1..g5 contiguous
constexpr int S_TMP = S_G1 + 5 * NY;
constexpr int S_DFDT = S_TMP + NY;
constexpr int S_PIV = S_DFDT + NY;         // pivot rows (stored as doubles)
constexpr int S_MP = S_PIV + NY;           // parameters
constexpr int S_CSE = S_MP + (NP_USED > 0 ? NP_USED : 1);
constexpr int kSmemDoublesPerThread = S_CSE + (ODESYS_NPCSE > 0 ? ODESYS_NPCSE : 1);

struct SCol {  // array-like view of one run of a thread's column: v[i] = col[(off + i) * 32]
    double *p;
    __device__ __forceinline__ double &operator[](int i) const { return p[i * kSB]; }
};
typedef OdeSysT<SCol> SSys;
__device__ __forceinline__ SCol scol(double *col, int off) { return SCol{col + off * kSB}; }
__device__ __forceinline__ SSys ssys(double *col) {
    SSys s;
    s.m_p = scol(col, S_MP);
    s.m_p_cse = scol(col, S_CSE);
    s.nfev = 0;
    s.njev = 0;
    return s;
}

struct SinkVec {
    SCol f;
    __device__ __forceinline__ void set(int i, double v) { f[i] = v; }
};
struct SinkW {  // W_ij = delta_ij/(gamma h) - J_ij, stored as the generated code produces J_ij
    SCol w, dfdt;
    double diag;
    __device__ __forceinline__ void setj(int i, int j, double v) { w[i * NY + j] = (i == j ? diag : 0.0) - v; }
    __device__ __forceinline__ void setdfdt(int i, double v) { dfdt[i] = v; }
};

// f(x, y) for the vector at ``yoff`` into the vector at ``foff``; one copy of the generated code
__device__ __noinline__ int s_rhs(double *col, double x, int yoff, int foff, int error_outside_bounds) {
    double yl[NY];
    const SCol y = scol(col, yoff);
#pragma unroll
    for (int i = 0; i < NY; ++i) yl[i] = y[i];
    SSys sys = ssys(col);
    SinkVec sink{scol(col, foff)};
    int st = sys.rhs_sink(x, yl, sink);
#if ODESYS_HAS_BOUNDS
    if (error_outside_bounds && sys.outside_bounds(yl)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NONNEGATIVE
#pragma unroll
    for (int i = 0; i < NY; ++i)
        if (yl[i] < 0) st = AnyODE::Status::recoverable_error;
#endif
    AnyODE::ignore(error_outside_bounds);
    return st;
}

__device__ __noinline__ int s_jac(double *col, double x, int yoff, double diag) {
    double yl[NY];
    const SCol y = scol(col, yoff);
#pragma unroll
    for (int i = 0; i < NY; ++i) yl[i] = y[i];
    SSys sys = ssys(col);
    SinkW sink{scol(col, S_W), scol(col, S_DFDT), diag};
    return sys.dense_jac_sink(x, yl, sink);
}

// In-place LU of W with partial pivoting (rows really move: LAPACK layout, like getrf_callback)
__device__ __noinline__ bool s_lu_factor(double *col) {
    const SCol w = scol(col, S_W), piv = scol(col, S_PIV);
    bool ok = true;
#pragma unroll 1
    for (int k = 0; k < NY; ++k) {
        int p = k;
        double amax = fabs(w[k * NY + k]);
#pragma unroll 2
        for (int i = k + 1; i < NY; ++i) {
            const double v = fabs(w[i * NY + k]);
            if (v > amax) {
                amax = v;
                p = i;
            }
        }
        piv[k] = (double)p;
        if (p != k) {
#pragma unroll 2
            for (int j = 0; j < NY; ++j) {
                const double u = w[k * NY + j];
                w[k * NY + j] = w[p * NY + j];
                w[p * NY + j] = u;
            }
        }
        ok = ok && (amax > 0.0) && (amax <= DBL_MAX);
        const double inv = fast_rcp<2>(w[k * NY + k]);
        w[k * NY + k] = inv;
#pragma unroll 1
        for (int i = k + 1; i < NY; ++i) {
            const double l = w[i * NY + k] * inv;
            w[i * NY + k] = l;
#pragma unroll 4
            for (int j = k + 1; j < NY; ++j) w[i * NY + j] = fma(-l, w[k * NY + j], w[i * NY + j]);
        }
    }
    return ok;
}

// Solves L U x = P b in place for the vector at ``boff`` (getrs_callback); two accumulators per dot product
__device__ __noinline__ void s_lu_solve(double *col, int boff) {
    const SCol w = scol(col, S_W), piv = scol(col, S_PIV), b = scol(col, boff);
#pragma unroll 1
    for (int k = 0; k < NY - 1; ++k) {
        const int p = (int)piv[k];
        if (p != k) {
            const double u = b[k];
            b[k] = b[p];
            b[p] = u;
        }
    }
#pragma unroll 1
    for (int i = 1; i < NY; ++i) {
        double s0 = b[i], s1 = 0.0;
        int k = 0;
#pragma unroll 2
        for (; k + 1 < i; k += 2) {
            s0 = fma(-w[i * NY + k], b[k], s0);
            s1 = fma(-w[i * NY + k + 1], b[k + 1], s1);
        }
        if (k < i) s0 = fma(-w[i * NY + k], b[k], s0);
        b[i] = s0 + s1;
    }
#pragma unroll 1
    for (int i = NY - 1; i >= 0; --i) {
        double s0 = b[i], s1 = 0.0;
        int j = i + 1;
#pragma unroll 2
        for (; j + 1 < NY; j += 2) {
            s0 = fma(-w[i * NY + j], b[j], s0);
            s1 = fma(-w[i * NY + j + 1], b[j + 1], s1);
        }
        if (j < NY) s0 = fma(-w[i * NY + j], b[j], s0);
        b[i] = (s0 + s1) * w[i * NY + i];
    }
}

// Rosenbrock coefficients as stage tables (same numbers as kRos4; uniform dynamic indexing from constant memory)
__constant__ double kRosA[6][5] = {
    {0, 0, 0, 0, 0},
    {0.1544000000000000e+01, 0, 0, 0, 0},
    {0.9466785280815826e+00, 0.2557011698983284e+00, 0, 0, 0},
    {0.3314825187068521e+01, 0.2896124015972201e+01, 0.9986419139977817e+00, 0, 0},
    {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950e+00, 0},
    {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950e+00, 1.0}};
__constant__ double kRosC[6][5] = {
    {0, 0, 0, 0, 0},
    {-0.5668800000000000e+01, 0, 0, 0, 0},
    {-0.2430093356833875e+01, -0.2063599157091915e+00, 0, 0, 0},
    {-0.1073529058151375e+00, -0.9594562251023355e+01, -0.2047028614809616e+02, 0, 0},
    {0.7496443313967647e+01, -0.1024680431464352e+02, -0.3399990352819905e+02, 0.1170890893206160e+02, 0},
    {0.8083246795921522e+01, -0.7981132988064893e+01, -0.3152159432874371e+02, 0.1631930543123136e+02,
     -0.6058818238834054e+01}};
__constant__ double kRosCt[6] = {0, 0.386, 0.21, 0.63, 1, 1};
__constant__ double kRosD[6] = {0.25, -0.1043, 0.1035, -0.3620000000000023e-01, 0, 0};
__constant__ double kRosDense[2][5] = {
    {0.1012623508344586e+02, -0.7487995877610167e+01, -0.3480091861555747e+02, -0.7992771707568823e+01,
     0.1025137723295662e+01},
    {-0.6762803392801253e+00, 0.6087714651680015e+01, 0.1643084320892478e+02, 0.2476722511418386e+02,
     -0.6594389125716872e+01}};

struct SmemRos4 {
    double h_old;
    float err_old;
    bool first, last_rejected;
    int nfev, njev;

    __device__ __forceinline__ void begin() {
        first = true;
        last_rejected = false;
        h_old = 0.0;
        err_old = 1.0f;
    }

    // Hairer's starting step (initial_step of integrators.cuh) on the shared-memory vectors
    __device__ __forceinline__ double auto_step(double *col, const KernelArgs &a, double t0, double span) {
        const SCol y = scol(col, S_Y), f0 = scol(col, S_F), y1 = scol(col, S_TMP), f1 = scol(col, S_YNEW);
        s_rhs(col, t0, S_Y, S_F, a.error_outside_bounds);
        nfev += 1;
        double d0 = 0.0, d1 = 0.0;
#pragma unroll 1
        for (int i = 0; i < NY; ++i) {
            const double sk = a.atol[i] + a.rtol * fabs(y[i]);
            d0 = fmax(d0, fabs(y[i]) / sk);
            d1 = fmax(d1, fabs(f0[i]) / sk);
        }
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, fabs(span));
        const double dir = span >= 0 ? 1.0 : -1.0;
#pragma unroll 1
        for (int i = 0; i < NY; ++i) y1[i] = fma(dir * h0, f0[i], y[i]);
        const int st = s_rhs(col, t0 + dir * h0, S_TMP, S_YNEW, a.error_outside_bounds);
        nfev += 1;
        double d2 = 0.0;
#pragma unroll 1
        for (int i = 0; i < NY; ++i) {
            const double sk = a.atol[i] + a.rtol * fabs(y[i]);
            d2 = fmax(d2, fabs(f1[i] - f0[i]) / sk);
        }
        d2 /= h0;
        if (st != 0 || !(d2 <= DBL_MAX)) d2 = 0.0;
        const double dmax = fmax(d1, d2);
        const double h1 = (dmax <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dmax, 0.25);
        return fmin(fmin(100.0 * h0, h1), fabs(span)) * dir;
    }

    // one trial step from S_Y to S_YNEW
    __device__ __forceinline__ int attempt(double *col, const KernelArgs &a, double t, double h, double &err,
                                           Counters &cnt) {
        const SCol y = scol(col, S_Y), ynew = scol(col, S_YNEW), f = scol(col, S_F), tmp = scol(col, S_TMP),
                   dfdt = scol(col, S_DFDT);
        int bad = 0;
        bad |= s_rhs(col, t, S_Y, S_F, a.error_outside_bounds);
        const double ih = fast_rcp<2>(h);
        bad |= s_jac(col, t, S_Y, 4.0 * ih /* 1/gamma */);
        nfev += 1;
        njev += 1;
        if (!s_lu_factor(col)) bad |= 1;
        cnt.nlu += 1;
        cnt.nsolve += 6;
#pragma unroll 1
        for (int s = 0; s < 6; ++s) {
            if (s > 0) {
                // tmp = y + sum_j a_sj g_j ; f = f(t + c_s h, tmp)
#pragma unroll 1
                for (int i = 0; i < NY; ++i) {
                    double acc = y[i];
                    for (int j = 0; j < s; ++j) acc = fma(kRosA[s][j], col[(S_G1 + j * NY + i) * kSB], acc);
                    tmp[i] = acc;
                }
                bad |= s_rhs(col, fma(kRosCt[s], h, t), S_TMP, S_F, a.error_outside_bounds);
                nfev += 1;
            }
            // rhs of the stage system: f + (1/h) sum_j c_sj g_j + h d_s f_t, solved in place
            const int dst = (s < 5) ? S_G1 + s * NY : S_F;
            const double hd = ODESYS_AUTONOMOUS ? 0.0 : h * kRosD[s];
#pragma unroll 1
            for (int i = 0; i < NY; ++i) {
                double acc = 0.0;
                for (int j = 0; j < s; ++j) acc = fma(kRosC[s][j], col[(S_G1 + j * NY + i) * kSB], acc);
                double v = fma(ih, acc, f[i]);
                if (!ODESYS_AUTONOMOUS) v = fma(hd, dfdt[i], v);
                col[(dst + i) * kSB] = v;
            }
            s_lu_solve(col, dst);
        }
        // S_F now holds the last increment = error estimate
        double e = 0.0;
        bool finite = true;
#pragma unroll 1
        for (int i = 0; i < NY; ++i) {
            const double xe = f[i];
            const double yn = tmp[i] + xe;
            ynew[i] = yn;
            finite = finite && (fabs(yn) <= DBL_MAX);
            const double sk = fma(a.rtol, fmax(fabs(y[i]), fabs(yn)), a.atol[i]) + 1e-300;
            e = fmax(e, fabs(xe) * fast_rcp<1>(sk));
        }
        err = e;
        if (!(e <= DBL_MAX) || !finite) bad |= 1;
        return bad;
    }

    __device__ __forceinline__ bool control(double err, double &hnext, double h) {
        constexpr float isafe = 1.0f / 0.9f, fac1 = 5.0f, fac2 = 1.0f / 6.0f;
        const float ef = (float)fmin(fmax(err, 1e-30), 1e30);
        const float lg = __log2f(ef);
        float fac = fmaxf(fac2, fminf(fac1, exp2f(0.25f * lg) * isafe));
        if (err <= 1.0) {
            if (!first) {
                const float ratio = (float)(h_old * fast_rcp<1>(h));
                float fp = ratio * exp2f(0.25f * (2.0f * lg - __log2f(err_old))) * isafe;
                fp = fmaxf(fac2, fminf(fac1, fp));
                fac = fmaxf(fac, fp);
            }
            double hn = h * (double)__frcp_rn(fac);
            first = false;
            h_old = h;
            err_old = fmaxf(0.01f, ef);
            if (last_rejected) hn = (h >= 0.0) ? fmin(hn, h) : fmax(hn, h);
            last_rejected = false;
            hnext = hn;
            return true;
        }
        last_rejected = true;
        hnext = h * (double)__frcp_rn(fac);
        return false;
    }

    // rosenbrock4_dense_output for component i at theta in (0, 1)
    __device__ __forceinline__ double dense(double *col, double theta, int i) {
        double c3 = 0.0, c4 = 0.0;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const double gj = col[(S_G1 + j * NY + i) * kSB];
            c3 = fma(kRosDense[0][j], gj, c3);
            c4 = fma(kRosDense[1][j], gj, c4);
        }
        const double s1 = 1.0 - theta;
        return col[(S_Y + i) * kSB] * s1 + theta * (col[(S_YNEW + i) * kSB] + s1 * (c3 + theta * c4));
    }
};

// ----------------------------------------------------------------------------------------------
// Persistent kernel (one warp per block).  Control flow = integrate_kernel (integrators.cuh).
// ----------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kSB) smem_ros4_kernel(const KernelArgs a) {
    extern __shared__ double s_smem[];  // [kSmemDoublesPerThread][32]
    double *col = s_smem + threadIdx.x;
    const SCol y = scol(col, S_Y), ynew = scol(col, S_YNEW), mp = scol(col, S_MP);
    const unsigned lane = threadIdx.x & 31u;
    const int64_t nthreads = (int64_t)gridDim.x * kSB;
    int64_t s = (int64_t)blockIdx.x * kSB + threadIdx.x - nthreads;

    SmemRos4 st;
    st.nfev = 0;
    st.njev = 0;
    double t = 0, h = 0, tend = 0, dir = 1, hmin = 0, hmax = INFINITY;
    int64_t steps_left = 0;
    int k_out = 0, restarts = 0, nfail = 0;
    Counters cnt = {0, 0, 0, 0};
    bool active = false, exhausted = false;

    for (;;) {
        // ---- (A) fetch a new system for every idle lane ---------------------------------------------
        if (a.work_counter != nullptr) {
            const unsigned need = __ballot_sync(FULL_MASK, !active && !exhausted);
            if (need) {
                const int leader = __ffs(need) - 1;
                unsigned long long base = 0;
                if ((int)lane == leader) base = atomicAdd(a.work_counter, (unsigned long long)__popc(need));
                base = __shfl_sync(FULL_MASK, base, leader);
                if (!active && !exhausted) s = (int64_t)base + __popc(need & ((1u << lane) - 1u));
            }
        } else if (!active && !exhausted) {
            s += nthreads;
        }
        if (!active && !exhausted) {
            if (s >= a.n) {
                exhausted = true;
            } else {
                bool fin = true;
#pragma unroll 1
                for (int i = 0; i < NY; ++i) {
                    const double v = a.y0[s * NY + i];
                    y[i] = v;
                    fin = fin && (fabs(v) <= DBL_MAX);
                }
#pragma unroll 1
                for (int i = 0; i < NP_USED; ++i) {
                    if (i >= ODESYS_NPARAMS && a.iv_from_y0) mp[i] = a.y0[s * NY + (i - ODESYS_NPARAMS)];
                    else mp[i] = a.params[s * a.params_ld + i];
                }
                {
                    SSys sys = ssys(col);
                    sys.setup();
                }
                cnt = {0, 0, 0, 0};
                st.nfev = 0;
                st.njev = 0;
                restarts = a.autorestart;
                nfail = fin ? 0 : (1 << 20);
                if (MODE == 0) {
                    const double *xo = a.xout + s * a.xout_stride;
                    t = xo[0];
                    tend = xo[a.nt - 1];
#pragma unroll 1
                    for (int i = 0; i < NY; ++i) {
                        if (a.yout_soa) a.yout[(int64_t)i * a.n + s] = y[i];
                        else a.yout[(s * a.nt) * NY + i] = y[i];
                    }
                } else {
                    t = a.x0[s];
                    tend = a.xend[s];
                    double *row = a.xyout + (s * a.max_rows) * (1 + NY);
                    row[0] = t;
#pragma unroll 1
                    for (int i = 0; i < NY; ++i) row[1 + i] = y[i];
                }
                k_out = 1;
                dir = (tend >= t) ? 1.0 : -1.0;
                hmin = a.dx_min ? a.dx_min[s] : 0.0;
                hmax = a.dx_max ? a.dx_max[s] : INFINITY;
                if (!(hmax > 0.0)) hmax = INFINITY;
                steps_left = a.mxsteps > 0 ? a.mxsteps : INT64_MAX;
                st.begin();
                h = a.dx0 ? a.dx0[s] : 0.0;
#if ODESYS_HAS_FIRST_STEP
                if (h == 0.0) {
                    double yl[NY];
#pragma unroll
                    for (int i = 0; i < NY; ++i) yl[i] = y[i];
                    SSys sys = ssys(col);
                    h = sys.get_dx0(t, yl);
                }
#endif
                if (h != 0.0) h = fabs(h) * dir;
                else h = st.auto_step(col, a, t, tend - t);
                active = true;
                if (tend == t && MODE == 0) {
                    for (; k_out < a.nt; ++k_out)
#pragma unroll 1
                        for (int i = 0; i < NY; ++i) {
                            if (a.yout_soa) a.yout[((int64_t)k_out * NY + i) * a.n + s] = y[i];
                            else a.yout[(s * a.nt + k_out) * NY + i] = y[i];
                        }
                }
            }
        }
        if (__all_sync(FULL_MASK, !active)) break;
        if (!active) continue;

        // ---- (B) termination / failure bookkeeping -----------------------------------------------------
        int status = -1;
        if (nfail >= (1 << 20)) status = ST_RHS;
        else if (t == tend) status = ST_OK;
        else if (steps_left <= 0) {
            if (restarts > 0) {
                restarts -= 1;
                steps_left = a.mxsteps;
                st.begin();
                h = st.auto_step(col, a, t, tend - t);
            } else {
                status = ST_MXSTEPS;
            }
        }

        if (status < 0) {
            // ---- (C) one trial step ----------------------------------------------------------------------
            double hh = h;
            if (fabs(hh) > hmax) hh = dir * hmax;
#if ODESYS_HAS_DX_MAX
            if (a.get_dx_max_factor > 0.0) {
                double yl[NY];
#pragma unroll
                for (int i = 0; i < NY; ++i) yl[i] = y[i];
                SSys sys = ssys(col);
                const double m = a.get_dx_max_factor * sys.get_dx_max(t, yl);
                if (fabs(hh) > m) hh = dir * m;
            }
#endif
            if (fabs(hh) < hmin) hh = dir * hmin;
            const double rem = tend - t;
            const bool lands = fabs(hh) >= fabs(rem) * (1.0 - 1e-13);
            double tnew = t + hh;
            if (lands) {
                hh = rem;
                tnew = tend;
            }
            double err = 0.0, hnext = hh;
            const int bad = st.attempt(col, a, t, hh, err, cnt);
            cnt.nsteps += 1;
            steps_left -= 1;
            bool ok;
            if (bad) {
                ok = false;
                hnext = 0.25 * hh;
                nfail += 1;
                if (nfail >= 12) nfail = 1 << 20;
                double ignored;
                st.control(2.0, ignored, hh);
            } else {
                ok = st.control(err, hnext, hh);
                if (ok) nfail = 0;
            }
            if (!ok) {
                cnt.nreject += 1;
                if (fabs(hnext) <= 16.0 * DBL_EPSILON * fabs(t) || fabs(hnext) == 0.0 || (hmin > 0.0 && fabs(hh) <= hmin))
                    status = ST_STEP_TOO_SMALL;
                h = hnext;
            } else {
                // ---- (D) accepted: emit outputs -----------------------------------------------------------
                if (MODE == 0) {
                    const double *xo = a.xout + s * a.xout_stride;
                    while (k_out < a.nt) {
                        const double xk = xo[k_out];
                        if ((xk - tnew) * dir > 0.0) break;
                        const bool at_end = (xk == tnew);
                        const double theta = (xk - t) / hh;
#pragma unroll 1
                        for (int i = 0; i < NY; ++i) {
                            const double v = at_end ? ynew[i] : st.dense(col, theta, i);
                            if (a.yout_soa) a.yout[((int64_t)k_out * NY + i) * a.n + s] = v;
                            else a.yout[(s * a.nt + k_out) * NY + i] = v;
                        }
                        k_out += 1;
                        steps_left = a.mxsteps > 0 ? a.mxsteps : INT64_MAX;
                    }
                } else {
                    if (k_out >= a.max_rows) {
                        status = ST_ROWS;
                    } else {
                        double *row = a.xyout + (s * a.max_rows + k_out) * (1 + NY);
                        row[0] = tnew;
#pragma unroll 1
                        for (int i = 0; i < NY; ++i) row[1 + i] = ynew[i];
                        k_out += 1;
                    }
                }
#pragma unroll 1
                for (int i = 0; i < NY; ++i) y[i] = ynew[i];
                t = tnew;
                h = hnext;
            }
        }

        if (status >= 0) {
            // ---- (E) retire the system ---------------------------------------------------------------------
            a.status[s] = status;
            if (MODE == 0) {
                a.nreached[s] = k_out;
                if (status != ST_OK) {
                    const double qnan = nan("");
                    for (int k = k_out; k < a.nt; ++k)
#pragma unroll 1
                        for (int i = 0; i < NY; ++i) {
                            if (a.yout_soa) a.yout[((int64_t)k * NY + i) * a.n + s] = qnan;
                            else a.yout[(s * a.nt + k) * NY + i] = qnan;
                        }
                }
            } else {
                a.nrows[s] = k_out;
            }
            if (a.counters) {
                a.counters[C_NSTEPS * a.counters_ld + s] = cnt.nsteps;
                a.counters[C_NFEV * a.counters_ld + s] = st.nfev;
                a.counters[C_NJEV * a.counters_ld + s] = st.njev;
                a.counters[C_NLU * a.counters_ld + s] = cnt.nlu;
                a.counters[C_NSOLVE * a.counters_ld + s] = cnt.nsolve;
                a.counters[C_NREJECT * a.counters_ld + s] = cnt.nreject;
            }
            active = false;
        }
    }
}

#endif  // ODESYS_HAS_JAC

}  // namespace pyodesys_cuda
