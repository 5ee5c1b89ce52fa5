// Batched stiff integrator for sm_100a: ONE SYSTEM PER LANE GROUP (G = 4, 8, 16 or 32 lanes of a warp).
//
// For ny beyond ~8 one thread cannot hold W = I/(gamma h) - J (ny*ny doubles) plus the stage vectors in
// registers (radiolysis, ny = 14: 255 registers + 4.4 KB of spills per thread, 2.5 % of the FP64 peak).
// Here the G lanes of a group share one system, the role the reference gives one OpenMP iteration of
// multi_predefined<OdeSys> (call sites _odeint_wrapper.pyx:163-166, _cvode_wrapper.pyx:339-345):
//   * vectors (y, stage increments, f) are distributed: variable i lives in lane i % G, slot i / G;
//   * the matrix is distributed by rows the same way, each lane keeps its R = ceil(ny/G) rows in registers;
//   * dense LU with partial pivoting (AnyODE::getrf_callback semantics, anyode_blasless.hpp:10-60) runs
//     across the group: pivot search = one REDUX over the high words of |a_ik| + a ballot, the pivot row
//     is staged through shared memory, every lane eliminates its own rows; pivoting is implicit (rows
//     never move, each row remembers its pivot order);
//   * the six triangular solves per Rosenbrock step broadcast one scalar per column with a shuffle;
//   * the error norm is a shuffle/REDUX reduction over the group;
//   * rhs / Jacobian: the generated expressions are evaluated by every lane of the group from the
//     state staged in shared memory (parameters live in shared memory too) and a "sink" keeps the
//     rows the lane owns (odesys_generated.h: rhs_sink / dense_jac_sink).
// A warp carries 32/G systems at once; groups fetch new systems from a global counter (work stealing).
// Stepper = the same Rosenbrock 4(3) as integrators.cuh (odeint ``rosenbrock4`` + controller + dense
// output); requires integrators.cuh to be included first.
#pragma once

namespace pyodesys_cuda {

using odesys_anyode::OdeSysRef;

#ifndef PYODESYS_GROUP_G
// smallest group whose lanes can keep their rows in registers (R * ny doubles per lane)
#define PYODESYS_GROUP_FITS(g) ((((NY + (g) - 1) / (g)) * NY) <= 40)
#define PYODESYS_GROUP_G (PYODESYS_GROUP_FITS(4) ? 4 : PYODESYS_GROUP_FITS(8) ? 8 : PYODESYS_GROUP_FITS(16) ? 16 : 32)
#endif

#if ODESYS_HAS_JAC

template <int G> struct Grp {
    static constexpr int R = (NY + G - 1) / G;
    unsigned mask;  // lanes of this group
    int gl;         // lane index inside the group
    int lane0;      // absolute lane of the group's first lane
    double *ys;     // [NY] staged state for rhs/jac evaluation
    double *prow;   // [NY] staged pivot row
    __device__ __forceinline__ bool owns(int r) const { return r * G + gl < NY; }
};

// shared memory per group, in doubles (odd: the 32/G groups of a warp hit distinct banks)
constexpr int kGroupSmemDoubles = (2 * NY + (NP_USED > 0 ? NP_USED : 1) + (ODESYS_NPCSE > 0 ? ODESYS_NPCSE : 1)) | 1;

template <int G> struct SinkF {
    int gl;
    double *f;
    __device__ __forceinline__ void set(int i, double v) {
        if (gl == i % G) f[i / G] = v;
    }
};
template <int G> struct SinkJ {
    int gl;
    double (*a)[NY];
    double *dfdt;
    __device__ __forceinline__ void setj(int i, int j, double v) {
        if (gl == i % G) a[i / G][j] = v;
    }
    __device__ __forceinline__ void setdfdt(int i, double v) {
        if (gl == i % G) dfdt[i / G] = v;
    }
};

// max over the group of non-negative floats (NaN wins, so callers see non-finite input)
template <int G> __device__ __forceinline__ float gmax_f(const Grp<G> &g, float v) {
    return __int_as_float(__reduce_max_sync(g.mask, __float_as_int(v)));
}
template <int G> __device__ __forceinline__ double gmax_d(const Grp<G> &g, double v) {
#pragma unroll
    for (int off = G / 2; off > 0; off /= 2) v = fmax(v, __shfl_xor_sync(g.mask, v, off));
    return v;
}

// The generated right-hand side of a network like C4 is several hundred straight-line instructions and a Rosenbrock
// step evaluates it at eight places; inlined eight times the step outgrows the instruction cache (ncu: no_instruction
// 1.5 stalls per issue, profiles/r1e_c4_group_inverse.md).  PYODESYS_GROUP_RHS_NOINLINE=1 keeps ONE copy: the state is
// already staged in shared memory, the rows a lane owns come back by value in registers.
#ifndef PYODESYS_GROUP_RHS_NOINLINE
#define PYODESYS_GROUP_RHS_NOINLINE (ODESYS_NY >= 9 && ODESYS_NINVAR == 0)
#endif
template <int G> struct RhsRows {
    double f[Grp<G>::R];
    int st;
};
template <int G>
__device__ __noinline__ RhsRows<G> g_rhs_shared_copy(double *ys, double *m_p, double *m_p_cse, int gl, double x,
                                                     int error_outside_bounds) {
    RhsRows<G> out;
#pragma unroll
    for (int r = 0; r < Grp<G>::R; ++r) out.f[r] = 0.0;
    OdeSysRef s;
    s.m_p = m_p;
    s.m_p_cse = m_p_cse;
    s.nfev = 0;
    s.njev = 0;
    SinkF<G> sink{gl, out.f};
    int st = s.rhs_sink(x, ys, sink);
#if ODESYS_HAS_BOUNDS
    if (error_outside_bounds && s.outside_bounds(ys)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NONNEGATIVE
#pragma unroll
    for (int i = 0; i < NY; ++i)
        if (ys[i] < 0) st = AnyODE::Status::recoverable_error;
#endif
    AnyODE::ignore(error_outside_bounds);
    out.st = st;
    return out;
}

template <int G> __device__ __forceinline__ int g_rhs(const Grp<G> &g, OdeSysRef &sys, const KernelArgs &a, double x,
                                                      const double *yv, double *f) {
    constexpr int R = Grp<G>::R;
    __syncwarp(g.mask);
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (g.owns(r)) g.ys[r * G + g.gl] = yv[r];
    __syncwarp(g.mask);
#if PYODESYS_GROUP_RHS_NOINLINE
    const RhsRows<G> out = g_rhs_shared_copy<G>(g.ys, sys.m_p, sys.m_p_cse, g.gl, x, a.error_outside_bounds);
#pragma unroll
    for (int r = 0; r < R; ++r) f[r] = out.f[r];
    sys.nfev += 1;
    return out.st;
#else
    SinkF<G> sink{g.gl, f};
    const int nfev0 = sys.nfev;
    int st = sys.rhs_sink(x, g.ys, sink);
    sys.nfev = nfev0 + 1;
#if ODESYS_HAS_BOUNDS
    if (a.error_outside_bounds && sys.outside_bounds(g.ys)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NINVAR > 0
    if (sys.invariant_violated(x, g.ys)) st = AnyODE::Status::recoverable_error;  // every lane: same y, same verdict
#endif
#if ODESYS_NONNEGATIVE
#pragma unroll
    for (int i = 0; i < NY; ++i)
        if (g.ys[i] < 0) st = AnyODE::Status::recoverable_error;
#endif
    return st;
#endif
}

// ---- dense LU across the group -------------------------------------------------------------------
// a[r][*]: rows owned by this lane.  ord[r]: pivot order of the row (NY while unused, -1 for padding rows).
// piv[k] = absolute lane | slot << 8 of the k-th pivot row.  The reciprocal pivot replaces the diagonal entry.
template <int G> __device__ __forceinline__ bool g_lu_factor(const Grp<G> &g, double (*a)[NY], int *ord, int *piv) {
    constexpr int R = Grp<G>::R;
    bool ok = true;
#pragma unroll
    for (int r = 0; r < R; ++r) ord[r] = g.owns(r) ? NY : -1;
#pragma unroll
    for (int k = 0; k < NY; ++k) {
        // pivot search: the high word of |a_ik| orders magnitudes to 2^-20 relative, which is all partial
        // pivoting needs; ties go to the lowest lane
        int best = -1, bslot = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int hi = (ord[r] == NY) ? (__double2hiint(a[r][k]) & 0x7fffffff) : -1;
            if (hi > best) {
                best = hi;
                bslot = r;
            }
        }
        const int mx = __reduce_max_sync(g.mask, best);
        const unsigned vote = __ballot_sync(g.mask, best == mx) & g.mask;
        const int pl = __ffs(vote) - 1;
        const int ps = __shfl_sync(g.mask, bslot, pl);
        ok = ok && (mx > 0) && (mx < 0x7ff00000);
        piv[k] = pl | (ps << 8);
        const bool mine = ((int)(threadIdx.x & 31u) == pl);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (mine && r == ps) {
                ord[r] = k;
#pragma unroll
                for (int j = k; j < NY; ++j) g.prow[j] = a[r][j];
            }
        }
        __syncwarp(g.mask);
        const double inv = fast_rcp<2>(g.prow[k]);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (mine && r == ps) a[r][k] = inv;
            const bool act = (ord[r] == NY);
            const double l = act ? a[r][k] * inv : 0.0;
            if (act) a[r][k] = l;
#pragma unroll
            for (int j = k + 1; j < NY; ++j) a[r][j] = fma(-l, g.prow[j], a[r][j]);
        }
        __syncwarp(g.mask);
    }
    return ok;
}

// Solves W x = b for the distributed right-hand side b[r] (equation layout); the solution comes back in
// b[r] in variable layout (both layouts are lane i % G, slot i / G).
template <int G> __device__ __forceinline__ void g_lu_solve(const Grp<G> &g, const double (*a)[NY], const int *ord,
                                                            const int *piv, double *b) {
    constexpr int R = Grp<G>::R;
    const int lane = threadIdx.x & 31u;
    // forward substitution in pivot order: rows pivoted later than k hold their multiplier in column k
#pragma unroll
    for (int k = 0; k < NY; ++k) {
        const int pl = piv[k] & 0xff, ps = piv[k] >> 8;
        double bk = b[0];
#pragma unroll
        for (int r = 1; r < R; ++r) bk = (ps == r) ? b[r] : bk;
        bk = __shfl_sync(g.mask, bk, pl);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double l = (ord[r] > k) ? a[r][k] : 0.0;
            b[r] = fma(-l, bk, b[r]);
        }
    }
    // back substitution: x_k = b_(pivot row k) / u_kk, then every earlier pivot row drops its u_ik x_k
    double x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) x[r] = 0.0;
#pragma unroll
    for (int k = NY - 1; k >= 0; --k) {
        const int pl = piv[k] & 0xff, ps = piv[k] >> 8;
        double xk = b[0] * a[0][k];
#pragma unroll
        for (int r = 1; r < R; ++r) xk = (ps == r) ? b[r] * a[r][k] : xk;
        xk = __shfl_sync(g.mask, xk, pl);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double u = (ord[r] >= 0 && ord[r] < k) ? a[r][k] : 0.0;
            b[r] = fma(-u, xk, b[r]);
        }
        if (g.gl == k % G) x[k / G] = xk;
    }
    AnyODE::ignore(lane);
#pragma unroll
    for (int r = 0; r < R; ++r) b[r] = x[r];
}

// ---- explicit inverse across the group ---------------------------------------------------------------------
// The six triangular solve pairs of one Rosenbrock step are 2 x NY strictly serial shuffle -> FMA links each; with
// two or three warps per scheduler that latency chain, not the FP64 pipe, set the speed of the kernel.  Instead W is
// inverted in place by Gauss-Jordan elimination with the same partial (row) pivoting (pivot = largest |a_ik| among
// the rows not used yet; rows never move, each row remembers the step it served as pivot in, ord[r]) and every
// solve becomes R dot products of length NY per lane: NY independent FMAs per row, no shuffles.
//   after the elimination  a[r][m] = (W^-1)[ord[r]][p_m],  p_m = row that was the pivot of step m
//   x_k = sum_m a[r][m] * b[p_m]  for the row r with ord[r] = k
// so the right-hand side is staged through shared memory in pivot order (bs[ord[r]] = b[r]) and the result
// comes back through shared memory in variable order (xs[ord[r]] = x).  Costs 2x the FMAs of the LU but removes
// ~170 dependent instructions per solve.
#ifndef PYODESYS_GROUP_INVERSE
#define PYODESYS_GROUP_INVERSE 1
#endif
template <int G> __device__ __forceinline__ bool g_invert(const Grp<G> &g, double (*a)[NY], int *ord) {
    constexpr int R = Grp<G>::R;
    bool ok = true;
#pragma unroll
    for (int r = 0; r < R; ++r) ord[r] = g.owns(r) ? NY : -1;
#pragma unroll
    for (int k = 0; k < NY; ++k) {
        int best = -1, bslot = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int hi = (ord[r] == NY) ? (__double2hiint(a[r][k]) & 0x7fffffff) : -1;
            if (hi > best) {
                best = hi;
                bslot = r;
            }
        }
        const int mx = __reduce_max_sync(g.mask, best);
        const unsigned vote = __ballot_sync(g.mask, best == mx) & g.mask;
        const int pl = __ffs(vote) - 1;
        const int ps = __shfl_sync(g.mask, bslot, pl);
        ok = ok && (mx >= 0x00100000) && (mx < 0x7ff00000);
        const bool mine = ((int)(threadIdx.x & 31u) == pl);
        // The owner only publishes its raw pivot row and clears it (a section that runs with one lane per group);
        // the scaling by 1/pivot is folded into the elimination factors that every lane forms anyway:
        //   pivot row:  0 + (1/p) row[j],  diagonal 1/p;   other rows:  a[r][j] - (l/p) row[j],  column k: -l/p
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (mine && r == ps) {
                ord[r] = k;
#pragma unroll
                for (int j = 0; j < NY; ++j) {
                    g.prow[j] = a[r][j];
                    a[r][j] = 0.0;
                }
            }
        }
        __syncwarp(g.mask);
        const double d = fast_rcp<2>(g.prow[k]);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double m = (ord[r] != k) ? -a[r][k] * d : d;
#pragma unroll
            for (int j = 0; j < NY; ++j) {
                if (j == k) continue;
                a[r][j] = fma(m, g.prow[j], a[r][j]);
            }
            a[r][k] = m;
        }
        __syncwarp(g.mask);
    }
    return ok;
}

// x = W^-1 b with the in-place inverse of g_invert; b distributed (slot r of lane gl = equation r*G + gl), the
// solution comes back in the same layout.  Uses g.prow / g.ys as staging (both are free between rhs evaluations).
template <int G> __device__ __forceinline__ void g_apply(const Grp<G> &g, const double (*a)[NY], const int *ord,
                                                         double *b) {
    constexpr int R = Grp<G>::R;
    __syncwarp(g.mask);
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (g.owns(r)) g.prow[ord[r]] = b[r];
    __syncwarp(g.mask);
    double acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.0;
#pragma unroll
    for (int m = 0; m < NY; ++m) {
        const double bm = g.prow[m];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = fma(a[r][m], bm, acc[r]);
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (g.owns(r)) g.ys[ord[r]] = acc[r];
    __syncwarp(g.mask);
#pragma unroll
    for (int r = 0; r < R; ++r) b[r] = g.owns(r) ? g.ys[r * G + g.gl] : 0.0;
}

// ---- Rosenbrock 4(3) on a lane group (same coefficients/controller as Stepper<M_ROS4>) -------------------
#define RC(name) kRos4[ros4c::name]
template <int G> struct GroupRos4 {
    static constexpr int R = Grp<G>::R;
    double g1[R], g2[R], g3[R], g4[R], g5[R];
    double h_old;
    float err_old;
    bool first, last_rejected;

    __device__ __forceinline__ void begin() {
        first = true;
        last_rejected = false;
        h_old = 0.0;
        err_old = 1.0f;
    }

    // Hairer's starting step (see initial_step in integrators.cuh), norms reduced over the group
    __device__ __forceinline__ double auto_step(const Grp<G> &g, OdeSysRef &sys, const KernelArgs &a, double t0,
                                                const double *y, double span) {
        double f0[R], y1[R], f1[R];
        g_rhs<G>(g, sys, a, t0, y, f0);
        double d0 = 0.0, d1 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!g.owns(r)) continue;
            const double sk = a.atol[r * G + g.gl] + a.rtol * fabs(y[r]);
            d0 = fmax(d0, fabs(y[r]) / sk);
            d1 = fmax(d1, fabs(f0[r]) / sk);
        }
        d0 = gmax_d<G>(g, d0);
        d1 = gmax_d<G>(g, d1);
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, fabs(span));
        const double dir = span >= 0 ? 1.0 : -1.0;
#pragma unroll
        for (int r = 0; r < R; ++r) y1[r] = fma(dir * h0, f0[r], y[r]);
        const int st = g_rhs<G>(g, sys, a, t0 + dir * h0, y1, f1);
        double d2 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!g.owns(r)) continue;
            const double sk = a.atol[r * G + g.gl] + a.rtol * fabs(y[r]);
            d2 = fmax(d2, fabs(f1[r] - f0[r]) / sk);
        }
        d2 = gmax_d<G>(g, d2) / h0;
        if (st != 0 || !(d2 <= DBL_MAX)) d2 = 0.0;
        const double dmax = fmax(d1, d2);
        const double h1 = (dmax <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dmax, 0.25);
        return fmin(fmin(100.0 * h0, h1), fabs(span)) * dir;
    }

    __device__ __forceinline__ int attempt(const Grp<G> &g, OdeSysRef &sys, const KernelArgs &a, double t,
                                           const double *y, double h, double *ynew, double &err, Counters &cnt) {
        double W[R][NY], dfdt[R], f[R], tmp[R], xerr[R];
        int ord[R], piv[NY];
        int bad = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            dfdt[r] = 0.0;
#pragma unroll
            for (int j = 0; j < NY; ++j) W[r][j] = 0.0;
        }
        bad |= g_rhs<G>(g, sys, a, t, y, f);
        {
            SinkJ<G> sink{g.gl, W, dfdt};
            const int njev0 = sys.njev;
            bad |= sys.dense_jac_sink(t, g.ys, sink);  // g.ys still holds y
            sys.njev = njev0 + 1;
        }
        const double ih = fast_rcp<2>(h);
        const double diag = 4.0 * ih;  // 1/gamma = 4
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int j = 0; j < NY; ++j) W[r][j] = ((r * G + g.gl == j) ? diag : 0.0) - W[r][j];
#if PYODESYS_GROUP_INVERSE
        if (!g_invert<G>(g, W, ord)) bad |= 1;
        AnyODE::ignore(piv);
#define G_SOLVE(vec) g_apply<G>(g, W, ord, vec)
#else
        if (!g_lu_factor<G>(g, W, ord, piv)) bad |= 1;
#define G_SOLVE(vec) g_lu_solve<G>(g, W, ord, piv, vec)
#endif
        cnt.nlu += 1;
        cnt.nsolve += 6;
#pragma unroll
        for (int r = 0; r < R; ++r) g1[r] = ODESYS_AUTONOMOUS ? f[r] : fma(h * RC(D1), dfdt[r], f[r]);
        G_SOLVE(g1);
#pragma unroll
        for (int r = 0; r < R; ++r) tmp[r] = fma(RC(A21), g1[r], y[r]);
        bad |= g_rhs<G>(g, sys, a, fma(RC(C2X), h, t), tmp, f);
        const double c21 = RC(C21) * ih;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g2[r] = fma(c21, g1[r], f[r]);
            if (!ODESYS_AUTONOMOUS) g2[r] = fma(h * RC(D2), dfdt[r], g2[r]);
        }
        G_SOLVE(g2);
#pragma unroll
        for (int r = 0; r < R; ++r) tmp[r] = fma(RC(A32), g2[r], fma(RC(A31), g1[r], y[r]));
        bad |= g_rhs<G>(g, sys, a, fma(RC(C3X), h, t), tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g3[r] = fma(ih, fma(RC(C32), g2[r], RC(C31) * g1[r]), f[r]);
            if (!ODESYS_AUTONOMOUS) g3[r] = fma(h * RC(D3), dfdt[r], g3[r]);
        }
        G_SOLVE(g3);
#pragma unroll
        for (int r = 0; r < R; ++r) tmp[r] = fma(RC(A43), g3[r], fma(RC(A42), g2[r], fma(RC(A41), g1[r], y[r])));
        bad |= g_rhs<G>(g, sys, a, fma(RC(C4X), h, t), tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g4[r] = fma(ih, fma(RC(C43), g3[r], fma(RC(C42), g2[r], RC(C41) * g1[r])), f[r]);
            if (!ODESYS_AUTONOMOUS) g4[r] = fma(h * RC(D4), dfdt[r], g4[r]);
        }
        G_SOLVE(g4);
#pragma unroll
        for (int r = 0; r < R; ++r)
            tmp[r] = fma(RC(A54), g4[r], fma(RC(A53), g3[r], fma(RC(A52), g2[r], fma(RC(A51), g1[r], y[r]))));
        bad |= g_rhs<G>(g, sys, a, t + h, tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r)
            g5[r] = fma(ih, fma(RC(C54), g4[r], fma(RC(C53), g3[r], fma(RC(C52), g2[r], RC(C51) * g1[r]))), f[r]);
        G_SOLVE(g5);
#pragma unroll
        for (int r = 0; r < R; ++r) tmp[r] += g5[r];
        bad |= g_rhs<G>(g, sys, a, t + h, tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r)
            xerr[r] = fma(ih, fma(RC(C65), g5[r], fma(RC(C64), g4[r], fma(RC(C63), g3[r],
                                  fma(RC(C62), g2[r], RC(C61) * g1[r])))), f[r]);
        G_SOLVE(xerr);
        double e = 0.0;
        bool finite = true;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            ynew[r] = tmp[r] + xerr[r];
            if (!g.owns(r)) continue;
            finite = finite && (fabs(ynew[r]) <= DBL_MAX);
            const double sk = fma(a.rtol, fmax(fabs(y[r]), fabs(ynew[r])), a.atol[r * G + g.gl]) + 1e-300;
            e = fmax(e, fabs(xerr[r]) * fast_rcp<1>(sk));
        }
        // shuffle/REDUX-reduced error norm; single precision is ample for step-size control
        const float ef = gmax_f<G>(g, (float)fmin(e, 3e38));
        err = (double)ef;
        if (!(ef <= 3e38f) || __any_sync(g.mask, !finite)) bad |= 1;
        return bad;
#undef G_SOLVE
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

    __device__ __forceinline__ void dense(double theta, const double *yold, const double *ynew, double *out) {
        const double s1 = 1.0 - theta;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double c3 = fma(RC(D25), g5[r], fma(RC(D24), g4[r], fma(RC(D23), g3[r], fma(RC(D22), g2[r], RC(D21) * g1[r]))));
            const double c4 = fma(RC(D35), g5[r], fma(RC(D34), g4[r], fma(RC(D33), g3[r], fma(RC(D32), g2[r], RC(D31) * g1[r]))));
            out[r] = yold[r] * s1 + theta * (ynew[r] + s1 * (c3 + theta * c4));
        }
    }
};
#undef RC

// ---- event functions on a lane group (see locate_roots in integrators.cuh for the scheme) --------------------
// The state is distributed over the lanes; g(x, y) is evaluated by every lane of the group from the state staged
// in shared memory, so all lanes see the same values and take the same decisions.
#if ODESYS_NROOTS > 0
template <int G> __device__ __forceinline__ void g_eval_roots(const Grp<G> &g, OdeSysRef &sys, double x,
                                                              const double *yv, double *gout) {
    constexpr int R = Grp<G>::R;
    __syncwarp(g.mask);
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (g.owns(r)) g.ys[r * G + g.gl] = yv[r];
    __syncwarp(g.mask);
    sys.roots(x, g.ys, gout);
    __syncwarp(g.mask);
}

template <int G, int MODE, class St>
__device__ __noinline__ bool g_locate_roots(const Grp<G> &g, St &st, OdeSysRef &sys, const KernelArgs &a, int64_t s,
                                            double t, double hh, double &tnew, const double *y, double *ynew,
                                            double *g_old, int &n_roots, int &k_out, double dir, int &status) {
    constexpr int R = Grp<G>::R;
    double g_lo[NROOTS], g_hi[NROOTS];
    g_eval_roots(g, sys, tnew, ynew, g_hi);
#pragma unroll
    for (int r = 0; r < NROOTS; ++r) g_lo[r] = g_old[r];
    double lo = 0.0;
    bool stop = false;
    const double ttol = 100.0 * DBL_EPSILON * (fabs(t) + fabs(hh)) / fabs(hh);
    for (int found = 0; found < 16 && !stop; ++found) {
        bool any = false;
#pragma unroll
        for (int r = 0; r < NROOTS; ++r) any = any || root_crossed(g_lo[r], g_hi[r]);
        if (!any) break;
        double ta = lo, tb = 1.0, ga[NROOTS], gb[NROOTS], yb[R];
#pragma unroll
        for (int r = 0; r < NROOTS; ++r) {
            ga[r] = g_lo[r];
            gb[r] = g_hi[r];
        }
#pragma unroll
        for (int r = 0; r < R; ++r) yb[r] = ynew[r];
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
            double ym[R], gm[NROOTS];
            st.dense(tm, y, ynew, ym);
            g_eval_roots(g, sys, fma(tm, hh, t), ym, gm);
            bool chg = false;
#pragma unroll
            for (int r = 0; r < NROOTS; ++r) chg = chg || root_crossed(ga[r], gm[r]);
            sideprev = side;
            if (chg) {
                tb = tm;
#pragma unroll
                for (int r = 0; r < NROOTS; ++r) gb[r] = gm[r];
#pragma unroll
                for (int r = 0; r < R; ++r) yb[r] = ym[r];
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
        const bool stored = n_roots < a.max_roots;
        if (stored) {
            double *ro = a.root_out + (s * a.max_roots + n_roots) * (1 + NY);
            if (g.gl == 0) {
                ro[0] = x_root;
                a.root_indices[s * a.max_roots + n_roots] = idx;
            }
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (g.owns(r)) ro[1 + r * G + g.gl] = yb[r];
        }
        n_roots += 1;
        if (MODE == 1 && a.return_on_root) {
            tnew = x_root;
#pragma unroll
            for (int r = 0; r < R; ++r) ynew[r] = yb[r];
            stop = true;
        } else if (MODE == 1 && !at_end && stored) {
            if (k_out >= a.max_rows) {
                status = ST_ROWS;
                stop = true;
            } else {
                double *row = a.xyout + (s * a.max_rows + k_out) * (1 + NY);
                if (g.gl == 0) row[0] = x_root;
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (g.owns(r)) row[1 + r * G + g.gl] = yb[r];
                k_out += 1;
            }
        }
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
// Persistent kernel, one system per lane group.  MODE 0: predefined output points, MODE 1: every
// accepted step.  Control flow mirrors integrate_kernel (integrators.cuh); all per-system control
// variables are replicated in the lanes of the group and stay identical by construction.
// ----------------------------------------------------------------------------------------------
#ifndef PYODESYS_GROUP_MIN_BLOCKS
#define PYODESYS_GROUP_MIN_BLOCKS 3
#endif
template <int G, int MODE, int BLOCK>
__global__ void __launch_bounds__(BLOCK, PYODESYS_GROUP_MIN_BLOCKS) group_ros4_kernel(const KernelArgs a) {
    constexpr int R = Grp<G>::R;
    constexpr int GS = kGroupSmemDoubles;
    extern __shared__ double g_smem[];
    const int lane = threadIdx.x & 31u;
    Grp<G> g;
    g.gl = lane % G;
    g.lane0 = lane - g.gl;
    g.mask = (G == 32) ? FULL_MASK : (((1u << (G & 31)) - 1u) << g.lane0);
    double *base = g_smem + (size_t)(threadIdx.x / G) * GS;
    g.ys = base;
    g.prow = base + NY;
    OdeSysRef sys;
    sys.m_p = base + 2 * NY;
    sys.m_p_cse = sys.m_p + (NP_USED > 0 ? NP_USED : 1);
    sys.nfev = 0;
    sys.njev = 0;

    const int64_t ngroups = (int64_t)gridDim.x * (BLOCK / G);
    int64_t s = (int64_t)blockIdx.x * (BLOCK / G) + threadIdx.x / G - ngroups;  // static schedule cursor

    GroupRos4<G> st;
    double y[R], t = 0, h = 0, tend = 0, dir = 1, hmin = 0, hmax = INFINITY;
    int64_t steps_left = 0;
    int k_out = 0, restarts = 0, nfail = 0;
    Counters cnt = {0, 0, 0, 0};
    bool active = false, exhausted = false;
#if ODESYS_NROOTS > 0
    double g_old[NROOTS];
    int n_roots = 0;
#endif
#pragma unroll
    for (int r = 0; r < R; ++r) y[r] = 0.0;

    for (;;) {
        __syncwarp();
        // ---- (A) fetch a new system for every idle group ---------------------------------------------
        if (!active && !exhausted) {
            if (a.work_counter != nullptr) {
                unsigned long long nxt = 0;
                if (g.gl == 0) nxt = atomicAdd(a.work_counter, 1ull);
                s = (int64_t)__shfl_sync(g.mask, nxt, g.lane0);
            } else {
                s += ngroups;
            }
            if (s >= a.n) {
                exhausted = true;
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) y[r] = g.owns(r) ? a.y0[s * NY + r * G + g.gl] : 0.0;
                __syncwarp(g.mask);
                for (int i = g.gl; i < NP_USED; i += G) {
                    if (i >= ODESYS_NPARAMS && a.iv_from_y0) sys.m_p[i] = a.y0[s * NY + (i - ODESYS_NPARAMS)];
                    else sys.m_p[i] = a.params[s * a.params_ld + i];
                }
                __syncwarp(g.mask);
                sys.setup();  // parameter-only sub-expressions: every lane writes the same values
                __syncwarp(g.mask);
                cnt = {0, 0, 0, 0};
                restarts = a.autorestart;
                nfail = 0;
                if (MODE == 0) {
                    const double *xo = a.xout + s * a.xout_stride;
                    t = xo[0];
                    tend = xo[a.nt - 1];
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (g.owns(r)) {
                            const int i = r * G + g.gl;
                            if (a.yout_soa) a.yout[(int64_t)i * a.n + s] = y[r];
                            else a.yout[(s * a.nt) * NY + i] = y[r];
                        }
                } else {
                    t = a.x0[s];
                    tend = a.xend[s];
                    double *row = a.xyout + (s * a.max_rows) * (1 + NY);
                    if (g.gl == 0) row[0] = t;
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (g.owns(r)) row[1 + r * G + g.gl] = y[r];
                }
                k_out = 1;
                dir = (tend >= t) ? 1.0 : -1.0;
                hmin = a.dx_min ? a.dx_min[s] : 0.0;
                hmax = a.dx_max ? a.dx_max[s] : INFINITY;
                if (!(hmax > 0.0)) hmax = INFINITY;
                steps_left = a.mxsteps > 0 ? a.mxsteps : INT64_MAX;
#if ODESYS_NINVAR > 0
                __syncwarp(g.mask);
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (g.owns(r)) g.ys[r * G + g.gl] = y[r];
                __syncwarp(g.mask);
                sys.init_invariants(a.max_invariant_violation, t, g.ys);
#endif
#if ODESYS_NROOTS > 0
                n_roots = 0;
                if (a.max_roots > 0) g_eval_roots(g, sys, t, y, g_old);
#endif
                st.begin();
                h = a.dx0 ? a.dx0[s] : 0.0;
#if ODESYS_HAS_FIRST_STEP
                if (h == 0.0) {
                    __syncwarp(g.mask);
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (g.owns(r)) g.ys[r * G + g.gl] = y[r];
                    __syncwarp(g.mask);
                    h = sys.get_dx0(t, g.ys);
                }
#endif
                if (h != 0.0) h = fabs(h) * dir;
                else h = st.auto_step(g, sys, a, t, y, tend - t);
                active = true;
                bool fin = true;
#pragma unroll
                for (int r = 0; r < R; ++r) fin = fin && (fabs(y[r]) <= DBL_MAX);
                if (__any_sync(g.mask, !fin)) nfail = 1 << 20;
                if (tend == t && MODE == 0) {
                    for (; k_out < a.nt; ++k_out) {
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            if (g.owns(r)) {
                                const int i = r * G + g.gl;
                                if (a.yout_soa) a.yout[((int64_t)k_out * NY + i) * a.n + s] = y[r];
                                else a.yout[(s * a.nt + k_out) * NY + i] = y[r];
                            }
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
                h = st.auto_step(g, sys, a, t, y, tend - t);
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
                __syncwarp(g.mask);
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (g.owns(r)) g.ys[r * G + g.gl] = y[r];
                __syncwarp(g.mask);
                const double m = a.get_dx_max_factor * sys.get_dx_max(t, g.ys);
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
            double ynew[R], err = 0.0, hnext = hh;
            const int bad = st.attempt(g, sys, a, t, y, hh, ynew, err, cnt);
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
                bool stop_at_root = false;
#if ODESYS_NROOTS > 0
                if (a.max_roots > 0)
                    stop_at_root = g_locate_roots<G, MODE>(g, st, sys, a, s, t, hh, tnew, y, ynew, g_old, n_roots,
                                                           k_out, dir, status);
#endif
                if (MODE == 0) {
                    const double *xo = a.xout + s * a.xout_stride;
                    while (k_out < a.nt) {
                        const double xk = xo[k_out];
                        if ((xk - tnew) * dir > 0.0) break;
                        double yo[R];
                        if (xk == tnew) {
#pragma unroll
                            for (int r = 0; r < R; ++r) yo[r] = ynew[r];
                        } else {
                            st.dense((xk - t) / hh, y, ynew, yo);
                        }
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            if (g.owns(r)) {
                                const int i = r * G + g.gl;
                                if (a.yout_soa) a.yout[((int64_t)k_out * NY + i) * a.n + s] = yo[r];
                                else a.yout[(s * a.nt + k_out) * NY + i] = yo[r];
                            }
                        k_out += 1;
                        steps_left = a.mxsteps > 0 ? a.mxsteps : INT64_MAX;
                    }
                } else {
                    if (k_out >= a.max_rows) {
                        status = ST_ROWS;
                    } else {
                        double *row = a.xyout + (s * a.max_rows + k_out) * (1 + NY);
                        if (g.gl == 0) row[0] = tnew;
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            if (g.owns(r)) row[1 + r * G + g.gl] = ynew[r];
                        k_out += 1;
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) y[r] = ynew[r];
                t = tnew;
                h = hnext;
                if (stop_at_root && status < 0) status = ST_OK;
            }
        }

        if (status >= 0) {
            // ---- (E) retire the system ---------------------------------------------------------------------
            if (MODE == 0 && status != ST_OK) {
                const double qnan = nan("");
                for (int k = k_out; k < a.nt; ++k)
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (g.owns(r)) {
                            const int i = r * G + g.gl;
                            if (a.yout_soa) a.yout[((int64_t)k * NY + i) * a.n + s] = qnan;
                            else a.yout[(s * a.nt + k) * NY + i] = qnan;
                        }
            }
            if (g.gl == 0) {
                a.status[s] = status;
                if (MODE == 0) a.nreached[s] = k_out;
                else a.nrows[s] = k_out;
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
            }
            sys.nfev = 0;
            sys.njev = 0;
            active = false;
        }
    }
}

#endif  // ODESYS_HAS_JAC

}  // namespace pyodesys_cuda
