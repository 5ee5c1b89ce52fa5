// Batched Rosenbrock integrator for sm_100a, ny ~ 9..32: BLOCK-COOPERATIVE.
//
//   * linear algebra: one system per lane group (G = 4, 8, 16 or 32 lanes), W = I/(gamma h) - J distributed by rows
//     in registers, in-place Gauss-Jordan inverse with partial pivoting + six matrix-vector products per step
//     (g_invert / g_apply of group_integrators.cuh; AnyODE::getrf_callback/getrs_callback semantics,
//     anyode_blasless.hpp:10-95);
//   * right-hand side and Jacobian: ONE SYSTEM PER LANE.  The 256 / G systems of a block step in lock-step; for every
//     evaluation the lane groups stage their state vectors in shared memory (Y[i][system]), one warp ("evaluator",
//     the role rotates over the warps of the block) evaluates the generated OdeSys::rhs / dense_jac_sink
//     (odesys_anyode_template.cpp:147-204,246-293) for all systems of the block - state in registers, parameters in
//     shared memory P[k][system] - and hands f / J back through shared memory.
//
// Why: in the lane-group kernel (group_integrators.cuh) every lane of a group evaluates ALL rows of f and J and keeps
// the R it owns - the generated expressions are straight-line code that only vectorises across systems.  On the
// radiolysis network (ny = 14, G = 8) that G-fold redundancy was 48 % of the executed instructions, the kernel
// spilled (368 M local loads per launch) and reached 5.4 % of the FP64 peak (profiles/r1g_c4_attribution.md); with
// 32-lane groups (ny = 28) it was 1.3 %.  Here an evaluation costs one warp-instruction per 32 systems.
//
// Everything per-system that steers the integration (t, h, controller memory, counters, output cursor) is replicated
// in the lanes of its group, as in group_ros4_kernel.  Supported: predefined and adaptive output, dx0/dx_min/dx_max
// arrays, autorestart, bounds / non-negativity guards.  Event functions, the invariant guard and the get_dx0 /
// get_dx_max expressions stay with the lane-group kernel (the shim routes such calls there).
// Requires integrators.cuh and group_integrators.cuh first.
#pragma once

namespace pyodesys_cuda {

#if ODESYS_HAS_JAC

// smallest group whose lanes keep their rows of W in at most 56 doubles (ny = 14: 4 lanes x 4 rows; ny = 28: 16 x 2)
#define PYODESYS_COOP_FITS(g) ((((NY + (g) - 1) / (g)) * NY) <= 56)
#ifndef PYODESYS_COOP_G
#define PYODESYS_COOP_G (PYODESYS_COOP_FITS(4) ? 4 : PYODESYS_COOP_FITS(8) ? 8 : PYODESYS_COOP_FITS(16) ? 16 : 32)
#endif
// Threads per block.  The systems of a block step in lock-step and every evaluation is a serial phase of ONE warp
// (the others wait at the barrier), so what keeps an SM busy is the number of independent blocks on it, not the size
// of a block: 128 threads, 2 blocks per SM at up to 255 registers (no spills).
#ifndef PYODESYS_COOP_BLOCK
#define PYODESYS_COOP_BLOCK 128
#endif
#ifndef PYODESYS_COOP_MIN_BLOCKS
#define PYODESYS_COOP_MIN_BLOCKS (256 / PYODESYS_COOP_BLOCK)
#endif

constexpr int kCoopBlock = PYODESYS_COOP_BLOCK;
constexpr int kCoopWarps = kCoopBlock / 32;

template <int S> struct SView {  // array-like view with a compile-time stride: v[i] = p[i * S]
    double *p;
    __device__ __forceinline__ double &operator[](int i) const { return p[i * S]; }
};

// Shared memory of one block (doubles).  "[i][system]" arrays have the odd pitch NSYSP: the evaluator (lane = system)
// reads consecutive words, the lanes of a group (lane = row i) hit distinct banks.
template <int G, int EXTRA = 0> struct CoopLayout {
    static constexpr int R = (NY + G - 1) / G;
    static constexpr int NSYS = kCoopBlock / G;    // systems per block
    static constexpr int NSYSP = NSYS | 1;
    static constexpr int NYP = NY | 1;              // row pitch of the Jacobian staging
    static constexpr int NE = (NSYS + 31) / 32;    // evaluator warps per evaluation
    static constexpr int NPU = NP_USED > 0 ? NP_USED : 1;
    static constexpr int NPC = ODESYS_NPCSE > 0 ? ODESYS_NPCSE : 1;
    static constexpr int oP = 0;                                     // parameters [k][system]
    static constexpr int oPC = oP + NPU * NSYS;                      // parameter-only sub-expressions [k][system]
    static constexpr int oY = oPC + NPC * NSYS;                      // staged state [i][system]
    static constexpr int oF = oY + NY * NSYSP;                       // f [i][system]
    static constexpr int oX = oF + NY * NSYSP;                       // abscissa [system]
    static constexpr int oDF = oX + NSYS;                            // dfdt [i][system] (non-autonomous systems)
    static constexpr int oJ = oDF + (ODESYS_AUTONOMOUS ? 0 : NY * NSYSP);  // Jacobian [(i * NYP + j)][system]
    static constexpr int NYE = NY + (NY & 1);                        // vectors that are moved 16 bytes at a time
    static constexpr int oScr = (oJ + NY * NYP * NSYSP + 1) & ~1;    // per system: two staging vectors [NYE] (16-byte aligned)
    static constexpr int oInt = oScr + 2 * NYE * NSYS;               // ints: status, fresh, request [NSYS] each
    static constexpr int oExtra = oInt + 2 * NSYS;                   // EXTRA doubles per system (stepper state)
    static constexpr int kDoubles = oExtra + EXTRA * NSYS;
    typedef odesys_anyode::OdeSysT<SView<NSYS>> ESys;
};

template <int G> struct CoopSinkJ {
    typedef CoopLayout<G> L;
    double *J, *DF;  // already offset by the evaluator's system
    __device__ __forceinline__ void setj(int i, int j, double v) { J[(i * L::NYP + j) * L::NSYSP] = v; }
    __device__ __forceinline__ void setdfdt(int i, double v) {
        if (!ODESYS_AUTONOMOUS) DF[i * L::NSYSP] = v;
    }
};

// f (and J) for one system of the block, state in registers, parameters / results in shared memory.  One out-of-line
// copy per variant (a Rosenbrock step has seven evaluation sites); only the evaluator lanes make the call, so the
// other warps pay neither for its code nor for its register save area.
template <int G> __device__ __noinline__ void coop_eval_rhs(int e, int error_outside_bounds) {
    typedef CoopLayout<G> L;
    extern __shared__ double c_smem[];
    double *sm = c_smem;
    typename L::ESys es;
    es.m_p = SView<L::NSYS>{sm + L::oP + e};
    es.m_p_cse = SView<L::NSYS>{sm + L::oPC + e};
    int *ist = reinterpret_cast<int *>(sm + L::oInt);
    if (ist[L::NSYS + e]) {  // parameters of a new system arrived: parameter-only sub-expressions (template:104-110)
        es.setup();
        ist[L::NSYS + e] = 0;
    }
    es.nfev = 0;
    es.njev = 0;
    double y[NY], f[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) y[i] = sm[L::oY + i * L::NSYSP + e];
    const double x = sm[L::oX + e];
    int st = es.rhs(x, y, f);
#if ODESYS_HAS_BOUNDS
    if (error_outside_bounds && es.outside_bounds(y)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NONNEGATIVE
#pragma unroll
    for (int i = 0; i < NY; ++i)
        if (y[i] < 0) st = AnyODE::Status::recoverable_error;
#endif
    AnyODE::ignore(error_outside_bounds);
#pragma unroll
    for (int i = 0; i < NY; ++i) sm[L::oF + i * L::NSYSP + e] = f[i];
    ist[e] = st;
}
// the Jacobian at the state that coop_eval_rhs just evaluated (one copy of each generated body in the kernel: the
// loop has to fit the instruction cache)
template <int G> __device__ __noinline__ void coop_eval_jac(int e) {
    typedef CoopLayout<G> L;
    extern __shared__ double c_smem[];
    double *sm = c_smem;
    typename L::ESys es;
    es.m_p = SView<L::NSYS>{sm + L::oP + e};
    es.m_p_cse = SView<L::NSYS>{sm + L::oPC + e};
    es.nfev = 0;
    es.njev = 0;
    double y[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) y[i] = sm[L::oY + i * L::NSYSP + e];
    CoopSinkJ<G> sink{sm + L::oJ + e, sm + L::oDF + e};
    const int st = es.dense_jac_sink(sm[L::oX + e], y, sink);
    int *ist = reinterpret_cast<int *>(sm + L::oInt);
    ist[e] |= st;
}

// ---- evaluation split over the four warps of a block ------------------------------------------------------------
// With one evaluator warp, an evaluation is a latency-bound serial phase (a single warp issues one dependent FP64
// instruction every few cycles while the other three wait at the barrier: 30 % of a block's time for Rosenbrock-4 on
// the radiolysis network, more for BDF).  The code generator deals the rows of f and the columns of J to
// ODESYS_EVAL_PARTS = 4 self-contained pieces (each with its own common-subexpression elimination; shared reaction
// rates are recomputed, so a piece is ~1/2 of f and ~1/4 of J); warp w evaluates piece w for all systems of the block.
#ifndef PYODESYS_COOP_SPLIT_EVAL
#define PYODESYS_COOP_SPLIT_EVAL 1
#endif
#if PYODESYS_COOP_SPLIT_EVAL && defined(ODESYS_EVAL_PARTS) && ODESYS_EVAL_PARTS == 4
#define PYODESYS_COOP_SPLIT 1
#else
#define PYODESYS_COOP_SPLIT 0
#endif
template <int G> struct CoopSplit {
    static constexpr bool value = PYODESYS_COOP_SPLIT && kCoopWarps == 4 && ((kCoopBlock / G) <= 32);
};
#if PYODESYS_COOP_SPLIT
template <int G> struct CoopSinkF {
    typedef CoopLayout<G> L;
    double *F;  // already offset by the evaluator's system
    __device__ __forceinline__ void set(int i, double v) { F[i * L::NSYSP] = v; }
};
// jac: 0 no, 1 yes, 2 where the system's request flag is set
template <int G, int P> __device__ __noinline__ void coop_eval_part(int e, int error_outside_bounds, int jac) {
    typedef CoopLayout<G> L;
    extern __shared__ double c_smem[];
    double *sm = c_smem;
    typename L::ESys es;
    es.m_p = SView<L::NSYS>{sm + L::oP + e};
    es.m_p_cse = SView<L::NSYS>{sm + L::oPC + e};
    double y[NY];
#pragma unroll
    for (int i = 0; i < NY; ++i) y[i] = sm[L::oY + i * L::NSYSP + e];
    const double x = sm[L::oX + e];
    int *ist = reinterpret_cast<int *>(sm + L::oInt);
    CoopSinkF<G> sf{sm + L::oF + e};
    int st = es.template rhs_sink_part<P>(x, y, sf);
    if (P == 0) {
#if ODESYS_HAS_BOUNDS
        if (error_outside_bounds && es.outside_bounds(y)) st = AnyODE::Status::recoverable_error;
#endif
#if ODESYS_NONNEGATIVE
#pragma unroll
        for (int i = 0; i < NY; ++i)
            if (y[i] < 0) st = AnyODE::Status::recoverable_error;
#endif
        ist[e] = st;
    }
    AnyODE::ignore(error_outside_bounds);
    if (jac == 1 || (jac == 2 && ist[2 * L::NSYS + e])) {
        CoopSinkJ<G> sink{sm + L::oJ + e, sm + L::oDF + e};
        es.template dense_jac_sink_part<P>(x, y, sink);
    }
}
#endif

__device__ __forceinline__ double *c_smem_base() {
    extern __shared__ double c_smem[];
    return c_smem;
}
// executed by every thread of the block between two block barriers; the evaluator role rotates over the warps
// (JAC: 0 right-hand side only, 1 + Jacobian, 2 + Jacobian for the systems whose request flag is set)
template <int G, int JAC> __device__ __forceinline__ long long coop_evaluate(int evalno, int error_outside_bounds) {
    typedef CoopLayout<G> L;
    __syncthreads();  // staged states (Y, X) are complete
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long body = 0;
#ifdef PYODESYS_COOP_PROFILE
    const long long t_body = clock64();
#endif
#if PYODESYS_COOP_SPLIT
    if (CoopSplit<G>::value) {
        if (lane < L::NSYS) {
            switch (warp) {  // (warp-uniform)
            case 0: coop_eval_part<G, 0>(lane, error_outside_bounds, JAC); break;
            case 1: coop_eval_part<G, 1>(lane, error_outside_bounds, JAC); break;
            case 2: coop_eval_part<G, 2>(lane, error_outside_bounds, JAC); break;
            default: coop_eval_part<G, 3>(lane, error_outside_bounds, JAC); break;
            }
        }
#ifdef PYODESYS_COOP_PROFILE
        body = clock64() - t_body;
#endif
        __syncthreads();  // results (F, J, status) are complete
        return body;
    }
#endif
    const int rel = (warp - (evalno % kCoopWarps) + kCoopWarps) % kCoopWarps;
    const int e = rel * 32 + lane;
    if (rel < L::NE && e < L::NSYS) {
        coop_eval_rhs<G>(e, error_outside_bounds);
        if (JAC == 1 || (JAC == 2 && reinterpret_cast<const int *>(c_smem_base() + L::oInt)[2 * L::NSYS + e]))
            coop_eval_jac<G>(e);
#ifdef PYODESYS_COOP_PROFILE
        body = clock64() - t_body;
#endif
    }
    __syncthreads();  // results (F, J, status) are complete
    return body;
}

// -DPYODESYS_COOP_PROFILE: per-phase SM cycles (clock64, summed over warps) are added to the eight 64-bit words at
// KernelArgs::root_out (the cooperative kernel locates no event functions): 0 evaluations incl. both barriers,
// 1 evaluator body, 2 inverse, 3 solves, 4 everything, 5 warp count
#ifdef PYODESYS_COOP_PROFILE
#define COOP_T0(var) const long long var = clock64()
#define COOP_ACC(slot, var) c.prof[slot] += clock64() - var
#else
#define COOP_T0(var)
#define COOP_ACC(slot, var)
#endif

template <int G> struct Coop {
    typedef CoopLayout<G> L;
    static constexpr int R = L::R;
    int sysl;     // local index of this lane's system inside the block
    int evalno;   // evaluations so far (block uniform): rotates the evaluator role over the warps
#ifdef PYODESYS_COOP_PROFILE
    long long prof[10];
#endif

    // stage the distributed vector yv, evaluate, bring back the rows this lane owns; returns the callback status
    template <int JAC>
    __device__ __forceinline__ int eval(const Grp<G> &g, const KernelArgs &a, double x, const double *yv, double *f,
                                        bool request = false) {
        extern __shared__ double c_smem[];
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (g.owns(r)) c_smem[L::oY + (r * G + g.gl) * L::NSYSP + sysl] = yv[r];
        if (g.gl == 0) {
            c_smem[L::oX + sysl] = x;
            if (JAC == 2) reinterpret_cast<int *>(c_smem + L::oInt)[2 * L::NSYS + sysl] = request ? 1 : 0;
        }
#ifdef PYODESYS_COOP_PROFILE
        const long long t_ev = clock64();
#endif
        const long long body = coop_evaluate<G, JAC>(evalno, a.error_outside_bounds);
        AnyODE::ignore(body);
#ifdef PYODESYS_COOP_PROFILE
        prof[0] += clock64() - t_ev;
        prof[1] += body;
#endif
        evalno += 1;
#pragma unroll
        for (int r = 0; r < R; ++r) f[r] = g.owns(r) ? c_smem[L::oF + (r * G + g.gl) * L::NSYSP + sysl] : 0.0;
        return reinterpret_cast<const int *>(c_smem + L::oInt)[sysl];
    }
};

// ---- in-place inverse for a block that steps in lock-step --------------------------------------------------------
// Same result layout as g_invert (group_integrators.cuh): after the elimination a[r][m] = (W^-1)[ord[r]][p_m].  What
// differs is how a pivot step is organised, because in the lane-group version the bookkeeping of a step (pivot search
// with per-group membermasks - the hardware serialises REDUX / SHFL over the distinct masks of a warp -, the owner
// copying its row to shared memory element by element under a predicate) was 3/4 of the executed instructions and the
// top stall site (profiles/r2c_coop14_by_line.txt):
//   * every lane of the warp is converged here, so all reductions are xor-shuffles with the full mask (offsets < G
//     stay inside a group) - no serialisation;
//   * THRESHOLD pivoting: the natural row k is the pivot whenever it is unused and |a_kk| >= max_i |a_ik| / 8 over
//     the unused rows (magnitudes compared on the high words), otherwise the row with the largest |a_ik|.  Relaxed
//     pivoting with threshold 1/8 bounds the multipliers by 8 (strict partial pivoting: by 1) - the standard trade of
//     sparse direct solvers.  W = I/(gamma h) - J of a kinetics network keeps its natural order almost always, and
//     then W^-1 comes out in variable order: the solves skip one of their two permutations through shared memory;
//   * the pivot row is published by its owner with 16-byte stores and read back by every lane with 16-byte
//     broadcast loads (a 64-bit shuffle costs two wavefront slots per double on the same data path).
template <int G> __device__ __forceinline__ int gmax_i_full(int v) {
#pragma unroll
    for (int off = G / 2; off > 0; off /= 2) v = max(v, __shfl_xor_sync(FULL_MASK, v, off));
    return v;
}
#ifndef PYODESYS_COOP_PIVOT_SHIFT
#define PYODESYS_COOP_PIVOT_SHIFT 3  /* natural pivot accepted down to 2^-3 of the column maximum */
#endif
// 16-byte shared-memory moves of a vector of NY doubles (buffer length NYE = NY rounded up to even, 16-byte aligned):
// the shared-memory / shuffle data path is what bounds this kernel (56 % of its peak in profiles/r2e_coop14.txt, the
// FP64 pipe 17 %), and a 16-byte access moves two doubles per wavefront slot
template <int N> __device__ __forceinline__ void lds_vec(const double *buf, double *v) {
    const double2 *b2 = reinterpret_cast<const double2 *>(buf);
#pragma unroll
    for (int j = 0; j < N / 2; ++j) {
        const double2 q = b2[j];
        v[2 * j] = q.x;
        v[2 * j + 1] = q.y;
    }
    if (N & 1) v[N - 1] = buf[N - 1];
}
template <int N> __device__ __forceinline__ void sts_vec(double *buf, const double *v) {
    double2 *b2 = reinterpret_cast<double2 *>(buf);
#pragma unroll
    for (int j = 0; j < N / 2; ++j) b2[j] = make_double2(v[2 * j], v[2 * j + 1]);
    if (N & 1) buf[N - 1] = v[N - 1];
}

// The pivot loop is ROLLED: the kernel is bound by instruction fetch, not by any execution pipe (ncu: the GPC-level
// instruction cache at 78 % of its peak request rate, 13 % of the SM instruction-cache requests missing, while FP64
// sat at 17 % and the shared-memory data path at 56 %; run time did not move with occupancy, block size or spills).
// Fully unrolled, the NY pivot steps were 54 KB of the 94 KB loop body.  Register arrays only allow literal indices,
// so every step works on PHYSICAL column 0 and shifts the columns left by one while it updates them
//     a[r][j-1] <- a[r][j] + m_r * prow[j]   (j = 1..NY-1),      a[r][NY-1] <- m_r  (the new column of the inverse)
// - after NY steps every logical column is back at its own position.
// Two more things keep the instruction count of a step down: the normalisation of the pivot rows is DEFERRED (a
// pivot row is left as it is - multiplier 0, a one in the new column - and every row is scaled by the reciprocal of
// its own pivot once, at the end; row operations commute with that scaling), and the loop runs slot by slot
// (unrolled, R copies of the body) over the lanes of a group (rolled), so that the natural pivot row a[ks] is a
// literal register row and needs no selects.
template <int G, int KS>
__device__ __forceinline__ void c_pivot_step(const Grp<G> &g, double (*a)[NY], int *ord, double *dinv, int k, int kl,
                                             bool &ok, bool &identity, bool part) {
    constexpr int R = Grp<G>::R;
    double *buf = (k & 1) ? g.ys : g.prow;  // two staging vectors in turn: one warp synchronisation per step
    int best = -1, bslot = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int hi = (ord[r] == NY) ? (__double2hiint(a[r][0]) & 0x7fffffff) : -1;
        if (hi > best) {
            best = hi;
            bslot = r;
        }
    }
    const int nat_here = (ord[KS] == NY) ? (__double2hiint(a[KS][0]) & 0x7fffffff) : -1;
    // key: magnitude class (top 24 of the 31 bits) | slot | inverted lane: the maximum is also the first lane on ties
    const int key = (best < 0) ? -1 : ((best & ~127) | (bslot << 5) | (G - 1 - g.gl));
    const int kmax = gmax_i_full<G>(key);
    const int mx = (kmax < 0) ? -1 : (kmax & ~127);
    ok = ok && (mx >= 0x00100000) && (mx < 0x7ff00000);
    const int nat = __shfl_sync(FULL_MASK, nat_here, g.lane0 + kl);
    // (a group that only takes part in the collectives - ``part`` false, see c_invert - never forces the slow path)
    const bool natural = !part || (nat >= 0 && nat + (PYODESYS_COOP_PIVOT_SHIFT << 20) >= mx);
    double prow[NY];
    bool piv[R];
    if (__all_sync(FULL_MASK, natural)) {
        const bool mine = (g.gl == kl);
#pragma unroll
        for (int r = 0; r < R; ++r) piv[r] = (r == KS) && mine;
        if (mine) {
#pragma unroll
            for (int j = 0; j < NY; ++j) buf[j] = a[KS][j];
        }
    } else {
        identity = false;
        const int pl = natural ? kl : G - 1 - (kmax & 31), ps = natural ? KS : (kmax >> 5) & 3;
        const bool mine = (g.gl == pl);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            piv[r] = mine && (ps == r);
            if (piv[r]) {
#pragma unroll
                for (int j = 0; j < NY; ++j) buf[j] = a[r][j];
            }
        }
    }
    __syncwarp();
    if (!part) return;  // (after the last collective of the step; uniform over the warp in the common case)
    lds_vec<NY>(buf, prow);
    const double d = fast_rcp<2>(prow[0]);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (piv[r]) {
            ord[r] = k;
            dinv[r] = d;
        }
        const double m = piv[r] ? 0.0 : -a[r][0] * d;
#pragma unroll
        for (int j = 1; j < NY; ++j) a[r][j - 1] = fma(m, prow[j], a[r][j]);
        a[r][NY - 1] = piv[r] ? 1.0 : m;
    }
}

// ``part`` = false: this group keeps the inverse it has (BDF: groups whose step and Jacobian did not change) and only
// takes part in the warp collectives of the elimination; its a / ord / identity are left alone.
template <int G>
__device__ __forceinline__ bool c_invert(const Grp<G> &g, double (*a)[NY], int *ord, bool &identity, bool part = true) {
    constexpr int R = Grp<G>::R;
    static_assert(R <= 4 && G <= 32, "pivot key packs the slot in 2 bits, the lane in 5");
    bool ok = true;
    bool ident = true;  // every pivot so far was the natural one: W^-1 comes out in variable order
    double dinv[R];
    int ordw[R];  // participants: the order under construction; others: all rows "used" (they publish nothing)
#pragma unroll
    for (int r = 0; r < R; ++r) {
        ordw[r] = (g.owns(r) && part) ? NY : -1;
        dinv[r] = 0.0;
    }
    // natural row k = slot ks, lane kl of the group: slots unrolled (literal register rows), lanes rolled
    if (R > 0) {
#pragma unroll 1
        for (int kl = 0; kl < G && kl < NY; ++kl) c_pivot_step<G, 0>(g, a, ordw, dinv, kl, kl, ok, ident, part);
    }
    if (R > 1) {
#pragma unroll 1
        for (int kl = 0; kl < G && G + kl < NY; ++kl) c_pivot_step<G, 1 % R>(g, a, ordw, dinv, G + kl, kl, ok, ident, part);
    }
    if (R > 2) {
#pragma unroll 1
        for (int kl = 0; kl < G && 2 * G + kl < NY; ++kl) c_pivot_step<G, 2 % R>(g, a, ordw, dinv, 2 * G + kl, kl, ok, ident, part);
    }
    if (R > 3) {
#pragma unroll 1
        for (int kl = 0; kl < G && 3 * G + kl < NY; ++kl) c_pivot_step<G, 3 % R>(g, a, ordw, dinv, 3 * G + kl, kl, ok, ident, part);
    }
    const bool all_ident = __all_sync(FULL_MASK, ident || !part);
    if (part) {
        // deferred normalisation: row p of the inverse is the accumulated row divided by its pivot
#pragma unroll
        for (int r = 0; r < R; ++r) {
            ord[r] = ordw[r];
#pragma unroll
            for (int j = 0; j < NY; ++j) a[r][j] *= dinv[r];
        }
        identity = all_ident;
    }
    return ok;
}

// x = W^-1 b with the in-place inverse of c_invert: a[r][m] = (W^-1)[ord[r]][p_m].  The right-hand side is staged in
// pivot order (bs[ord[r]] = b[r]), every lane reads all of it (broadcast, 16 bytes at a time) and forms the R dot
// products of its rows; the results are in variable order already when every pivot was the natural one (identity),
// otherwise they go through shared memory once more.
template <int G> __device__ __forceinline__ void c_apply(const Grp<G> &g, const double (*a)[NY], const int *ord,
                                                         bool identity, double *b) {
    constexpr int R = Grp<G>::R;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (g.owns(r) && (unsigned)ord[r] < (unsigned)NY) g.prow[ord[r]] = b[r];  // (a group without a valid inverse takes part in the collective only)
    __syncwarp();
    double bs[NY], acc[R];
    lds_vec<NY>(g.prow, bs);
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.0;
#pragma unroll
    for (int m = 0; m < NY; ++m) {
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = fma(a[r][m], bs[m], acc[r]);
    }
    // (warp vote: the groups of a warp may hold inverses from different eliminations - the BDF stepper keeps them
    // across steps - and the permuting path below synchronises the whole warp)
    if (__all_sync(FULL_MASK, identity)) {
#pragma unroll
        for (int r = 0; r < R; ++r) b[r] = g.owns(r) ? acc[r] : 0.0;
    } else {
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (g.owns(r) && (unsigned)ord[r] < (unsigned)NY) g.ys[ord[r]] = acc[r];
        __syncwarp();
#pragma unroll
        for (int r = 0; r < R; ++r) b[r] = g.owns(r) ? g.ys[r * G + g.gl] : 0.0;
    }
}

#define RC(name) kRos4[ros4c::name]
template <int G> struct CoopRos4 {
    typedef CoopLayout<G> L;
    static constexpr int R = L::R;
    double g1[R], g2[R], g3[R], g4[R], g5[R];  // stage increments of the last attempt (dense output needs them)
    double h_old;
    float err_old;
    bool first, last_rejected;

    __device__ __forceinline__ void begin() {
        first = true;
        last_rejected = false;
        h_old = 0.0;
        err_old = 1.0f;
    }

    static constexpr bool kAlwaysInit = false;  // init() only chooses a starting step
    static constexpr int kEpoch = 0;  // no start grid (see CoopBdf)
#ifdef PYODESYS_COOP_TRACE
    __device__ __forceinline__ double trace_order() const { return 4.0; }
    __device__ __forceinline__ double trace_flags() const { return 0.0; }
    __device__ __forceinline__ double trace_extra(int) const { return 0.0; }
#endif
    static constexpr int kExtraDoubles = 0;
    // Hairer's starting step (initial_step in integrators.cuh); executed by every thread of the block (two
    // evaluation slots), the result is used by the groups that asked for it
    __device__ __forceinline__ double init(Coop<G> &c, const Grp<G> &g, const KernelArgs &a, double t0, const double *y,
                                           double span, bool want, bool have_h, int &nfev, int &njev) {
        AnyODE::ignore(njev);
        AnyODE::ignore(have_h);
        int used = 0;
        const double hs = auto_step(c, g, a, t0, y, span, used);
        if (want) nfev += used;
        return hs;
    }
    __device__ __forceinline__ double auto_step(Coop<G> &c, const Grp<G> &g, const KernelArgs &a, double t0,
                                                const double *y, double span, int &nfev) {
        double f0[R], y1[R], f1[R];
        c.template eval<0>(g, a, t0, y, f0);
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
        if (!(h0 > 0.0)) h0 = 1e-6;  // idle group (span 0): keep the arithmetic finite
        const double dir = span >= 0 ? 1.0 : -1.0;
#pragma unroll
        for (int r = 0; r < R; ++r) y1[r] = fma(dir * h0, f0[r], y[r]);
        const int st = c.template eval<0>(g, a, t0 + dir * h0, y1, f1);
        nfev += 2;
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
        const double h1 = (dmax <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : fast_pow(0.01 * fast_rcp<1>(dmax), 0.25f);
        return fmin(fmin(100.0 * h0, h1), fabs(span)) * dir;
    }

    __device__ __forceinline__ int attempt(Coop<G> &c, const Grp<G> &g, const KernelArgs &a, double t, const double *y,
                                           double h, double *ynew, double &err, Counters &cnt, int &nfev, int &njev,
                                           bool live) {
        extern __shared__ double c_smem[];
        if (live) {
            nfev += 6;
            njev += 1;
        }
        double W[R][NY], dfdt[R], f[R], tmp[R], b[R];
        int ord[R];
        bool identity;
        int bad = c.template eval<1>(g, a, t, y, f);
        const double ih = fast_rcp<2>(h);
        const double diag = 4.0 * ih;  // 1/gamma = 4
        // W = I/(gamma h) - J: the rows this lane owns, from the Jacobian the evaluator left in shared memory
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int i = g.owns(r) ? r * G + g.gl : 0;
            const double *row = c_smem + L::oJ + (i * L::NYP) * L::NSYSP + c.sysl;
#pragma unroll
            for (int j = 0; j < NY; ++j) {
                const double v = row[j * L::NSYSP];
                W[r][j] = g.owns(r) ? ((i == j) ? diag - v : -v) : 0.0;
            }
            dfdt[r] = (!ODESYS_AUTONOMOUS && g.owns(r)) ? c_smem[L::oDF + i * L::NSYSP + c.sysl] : 0.0;
        }
        {
            COOP_T0(t_inv);
            if (!c_invert<G>(g, W, ord, identity)) bad |= 1;
            COOP_ACC(2, t_inv);
        }
#ifdef PYODESYS_COOP_PROFILE
#define G_SOLVE(vec) { COOP_T0(t_s); c_apply<G>(g, W, ord, identity, vec); COOP_ACC(3, t_s); }
#else
#define G_SOLVE(vec) c_apply<G>(g, W, ord, identity, vec)
#endif
        cnt.nlu += 1;
        cnt.nsolve += 6;
#pragma unroll
        for (int r = 0; r < R; ++r) b[r] = ODESYS_AUTONOMOUS ? f[r] : fma(h * RC(D1), dfdt[r], f[r]);
        G_SOLVE(b);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g1[r] = b[r];
            tmp[r] = fma(RC(A21), b[r], y[r]);
        }
        bad |= c.template eval<0>(g, a, fma(RC(C2X), h, t), tmp, f);
        const double c21 = RC(C21) * ih;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            b[r] = fma(c21, g1[r], f[r]);
            if (!ODESYS_AUTONOMOUS) b[r] = fma(h * RC(D2), dfdt[r], b[r]);
        }
        G_SOLVE(b);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g2[r] = b[r];
            tmp[r] = fma(RC(A32), b[r], fma(RC(A31), g1[r], y[r]));
        }
        bad |= c.template eval<0>(g, a, fma(RC(C3X), h, t), tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            b[r] = fma(ih, fma(RC(C32), g2[r], RC(C31) * g1[r]), f[r]);
            if (!ODESYS_AUTONOMOUS) b[r] = fma(h * RC(D3), dfdt[r], b[r]);
        }
        G_SOLVE(b);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g3[r] = b[r];
            tmp[r] = fma(RC(A43), b[r], fma(RC(A42), g2[r], fma(RC(A41), g1[r], y[r])));
        }
        bad |= c.template eval<0>(g, a, fma(RC(C4X), h, t), tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            b[r] = fma(ih, fma(RC(C43), g3[r], fma(RC(C42), g2[r], RC(C41) * g1[r])), f[r]);
            if (!ODESYS_AUTONOMOUS) b[r] = fma(h * RC(D4), dfdt[r], b[r]);
        }
        G_SOLVE(b);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g4[r] = b[r];
            tmp[r] = fma(RC(A54), b[r], fma(RC(A53), g3[r], fma(RC(A52), g2[r],
                                                                          fma(RC(A51), g1[r], y[r]))));
        }
        bad |= c.template eval<0>(g, a, t + h, tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r)
            b[r] = fma(ih, fma(RC(C54), g4[r], fma(RC(C53), g3[r],
                               fma(RC(C52), g2[r], RC(C51) * g1[r]))), f[r]);
        G_SOLVE(b);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            g5[r] = b[r];
            tmp[r] += b[r];
        }
        bad |= c.template eval<0>(g, a, t + h, tmp, f);
#pragma unroll
        for (int r = 0; r < R; ++r)
            b[r] = fma(ih, fma(RC(C65), g5[r], fma(RC(C64), g4[r], fma(RC(C63), g3[r],
                               fma(RC(C62), g2[r], RC(C61) * g1[r])))), f[r]);
        G_SOLVE(b);
#undef G_SOLVE
        double e = 0.0;
        bool finite = true;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            ynew[r] = tmp[r] + b[r];
            if (!g.owns(r)) continue;
            finite = finite && (fabs(ynew[r]) <= DBL_MAX);
            const double sk = fma(a.rtol, fmax(fabs(y[r]), fabs(ynew[r])), a.atol[r * G + g.gl]) + 1e-300;
            e = fmax(e, fabs(b[r]) * fast_rcp<1>(sk));
        }
        const float ef = gmax_f<G>(g, (float)fmin(e, 3e38));
        err = (double)ef;
        if (!(ef <= 3e38f) || __any_sync(g.mask, !finite)) bad |= 1;
        return bad;
    }

    // rosenbrock4_controller::try_step (same rule as GroupRos4::control)
    __device__ __forceinline__ bool control(const Coop<G> &, const Grp<G> &, double err, double &hnext, double h) {
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

    __device__ __forceinline__ void dense(const Coop<G> &c, const Grp<G> &g, double theta, double, const double *yold,
                                          const double *ynew, double *out) {
        AnyODE::ignore(c);
        AnyODE::ignore(g);
        const double s1 = 1.0 - theta;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double q1 = g1[r], q2 = g2[r], q3 = g3[r], q4 = g4[r], q5 = g5[r];
            const double c3 = fma(RC(D25), q5, fma(RC(D24), q4, fma(RC(D23), q3, fma(RC(D22), q2, RC(D21) * q1))));
            const double c4 = fma(RC(D35), q5, fma(RC(D34), q4, fma(RC(D33), q3, fma(RC(D32), q2, RC(D31) * q1))));
            out[r] = yold[r] * s1 + theta * (ynew[r] + s1 * (c3 + theta * c4));
        }
    }
};
#undef RC

// ----------------------------------------------------------------------------------------------
// Persistent kernel.  MODE 0: predefined output points, MODE 1: every accepted step.  One iteration of the outer
// loop = one Rosenbrock step attempt of every system of the block (lock-step), preceded - when some group just
// loaded a system without a given first step - by the two evaluations of the starting-step heuristic.
// ----------------------------------------------------------------------------------------------
template <class St, int G, int MODE> __device__ __forceinline__ void coop_kernel_body(const KernelArgs &a) {
    typedef CoopLayout<G, St::kExtraDoubles> L;
    constexpr int R = L::R;
    extern __shared__ double c_smem[];
    const int lane = threadIdx.x & 31u;
    Grp<G> g;
    g.gl = lane % G;
    g.lane0 = lane - g.gl;
    g.mask = (G == 32) ? FULL_MASK : (((1u << (G & 31)) - 1u) << g.lane0);
    Coop<G> c;
    c.sysl = threadIdx.x / G;
    c.evalno = 0;
#ifdef PYODESYS_COOP_PROFILE
    for (int k = 0; k < 10; ++k) c.prof[k] = 0;
    const long long t_all = clock64();
#endif
    g.prow = c_smem + L::oScr + (size_t)c.sysl * 2 * L::NYE;
    g.ys = g.prow + L::NYE;
    int *ist = reinterpret_cast<int *>(c_smem + L::oInt);
    for (int k = threadIdx.x; k < 4 * L::NSYS; k += kCoopBlock) ist[k] = 0;
    for (int k = threadIdx.x; k < L::oScr - L::oP; k += kCoopBlock) c_smem[L::oP + k] = 0.0;  // finite garbage for idle slots
    __syncthreads();

    const int64_t ngroups = (int64_t)gridDim.x * L::NSYS;
    int64_t s = (int64_t)blockIdx.x * L::NSYS + c.sysl - ngroups;  // static schedule cursor

    St st;
    st.begin();
    double y[R], t = 0, h = 1.0, tend = 0, dir = 1, hmin = 0, hmax = INFINITY;
    bool have_h = false;
    int64_t steps_left = 0;
    int k_out = 0, restarts = 0, nfail = 0, nfev = 0, njev = 0;
    Counters cnt = {0, 0, 0, 0};
    bool active = false, exhausted = false, need_h = false;
#pragma unroll
    for (int r = 0; r < R; ++r) y[r] = 0.0;

#ifdef PYODESYS_COOP_DEBUG_ITERS
    long long dbg_iters = 0;
#endif
    int epoch_phase = 0;  // St::kEpoch > 0: idle groups draw their next system every kEpoch-th iteration only
    for (;;) {
#ifdef PYODESYS_COOP_DEBUG_ITERS
        if (++dbg_iters > PYODESYS_COOP_DEBUG_ITERS) break;  // debugging aid: tells a livelock from a deadlock
#endif
        bool may_fetch = true;
        if (St::kEpoch > 0) {
            may_fetch = epoch_phase == 0;
            epoch_phase = epoch_phase + 1 == St::kEpoch ? 0 : epoch_phase + 1;
        }
        COOP_T0(t_top);
        // ---- (B) + (E): termination test of the running system, retire ----------------------------------------
        if (active) {
            int status = -1;
            if (nfail >= (1 << 20)) status = ST_RHS;
            else if (nfail == -1) status = ST_STEP_TOO_SMALL;
            else if (nfail == -2) status = ST_ROWS;
            else if (t == tend) status = ST_OK;
            else if (steps_left <= 0) {
                if (restarts > 0) {  // autorestart (_cvode_wrapper.pyx:90): new step budget, re-initialised stepper
                    restarts -= 1;
                    steps_left = a.mxsteps;
                    st.begin();
                    need_h = true;
                    have_h = false;
                } else {
                    status = ST_MXSTEPS;
                }
            }
            if (status >= 0) {
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
                    if (a.counters) {
                        a.counters[C_NSTEPS * a.counters_ld + s] = cnt.nsteps;
                        a.counters[C_NFEV * a.counters_ld + s] = nfev;
                        a.counters[C_NJEV * a.counters_ld + s] = njev;
                        a.counters[C_NLU * a.counters_ld + s] = cnt.nlu;
                        a.counters[C_NSOLVE * a.counters_ld + s] = cnt.nsolve;
                        a.counters[C_NREJECT * a.counters_ld + s] = cnt.nreject;
                    }
                }
                active = false;
            }
        }
        // ---- (A) fetch a new system for an idle group ------------------------------------------------------------
        if (may_fetch && !active && !exhausted) {
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
                for (int i = g.gl; i < NP_USED; i += G) {
                    double v;
                    if (i >= ODESYS_NPARAMS && a.iv_from_y0) v = a.y0[s * NY + (i - ODESYS_NPARAMS)];
                    else v = a.params[s * a.params_ld + i];
                    c_smem[L::oP + i * L::NSYS + c.sysl] = v;
                }
                if (CoopSplit<G>::value) {
                    // parameter-only sub-expressions (template:104-110), once per system, by the first lane of the
                    // group: with four evaluating warps nobody could own the "first evaluation" of a slot
#if ODESYS_NPCSE > 0
                    __syncwarp(g.mask);
                    if (g.gl == 0) {
                        typename L::ESys es;
                        es.m_p = SView<L::NSYS>{c_smem + L::oP + c.sysl};
                        es.m_p_cse = SView<L::NSYS>{c_smem + L::oPC + c.sysl};
                        es.setup();
                    }
#endif
                } else if (g.gl == 0) {
                    ist[L::NSYS + c.sysl] = 1;  // the evaluator runs OdeSys::setup() for this slot
                }
                cnt = {0, 0, 0, 0};
                nfev = 0;
                njev = 0;
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
                st.begin();
                h = a.dx0 ? a.dx0[s] : 0.0;
                have_h = (h != 0.0);
                need_h = St::kAlwaysInit || !have_h;
                if (have_h) h = fabs(h) * dir;
                active = true;
                bool fin = true;
#pragma unroll
                for (int r = 0; r < R; ++r) fin = fin && (fabs(y[r]) <= DBL_MAX);
                if (__any_sync(g.mask, !fin)) nfail = 1 << 20;
                if (tend == t) {
                    need_h = false;
                    if (MODE == 0) {  // zero-length span: every requested point equals the start
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
        }
        COOP_ACC(6, t_top);
        COOP_T0(t_bar);
        if (__syncthreads_and(!active && exhausted)) break;
        COOP_ACC(7, t_bar);
        // a system that is about to retire (finished span, failure) idles through this iteration
        const bool live = active && t != tend && nfail >= 0 && nfail < (1 << 20);

        // ---- starting step size where none was given: two evaluation slots, taken by the whole block ----------------
        if (__syncthreads_or(live && need_h)) {
            const bool want = live && need_h;
            const double hs = st.init(c, g, a, t, y, live ? tend - t : 0.0, want, have_h, nfev, njev);
            if (want && !have_h) h = hs;
            need_h = false;
        }

        // ---- (C) one trial step of every system of the block -----------------------------------------------------------
        COOP_T0(t_att);
        double hh = live ? h : 1.0;
        if (fabs(hh) > hmax) hh = dir * hmax;
        if (fabs(hh) < hmin) hh = dir * hmin;
        const double rem = tend - t;
        const bool lands = live && fabs(hh) >= fabs(rem) * (1.0 - 1e-13);
        double tnew = t + hh;
        if (lands) {
            hh = rem;
            tnew = tend;
        }
        double ynew[R], err = 0.0, hnext = hh;
        Counters scratch = cnt;
        const int bad = st.attempt(c, g, a, t, y, hh, ynew, err, live ? cnt : scratch, nfev, njev, live);
        COOP_ACC(8, t_att);
        if (!live) continue;
        COOP_T0(t_post);
        cnt.nsteps += 1;
        steps_left -= 1;
        bool ok;
        if (bad) {
            ok = false;
            hnext = 0.25 * hh;
            nfail += 1;
            if (nfail >= 12) nfail = 1 << 20;
            double ignored;
            st.control(c, g, 2.0, ignored, hh);
        } else {
            ok = st.control(c, g, err, hnext, hh);
            if (ok) nfail = 0;
        }
#ifdef PYODESYS_COOP_TRACE
        // debugging aid: one record of 12 doubles per attempt of system 0 at KernelArgs::root_out
        if (s == 0 && g.gl == 0 && cnt.nsteps <= 4000) {
            double *tr = a.root_out + 12 * (cnt.nsteps - 1);
            tr[0] = t; tr[1] = hh; tr[2] = err; tr[3] = ok ? 1.0 : 0.0; tr[4] = bad; tr[5] = st.trace_order();
            tr[6] = hnext; tr[7] = st.trace_flags();
            for (int q = 0; q < 4; ++q) tr[8 + q] = st.trace_extra(q);
        }
#endif
        if (!ok) {
            cnt.nreject += 1;
            if (fabs(hnext) <= 16.0 * DBL_EPSILON * fabs(t) || fabs(hnext) == 0.0 || (hmin > 0.0 && fabs(hh) <= hmin))
                nfail = -1;  // ST_STEP_TOO_SMALL at the top of the next iteration
            h = hnext;
        } else {
            // ---- (D) accepted: emit outputs -----------------------------------------------------------------------
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
                        st.dense(c, g, (xk - t) / hh, hh, y, ynew, yo);
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
                    nfail = -2;  // ST_ROWS
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
        }
        COOP_ACC(9, t_post);
    }
#ifdef PYODESYS_COOP_PROFILE
    if (lane == 0) {
        unsigned long long *out = reinterpret_cast<unsigned long long *>(a.root_out);
        atomicAdd(out + 0, (unsigned long long)c.prof[0]);
        atomicAdd(out + 1, (unsigned long long)c.prof[1]);
        atomicAdd(out + 2, (unsigned long long)c.prof[2]);
        atomicAdd(out + 3, (unsigned long long)c.prof[3]);
        atomicAdd(out + 4, (unsigned long long)(clock64() - t_all));
        atomicAdd(out + 5, 1ull);
        for (int k = 6; k < 10; ++k) atomicAdd(out + k, (unsigned long long)c.prof[k]);
    }
#endif
}

template <int G, int MODE>
__global__ void __launch_bounds__(kCoopBlock, PYODESYS_COOP_MIN_BLOCKS) coop_ros4_kernel(const KernelArgs a) {
    coop_kernel_body<CoopRos4<G>, G, MODE>(a);
}

#endif  // ODESYS_HAS_JAC

}  // namespace pyodesys_cuda
