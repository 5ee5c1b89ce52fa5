// C-ABI shim of one generated ODE system (see include/pyodesys_cuda.h for the contract).
//
// Plays the role of the Cython batch wrappers of the reference
// (pyodesys/native/sources/_odeint_wrapper.pyx:37-177, _cvode_wrapper.pyx:78-365): validate,
// stage the row-major arrays, run the batch, hand back per-system status/counters.  Where the
// wrappers ``new`` one OdeSys per row under the GIL (_odeint_wrapper.pyx:158-160) and fan out
// with OpenMP, this shim copies the arrays to HBM once and launches one persistent kernel.
// Compiled once per system:  nvcc -gencode arch=compute_100a,code=sm_100a -shared ...
#include "odesys_generated.h"

#include "integrators.cuh"
// BDF keeps 8 difference vectors + J + LU per lane in registers
#define ODESYS_THREAD_BDF (ODESYS_HAS_JAC && ODESYS_NY <= 8)
#if ODESYS_THREAD_BDF
#include "bdf_integrator.cuh"
#endif
// Kernel families are only compiled where they can matter (the fully unrolled register code explodes for large ny):
#define ODESYS_THREAD_ROS4 (ODESYS_HAS_JAC && ODESYS_NY <= 16)
#define ODESYS_SMEM_KERNEL (ODESYS_HAS_JAC && ODESYS_NY >= 9)
#define ODESYS_GROUP_KERNEL (ODESYS_HAS_JAC && ODESYS_NY >= 6)
#if ODESYS_SMEM_KERNEL
#include "smem_integrators.cuh"
#endif
#if ODESYS_GROUP_KERNEL
#include "group_integrators.cuh"
#endif
// block-cooperative kernel: lane groups for the linear algebra, one system per lane for f and J (default for ny > 8)
#define ODESYS_COOP_KERNEL (ODESYS_GROUP_KERNEL && ODESYS_NY >= 9)
#if ODESYS_COOP_KERNEL
#include "coop_integrators.cuh"
#include "coop_bdf.cuh"
#endif
#include "pyodesys_cuda.h"

#include <algorithm>
#include <exception>
#include <ctype.h>
#include <sched.h>
#include <string>
#include <vector>
#include <atomic>
#include <thread>
#include <chrono>
#include <mutex>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

using namespace pyodesys_cuda;

namespace {

thread_local char g_err[512] = "";
// systems with status != 0 in the last integrate call of this thread, where the library counted them while it waited
// for the device anyway (-1: not counted; the caller screens the status array itself)
thread_local int64_t g_last_failed = -1;
std::atomic<int64_t> g_launches{0};

int fail(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return -1;
}

#define CU_CHECK(expr)                                                                        \
    do {                                                                                      \
        cudaError_t e_ = (expr);                                                              \
        if (e_ != cudaSuccess) return fail("%s failed: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

constexpr int kMaxDev = 64;
constexpr int kMaxSlices = 64;
#ifndef PYODESYS_BLOCK
#define PYODESYS_BLOCK 128
#endif
constexpr int kBlock = PYODESYS_BLOCK;

enum Buf { B_Y0, B_PARAMS, B_XOUT, B_XEND, B_DX0, B_DXMIN, B_DXMAX, B_YOUT, B_NREACHED, B_STATUS, B_COUNTERS, B_WORK, B_ROOTOUT, B_ROOTIDX, B_NROOTS, B_READY, B_SLICEDONE, B_OFFSETS, B_BLOCKSUMS, B_COMPACT, NBUF };

struct DevCtx {
    std::mutex mtx;
    bool ready = false;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;  // second lane of the chunk pipeline (host-pointer calls)
    cudaEvent_t ev[4];
    cudaEvent_t ev_k0, ev_k1, ev_fork, ev_join;
    cudaEvent_t ev_slice[kMaxSlices] = {};  // streaming calls: "the results of slice j are on the host" (created lazily)
    void *buf[NBUF] = {};
    size_t cap[NBUF] = {};
    int sm_count = 0;
    // streaming calls: pinned table of "systems ready" values (source of the 8-byte flag copies) and the
    // host-mapped completion flags the kernel sets per output slice
    unsigned long long *h_ready = nullptr;
    unsigned int *h_flags = nullptr, *d_flags = nullptr;
    int64_t compact_rows = 0;  // rows left in buf[B_COMPACT] by the last compacting adaptive call
};
#ifndef PYODESYS_CUDA_SLICE_MB_DEFAULT
#define PYODESYS_CUDA_SLICE_MB_DEFAULT 8.0
#endif
constexpr int kLanes = 2;  // independent call contexts per device (options.ctx_lane): two host threads can keep one GPU busy
DevCtx g_ctx[kMaxDev][kLanes];
int lane_of(const pyodesys_cuda_options *opt) { return (opt && opt->ctx_lane > 0 && opt->ctx_lane < kLanes) ? opt->ctx_lane : 0; }

int ctx_init(DevCtx &c, int device) {
    CU_CHECK(cudaSetDevice(device));
    if (c.ready) return 0;
    cudaDeviceProp prop;
    CU_CHECK(cudaGetDeviceProperties(&prop, device));
    c.sm_count = prop.multiProcessorCount;
    CU_CHECK(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    CU_CHECK(cudaStreamCreateWithFlags(&c.stream2, cudaStreamNonBlocking));
    for (auto &e : c.ev) CU_CHECK(cudaEventCreate(&e));
    CU_CHECK(cudaEventCreate(&c.ev_k0));
    CU_CHECK(cudaEventCreate(&c.ev_k1));
    CU_CHECK(cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming));
    CU_CHECK(cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming));
    CU_CHECK(cudaHostAlloc((void **)&c.h_ready, kMaxSlices * sizeof(unsigned long long), cudaHostAllocPortable));
    // [kMaxSlices] completion flags + [1] abort word (kernel <-> host, see wait_inputs in integrators.cuh)
    CU_CHECK(cudaHostAlloc((void **)&c.h_flags, (kMaxSlices + 1) * sizeof(unsigned int),
                           cudaHostAllocPortable | cudaHostAllocMapped));
    CU_CHECK(cudaHostGetDevicePointer((void **)&c.d_flags, c.h_flags, 0));
    c.ready = true;
    return 0;
}

int ensure(DevCtx &c, Buf b, size_t bytes) {
    if (bytes == 0) bytes = 8;
    if (c.cap[b] >= bytes) return 0;
    if (c.buf[b]) CU_CHECK(cudaFree(c.buf[b]));
    c.buf[b] = nullptr;
    c.cap[b] = 0;
    CU_CHECK(cudaMalloc(&c.buf[b], bytes));
    c.cap[b] = bytes;
    return 0;
}

template <int METHOD, int MODE> int launch(DevCtx &c, const KernelArgs &a, cudaStream_t stream, int blocks_per_sm) {
    auto kern = integrate_kernel<METHOD, MODE, kBlock, false>;
#if ODESYS_NINVAR > 0
    if (a.max_invariant_violation != 0.0) kern = integrate_kernel<METHOD, MODE, kBlock, true>;
#endif
    // steppers whose state is addressed dynamically (BDF difference table) keep it in shared memory
    const size_t smem = (size_t)Stepper<METHOD>::kSharedDoublesPerThread * kBlock * sizeof(double);
    if (smem > 48 * 1024) CU_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kBlock, smem));
    if (occ < 1) occ = 1;
    if (blocks_per_sm > 0 && blocks_per_sm < occ) occ = blocks_per_sm;
    // persistent grid: a multiple of the SM count, never more blocks than there is work
    int64_t grid = (int64_t)c.sm_count * occ;
    const int64_t need = (a.n + kBlock - 1) / kBlock;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kBlock, smem, stream>>>(a);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    return 0;
}

#if ODESYS_GROUP_KERNEL
constexpr int kGroup = PYODESYS_GROUP_G;
template <int MODE> int launch_group(DevCtx &c, const KernelArgs &a, cudaStream_t stream, int blocks_per_sm) {
    auto kern = group_ros4_kernel<kGroup, MODE, kBlock>;
    const size_t smem = (size_t)(kBlock / kGroup) * kGroupSmemDoubles * sizeof(double);
    if (smem > 48 * 1024) CU_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kBlock, smem));
    if (occ < 1) occ = 1;
    if (blocks_per_sm > 0 && blocks_per_sm < occ) occ = blocks_per_sm;
    int64_t grid = (int64_t)c.sm_count * occ;
    const int64_t need = (a.n + (kBlock / kGroup) - 1) / (kBlock / kGroup);
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kBlock, smem, stream>>>(a);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    return 0;
}
#endif

#if ODESYS_COOP_KERNEL
constexpr int kCoopG = PYODESYS_COOP_G;
template <int MODE> int launch_coop(DevCtx &c, const KernelArgs &a, cudaStream_t stream, int blocks_per_sm) {
    auto kern = coop_ros4_kernel<kCoopG, MODE>;
    typedef CoopLayout<kCoopG> L;
    const size_t smem = (size_t)L::kDoubles * sizeof(double);
    if (smem > 227 * 1024) return fail("ny = %d needs %zu bytes of shared memory per block in the cooperative kernel; use kernel='group'", NY, smem);
    CU_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kCoopBlock, smem));
    if (occ < 1) occ = 1;
    if (blocks_per_sm > 0 && blocks_per_sm < occ) occ = blocks_per_sm;
    int64_t grid = (int64_t)c.sm_count * occ;
    const int64_t need = (a.n + L::NSYS - 1) / L::NSYS;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kCoopBlock, smem, stream>>>(a);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    return 0;
}
#endif

#if ODESYS_COOP_KERNEL
constexpr int kCoopBdfG = PYODESYS_COOP_BDF_G;
template <int MODE> int launch_coop_bdf(DevCtx &c, const KernelArgs &a, cudaStream_t stream, int blocks_per_sm) {
    auto kern = coop_bdf_kernel<kCoopBdfG, MODE>;
    typedef CoopBdf<kCoopBdfG>::L L;
    const size_t smem = (size_t)L::kDoubles * sizeof(double);
    if (smem > 227 * 1024) return fail("ny = %d needs %zu bytes of shared memory per block in the cooperative BDF kernel", NY, smem);
    CU_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kCoopBlock, smem));
    if (occ < 1) occ = 1;
    if (blocks_per_sm > 0 && blocks_per_sm < occ) occ = blocks_per_sm;
    int64_t grid = (int64_t)c.sm_count * occ;
    const int64_t need = (a.n + L::NSYS - 1) / L::NSYS;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kCoopBlock, smem, stream>>>(a);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    return 0;
}
#endif

#if ODESYS_SMEM_KERNEL
// one system per thread, W / vectors / parameters in shared memory, rolled loops (one warp per block)
template <int MODE> int launch_smem(DevCtx &c, const KernelArgs &a, cudaStream_t stream, int blocks_per_sm) {
    auto kern = smem_ros4_kernel<MODE>;
    const size_t smem = (size_t)kSmemDoublesPerThread * kSB * sizeof(double);
    if (smem > 227 * 1024) return fail("ny = %d needs %zu bytes of shared memory per warp; use kernel='group'", NY, smem);
    CU_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kSB, smem));
    if (occ < 1) occ = 1;
    if (blocks_per_sm > 0 && blocks_per_sm < occ) occ = blocks_per_sm;
    int64_t grid = (int64_t)c.sm_count * occ;
    const int64_t need = (a.n + kSB - 1) / kSB;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kSB, smem, stream>>>(a);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    return 0;
}
#endif

// Which kernel family integrates this call: 1 = one system per thread in registers, 2 = one system per lane
// group, 3 = one system per thread with its arrays in shared memory.  Auto rule from measurements on B200
// (radiolysis ny = 14, ms per 1e5 systems: lane group with in-place inverse 85, registers 104-140 (4.3 KB of spills
// per thread), shared memory 198; profiles/r1_c4_kernels.md, r1e_c4_group_inverse.md): registers up to ny = 8, the
// lane-group kernel above.
int pick_kernel(int method, const pyodesys_cuda_options *opt) {
    const int want = opt ? opt->kernel : 0;
    if (want == 1 || method != PYODESYS_CUDA_ROSENBROCK4) return (want == 2 || want == 3 || want == 4) ? -1 : 1;
#if ODESYS_SMEM_KERNEL
    if (want == 3) return 3;
#endif
#if ODESYS_COOP_KERNEL
    {
        // what the cooperative kernel does not carry stays with the lane-group kernel
        const bool extras = opt && ((ODESYS_NROOTS > 0 && opt->max_roots > 0) || opt->max_invariant_violation != 0.0 ||
                                    (ODESYS_HAS_DX_MAX && opt->get_dx_max_factor > 0.0));
        if (want == 4) return (extras || ODESYS_HAS_FIRST_STEP) ? -1 : 4;
        if (want == 0 && NY > 8 && !extras && !ODESYS_HAS_FIRST_STEP) return 4;
    }
#endif
#if ODESYS_GROUP_KERNEL
    if (want == 2) return 2;
    if (want == 0 && NY > 8) return 2;
#endif
    return (want == 0 || want == 1) ? 1 : -1;
}

template <int MODE> int dispatch(DevCtx &c, const KernelArgs &a, int method, cudaStream_t stream, int bps,
                                 const pyodesys_cuda_options *opt) {
    const int family = pick_kernel(method, opt);
    if (family < 0)
        return fail("kernel=%d is not available for this system/method (lane-group kernel: rosenbrock4, Jacobian, ny >= 6; "
                    "shared-memory thread kernel and block-cooperative kernel: rosenbrock4, Jacobian, ny >= 9; the "
                    "cooperative kernel carries no event functions / invariant guard / get_dx0 / get_dx_max)",
                    opt ? opt->kernel : 0);
#if ODESYS_GROUP_KERNEL
    if (family == 2) return launch_group<MODE>(c, a, stream, bps);
#endif
#if ODESYS_COOP_KERNEL
    if (family == 4) return launch_coop<MODE>(c, a, stream, bps);
#endif
#if ODESYS_SMEM_KERNEL
    if (family == 3) return launch_smem<MODE>(c, a, stream, bps);
#endif
    switch (method) {
    case PYODESYS_CUDA_RK4: return launch<M_RK4, MODE>(c, a, stream, bps);
    case PYODESYS_CUDA_DOPRI5: return launch<M_DOPRI5, MODE>(c, a, stream, bps);
#if ODESYS_THREAD_ROS4
    case PYODESYS_CUDA_ROSENBROCK4: return launch<M_ROS4, MODE>(c, a, stream, bps);
#elif ODESYS_HAS_JAC
    case PYODESYS_CUDA_ROSENBROCK4: return fail("the register Rosenbrock kernel is limited to ny <= 16; use kernel 0 (auto) or 2");
#else
    case PYODESYS_CUDA_ROSENBROCK4: return fail("rosenbrock4 needs a Jacobian but the system was generated with jac=False");
#endif
#if ODESYS_THREAD_BDF
    case PYODESYS_CUDA_BDF:
#if ODESYS_HAS_JTIMES
        if (a.linear_solver == PYODESYS_CUDA_LINSOL_GMRES) return launch<M_BDF_KRYLOV, MODE>(c, a, stream, bps);
#endif
        return launch<M_BDF, MODE>(c, a, stream, bps);
#elif ODESYS_COOP_KERNEL
    case PYODESYS_CUDA_BDF: {
        // ny > 8: the block-cooperative BDF (lane groups for the linear algebra, one system per lane for f and J)
        const bool extras = opt && ((ODESYS_NROOTS > 0 && opt->max_roots > 0) || opt->max_invariant_violation != 0.0 ||
                                    (ODESYS_HAS_DX_MAX && opt->get_dx_max_factor > 0.0) ||
                                    opt->linear_solver != PYODESYS_CUDA_LINSOL_DENSE);
        if (extras || ODESYS_HAS_FIRST_STEP)
            return fail("method bdf for ny > 8 runs on the block-cooperative kernel, which carries no event functions / "
                        "invariant guard / get_dx0 / get_dx_max / matrix-free solver");
        return launch_coop_bdf<MODE>(c, a, stream, bps);
    }
#elif ODESYS_HAS_JAC
    case PYODESYS_CUDA_BDF: return fail("the BDF kernel keeps its difference table in registers and is limited to ny <= 8; use rosenbrock4");
#else
    case PYODESYS_CUDA_BDF: return fail("bdf needs a Jacobian but the system was generated with jac=False");
#endif
    default: return fail("unknown method id %d", method);
    }
}

struct Common {
    int device;
    int64_t n;
    const double *y0, *params, *atol;
    int natol;
    double rtol;
    const double *dx0, *dx_min, *dx_max;
    int64_t mxsteps;
    int method, autorestart;
    int32_t *status, *counters;
    double *timings;
    int ptr_kind;
    const pyodesys_cuda_options *opt;
};

int check_common(const Common &q) {
    if (q.device < 0 || q.device >= kMaxDev) return fail("invalid device %d", q.device);
    if (q.n < 0) return fail("negative batch size");
    if (q.natol != 1 && q.natol != NY) return fail("atol needs to be of length 1 or %d", NY);
    if (!q.atol) return fail("atol is NULL");
    if (!q.y0 && q.n > 0) return fail("y0 is NULL");
    const bool ivy = q.opt && q.opt->iv_from_y0;
    const int pcols = ivy ? ODESYS_NPARAMS : NP_TOT;
    if (!q.params && pcols > 0 && q.n > 0) return fail("params is NULL but the system has %d parameters", pcols);
    if (q.opt && q.opt->params_ld != 0 && q.opt->params_ld < pcols) return fail("params_ld %d < %d columns", q.opt->params_ld, pcols);
    if (!q.status && q.n > 0) return fail("status is NULL");
    if (q.ptr_kind != 0 && q.ptr_kind != 1) return fail("ptr_kind must be 0 (host) or 1 (device)");
    if (!(q.rtol >= 0.0)) return fail("rtol must be non-negative");
    if (q.opt && q.opt->counters_ld != 0 && q.opt->counters_ld < q.n) return fail("counters_ld %lld < n", (long long)q.opt->counters_ld);
    if (ODESYS_NROOTS > 0 && q.opt && q.opt->max_roots > 0) {
        if (!q.opt->root_out || !q.opt->root_indices || !q.opt->nroots_found)
            return fail("max_roots > 0 needs root_out, root_indices and nroots_found");
        if (pick_kernel(q.method, q.opt) == 3)
            return fail("the shared-memory thread kernel does not locate event functions: pass kernel=1 or kernel=2");
    }
    if (q.opt && q.opt->linear_solver != PYODESYS_CUDA_LINSOL_DENSE) {
        if (q.opt->linear_solver != PYODESYS_CUDA_LINSOL_GMRES) return fail("unknown linear_solver %d", q.opt->linear_solver);
        if (!ODESYS_HAS_JTIMES) return fail("linear_solver gmres needs a system generated with jtimes=True");
        if (q.method != PYODESYS_CUDA_BDF || !ODESYS_THREAD_BDF)
            return fail("linear_solver gmres is the matrix-free Newton iteration of method bdf (ny <= 8)");
    }
    if (q.opt && q.opt->max_invariant_violation != 0.0) {
        if (ODESYS_NINVAR == 0) return fail("max_invariant_violation given but the system defines no invariants");
        if (pick_kernel(q.method, q.opt) == 3) return fail("the shared-memory thread kernel has no invariant guard");
    }
    return 0;
}

void fill_args(KernelArgs &a, const Common &q) {
    memset(&a, 0, sizeof(a));
    a.n = q.n;
    a.counters_ld = (q.ptr_kind == 1 && q.opt && q.opt->counters_ld > 0) ? q.opt->counters_ld : q.n;
    // calibrated local-error tolerances: kappa(method) * (atol, rtol), see odesys_prelude.h
    for (int i = 0; i < NY; ++i) a.atol[i] = odesys_kappa(q.method) * q.atol[q.natol == 1 ? 0 : i];
    a.rtol = odesys_effective_rtol(q.method, q.rtol);
    a.mxsteps = q.mxsteps;
    a.autorestart = q.autorestart;
    a.iv_from_y0 = (q.opt && q.opt->iv_from_y0) ? 1 : 0;
    a.params_ld = (q.opt && q.opt->params_ld > 0) ? q.opt->params_ld : (a.iv_from_y0 ? ODESYS_NPARAMS : NP_TOT);
    a.get_dx_max_factor = q.opt ? q.opt->get_dx_max_factor : 0.0;
    a.error_outside_bounds = q.opt ? q.opt->error_outside_bounds : 0;
    a.max_invariant_violation = q.opt ? q.opt->max_invariant_violation : 0.0;
    a.linear_solver = q.opt ? q.opt->linear_solver : 0;
#if ODESYS_NROOTS > 0
    if (q.opt && q.opt->max_roots > 0) {
        a.max_roots = q.opt->max_roots;
        a.return_on_root = q.opt->return_on_root;
    }
#endif
}

// row pitch in bytes of the caller's counters array: [NCOUNT, counters_ld] (options) or dense [NCOUNT, n]
size_t host_counters_pitch(const Common &q) {
    const int64_t ld = (q.opt && q.opt->counters_ld > 0) ? q.opt->counters_ld : q.n;
    return (size_t)ld * 4;
}

// event-function outputs requested and available?
bool want_roots(const Common &q) { return ODESYS_NROOTS > 0 && q.opt && q.opt->max_roots > 0; }

bool want_dynamic(const Common &q) {
    const int sched = q.opt ? q.opt->schedule : 0;
    if (sched == 1) return false;
    if (sched == 2) return true;
    return q.method != PYODESYS_CUDA_RK4;  // adaptive steppers have data dependent step counts
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Stage an optional host array; returns the device pointer (nullptr when src is nullptr).
int stage(DevCtx &c, Buf b, const void *src, size_t bytes, const void **dst) {
    *dst = nullptr;
    if (!src) return 0;
    if (ensure(c, b, bytes)) return -1;
    CU_CHECK(cudaMemcpyAsync(c.buf[b], src, bytes, cudaMemcpyHostToDevice, c.stream));
    *dst = c.buf[b];
    return 0;
}

// ---- chunk pipeline of the host-pointer calls ------------------------------------------------------------------
// The batch is cut into up to kMaxChunks index ranges that alternate between two streams: the H2D copy of chunk
// i+1 and the D2H copy of chunk i-1 overlap the kernel of chunk i, and because consecutive kernels sit on different
// streams the blocks of chunk i+1 move onto the SMs as the persistent blocks of chunk i drain (the tail of one
// chunk is filled by the head of the next).  Device buffers span the whole batch; a chunk is a window into them.
// How many chunks: every chunk boundary costs load balance (measured on C3: ~0.6 ms per boundary, a chunk's
// persistent blocks drain while the next chunk's move in), every additional chunk hides more of the copies
// (exposed copy time ~ T_copy / k).  Minimising T_copy / k + (k - 1) * 0.6 ms with T_copy = bytes / 50 GB/s gives
// k = sqrt(bytes / 30 MB): C3 (1e6 systems, 128 MB) -> 2 chunks (17.15 -> 16.64 ms per call; 3 chunks 16.9, 4 chunks
// 17.2, 8 chunks 21), C2 (1e6 trajectories x 100 samples, 2.5 GB, D2H 47 ms against a 33 ms kernel) -> 9 chunks.
constexpr int kMaxChunks = 16;
constexpr int64_t kMinChunk = 32768;
constexpr double kBytesPerBoundary = 30e6;

struct ChunkPlan {
    int nchunks;
    int64_t per;
    int64_t lo(int k) const { return (int64_t)k * per; }
    int64_t len(int k, int64_t n) const { return (k + 1) * per > n ? n - k * per : per; }
};

ChunkPlan plan_chunks(int64_t n, size_t bytes_per_system) {
    ChunkPlan p;
    p.nchunks = 1;
    p.per = n;
    // experiment knob: PYODESYS_CUDA_CHUNKS=k forces k equal chunks (k <= kMaxChunks)
    if (const char *env = getenv("PYODESYS_CUDA_CHUNKS")) {
        const int k = atoi(env);
        if (k >= 1 && k <= kMaxChunks && n >= k) {
            p.per = ((n + k - 1) / k + 1023) / 1024 * 1024;
            p.nchunks = (int)((n + p.per - 1) / p.per);
            return p;
        }
    }
    int k = (int)(sqrt((double)n * (double)bytes_per_system / kBytesPerBoundary) + 0.5);
    if (k > kMaxChunks) k = kMaxChunks;
    if (k <= 1) return p;
    int64_t per = (n + k - 1) / k;
    if (per < kMinChunk) per = kMinChunk;
    per = (per + 1023) / 1024 * 1024;
    if (per >= n) return p;
    p.per = per;
    p.nchunks = (int)((n + per - 1) / per);
    return p;
}

// device window (cap check + pointer) without a copy
int reserve(DevCtx &c, Buf b, const void *src, size_t bytes) {
    if (!src) return 0;
    return ensure(c, b, bytes);
}
int h2d(DevCtx &c, Buf b, const void *src, size_t elem_bytes, int64_t lo, int64_t len, cudaStream_t st) {
    if (!src || len <= 0) return 0;
    CU_CHECK(cudaMemcpyAsync((char *)c.buf[b] + (size_t)lo * elem_bytes, (const char *)src + (size_t)lo * elem_bytes,
                             (size_t)len * elem_bytes, cudaMemcpyHostToDevice, st));
    return 0;
}
int d2h(DevCtx &c, Buf b, void *dst, size_t elem_bytes, int64_t lo, int64_t len, cudaStream_t st) {
    if (!dst || len <= 0) return 0;
    CU_CHECK(cudaMemcpyAsync((char *)dst + (size_t)lo * elem_bytes, (const char *)c.buf[b] + (size_t)lo * elem_bytes,
                             (size_t)len * elem_bytes, cudaMemcpyDeviceToHost, st));
    return 0;
}
template <class T> const T *dev_or_null(DevCtx &c, Buf b, const void *src, int64_t lo, size_t elems_per_row = 1) {
    return src ? (const T *)c.buf[b] + (size_t)lo * elems_per_row : nullptr;
}

// timings of a pipelined call: h2d = start .. first kernel start, kernel = first kernel start .. last kernel end
// (copies of the neighbouring chunks run underneath), d2h = last kernel end .. everything on the host
int pipeline_timings(DevCtx &c, double *timings, double t_begin) {
    if (!timings) return 0;
    float ms = 0;
    CU_CHECK(cudaEventElapsedTime(&ms, c.ev[0], c.ev_k0)); timings[PYODESYS_CUDA_T_H2D] = ms * 1e-3;
    CU_CHECK(cudaEventElapsedTime(&ms, c.ev_k0, c.ev_k1)); timings[PYODESYS_CUDA_T_KERNEL] = ms * 1e-3;
    CU_CHECK(cudaEventElapsedTime(&ms, c.ev_k1, c.ev[3])); timings[PYODESYS_CUDA_T_D2H] = ms * 1e-3;
    timings[PYODESYS_CUDA_T_TOTAL] = now_s() - t_begin;
    return 0;
}

// ---- streaming host-pointer call (predefined mode, one-system-per-thread kernels, work stealing) ----------------
// ONE persistent launch for the whole batch.  The inputs of the first slice are copied in front of the kernel; the
// remaining slices follow on the second stream while the kernel runs, each one published through an 8-byte copy
// into ``ready`` that the work-stealing leader lane waits for before it hands out indices beyond it.  Every retired
// system counts into its output slice; the last one raises a host-mapped flag and the host thread (which is inside
// this synchronous call anyway) then queues the D2H of that slice.  Compared with cutting the batch into separate
// launches there is no drain/fill at chunk boundaries, and only the first H2D slice and the last D2H slice are
// exposed (C3: 17.6 -> see DESIGN.md section 3).
struct SlicePlan {
    int64_t len;
    int n;
    int64_t lo(int j) const { return (int64_t)j * len; }
    int64_t size(int j, int64_t total) const { return (j + 1) * len > total ? total - j * len : len; }
};

// bytes of output per slice of the streaming launch: what stays exposed at the end of a call is the device-to-host
// copy of the slices that complete last (PYODESYS_CUDA_SLICE_MB overrides)
size_t slice_bytes() {
    static const size_t v = [] {
        const char *e = getenv("PYODESYS_CUDA_SLICE_MB");
        const double mb = e ? atof(e) : 0.0;
        return (size_t)((mb >= 0.25 && mb <= 1024.0 ? mb : PYODESYS_CUDA_SLICE_MB_DEFAULT) * (1u << 20));
    }();
    return v;
}

SlicePlan plan_slices(int64_t n, size_t out_bytes_per_system) {
    int64_t len = (int64_t)(slice_bytes() / (out_bytes_per_system ? out_bytes_per_system : 1));
    const int64_t lo = std::max<int64_t>(16384, (n + kMaxSlices - 1) / kMaxSlices), hi = (n + 7) / 8;
    if (len > hi) len = hi;
    if (len < lo) len = lo;
    len = (len + 1023) / 1024 * 1024;
    SlicePlan p;
    p.len = len;
    p.n = (int)((n + len - 1) / len);
    return p;
}

// page-locked (cudaHostAlloc / cudaHostRegister) host memory?
bool is_pinned(const void *ptr) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost;
}

// wall-clock limit a streaming kernel waits for one input slice (generous: a profiler, a debugger, MPS time slicing
// or a pageable copy from swapped memory may hold the copies back for seconds)
unsigned long long wait_timeout_ns() {
    double s = 60.0;
    if (const char *env = getenv("PYODESYS_CUDA_WAIT_TIMEOUT_S")) {
        const double v = atof(env);
        if (v > 0.0) s = v;
    }
    return (unsigned long long)(s * 1e9);
}

bool streaming_enabled() {
    const char *env = getenv("PYODESYS_CUDA_STREAMING");
    return !(env && env[0] == '0');
}


// ---- compaction of the ragged adaptive output ------------------------------------------------------
// The kernels write every accepted step into bounded per-system buffers xyout[n, max_rows, 1 + ny]; what the caller
// gets - like the reference's cvode wrapper, which hands back exactly ``result[idx].first + 1`` rows per system
// (_cvode_wrapper.pyx:198-205) - is ONE contiguous array of the rows that were written plus offsets[n + 1]:
// exclusive scan of nrows (three small kernels), one warp per system copies its rows (coalesced), and only
// sum(nrows) * (1 + ny) * 8 bytes cross PCIe instead of n * max_rows * (1 + ny) * 8.
constexpr int kScanTile = 2048, kScanBlock = 256, kScanPer = kScanTile / kScanBlock;
__device__ __forceinline__ long long block_exclusive_scan(long long v, long long *total) {
    __shared__ long long warp_sums[kScanBlock / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    long long inc = v;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const long long u = __shfl_up_sync(FULL_MASK, inc, off);
        if (lane >= off) inc += u;
    }
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    long long base = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < kScanBlock / 32; ++k) {
        const long long ws = warp_sums[k];
        if (k < w) base += ws;
        tot += ws;
    }
    __syncthreads();
    *total = tot;
    return base + inc - v;
}
__global__ void scan_tile_sums(const int32_t *nrows, int64_t n, long long *tile_sums) {
    const int64_t lo = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPer;
    long long v = 0;
#pragma unroll
    for (int k = 0; k < kScanPer; ++k)
        if (lo + k < n) v += nrows[lo + k];
    long long tot;
    block_exclusive_scan(v, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}
__global__ void scan_tiles(long long *tile_sums, int ntiles) {  // one block: exclusive scan in place, carry over passes
    long long carry = 0;
    for (int base = 0; base < ntiles; base += kScanBlock) {
        const int i = base + threadIdx.x;
        const long long v = i < ntiles ? tile_sums[i] : 0;
        long long tot;
        const long long ex = block_exclusive_scan(v, &tot);
        if (i < ntiles) tile_sums[i] = carry + ex;
        carry += tot;
    }
}
__global__ void scan_offsets(const int32_t *nrows, int64_t n, const long long *tile_sums, long long *offsets) {
    const int64_t lo = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanPer;
    long long v[kScanPer], sum = 0;
#pragma unroll
    for (int k = 0; k < kScanPer; ++k) {
        v[k] = lo + k < n ? nrows[lo + k] : 0;
        sum += v[k];
    }
    long long tot;
    long long run = tile_sums[blockIdx.x] + block_exclusive_scan(sum, &tot);
#pragma unroll
    for (int k = 0; k < kScanPer; ++k) {
        if (lo + k < n) offsets[lo + k] = run;
        run += v[k];
        if (lo + k == n - 1) offsets[n] = run;
    }
}
__global__ void compact_rows(const double *xyout, const long long *offsets, int64_t n, int64_t max_rows, int rowlen,
                             double *out) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < n; s += nwarps) {
        const long long o = offsets[s], cnt = (offsets[s + 1] - o) * rowlen;
        const double *src = xyout + s * max_rows * rowlen;
        double *dst = out + o * rowlen;
        for (long long i = lane; i < cnt; i += 32) dst[i] = src[i];
    }
}


// ---- compaction without any allocation (pyodesys_cuda_compact_rows_noalloc) -----------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  The scratch of the scan - one partial sum per tile of
// kScanTile systems - lives in this file-scope array: kNoAllocMaxTiles * 8 bytes = 64 KiB that every GPU which
// loads the library holds until exit.  In HOST code ``g_noalloc_tile_sums`` is only a stand-in for the symbol; the
// device address is taken once per device with cudaGetSymbolAddress (noalloc_scratch below) and passed to the kernels.
constexpr int kNoAllocMaxTiles = PYODESYS_CUDA_NOALLOC_MAX_TILES;
constexpr int64_t kNoAllocMaxSystems = (int64_t)kNoAllocMaxTiles * kScanTile;  // 16 777 216
static_assert(kNoAllocMaxSystems == PYODESYS_CUDA_NOALLOC_MAX_SYSTEMS, "header and scan tile size disagree");
static __device__ __align__(256) long long g_noalloc_tile_sums[kNoAllocMaxTiles];

// compact_rows with a bound: rows that would land beyond ``capacity`` rows of ``out`` are not written (the caller
// sees offsets[n] > capacity and knows)
__global__ void compact_rows_bounded(const double *xyout, const long long *offsets, int64_t n, int64_t max_rows,
                                     int rowlen, double *out, long long capacity) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < n; s += nwarps) {
        const long long o = offsets[s];
        long long e = offsets[s + 1];
        if (e > capacity) e = capacity;
        const long long cnt = (e - o) * rowlen;
        const double *src = xyout + s * max_rows * rowlen;
        double *dst = out + o * rowlen;
        for (long long i = lane; i < cnt; i += 32) dst[i] = src[i];
    }
}

struct NoAllocDev {
    long long *scratch = nullptr;  // device address of g_noalloc_tile_sums ON THIS DEVICE
    int sm_count = 0;
};
NoAllocDev g_noalloc[kMaxDev];
std::mutex g_noalloc_mtx;

// device address of the scratch array on ``device`` (made current here); looked up once per device
int noalloc_scratch(int device, NoAllocDev **out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mtx);
    CU_CHECK(cudaSetDevice(device));
    NoAllocDev &d = g_noalloc[device];
    if (!d.scratch) {
        long long *ptr = nullptr;
        CU_CHECK(cudaGetSymbolAddress((void **)&ptr, g_noalloc_tile_sums));
        if (!ptr) return fail("cudaGetSymbolAddress returned a null address for the scratch array on device %d", device);
        CU_CHECK(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
        d.scratch = ptr;
    }
    *out = &d;
    return 0;
}

}  // namespace

extern "C" {

int pyodesys_cuda_info(int *ny, int *np_tot, int *npcse, int *flops_rhs, int *flops_jac) {
    if (ny) *ny = NY;
    if (np_tot) *np_tot = NP_TOT;
    if (npcse) *npcse = ODESYS_NPCSE;
    if (flops_rhs) *flops_rhs = ODESYS_FLOPS_RHS;
    if (flops_jac) *flops_jac = ODESYS_FLOPS_JAC;
    return 0;
}

int64_t pyodesys_cuda_query(const char *key) {
    if (!key) return -1;
    if (strcmp(key, "last_failed_systems") == 0) return g_last_failed;
    struct KV { const char *k; int64_t v; };
    const KV kv[] = {
        {"abi_version", PYODESYS_CUDA_ABI_VERSION}, {"ny", NY}, {"nparams", ODESYS_NPARAMS}, {"np_tot", NP_TOT},
        {"np_used", NP_USED}, {"npcse", ODESYS_NPCSE}, {"has_jac", ODESYS_HAS_JAC}, {"autonomous", ODESYS_AUTONOMOUS},
        {"nroots", ODESYS_NROOTS}, {"nnz", ODESYS_NNZ}, {"has_jtimes", ODESYS_HAS_JTIMES}, {"ninvar", ODESYS_NINVAR}, {"flops_rhs", ODESYS_FLOPS_RHS}, {"flops_jac", ODESYS_FLOPS_JAC},
        {"flops_pcse", ODESYS_FLOPS_PCSE}, {"jac_nnz", ODESYS_JAC_NNZ}, {"block_size", kBlock},
        {"ncounters", NCOUNT}, {"has_group_kernel", ODESYS_GROUP_KERNEL}, {"has_coop_kernel", ODESYS_COOP_KERNEL},
#if ODESYS_COOP_KERNEL
        {"coop_group_size", kCoopG}, {"coop_systems_per_block", CoopLayout<kCoopG>::NSYS},
        {"coop_smem_bytes", (int64_t)CoopLayout<kCoopG>::kDoubles * 8},
#endif
        {"has_smem_kernel", ODESYS_SMEM_KERNEL}, {"has_bdf", ODESYS_THREAD_BDF || ODESYS_COOP_KERNEL},
#if ODESYS_SMEM_KERNEL
        {"smem_doubles_per_thread", kSmemDoublesPerThread},
#endif
#if ODESYS_GROUP_KERNEL
        {"group_size", kGroup}, {"group_smem_doubles", kGroupSmemDoubles},
#endif
        {"has_first_step", ODESYS_HAS_FIRST_STEP}, {"has_dx_max", ODESYS_HAS_DX_MAX},
    };
    for (const auto &e : kv)
        if (strcmp(e.k, key) == 0) return e.v;
    // registers per thread of a kernel variant: "regs:<method>:<mode>" (needs a CUDA device)
    int m = -1, mode = -1;
    if (sscanf(key, "regs:%d:%d", &m, &mode) == 2) {
        cudaFuncAttributes fa;
        cudaError_t e = cudaErrorInvalidValue;
        if (m == PYODESYS_CUDA_RK4) e = mode ? cudaFuncGetAttributes(&fa, integrate_kernel<M_RK4, 1, kBlock>) : cudaFuncGetAttributes(&fa, integrate_kernel<M_RK4, 0, kBlock>);
        if (m == PYODESYS_CUDA_DOPRI5) e = mode ? cudaFuncGetAttributes(&fa, integrate_kernel<M_DOPRI5, 1, kBlock>) : cudaFuncGetAttributes(&fa, integrate_kernel<M_DOPRI5, 0, kBlock>);
#if ODESYS_THREAD_ROS4
        if (m == PYODESYS_CUDA_ROSENBROCK4) e = mode ? cudaFuncGetAttributes(&fa, integrate_kernel<M_ROS4, 1, kBlock>) : cudaFuncGetAttributes(&fa, integrate_kernel<M_ROS4, 0, kBlock>);
#endif
#if ODESYS_THREAD_BDF
        if (m == PYODESYS_CUDA_BDF) e = mode ? cudaFuncGetAttributes(&fa, integrate_kernel<M_BDF, 1, kBlock>) : cudaFuncGetAttributes(&fa, integrate_kernel<M_BDF, 0, kBlock>);
#endif
        if (e != cudaSuccess) { cudaGetLastError(); return -1; }
        return fa.numRegs;
    }
    return -1;
}

int64_t pyodesys_cuda_launch_count(void) { return g_launches.load(); }

void *pyodesys_cuda_host_alloc(size_t bytes) {
    void *ptr = nullptr;
    if (bytes == 0) bytes = 8;
    if (cudaHostAlloc(&ptr, bytes, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        fail("cudaHostAlloc(%zu) failed", bytes);
        return nullptr;
    }
    return ptr;
}

void pyodesys_cuda_host_free(void *ptr) {
    if (ptr) cudaFreeHost(ptr);
}

const char *pyodesys_cuda_last_error(void) { return g_err; }

static int integrate_predefined_impl(int device, int64_t n, int nt, const double *y0, const double *xout,
                                       const double *params, const double *atol, int natol, double rtol,
                                       const double *dx0, const double *dx_min, const double *dx_max, int64_t mxsteps,
                                       int method, int autorestart, int return_on_error, double *yout,
                                       int32_t *nreached, int32_t *status, int32_t *counters, double *timings,
                                       int ptr_kind, const pyodesys_cuda_options *opt) {
    (void)return_on_error;  // per-system failures are always reported through status/nreached
    const double t_begin = now_s();
    g_last_failed = -1;
    Common q{device, n, y0, params, atol, natol, rtol, dx0, dx_min, dx_max, mxsteps, method, autorestart,
             status, counters, timings, ptr_kind, opt};
    if (check_common(q)) return -1;
    if (nt < 1) return fail("need at least one output point (nt=%d)", nt);
    if ((!xout || !yout || !nreached) && n > 0) return fail("xout/yout/nreached must not be NULL");
    if (timings) for (int i = 0; i < PYODESYS_CUDA_NTIMINGS; ++i) timings[i] = 0.0;
    if (n == 0) return 0;
    DevCtx &c = g_ctx[device][lane_of(opt)];
    std::lock_guard<std::mutex> lock(c.mtx);
    if (ctx_init(c, device)) return -1;

    const bool shared_x = opt && opt->xout_shared;
    KernelArgs a;
    fill_args(a, q);
    a.nt = nt;
    a.xout_stride = shared_x ? 0 : nt;
    a.yout_soa = (opt && ptr_kind == 1) ? opt->yout_soa : 0;
    cudaStream_t stream = c.stream;
    if (want_dynamic(q)) {
        if (ensure(c, B_WORK, kMaxChunks * sizeof(unsigned long long))) return -1;
        a.work_counter = (unsigned long long *)c.buf[B_WORK];
    }
    const int bps = opt ? opt->blocks_per_sm : 0;

    if (ptr_kind == 1) {
        if (opt && opt->stream) stream = (cudaStream_t)opt->stream;
        a.y0 = y0; a.params = params; a.xout = xout; a.dx0 = dx0; a.dx_min = dx_min; a.dx_max = dx_max;
        a.yout = yout; a.nreached = nreached; a.status = status; a.counters = counters;
        if (want_roots(q)) { a.root_out = opt->root_out; a.root_indices = opt->root_indices; a.nroots_found = opt->nroots_found; }
#if defined(PYODESYS_COOP_PROFILE) || defined(PYODESYS_COOP_TRACE)
        if (opt) a.root_out = opt->root_out;  // profiling / tracing build: phase counters or the step trace land here
#endif
        if (a.work_counter) CU_CHECK(cudaMemsetAsync(a.work_counter, 0, sizeof(unsigned long long), stream));
        CU_CHECK(cudaEventRecord(c.ev[0], stream));
        if (dispatch<0>(c, a, method, stream, bps, opt)) return -1;
        CU_CHECK(cudaEventRecord(c.ev[1], stream));
        CU_CHECK(cudaStreamSynchronize(stream));
        float ms = 0;
        CU_CHECK(cudaEventElapsedTime(&ms, c.ev[0], c.ev[1]));
        if (timings) { timings[PYODESYS_CUDA_T_KERNEL] = ms * 1e-3; timings[PYODESYS_CUDA_T_TOTAL] = now_s() - t_begin; }
        return 0;
    }

    // ---- host pointers: chunked H2D / kernel / D2H pipeline on two streams ---------------------------------
    const size_t nn = (size_t)n;
    const size_t cnt_pitch = host_counters_pitch(q);
    const size_t prow = (size_t)a.params_ld;
    if (reserve(c, B_Y0, y0, nn * NY * 8) || reserve(c, B_PARAMS, prow ? params : nullptr, nn * prow * 8) ||
        reserve(c, B_XOUT, xout, (shared_x ? 1 : nn) * nt * 8) || reserve(c, B_DX0, dx0, nn * 8) ||
        reserve(c, B_DXMIN, dx_min, nn * 8) || reserve(c, B_DXMAX, dx_max, nn * 8) ||
        ensure(c, B_YOUT, nn * nt * NY * 8) || ensure(c, B_NREACHED, nn * 4) || ensure(c, B_STATUS, nn * 4) ||
        (counters && ensure(c, B_COUNTERS, nn * NCOUNT * 4)) ||
        ensure(c, B_WORK, kMaxChunks * sizeof(unsigned long long)))
        return -1;
    const bool roots = want_roots(q);
    const size_t root_row = roots ? (size_t)a.max_roots * (1 + NY) * 8 : 0;
    if (roots && (ensure(c, B_ROOTOUT, nn * root_row) || ensure(c, B_ROOTIDX, nn * a.max_roots * 4) ||
                  ensure(c, B_NROOTS, nn * 4)))
        return -1;
    const bool dynamic = a.work_counter != nullptr;
    if (dynamic && !roots && n >= 131072 && pick_kernel(method, opt) == 1 && streaming_enabled()) {
        // ---- streaming: one persistent launch, sliced copies underneath ------------------------------------
        const SlicePlan sp = plan_slices(n, (size_t)nt * NY * 8 + 8 + (counters ? NCOUNT * 4 : 0));
        if (ensure(c, B_READY, 8) || ensure(c, B_SLICEDONE, kMaxSlices * 4)) return -1;
        for (int j = 0; j < sp.n; ++j) {
            c.h_ready[j] = (unsigned long long)std::min<int64_t>(n, sp.lo(j) + sp.len);
            c.h_flags[j] = 0;
        }
        volatile unsigned int *h_abort = c.h_flags + kMaxSlices;
        *h_abort = 0;
        auto copy_in = [&](int j, cudaStream_t st) -> int {
            const int64_t lo = sp.lo(j), len = sp.size(j, n);
            if (h2d(c, B_Y0, y0, NY * 8, lo, len, st) || h2d(c, B_PARAMS, prow ? params : nullptr, prow * 8, lo, len, st) ||
                (!shared_x && h2d(c, B_XOUT, xout, (size_t)nt * 8, lo, len, st)) || h2d(c, B_DX0, dx0, 8, lo, len, st) ||
                h2d(c, B_DXMIN, dx_min, 8, lo, len, st) || h2d(c, B_DXMAX, dx_max, 8, lo, len, st))
                return -1;
            CU_CHECK(cudaMemcpyAsync(c.buf[B_READY], c.h_ready + j, 8, cudaMemcpyHostToDevice, st));
            return 0;
        };
        CU_CHECK(cudaEventRecord(c.ev[0], c.stream));
        CU_CHECK(cudaMemsetAsync(c.buf[B_WORK], 0, sizeof(unsigned long long), c.stream));
        CU_CHECK(cudaMemsetAsync(c.buf[B_SLICEDONE], 0, kMaxSlices * 4, c.stream));
        if (shared_x) CU_CHECK(cudaMemcpyAsync(c.buf[B_XOUT], xout, (size_t)nt * 8, cudaMemcpyHostToDevice, c.stream));
        if (copy_in(0, c.stream)) return -1;
        CU_CHECK(cudaEventRecord(c.ev_fork, c.stream));
        CU_CHECK(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0));
        KernelArgs ak = a;
        ak.y0 = (const double *)c.buf[B_Y0];
        ak.params = prow ? (const double *)c.buf[B_PARAMS] : nullptr;
        ak.xout = (const double *)c.buf[B_XOUT];
        ak.dx0 = dev_or_null<double>(c, B_DX0, dx0, 0);
        ak.dx_min = dev_or_null<double>(c, B_DXMIN, dx_min, 0);
        ak.dx_max = dev_or_null<double>(c, B_DXMAX, dx_max, 0);
        ak.yout = (double *)c.buf[B_YOUT];
        ak.nreached = (int32_t *)c.buf[B_NREACHED];
        ak.status = (int32_t *)c.buf[B_STATUS];
        ak.counters = counters ? (int32_t *)c.buf[B_COUNTERS] : nullptr;
        ak.work_counter = (unsigned long long *)c.buf[B_WORK];
        ak.ready = (const unsigned long long *)c.buf[B_READY];
        ak.slice_done = (unsigned int *)c.buf[B_SLICEDONE];
        ak.slice_flags = c.d_flags;
        ak.slice_len = sp.len;
        ak.abort_flag = c.d_flags + kMaxSlices;
        ak.wait_timeout_ns = wait_timeout_ns();
        // The remaining input slices are queued by a helper thread that starts BEFORE the launch.  Copies from
        // pageable memory block the calling thread while the driver stages them, so queued from this thread they
        // would either delay the launch (all input copied first) or, queued after it, never be issued under a tool
        // that makes launches synchronous (ncu) - the kernel would wait for its inputs forever.
        struct Feeder {
            std::thread th;
            int rc = 0;
            char err[sizeof(g_err)] = "";
            ~Feeder() { if (th.joinable()) th.join(); }
        } feeder;
        // Page-locked inputs: the copies are asynchronous for real, so they are simply queued here, in front of the
        // launch (also what a profiler needs: ncu serialises CUDA calls behind a running kernel, and a kernel that
        // waits for copies issued during its own execution would never finish).  Pageable inputs: helper thread;
        // profile such calls with PYODESYS_CUDA_STREAMING=0.
        const bool pinned_inputs = is_pinned(y0) && (!prow || is_pinned(params)) && (shared_x || is_pinned(xout)) &&
                                   (!dx0 || is_pinned(dx0)) && (!dx_min || is_pinned(dx_min)) &&
                                   (!dx_max || is_pinned(dx_max));
        if (sp.n > 1 && pinned_inputs) {
            for (int j = 1; j < sp.n; ++j)
                if (copy_in(j, c.stream2)) return -1;
        } else if (sp.n > 1)
            feeder.th = std::thread([&]() {
                // a failure here raises the abort word at once: the kernel must not wait for inputs that never come
                if (cudaSetDevice(device) != cudaSuccess) { feeder.rc = -1; *h_abort = 1; snprintf(feeder.err, sizeof(feeder.err), "cudaSetDevice failed in the copy thread"); return; }
                for (int j = 1; j < sp.n; ++j)
                    if (copy_in(j, c.stream2)) {
                        feeder.rc = -1;
                        *h_abort = 1;
                        memcpy(feeder.err, g_err, sizeof(feeder.err));  // thread-local text of THIS thread
                        return;
                    }
            });
        CU_CHECK(cudaEventRecord(c.ev_k0, c.stream));
        if (dispatch<0>(c, ak, method, c.stream, bps, opt)) return -1;
        CU_CHECK(cudaEventRecord(c.ev_k1, c.stream));
        if (feeder.th.joinable()) feeder.th.join();  // D2H copies below go to the same stream, behind the inputs
        if (feeder.rc != 0) return fail("%s", feeder.err);
        // While it waits for the device, this thread also screens the status words that have already arrived for
        // failed systems (the caller's first question after the call; a pass over n words costs 0.25 ms per 1e6
        // systems when it is done afterwards).
        int counted = 0, queued = 0;
        int64_t n_failed = 0;
        auto count_arrived = [&](bool wait) {
            while (counted < queued) {
                const cudaError_t q = wait ? cudaEventSynchronize(c.ev_slice[counted]) : cudaEventQuery(c.ev_slice[counted]);
                if (q != cudaSuccess) {
                    if (q == cudaErrorNotReady) cudaGetLastError();
                    return;
                }
                const int64_t lo = sp.lo(counted), hi = lo + sp.size(counted, n);
                int64_t bad = 0;
                for (int64_t i = lo; i < hi; ++i) bad += status[i] != 0;
                n_failed += bad;
                counted += 1;
            }
        };
        for (int j = 0; j < sp.n && !*h_abort; ++j) {
            // wait for the kernel to complete slice j (or to end: then every slice is complete, or it failed).
            // A short spin catches flags that are about to flip; after that the thread sleeps between polls
            // instead of burning a core per device for the length of the kernel.
            volatile unsigned int *flag = c.h_flags + j;
            for (unsigned spin = 0; !*flag; ++spin) {
                if (spin < 512u) continue;
                const cudaError_t q = cudaStreamQuery(c.stream);
                if (q == cudaSuccess) break;
                if (q != cudaErrorNotReady) return fail("streaming kernel failed: %s", cudaGetErrorString(q));
                count_arrived(false);
                std::this_thread::sleep_for(std::chrono::microseconds(spin < 4096u ? 20 : 100));
            }
            std::atomic_thread_fence(std::memory_order_acquire);
            const int64_t lo = sp.lo(j), len = sp.size(j, n);
            if (d2h(c, B_YOUT, yout, (size_t)nt * NY * 8, lo, len, c.stream2) ||
                d2h(c, B_NREACHED, nreached, 4, lo, len, c.stream2) || d2h(c, B_STATUS, status, 4, lo, len, c.stream2))
                return -1;
            if (counters)
                CU_CHECK(cudaMemcpy2DAsync(counters + lo, cnt_pitch, (const int32_t *)c.buf[B_COUNTERS] + lo, nn * 4,
                                           (size_t)len * 4, NCOUNT, cudaMemcpyDeviceToHost, c.stream2));
            if (!c.ev_slice[j]) CU_CHECK(cudaEventCreateWithFlags(&c.ev_slice[j], cudaEventDisableTiming));
            CU_CHECK(cudaEventRecord(c.ev_slice[j], c.stream2));
            queued = j + 1;
        }
        CU_CHECK(cudaEventRecord(c.ev_join, c.stream2));
        CU_CHECK(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
        CU_CHECK(cudaEventRecord(c.ev[3], c.stream));
        if (!*h_abort && queued == sp.n) count_arrived(true);  // the last slices, as their copies complete
        CU_CHECK(cudaStreamSynchronize(c.stream));
        if (!*h_abort && counted == sp.n) g_last_failed = n_failed;
        if (*h_abort)
            return fail("streaming call aborted: %s (the CUDA context is intact; PYODESYS_CUDA_STREAMING=0 selects the "
                        "chunked path, PYODESYS_CUDA_WAIT_TIMEOUT_S the limit)",
                        *h_abort == 2 ? "the kernel waited longer than the time-out for its next input slice"
                                      : "the input copy thread failed");
        return pipeline_timings(c, timings, t_begin);
    }
    const ChunkPlan plan = plan_chunks(n, ((size_t)NY + prow + (shared_x ? 0 : nt) + (size_t)nt * NY) * 8 + 32);
    cudaStream_t lanes[2] = {c.stream, c.stream2};
    CU_CHECK(cudaEventRecord(c.ev[0], c.stream));
    CU_CHECK(cudaMemsetAsync(c.buf[B_WORK], 0, kMaxChunks * sizeof(unsigned long long), c.stream));
    if (shared_x) CU_CHECK(cudaMemcpyAsync(c.buf[B_XOUT], xout, (size_t)nt * 8, cudaMemcpyHostToDevice, c.stream));
    CU_CHECK(cudaEventRecord(c.ev_fork, c.stream));
    CU_CHECK(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0));
    for (int k = 0; k < plan.nchunks; ++k) {
        cudaStream_t st = lanes[k & 1];
        const int64_t lo = plan.lo(k), len = plan.len(k, n);
        if (h2d(c, B_Y0, y0, NY * 8, lo, len, st) || h2d(c, B_PARAMS, prow ? params : nullptr, prow * 8, lo, len, st) ||
            (!shared_x && h2d(c, B_XOUT, xout, (size_t)nt * 8, lo, len, st)) || h2d(c, B_DX0, dx0, 8, lo, len, st) ||
            h2d(c, B_DXMIN, dx_min, 8, lo, len, st) || h2d(c, B_DXMAX, dx_max, 8, lo, len, st))
            return -1;
        KernelArgs ak = a;
        ak.n = len;
        ak.y0 = (const double *)c.buf[B_Y0] + (size_t)lo * NY;
        ak.params = prow ? (const double *)c.buf[B_PARAMS] + (size_t)lo * prow : nullptr;
        ak.xout = (const double *)c.buf[B_XOUT] + (shared_x ? 0 : (size_t)lo * nt);
        ak.dx0 = dev_or_null<double>(c, B_DX0, dx0, lo);
        ak.dx_min = dev_or_null<double>(c, B_DXMIN, dx_min, lo);
        ak.dx_max = dev_or_null<double>(c, B_DXMAX, dx_max, lo);
        ak.yout = (double *)c.buf[B_YOUT] + (size_t)lo * nt * NY;
        ak.nreached = (int32_t *)c.buf[B_NREACHED] + lo;
        ak.status = (int32_t *)c.buf[B_STATUS] + lo;
        ak.counters = counters ? (int32_t *)c.buf[B_COUNTERS] + lo : nullptr;
        ak.counters_ld = n;
        ak.work_counter = dynamic ? (unsigned long long *)c.buf[B_WORK] + k : nullptr;
        if (roots) {
            ak.root_out = (double *)c.buf[B_ROOTOUT] + (size_t)lo * a.max_roots * (1 + NY);
            ak.root_indices = (int32_t *)c.buf[B_ROOTIDX] + (size_t)lo * a.max_roots;
            ak.nroots_found = (int32_t *)c.buf[B_NROOTS] + lo;
        }
        if (k == 0) CU_CHECK(cudaEventRecord(c.ev_k0, st));
        if (dispatch<0>(c, ak, method, st, bps, opt)) return -1;
        if (k == plan.nchunks - 1) CU_CHECK(cudaEventRecord(c.ev_k1, st));
        if (d2h(c, B_YOUT, yout, (size_t)nt * NY * 8, lo, len, st) || d2h(c, B_NREACHED, nreached, 4, lo, len, st) ||
            d2h(c, B_STATUS, status, 4, lo, len, st))
            return -1;
        if (counters)
            CU_CHECK(cudaMemcpy2DAsync(counters + lo, cnt_pitch, (const int32_t *)c.buf[B_COUNTERS] + lo, nn * 4,
                                       (size_t)len * 4, NCOUNT, cudaMemcpyDeviceToHost, st));
        if (roots && (d2h(c, B_ROOTOUT, opt->root_out, root_row, lo, len, st) ||
                      d2h(c, B_ROOTIDX, opt->root_indices, (size_t)a.max_roots * 4, lo, len, st) ||
                      d2h(c, B_NROOTS, opt->nroots_found, 4, lo, len, st)))
            return -1;
    }
    CU_CHECK(cudaEventRecord(c.ev_join, c.stream2));
    CU_CHECK(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
    CU_CHECK(cudaEventRecord(c.ev[3], c.stream));
    CU_CHECK(cudaStreamSynchronize(c.stream));
    return pipeline_timings(c, timings, t_begin);
}

static int integrate_adaptive_impl(int device, int64_t n, const double *y0, const double *x0, const double *xend,
                                     const double *params, const double *atol, int natol, double rtol,
                                     const double *dx0, const double *dx_min, const double *dx_max, int64_t mxsteps,
                                     int method, int autorestart, int return_on_error, int64_t max_rows,
                                     double *xyout, int32_t *nrows, int32_t *status, int32_t *counters,
                                     double *timings, int ptr_kind, const pyodesys_cuda_options *opt,
                                     int64_t *compact_offsets = nullptr) {
    (void)return_on_error;
    const double t_begin = now_s();
    Common q{device, n, y0, params, atol, natol, rtol, dx0, dx_min, dx_max, mxsteps, method, autorestart,
             status, counters, timings, ptr_kind, opt};
    if (check_common(q)) return -1;
    if (max_rows < 2) return fail("max_rows_per_system must be at least 2");
    const bool compact = compact_offsets != nullptr;  // rows stay on the device, compacted; see fetch_compact_rows
    if (compact && ptr_kind != 0) return fail("the compacting call takes host pointers");
    if ((!x0 || !xend || (!compact && (!xyout || !nrows))) && n > 0) return fail("x0/xend/xyout/nrows must not be NULL");
    if (timings) for (int i = 0; i < PYODESYS_CUDA_NTIMINGS; ++i) timings[i] = 0.0;
    if (n == 0) return 0;
    DevCtx &c = g_ctx[device][lane_of(opt)];
    std::lock_guard<std::mutex> lock(c.mtx);
    if (ctx_init(c, device)) return -1;

    KernelArgs a;
    fill_args(a, q);
    a.max_rows = max_rows;
    cudaStream_t stream = c.stream;
    if (want_dynamic(q)) {
        if (ensure(c, B_WORK, kMaxChunks * sizeof(unsigned long long))) return -1;
        a.work_counter = (unsigned long long *)c.buf[B_WORK];
    }
    const int bps = opt ? opt->blocks_per_sm : 0;
    const size_t nn = (size_t)n;
    const size_t out_bytes = nn * (size_t)max_rows * (1 + NY) * 8;

    if (ptr_kind == 1) {
        if (opt && opt->stream) stream = (cudaStream_t)opt->stream;
        a.y0 = y0; a.params = params; a.x0 = x0; a.xend = xend; a.dx0 = dx0; a.dx_min = dx_min; a.dx_max = dx_max;
        a.xyout = xyout; a.nrows = nrows; a.status = status; a.counters = counters;
        if (want_roots(q)) { a.root_out = opt->root_out; a.root_indices = opt->root_indices; a.nroots_found = opt->nroots_found; }
        if (a.work_counter) CU_CHECK(cudaMemsetAsync(a.work_counter, 0, sizeof(unsigned long long), stream));
        CU_CHECK(cudaEventRecord(c.ev[0], stream));
        if (dispatch<1>(c, a, method, stream, bps, opt)) return -1;
        CU_CHECK(cudaEventRecord(c.ev[1], stream));
        CU_CHECK(cudaStreamSynchronize(stream));
        float ms = 0;
        CU_CHECK(cudaEventElapsedTime(&ms, c.ev[0], c.ev[1]));
        if (timings) { timings[PYODESYS_CUDA_T_KERNEL] = ms * 1e-3; timings[PYODESYS_CUDA_T_TOTAL] = now_s() - t_begin; }
        return 0;
    }

    const size_t prow = (size_t)a.params_ld;
    const size_t cnt_pitch = host_counters_pitch(q);
    const size_t row_bytes = (size_t)max_rows * (1 + NY) * 8;
    if (reserve(c, B_Y0, y0, nn * NY * 8) || reserve(c, B_PARAMS, prow ? params : nullptr, nn * prow * 8) ||
        ensure(c, B_XOUT, nn * 8) || ensure(c, B_XEND, nn * 8) || reserve(c, B_DX0, dx0, nn * 8) ||
        reserve(c, B_DXMIN, dx_min, nn * 8) || reserve(c, B_DXMAX, dx_max, nn * 8) || ensure(c, B_YOUT, out_bytes) ||
        ensure(c, B_NREACHED, nn * 4) || ensure(c, B_STATUS, nn * 4) ||
        (counters && ensure(c, B_COUNTERS, nn * NCOUNT * 4)) ||
        ensure(c, B_WORK, kMaxChunks * sizeof(unsigned long long)))
        return -1;
    const bool roots = want_roots(q);
    const size_t root_row = roots ? (size_t)a.max_roots * (1 + NY) * 8 : 0;
    if (roots && (ensure(c, B_ROOTOUT, nn * root_row) || ensure(c, B_ROOTIDX, nn * a.max_roots * 4) ||
                  ensure(c, B_NROOTS, nn * 4)))
        return -1;
    const ChunkPlan plan = plan_chunks(n, ((size_t)NY + prow + 2) * 8 + row_bytes + 32);
    const bool dynamic = a.work_counter != nullptr;
    cudaStream_t lanes[2] = {c.stream, c.stream2};
    CU_CHECK(cudaEventRecord(c.ev[0], c.stream));
    CU_CHECK(cudaMemsetAsync(c.buf[B_WORK], 0, kMaxChunks * sizeof(unsigned long long), c.stream));
    CU_CHECK(cudaEventRecord(c.ev_fork, c.stream));
    CU_CHECK(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0));
    for (int k = 0; k < plan.nchunks; ++k) {
        cudaStream_t st = lanes[k & 1];
        const int64_t lo = plan.lo(k), len = plan.len(k, n);
        if (h2d(c, B_Y0, y0, NY * 8, lo, len, st) || h2d(c, B_PARAMS, prow ? params : nullptr, prow * 8, lo, len, st) ||
            h2d(c, B_XOUT, x0, 8, lo, len, st) || h2d(c, B_XEND, xend, 8, lo, len, st) ||
            h2d(c, B_DX0, dx0, 8, lo, len, st) || h2d(c, B_DXMIN, dx_min, 8, lo, len, st) ||
            h2d(c, B_DXMAX, dx_max, 8, lo, len, st))
            return -1;
        KernelArgs ak = a;
        ak.n = len;
        ak.y0 = (const double *)c.buf[B_Y0] + (size_t)lo * NY;
        ak.params = prow ? (const double *)c.buf[B_PARAMS] + (size_t)lo * prow : nullptr;
        ak.x0 = (const double *)c.buf[B_XOUT] + lo;
        ak.xend = (const double *)c.buf[B_XEND] + lo;
        ak.dx0 = dev_or_null<double>(c, B_DX0, dx0, lo);
        ak.dx_min = dev_or_null<double>(c, B_DXMIN, dx_min, lo);
        ak.dx_max = dev_or_null<double>(c, B_DXMAX, dx_max, lo);
        ak.xyout = (double *)c.buf[B_YOUT] + (size_t)lo * max_rows * (1 + NY);
        ak.nrows = (int32_t *)c.buf[B_NREACHED] + lo;
        ak.status = (int32_t *)c.buf[B_STATUS] + lo;
        ak.counters = counters ? (int32_t *)c.buf[B_COUNTERS] + lo : nullptr;
        ak.counters_ld = n;
        ak.work_counter = dynamic ? (unsigned long long *)c.buf[B_WORK] + k : nullptr;
        if (roots) {
            ak.root_out = (double *)c.buf[B_ROOTOUT] + (size_t)lo * a.max_roots * (1 + NY);
            ak.root_indices = (int32_t *)c.buf[B_ROOTIDX] + (size_t)lo * a.max_roots;
            ak.nroots_found = (int32_t *)c.buf[B_NROOTS] + lo;
        }
        if (k == 0) CU_CHECK(cudaEventRecord(c.ev_k0, st));
        if (dispatch<1>(c, ak, method, st, bps, opt)) return -1;
        if (k == plan.nchunks - 1) CU_CHECK(cudaEventRecord(c.ev_k1, st));
        if ((!compact && (d2h(c, B_YOUT, xyout, row_bytes, lo, len, st) || d2h(c, B_NREACHED, nrows, 4, lo, len, st))) ||
            d2h(c, B_STATUS, status, 4, lo, len, st))
            return -1;
        if (counters)
            CU_CHECK(cudaMemcpy2DAsync(counters + lo, cnt_pitch, (const int32_t *)c.buf[B_COUNTERS] + lo, nn * 4,
                                       (size_t)len * 4, NCOUNT, cudaMemcpyDeviceToHost, st));
        if (roots && (d2h(c, B_ROOTOUT, opt->root_out, root_row, lo, len, st) ||
                      d2h(c, B_ROOTIDX, opt->root_indices, (size_t)a.max_roots * 4, lo, len, st) ||
                      d2h(c, B_NROOTS, opt->nroots_found, 4, lo, len, st)))
            return -1;
    }
    CU_CHECK(cudaEventRecord(c.ev_join, c.stream2));
    CU_CHECK(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
    if (compact) {
        const int ntiles = (int)((n + kScanTile - 1) / kScanTile);
        if (ensure(c, B_OFFSETS, (nn + 1) * 8) || ensure(c, B_BLOCKSUMS, (size_t)ntiles * 8)) return -1;
        const int32_t *d_nrows = (const int32_t *)c.buf[B_NREACHED];
        long long *d_off = (long long *)c.buf[B_OFFSETS], *d_tiles = (long long *)c.buf[B_BLOCKSUMS];
        scan_tile_sums<<<ntiles, kScanBlock, 0, c.stream>>>(d_nrows, n, d_tiles);
        scan_tiles<<<1, kScanBlock, 0, c.stream>>>(d_tiles, ntiles);
        scan_offsets<<<ntiles, kScanBlock, 0, c.stream>>>(d_nrows, n, d_tiles, d_off);
        g_launches.fetch_add(3);
        CU_CHECK(cudaGetLastError());
        CU_CHECK(cudaMemcpyAsync(compact_offsets, d_off, (nn + 1) * 8, cudaMemcpyDeviceToHost, c.stream));
        CU_CHECK(cudaStreamSynchronize(c.stream));  // the total decides the size of the compact buffer
        const int64_t total = compact_offsets[n];
        c.compact_rows = total;
        if (total > 0) {
            if (ensure(c, B_COMPACT, (size_t)total * (1 + NY) * 8)) return -1;
            const int blocks = (int)std::min<int64_t>((n + 7) / 8, (int64_t)c.sm_count * 8);
            compact_rows<<<blocks, 256, 0, c.stream>>>((const double *)c.buf[B_YOUT], d_off, n, max_rows, 1 + NY,
                                                       (double *)c.buf[B_COMPACT]);
            g_launches.fetch_add(1);
            CU_CHECK(cudaGetLastError());
        }
    }
    CU_CHECK(cudaEventRecord(c.ev[3], c.stream));
    CU_CHECK(cudaStreamSynchronize(c.stream));
    return pipeline_timings(c, timings, t_begin);
}

// A call that fails half way (bad argument found late, CUDA error) must not leave copies or kernels of its own
// in flight on the library streams: the caller is free to release its buffers as soon as the call returns.
static void drain_after_failure(int device) {
    if (device < 0 || device >= kMaxDev) return;
    char keep[sizeof(g_err)];
    memcpy(keep, g_err, sizeof(keep));  // keep the first error text
    for (int lane = 0; lane < kLanes; ++lane) {
        DevCtx &c = g_ctx[device][lane];
        if (!c.mtx.try_lock()) continue;  // the other lane is inside a call of its own: not ours to drain
        std::lock_guard<std::mutex> lock(c.mtx, std::adopt_lock);
        if (!c.ready) continue;
        cudaSetDevice(device);
        cudaStreamSynchronize(c.stream);
        cudaStreamSynchronize(c.stream2);
        cudaGetLastError();
    }
    memcpy(g_err, keep, sizeof(keep));
}

int pyodesys_cuda_integrate_predefined(int device, int64_t n, int nt, const double *y0, const double *xout,
                                       const double *params, const double *atol, int natol, double rtol,
                                       const double *dx0, const double *dx_min, const double *dx_max, int64_t mxsteps,
                                       int method, int autorestart, int return_on_error, double *yout,
                                       int32_t *nreached, int32_t *status, int32_t *counters, double *timings,
                                       int ptr_kind, const pyodesys_cuda_options *opt) {
    const int rc = integrate_predefined_impl(device, n, nt, y0, xout, params, atol, natol, rtol, dx0, dx_min, dx_max,
                                             mxsteps, method, autorestart, return_on_error, yout, nreached, status,
                                             counters, timings, ptr_kind, opt);
    if (rc != 0) drain_after_failure(device);
    return rc;
}

// ---- several devices from one call -------------------------------------------------------------
// The batch is cut into chunks of ``chunk`` systems; ``lanes`` host threads per device (each with its own call
// context: streams + staging buffers) draw chunk numbers from one atomic counter - whichever lane is free takes the
// next chunk, so devices and lanes that run ahead (stiffness varies with the parameters) simply take more - and run
// the ordinary host-pointer path on their window of the caller's arrays.  Every thread is bound to the CPUs of its
// GPU's NUMA node.  No collective, no gather: the device-to-host copies land in the caller's arrays.
static bool local_cpus_of(int device, cpu_set_t *set) {
    char bus[32] = "";
    if (cudaDeviceGetPCIBusId(bus, (int)sizeof(bus), device) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    for (char *q = bus; *q; ++q) *q = (char)tolower(*q);
    char path[128];
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/local_cpulist", bus);
    FILE *fh = fopen(path, "r");
    if (!fh) return false;
    char line[4096] = "";
    const bool got = fgets(line, sizeof(line), fh) != nullptr;
    fclose(fh);
    if (!got) return false;
    CPU_ZERO(set);
    int count = 0;
    char *save = nullptr;  // (strtok_r: several device threads parse their lists at the same time)
    for (char *tok = strtok_r(line, ",\n", &save); tok; tok = strtok_r(nullptr, ",\n", &save)) {
        int lo = 0, hi = 0;
        const int k = sscanf(tok, "%d-%d", &lo, &hi);
        if (k < 1) continue;
        if (k == 1) hi = lo;
        for (int cpu = lo; cpu <= hi && cpu < CPU_SETSIZE; ++cpu) {
            CPU_SET(cpu, set);
            ++count;
        }
    }
    return count > 0;
}

int pyodesys_cuda_integrate_predefined_multi(const int *devices, int ndev, int lanes, int64_t chunk, int64_t n, int nt,
                                             const double *y0, const double *xout, const double *params,
                                             const double *atol, int natol, double rtol, const double *dx0,
                                             const double *dx_min, const double *dx_max, int64_t mxsteps, int method,
                                             int autorestart, int return_on_error, double *yout, int32_t *nreached,
                                             int32_t *status, int32_t *counters, double *busy_s, int32_t *chunks_taken,
                                             int64_t *nfailed, const pyodesys_cuda_options *opt) {
    if (nfailed) *nfailed = 0;
    if (!devices || ndev < 1 || ndev > kMaxDev) return fail("invalid device list (ndev=%d)", ndev);
    if (lanes < 1 || lanes > kLanes) return fail("lanes must be 1..%d, got %d", kLanes, lanes);
    if (chunk < 1 || n < 0 || nt < 1) return fail("invalid sizes (n=%lld, chunk=%lld, nt=%d)", (long long)n, (long long)chunk, nt);
    if (opt && (opt->max_roots > 0 || opt->root_out)) return fail("event functions are not routed through the multi-device entry point");
    for (int d = 0; d < ndev; ++d)
        if (devices[d] < 0 || devices[d] >= kMaxDev) return fail("invalid device %d", devices[d]);
    const int nworkers = ndev * lanes;
    for (int d = 0; d < ndev; ++d) {
        if (busy_s) busy_s[d] = 0.0;
        if (chunks_taken) chunks_taken[d] = 0;
    }
    if (n == 0) return 0;
    pyodesys_cuda_options base;
    memset(&base, 0, sizeof(base));
    if (opt) base = *opt;
    const bool shared_x = base.xout_shared != 0;
    const int64_t prow = base.params_ld > 0 ? base.params_ld : (base.iv_from_y0 ? ODESYS_NPARAMS : NP_TOT);
    const int64_t nchunks = (n + chunk - 1) / chunk;
    std::atomic<int64_t> next{0};
    std::atomic<int> failed{0};
    std::vector<std::string> errs((size_t)nworkers);
    std::vector<double> busy((size_t)nworkers, 0.0);
    std::vector<int> taken((size_t)nworkers, 0);
    std::atomic<int64_t> bad_systems{0};
    const double t_entry = now_s();
    const bool trace = getenv("PYODESYS_CUDA_MULTI_TRACE") != nullptr;
    auto work = [&](int w) {
        const int dev = devices[w / lanes], lane = w % lanes;
        const double t_thread = now_s();
        cpu_set_t cpus;
        if (local_cpus_of(dev, &cpus)) sched_setaffinity(0, sizeof(cpus), &cpus);  // this thread only
        const double t_bound = now_s();
        double t_first = 0.0;
        pyodesys_cuda_options o = base;
        o.ctx_lane = lane;
        o.counters_ld = n;  // the caller's [NCOUNTERS, n] array: this call fills a column window
        while (!failed.load(std::memory_order_relaxed)) {
            const int64_t k = next.fetch_add(1);
            if (k >= nchunks) break;
            const int64_t lo = k * chunk, len = std::min<int64_t>(chunk, n - lo);
            double tm[PYODESYS_CUDA_NTIMINGS] = {0, 0, 0, 0};
            const int rc = integrate_predefined_impl(
                dev, len, nt, y0 + lo * NY, shared_x ? xout : xout + lo * nt, params ? params + lo * prow : nullptr, atol,
                natol, rtol, dx0 ? dx0 + lo : nullptr, dx_min ? dx_min + lo : nullptr, dx_max ? dx_max + lo : nullptr,
                mxsteps, method, autorestart, return_on_error, yout + lo * (int64_t)nt * NY, nreached + lo, status + lo,
                counters ? counters + lo : nullptr, tm, 0, &o);
            if (rc != 0) {
                errs[(size_t)w] = g_err;  // thread-local text of THIS thread
                failed.store(1);
                break;
            }
            busy[(size_t)w] += tm[PYODESYS_CUDA_T_TOTAL];
            taken[(size_t)w] += 1;
            if (taken[(size_t)w] == 1) t_first = now_s();
            if (nfailed) {  // the caller's screen for failed systems, done here by all workers in parallel
                int64_t bad = 0;
                for (int64_t i = lo; i < lo + len; ++i) bad += status[i] != 0;
                if (bad) bad_systems.fetch_add(bad);
            }
        }
        if (trace)
            fprintf(stderr, "multi: worker %d (device %d lane %d): started +%.3f ms, bound +%.3f, first chunk done +%.3f, "
                    "exit +%.3f, busy %.3f ms, %d chunks\n", w, dev, lane, 1e3 * (t_thread - t_entry), 1e3 * (t_bound - t_entry),
                    1e3 * (t_first - t_entry), 1e3 * (now_s() - t_entry), 1e3 * busy[(size_t)w], taken[(size_t)w]);
    };
    std::vector<std::thread> threads;
    try {
        for (int w = 1; w < nworkers; ++w) threads.emplace_back(work, w);
    } catch (const std::exception &e) {  // no more threads: the workers that exist (and this thread) do all the chunks
        if (trace) fprintf(stderr, "multi: %s after %zu threads\n", e.what(), threads.size());
    }
    cpu_set_t mine;
    const bool have_mine = sched_getaffinity(0, sizeof(mine), &mine) == 0;
    work(0);  // the calling thread is worker 0; its CPU binding is restored afterwards
    if (have_mine) sched_setaffinity(0, sizeof(mine), &mine);
    for (auto &t : threads) t.join();
    for (int w = 0; w < nworkers; ++w) {
        if (busy_s) busy_s[w / lanes] += busy[(size_t)w];
        if (chunks_taken) chunks_taken[w / lanes] += taken[(size_t)w];
    }
    if (nfailed) *nfailed = bad_systems.load();
    if (failed.load()) {
        for (int d = 0; d < ndev; ++d) drain_after_failure(devices[d]);
        for (const auto &e : errs)
            if (!e.empty()) return fail("%s", e.c_str());
        return fail("a device thread failed");
    }
    return 0;
}

int pyodesys_cuda_integrate_adaptive(int device, int64_t n, const double *y0, const double *x0, const double *xend,
                                     const double *params, const double *atol, int natol, double rtol,
                                     const double *dx0, const double *dx_min, const double *dx_max, int64_t mxsteps,
                                     int method, int autorestart, int return_on_error, int64_t max_rows,
                                     double *xyout, int32_t *nrows, int32_t *status, int32_t *counters,
                                     double *timings, int ptr_kind, const pyodesys_cuda_options *opt) {
    const int rc = integrate_adaptive_impl(device, n, y0, x0, xend, params, atol, natol, rtol, dx0, dx_min, dx_max,
                                           mxsteps, method, autorestart, return_on_error, max_rows, xyout, nrows,
                                           status, counters, timings, ptr_kind, opt);
    if (rc != 0) drain_after_failure(device);
    return rc;
}

int pyodesys_cuda_integrate_adaptive_compact(int device, int64_t n, const double *y0, const double *x0,
                                             const double *xend, const double *params, const double *atol, int natol,
                                             double rtol, const double *dx0, const double *dx_min, const double *dx_max,
                                             int64_t mxsteps, int method, int autorestart, int return_on_error,
                                             int64_t max_rows, int64_t *offsets, int32_t *status, int32_t *counters,
                                             double *timings, const pyodesys_cuda_options *opt) {
    if (!offsets) return fail("offsets must not be NULL");
    if (n == 0) { offsets[0] = 0; return 0; }
    const int rc = integrate_adaptive_impl(device, n, y0, x0, xend, params, atol, natol, rtol, dx0, dx_min, dx_max,
                                           mxsteps, method, autorestart, return_on_error, max_rows, nullptr, nullptr,
                                           status, counters, timings, 0, opt, offsets);
    if (rc != 0) drain_after_failure(device);
    return rc;
}

int pyodesys_cuda_fetch_compact_rows(int device, int ctx_lane, double *rows, int64_t nrows_total) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
    DevCtx &c = g_ctx[device][(ctx_lane > 0 && ctx_lane < kLanes) ? ctx_lane : 0];
    std::lock_guard<std::mutex> lock(c.mtx);
    if (!c.ready) return fail("no compacted rows on device %d", device);
    if (nrows_total != c.compact_rows) return fail("the last compacting call left %lld rows, %lld asked for", (long long)c.compact_rows, (long long)nrows_total);
    if (nrows_total == 0) return 0;
    if (!rows) return fail("rows is NULL");
    CU_CHECK(cudaSetDevice(device));
    CU_CHECK(cudaMemcpyAsync(rows, c.buf[B_COMPACT], (size_t)nrows_total * (1 + NY) * 8, cudaMemcpyDeviceToHost, c.stream));
    CU_CHECK(cudaStreamSynchronize(c.stream));
    return 0;
}

// See include/pyodesys_cuda.h.  Everything is queued on ``stream``; nothing is allocated, nothing is synchronised.
int pyodesys_cuda_compact_rows_noalloc(int device, int64_t n, int64_t max_rows, const double *xyout_dev,
                                       const int32_t *nrows_dev, int64_t *offsets_dev, double *rows_dev,
                                       int64_t rows_capacity, void *stream) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
    if (n < 0 || max_rows < 1 || rows_capacity < 0) return fail("invalid sizes (n=%lld, max_rows=%lld, rows_capacity=%lld)", (long long)n, (long long)max_rows, (long long)rows_capacity);
    if (n > kNoAllocMaxSystems)
        return fail("pyodesys_cuda_compact_rows_noalloc handles at most %lld systems per call (static scratch of %d tile "
                    "sums), got %lld: split the batch or use pyodesys_cuda_integrate_adaptive_compact",
                    (long long)kNoAllocMaxSystems, kNoAllocMaxTiles, (long long)n);
    if (!offsets_dev) return fail("offsets_dev must not be NULL");
    if (n > 0 && (!xyout_dev || !nrows_dev)) return fail("xyout_dev/nrows_dev must not be NULL");
    if (rows_capacity > 0 && !rows_dev) return fail("rows_dev is NULL with rows_capacity=%lld", (long long)rows_capacity);
    NoAllocDev *d = nullptr;
    if (noalloc_scratch(device, &d)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    long long *d_off = reinterpret_cast<long long *>(offsets_dev);
    if (n == 0) {
        CU_CHECK(cudaMemsetAsync(d_off, 0, 8, st));
        return 0;
    }
    const int ntiles = (int)((n + kScanTile - 1) / kScanTile);
    // the array keeps what the last call left: the part this call uses starts from zero
    CU_CHECK(cudaMemsetAsync(d->scratch, 0, (size_t)ntiles * 8, st));
    scan_tile_sums<<<ntiles, kScanBlock, 0, st>>>(nrows_dev, n, d->scratch);
    scan_tiles<<<1, kScanBlock, 0, st>>>(d->scratch, ntiles);
    scan_offsets<<<ntiles, kScanBlock, 0, st>>>(nrows_dev, n, d->scratch, d_off);
    g_launches.fetch_add(3);
    CU_CHECK(cudaGetLastError());
    if (rows_capacity > 0) {
        const int blocks = (int)std::min<int64_t>((n + 7) / 8, (int64_t)d->sm_count * 8);
        compact_rows_bounded<<<blocks, 256, 0, st>>>(xyout_dev, d_off, n, max_rows, 1 + NY, rows_dev,
                                                    (long long)rows_capacity);
        g_launches.fetch_add(1);
        CU_CHECK(cudaGetLastError());
    }
    return 0;
}

// Free every device buffer, stream and pinned table this library holds for ``device`` (they are created lazily and
// grow to the largest batch seen).  The next call re-creates what it needs.
int pyodesys_cuda_release(int device) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
    for (int lane = 0; lane < kLanes; ++lane) {
        DevCtx &c = g_ctx[device][lane];
        std::lock_guard<std::mutex> lock(c.mtx);
        if (!c.ready) continue;
        CU_CHECK(cudaSetDevice(device));
        cudaStreamSynchronize(c.stream);
        cudaStreamSynchronize(c.stream2);
        for (int b = 0; b < NBUF; ++b) {
            if (c.buf[b]) cudaFree(c.buf[b]);
            c.buf[b] = nullptr;
            c.cap[b] = 0;
        }
        cudaFreeHost(c.h_ready);
        cudaFreeHost(c.h_flags);
        c.h_ready = nullptr;
        c.h_flags = c.d_flags = nullptr;
        for (auto &e : c.ev) cudaEventDestroy(e);
        for (auto &e : c.ev_slice)
            if (e) { cudaEventDestroy(e); e = nullptr; }
        cudaEventDestroy(c.ev_k0);
        cudaEventDestroy(c.ev_k1);
        cudaEventDestroy(c.ev_fork);
        cudaEventDestroy(c.ev_join);
        cudaStreamDestroy(c.stream);
        cudaStreamDestroy(c.stream2);
        c.stream = c.stream2 = nullptr;
        c.ready = false;
        cudaGetLastError();
    }
    return 0;
}

// PCI bus id ("0000:1b:00.0") of a CUDA device: lets the host side bind the thread that drives a GPU to the CPUs of
// the GPU's NUMA node (/sys/bus/pci/devices/<id>/local_cpulist).
int pyodesys_cuda_device_pci_bus_id(int device, char *buf, int len) {
    if (!buf || len < 16) return fail("pci bus id buffer too small");
    CU_CHECK(cudaDeviceGetPCIBusId(buf, len, device));
    return 0;
}

int pyodesys_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// scratch device buffers of the verification hooks: freed on every return path
extern "C++" {
struct Scratch {
    std::vector<void *> ptrs;
    ~Scratch() { for (void *p : ptrs) cudaFree(p); }
    template <class T> T *get(size_t count) {
        void *p = nullptr;
        if (cudaMalloc(&p, (count ? count : 1) * sizeof(T)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        ptrs.push_back(p);
        return (T *)p;
    }
};
}
#define SCRATCH(var, T, count) T *var = scratch.get<T>(count); if (!var) return fail("out of device memory in a verification hook")

int pyodesys_cuda_eval(int device, int64_t n, const double *x, const double *y, const double *params, double *f,
                       double *jac, double *dfdt) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
    if (n <= 0) return 0;
    CU_CHECK(cudaSetDevice(device));
    const size_t nn = (size_t)n;
    Scratch scratch;
    SCRATCH(dx, double, nn);
    SCRATCH(dy, double, nn * NY);
    SCRATCH(dp, double, nn * NP_TOT);
    SCRATCH(df, double, nn * NY);
    double *dj = nullptr, *dd = nullptr;
    if (jac) { SCRATCH(t, double, nn * NY * NY); dj = t; }
    if (jac && dfdt) { SCRATCH(t, double, nn * NY); dd = t; }
    CU_CHECK(cudaMemcpy(dx, x, nn * 8, cudaMemcpyHostToDevice));
    CU_CHECK(cudaMemcpy(dy, y, nn * NY * 8, cudaMemcpyHostToDevice));
    if (NP_TOT) CU_CHECK(cudaMemcpy(dp, params, nn * NP_TOT * 8, cudaMemcpyHostToDevice));
    eval_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, dx, dy, dp, df, dj, dd);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaMemcpy(f, df, nn * NY * 8, cudaMemcpyDeviceToHost));
    if (dj) CU_CHECK(cudaMemcpy(jac, dj, nn * NY * NY * 8, cudaMemcpyDeviceToHost));
    if (dd) CU_CHECK(cudaMemcpy(dfdt, dd, nn * NY * 8, cudaMemcpyDeviceToHost));
    return 0;
}

int pyodesys_cuda_eval_jtimes(int device, int64_t n, const double *x, const double *y, const double *params,
                              const double *v, double *Jv) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
#if !ODESYS_HAS_JTIMES
    (void)n; (void)x; (void)y; (void)params; (void)v; (void)Jv;
    return fail("this system was generated without a Jacobian-vector product (jtimes=True)");
#else
    if (n <= 0) return 0;
    CU_CHECK(cudaSetDevice(device));
    const size_t nn = (size_t)n;
    Scratch scratch;
    SCRATCH(dx, double, nn);
    SCRATCH(dy, double, nn * NY);
    SCRATCH(dp, double, nn * NP_TOT);
    SCRATCH(dv, double, nn * NY);
    SCRATCH(dout, double, nn * NY);
    CU_CHECK(cudaMemcpy(dx, x, nn * 8, cudaMemcpyHostToDevice));
    CU_CHECK(cudaMemcpy(dy, y, nn * NY * 8, cudaMemcpyHostToDevice));
    CU_CHECK(cudaMemcpy(dv, v, nn * NY * 8, cudaMemcpyHostToDevice));
    if (NP_TOT) CU_CHECK(cudaMemcpy(dp, params, nn * NP_TOT * 8, cudaMemcpyHostToDevice));
    eval_jtimes_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, dx, dy, dp, dv, dout);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaMemcpy(Jv, dout, nn * NY * 8, cudaMemcpyDeviceToHost));
    return 0;
#endif
}

int pyodesys_cuda_eval_sparse_jac(int device, int64_t n, const double *x, const double *y, const double *params,
                                  double *data, int32_t *colptrs, int32_t *rowvals) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
#if ODESYS_NNZ < 0
    (void)n; (void)x; (void)y; (void)params; (void)data; (void)colptrs; (void)rowvals;
    return fail("this system stores its Jacobian densely (nnz = -1); generate it with sparse=True");
#else
    if (n <= 0) return 0;
    CU_CHECK(cudaSetDevice(device));
    const size_t nn = (size_t)n;
    Scratch scratch;
    SCRATCH(dx, double, nn);
    SCRATCH(dy, double, nn * NY);
    SCRATCH(dp, double, nn * NP_TOT);
    SCRATCH(dd, double, nn * ODESYS_NNZ);
    SCRATCH(dc, int32_t, NY + 1);
    SCRATCH(dr, int32_t, ODESYS_NNZ);
    CU_CHECK(cudaMemcpy(dx, x, nn * 8, cudaMemcpyHostToDevice));
    CU_CHECK(cudaMemcpy(dy, y, nn * NY * 8, cudaMemcpyHostToDevice));
    if (NP_TOT) CU_CHECK(cudaMemcpy(dp, params, nn * NP_TOT * 8, cudaMemcpyHostToDevice));
    eval_sparse_jac_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, dx, dy, dp, dd, dc, dr);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    if (ODESYS_NNZ > 0) {
        CU_CHECK(cudaMemcpy(data, dd, nn * ODESYS_NNZ * 8, cudaMemcpyDeviceToHost));
        CU_CHECK(cudaMemcpy(rowvals, dr, ODESYS_NNZ * 4, cudaMemcpyDeviceToHost));
    }
    CU_CHECK(cudaMemcpy(colptrs, dc, (NY + 1) * 4, cudaMemcpyDeviceToHost));
    return 0;
#endif
}

int pyodesys_cuda_lu_solve(int device, int64_t n, const double *A, const double *b, double *x, int32_t *ok) {
    if (device < 0 || device >= kMaxDev) return fail("invalid device %d", device);
#if ODESYS_NY > 16
    (void)n; (void)A; (void)b; (void)x; (void)ok;
    return fail("the register LU hook is limited to ny <= 16");
#else
    if (n <= 0) return 0;
    CU_CHECK(cudaSetDevice(device));
    const size_t nn = (size_t)n;
    Scratch scratch;
    SCRATCH(dA, double, nn * NY * NY);
    SCRATCH(db, double, nn * NY);
    SCRATCH(dx, double, nn * NY);
    SCRATCH(dok, int32_t, nn);
    CU_CHECK(cudaMemcpy(dA, A, nn * NY * NY * 8, cudaMemcpyHostToDevice));
    CU_CHECK(cudaMemcpy(db, b, nn * NY * 8, cudaMemcpyHostToDevice));
    lu_kernel<<<(unsigned)((n + 127) / 128), 128>>>(n, dA, db, dx, dok);
    g_launches.fetch_add(1);
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaMemcpy(x, dx, nn * NY * 8, cudaMemcpyDeviceToHost));
    CU_CHECK(cudaMemcpy(ok, dok, nn * 4, cudaMemcpyDeviceToHost));
    return 0;
#endif
}

}  // extern "C"
