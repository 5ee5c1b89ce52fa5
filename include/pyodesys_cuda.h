/*
 * pyodesys_cuda.h - C ABI of one compiled ODE system (one shared library per system).
 *
 * A library exporting these symbols is what pyodesys' native layer would bind for a
 * ``cuda`` back end, next to the Cython modules ``_cvode_wrapper`` / ``_odeint_wrapper`` /
 * ``_gsl_wrapper``.  Each entry point names the reference interface it replaces
 * (paths relative to the pyodesys source tree):
 *
 *   pyodesys_cuda_integrate_predefined  <->  integrate_predefined(y0, xout, params, atol, rtol,
 *        dx0, dx_max, mxsteps, method, autorestart, return_on_error, special_settings)
 *        pyodesys/native/sources/_odeint_wrapper.pyx:111-177, _cvode_wrapper.pyx:238-365,
 *        _gsl_wrapper.pyx:117-194  (which call multi_predefined<OdeSys>, :163-166)
 *   pyodesys_cuda_integrate_adaptive    <->  integrate_adaptive(y0, x0, xend, params, ...)
 *        _odeint_wrapper.pyx:37-108, _cvode_wrapper.pyx:78-235, _gsl_wrapper.pyx:37-114
 *        (multi_adaptive<OdeSys>, :89-92)
 *   pyodesys_cuda_info                  <->  OdeSys::get_ny()/get_nnz()/get_nroots()
 *        pyodesys/native/sources/odesys_anyode_template.cpp:127-145
 *   per-system ``status`` / ``counters``  <->  AnyODE::Status and AnyODE::Info
 *        external/anyode/include/anyode/anyode.hpp:25-92,125
 *
 * Conventions
 *   - every array is float64 / int32, C-contiguous, row-major exactly as the Cython wrappers
 *     receive them: y0[n,ny], params[n,np_tot], xout[n,nt] (or one shared row, see options),
 *     yout[n,nt,ny];
 *   - the caller owns every buffer; nothing returned needs to be freed;
 *   - ptr_kind 0: all array arguments are host pointers (pageable or pinned), the library
 *     stages them through device memory (H2D, kernels, D2H inside the call);
 *     ptr_kind 1: all array arguments (except atol) are device pointers on ``device`` and
 *     stay resident; the call returns after the kernels completed;
 *   - return value 0 = OK, <0 = argument or CUDA error (text via pyodesys_cuda_last_error);
 *     failure of an individual system is NOT an error return: it is status[i] != 0 and
 *     (predefined mode) nreached[i] < nt, like ``return_on_error`` in the reference;
 *   - one in-flight call per (device, options.ctx_lane) per library; distinct devices - and the two lanes of one
 *     device - may be driven from distinct host threads concurrently.
 */
#ifndef PYODESYS_CUDA_H
#define PYODESYS_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PYODESYS_CUDA_ABI_VERSION 4
#define PYODESYS_CUDA_LINSOL_DENSE 0
#define PYODESYS_CUDA_LINSOL_GMRES 1

/* integration methods (``method=`` keyword of the reference wrappers) */
#define PYODESYS_CUDA_RK4 0        /* fixed step classical Runge-Kutta (pyodesys/integrators.py:17-61) */
#define PYODESYS_CUDA_DOPRI5 1     /* odeint ``dopri5``: Dormand-Prince 5(4), dense output */
#define PYODESYS_CUDA_ROSENBROCK4 2 /* odeint ``rosenbrock4`` (default of _odeint_wrapper.pyx:43,116) */
#define PYODESYS_CUDA_BDF 3        /* variable step BDF + Newton, order <= 5 (cvode ``bdf`` role) */

/* per-system status codes written to ``status`` */
#define PYODESYS_CUDA_OK 0
#define PYODESYS_CUDA_ERR_MXSTEPS 1       /* more than mxsteps steps were needed */
#define PYODESYS_CUDA_ERR_STEP_TOO_SMALL 2 /* step size underflow (|h| <= dx_min or eps*|x|) */
#define PYODESYS_CUDA_ERR_RHS 3           /* rhs/jac kept returning an error or produced NaN/Inf */
#define PYODESYS_CUDA_ERR_ROWS 4          /* adaptive mode: more accepted steps than max_rows */

/* counters[k*n + i], k = */
#define PYODESYS_CUDA_NCOUNTERS 6
#define PYODESYS_CUDA_C_NSTEPS 0
#define PYODESYS_CUDA_C_NFEV 1
#define PYODESYS_CUDA_C_NJEV 2
#define PYODESYS_CUDA_C_NLU 3
#define PYODESYS_CUDA_C_NSOLVE 4
#define PYODESYS_CUDA_C_NREJECT 5

/* timings[] (seconds) written by the integrate calls when non-NULL */
#define PYODESYS_CUDA_NTIMINGS 4
#define PYODESYS_CUDA_T_KERNEL 0 /* CUDA-event time of the integration kernel(s) */
#define PYODESYS_CUDA_T_H2D 1
#define PYODESYS_CUDA_T_D2H 2
#define PYODESYS_CUDA_T_TOTAL 3 /* host wall clock of the whole call */

/* optional knobs; pass NULL for defaults.  Mirrors the trailing constructor arguments of
 * the generated OdeSys (odesys_anyode_template.cpp:92-98) plus launch controls. */
typedef struct pyodesys_cuda_options {
    double get_dx_max_factor;        /* <=0: ignore OdeSys::get_dx_max (default 0) */
    double max_invariant_violation;  /* != 0: rhs returns a recoverable error (step retried with h/4) when an
                                        invariant of the system leaves |I - I(x0)| <= v (v > 0: absolute; v < 0:
                                        |v I(x0)| - v), odesys_anyode_template.cpp:112-123,191-198 */
    int32_t error_outside_bounds;    /* nonzero: rhs outside bounds -> recoverable error */
    int32_t xout_shared;             /* nonzero: ``xout`` is ONE row [nt] shared by all systems */
    int32_t schedule;                /* 0 auto, 1 static (grid-stride), 2 dynamic (work stealing) */
    int32_t block_size;              /* 0 = library default */
    int32_t blocks_per_sm;           /* 0 = library default (persistent grid = SMs * this) */
    int32_t yout_soa;                /* nonzero (ptr_kind 1 only): yout laid out [nt,ny,n] */
    void *stream;                    /* cudaStream_t to launch on (ptr_kind 1); NULL = library stream */
    /* ABI v2.  The reference appends the initial values to every parameter row on the host
     * (``append_iv``: ``numpy.concatenate((p, y0), axis=-1)``, pyodesys/core.py:410-411; consumed by
     * ``m_p.assign(params, params + len(params) + ny)``, odesys_anyode_template.cpp:102).  With
     * iv_from_y0 != 0 ``params`` holds only the user's np columns ([n, np], or [n, params_ld]) and the
     * kernels take m_p[np + i] from y0[i] themselves: no O(N) host concatenation, ny*8 B less H2D per system. */
    int32_t iv_from_y0;
    int32_t params_ld;               /* leading dimension of ``params`` in doubles; 0 = dense rows */
    int32_t kernel;                  /* Rosenbrock kernel family: 0 auto (1 for ny <= 8, else 4, or 2 where 4 lacks a
                                        feature), 1 one system per thread in registers, 2 one system per lane group
                                        (ny >= 6), 3 one system per thread with W / vectors / parameters in shared
                                        memory (ny >= 9), 4 block-cooperative (ny >= 9): lane groups for the linear
                                        algebra, one system per lane for f and J */
    /* ABI v3.  Event functions: OdeSys::roots (odesys_anyode_template.cpp:377-396), consumed by the cvode wrapper
     * (``return_on_root``, ``root_indices``, ``root_out``: _cvode_wrapper.pyx:89,206-213,354-355).  Sign changes of
     * the nroots functions are located on the dense output of each accepted step to 100 eps (|x| + |h|) (thread and
     * lane-group kernels).  The three
     * arrays follow ptr_kind like every other array; max_roots == 0 (or a system without roots) switches it off. */
    int32_t return_on_root;          /* adaptive mode: the first root ends the integration; its row is the last one */
    int32_t max_roots;               /* capacity per system of root_out / root_indices */
    int32_t linear_solver;           /* method bdf: PYODESYS_CUDA_LINSOL_DENSE (0, default: LU of I - c J in registers;
                                        what cvode's dense and klu solvers do for ny <= 8) or
                                        PYODESYS_CUDA_LINSOL_GMRES (1): matrix-free Newton-Krylov on OdeSys::jtimes
                                        (cvode ``linear_solver='gmres', with_jtimes=True``, _cvode_wrapper.pyx:88-92;
                                        test_cvode.py:198-217).  No Jacobian is formed (njev stays 0) and
                                        counters[PYODESYS_CUDA_C_NLU] then counts the J v products (``njvev``).
                                        (This slot was ``reserved`` = 0 up to ABI v3: old callers get the default.) */
    double *root_out;                /* [n, max_roots, 1 + ny]: rows (x_root, y(x_root)) */
    int32_t *root_indices;           /* [n, max_roots]: adaptive: the row of xyout that holds the root;
                                        predefined: index of the first output point at or behind the root */
    int32_t *nroots_found;           /* [n]: roots met (may exceed max_roots; only the first max_roots are stored) */
    int64_t counters_ld;             /* row length of ``counters`` ([NCOUNTERS, counters_ld]); 0 = n.  Lets several
                                        calls (one per device) fill column ranges of ONE array without a gather */
    /* ABI v4 */
    int32_t ctx_lane;                /* 0 or 1: which of the two independent call contexts (streams + staging buffers)
                                        of ``device`` serves this call.  Calls on different lanes of one device may
                                        run concurrently from two host threads: the persistent blocks of one call
                                        move onto the SMs as those of the other drain (multi-GPU driver) */
    int32_t reserved4;
} pyodesys_cuda_options;

/* Compile-time facts of the system this library was generated for. Any pointer may be NULL. */
int pyodesys_cuda_info(int *ny, int *np_tot, int *npcse, int *flops_rhs, int *flops_jac);

/* abi version, warp-per-system flag, etc.: key/value lookup; returns -1 for unknown keys.  One key is dynamic:
 * "last_failed_systems" = number of systems with status != 0 in the last integrate call of the CALLING THREAD where
 * the library counted them while it waited for the device (streaming predefined calls), else -1 - the role of the
 * ``success`` scan the reference's wrappers do per system (_odeint_wrapper.pyx:169-173). */
int64_t pyodesys_cuda_query(const char *key);

int pyodesys_cuda_integrate_predefined(
    int device, int64_t n, int nt,
    const double *y0,      /* [n, ny] */
    const double *xout,    /* [n, nt]  (or [nt] when options->xout_shared) */
    const double *params,  /* [n, np_tot] (NULL allowed when np_tot == 0) */
    const double *atol, int natol, /* host; natol == 1 or ny */
    double rtol,
    const double *dx0,     /* [n] or NULL (0.0 = choose automatically) */
    const double *dx_min,  /* [n] or NULL */
    const double *dx_max,  /* [n] or NULL (inf) */
    int64_t mxsteps, int method, int autorestart, int return_on_error,
    double *yout,          /* [n, nt, ny] */
    int32_t *nreached,     /* [n] number of rows of yout that are valid */
    int32_t *status,       /* [n] */
    int32_t *counters,     /* [PYODESYS_CUDA_NCOUNTERS, n] or NULL */
    double *timings,       /* [PYODESYS_CUDA_NTIMINGS] host, or NULL */
    int ptr_kind, const pyodesys_cuda_options *options);

/* The predefined-mode call over SEVERAL devices (host pointers; ABI v4).  Role in the reference: the OpenMP loop of
 * multi_predefined (call sites _cvode_wrapper.pyx:339-345, _odeint_wrapper.pyx:163-166) - a bag of tasks over the
 * systems, here over chunks of ``chunk`` systems and over devices: ``lanes`` (1 or 2) host threads per entry of
 * ``devices`` draw chunk numbers from one atomic counter and run the single-device path on their window of the
 * caller's arrays (counters: column windows of [NCOUNTERS, n]); each thread is bound to the CPUs of its GPU's NUMA
 * node.  No collective, no gather.  busy_s[ndev] / chunks_taken[ndev] (may be NULL) report the schedule.  Results
 * are those of the single-device call bit for bit.  Event functions are not routed through this entry point. */
int pyodesys_cuda_integrate_predefined_multi(
    const int *devices, int ndev, int lanes, int64_t chunk, int64_t n, int nt,
    const double *y0, const double *xout, const double *params, const double *atol, int natol, double rtol,
    const double *dx0, const double *dx_min, const double *dx_max, int64_t mxsteps, int method, int autorestart,
    int return_on_error, double *yout, int32_t *nreached, int32_t *status, int32_t *counters,
    double *busy_s, int32_t *chunks_taken, int64_t *nfailed /* systems with status != 0, or NULL */,
    const pyodesys_cuda_options *options);

int pyodesys_cuda_integrate_adaptive(
    int device, int64_t n,
    const double *y0, const double *x0 /* [n] */, const double *xend /* [n] */,
    const double *params, const double *atol, int natol, double rtol,
    const double *dx0, const double *dx_min, const double *dx_max,
    int64_t mxsteps, int method, int autorestart, int return_on_error,
    int64_t max_rows_per_system,
    double *xyout,         /* [n, max_rows_per_system, 1 + ny]: rows of (x, y...) */
    int32_t *nrows,        /* [n] rows written per system (first row is (x0, y0)) */
    int32_t *status, int32_t *counters, double *timings,
    int ptr_kind, const pyodesys_cuda_options *options);

/* Adaptive mode with COMPACTED output (host pointers; ABI v4).  Same integration as
 * pyodesys_cuda_integrate_adaptive, but the rows stay on the device: an exclusive scan of the per-system row counts
 * gives offsets[n + 1] (written to the host array), the rows that were actually written are packed into ONE
 * contiguous device array [offsets[n], 1 + ny] and pyodesys_cuda_fetch_compact_rows copies exactly those to the
 * caller - the wrapper of the reference hands back ``result[idx].first + 1`` rows per system
 * (_cvode_wrapper.pyx:198-205), not the whole preallocated buffer.  System i owns rows offsets[i] .. offsets[i+1]-1;
 * with event functions root_indices count rows of system i from its first row.  ``max_rows_per_system`` only bounds
 * the device-side scratch (status PYODESYS_CUDA_ERR_ROWS when a system needs more).  The fetch must follow its
 * compacting call on the same (device, options.ctx_lane) before the next integrate call on that lane. */
int pyodesys_cuda_integrate_adaptive_compact(
    int device, int64_t n, const double *y0, const double *x0, const double *xend, const double *params,
    const double *atol, int natol, double rtol, const double *dx0, const double *dx_min, const double *dx_max,
    int64_t mxsteps, int method, int autorestart, int return_on_error, int64_t max_rows_per_system,
    int64_t *offsets,      /* host [n + 1] */
    int32_t *status, int32_t *counters, double *timings, const pyodesys_cuda_options *options);
int pyodesys_cuda_fetch_compact_rows(int device, int ctx_lane, double *rows /* host [offsets[n], 1 + ny] */,
                                     int64_t nrows_total);

/* Compaction of a device-resident adaptive result WITHOUT ANY ALLOCATION.  This entry point has no counterpart in
 * the reference: it was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the compaction stage of
 * pyodesys_cuda_integrate_adaptive_compact - exclusive scan of nrows, then one warp per system packs its rows - on
 * buffers the caller owns: input xyout_dev[n, max_rows_per_system, 1 + ny] and nrows_dev[n] exactly as
 * pyodesys_cuda_integrate_adaptive leaves them with ptr_kind = 1; output offsets_dev[n + 1] (device) and
 * rows_dev[rows_capacity, 1 + ny] (device).  Rows beyond rows_capacity are not written: offsets_dev[n] is the
 * number of rows the result has, compare it with rows_capacity after synchronising (rows_capacity = 0 computes the
 * offsets only; rows_dev may then be NULL).  All work is queued on ``stream`` (a cudaStream_t, NULL = the default
 * stream) and the call returns without synchronising; it calls no cudaMalloc / cudaFree / allocating routine.
 * The scratch of the scan (one partial sum per 2048 systems) is a static __device__ array of
 * PYODESYS_CUDA_NOALLOC_MAX_TILES int64 = 64 KiB that every GPU which loads the library holds until exit, hence
 * n <= PYODESYS_CUDA_NOALLOC_MAX_SYSTEMS per call (larger n: error -1 with a message that says so).  Every call on a
 * device shares that array (its used part is cleared at the start of each call): calls that overlap on the GPU -
 * different streams, several host threads - race; serialise them on one stream. */
#define PYODESYS_CUDA_NOALLOC_MAX_TILES 8192
#define PYODESYS_CUDA_NOALLOC_MAX_SYSTEMS (8192LL * 2048LL)
int pyodesys_cuda_compact_rows_noalloc(int device, int64_t n, int64_t max_rows_per_system, const double *xyout_dev,
                                       const int32_t *nrows_dev, int64_t *offsets_dev, double *rows_dev,
                                       int64_t rows_capacity, void *stream);

/* Verification hooks (host pointers): evaluate the generated OdeSys::rhs / dense_jac_rmaj
 * (odesys_anyode_template.cpp:147-204,246-293) for a batch, and run the in-register LU that
 * the Rosenbrock stepper uses (role of AnyODE::getrf_callback/getrs_callback,
 * external/anyode/include/anyode/anyode_blasless.hpp:10-95) on a batch of ny x ny systems.
 * x[n], y[n,ny], params[n,np_tot] -> f[n,ny], jac[n,ny,ny] row-major (may be NULL), dfdt[n,ny]
 * (may be NULL);  A[n,ny,ny] row-major, b[n,ny] -> x[n,ny], ok[n] (0 = singular). */
int pyodesys_cuda_eval(int device, int64_t n, const double *x, const double *y, const double *params,
                       double *f, double *jac, double *dfdt);
int pyodesys_cuda_lu_solve(int device, int64_t n, const double *A, const double *b, double *x, int32_t *ok);

/* The two alternative Jacobian forms of the generated OdeSys, evaluated for a batch (host pointers):
 *  - OdeSys::jtimes (odesys_anyode_template.cpp:214-243; expressions from SymbolicSys.get_jtimes,
 *    pyodesys/symbolic.py:652-665): v[n,ny] -> Jv[n,ny] = J(x, y) v without forming J.  Systems generated
 *    without ``jtimes=True`` return an error (the template returns unrecoverable_error, :239-243);
 *    ``pyodesys_cuda_query("has_jtimes")`` tells.
 *  - OdeSys::sparse_jac_csc / get_nnz (template:311-344,131-133; layout of be.sparse_jacobian_csc,
 *    pyodesys/symbolic.py:641): data[n,nnz] in compressed-sparse-column order, colptrs[ny+1] and rowvals[nnz]
 *    (shared by all systems, written once).  Systems generated without ``sparse=True`` have nnz = -1
 *    (``pyodesys_cuda_query("nnz")``) and return an error.
 * The integrator kernels of a ``sparse=True`` system assemble their dense ny x ny iteration matrix from exactly
 * these nnz expressions (structural zeros are compile-time constants). */
int pyodesys_cuda_eval_jtimes(int device, int64_t n, const double *x, const double *y, const double *params,
                              const double *v, double *Jv);
int pyodesys_cuda_eval_sparse_jac(int device, int64_t n, const double *x, const double *y, const double *params,
                                  double *data, int32_t *colptrs, int32_t *rowvals);

/* Page-locked host memory for inputs/outputs of the ptr_kind 0 calls (portable across devices).  D2H/H2D
 * of pageable numpy memory runs at ~6 GB/s through the driver's staging buffer; pinned memory reaches
 * PCIe speed.  The reference wrappers allocate ``yout`` themselves (numpy.empty, _odeint_wrapper.pyx:162);
 * the binding uses this allocator for the same arrays.  Returns NULL on failure. */
void *pyodesys_cuda_host_alloc(size_t bytes);
void pyodesys_cuda_host_free(void *ptr);

/* Free the device buffers, streams and pinned tables the library keeps for ``device`` (created lazily, sized for the
 * largest batch seen; the role of ``del systems[idx]`` / ``free`` at the end of the Cython wrappers,
 * _cvode_wrapper.pyx:227-233, for state that here outlives a call).  The next call re-creates what it needs. */
int pyodesys_cuda_release(int device);

/* PCI bus id of CUDA device ``device`` as text ("0000:1b:00.0", len >= 16): the host-side multi-GPU driver binds the
 * thread that feeds a GPU to the CPUs of the GPU's NUMA node.  pyodesys_cuda_device_count: visible CUDA devices. */
int pyodesys_cuda_device_pci_bus_id(int device, char *buf, int len);
int pyodesys_cuda_device_count(void);

/* Number of kernel launches issued by this library since it was loaded (bench bookkeeping). */
int64_t pyodesys_cuda_launch_count(void);

/* Thread-local, valid until the next failing call on the same thread. */
const char *pyodesys_cuda_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* PYODESYS_CUDA_H */
