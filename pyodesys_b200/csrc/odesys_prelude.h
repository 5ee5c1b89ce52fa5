// Shared prelude of every generated ODE-system header.
//
// Provides the few names the generated ``struct odesys_anyode::OdeSys`` (and user supplied
// ``namespace_override`` snippets written for the reference's template) rely on:
//   AnyODE::Status   - callback return codes, same values as
//                      external/anyode/include/anyode/anyode.hpp:125
//   AnyODE::ignore   - silences unused-argument warnings (anyode.hpp:20-22)
// Compiles under nvcc (device + host) and under plain g++ (CPU oracle).
#pragma once
#include <math.h>
#ifdef __CUDACC__
#define ODESYS_FN __host__ __device__ __forceinline__
#define ODESYS_RESTRICT __restrict__
#else
#define ODESYS_FN inline
#define ODESYS_RESTRICT __restrict__
#endif

// Tolerance calibration: a stepper controls the LOCAL error of one step, the parity rule of this
// back end (BASELINE.json north_star) bounds the GLOBAL error: |y - y_ref| <= max(10 rtol |y|, 10 atol) at every
// output point.  Each adaptive stepper therefore runs its error test at kappa * (atol, rtol), kappa calibrated on the
// BASELINE configurations against tight-tolerance solutions of the reference's ODESys + LSODA
// (tests/golden/make_golden_r2.py; achieved ratios: profiles/r2_parity.json):
//   Dormand-Prince 5(4): 0.01  (Lorenz at t = 2 sits at 55 bands with kappa = 1, 0.5 with 0.01; van der Pol 1.7 -> 0.02)
//   Rosenbrock-4:        1     (Robertson 0.15, radiolysis 0.10 of the band)
//   BDF 1-5:             0.1   (Robertson 0.6 / radiolysis 1.3 of the band with 0.25)
// never below 100 eps relative (a request of rtol <= 2.2e-14 is used as given).  The CPU oracle applies the same
// table, so "same inputs, same tolerances" means the same work on both sides.
#ifndef ODESYS_KAPPA_DOPRI5
#define ODESYS_KAPPA_DOPRI5 0.01
#endif
#ifndef ODESYS_KAPPA_ROS4
#define ODESYS_KAPPA_ROS4 1.0
#endif
#ifndef ODESYS_KAPPA_BDF
#define ODESYS_KAPPA_BDF 0.1
#endif
ODESYS_FN double odesys_kappa(int method) {  // method ids of include/pyodesys_cuda.h: 0 rk4, 1 dopri5, 2 rosenbrock4, 3 bdf
    return method == 1 ? ODESYS_KAPPA_DOPRI5 : method == 2 ? ODESYS_KAPPA_ROS4 : method == 3 ? ODESYS_KAPPA_BDF : 1.0;
}
ODESYS_FN double odesys_effective_rtol(int method, double rtol) {
    const double floor_ = rtol < 2.220446049250313e-14 ? rtol : 2.220446049250313e-14;
    const double r = odesys_kappa(method) * rtol;
    return r > floor_ ? r : floor_;
}

// BDF: order / step-size selection happens on a fixed grid of step attempts ("epochs"), not whenever order + 1
// equal steps have accumulated: after attempt number k of a system (counted from the start of the integration) a
// new order and step are chosen only if k is a multiple of ODESYS_BDF_EPOCH and at least order + 1 equal steps were
// taken (CVODE likewise waits - its ``qwait`` - before it reconsiders order and step; waiting longer is always
// allowed).  The Jacobian is re-evaluated with every such selection (the iteration matrix is re-formed then
// anyway).  Why a grid: on the GPU the systems of a warp (thread kernel) or block (cooperative kernel) start on
// the same grid, so the expensive, formerly divergent parts of a step - rescaling the difference table,
// Jacobian, factorisation - fall into the same iteration for all of them.  The CPU oracle applies the same rule.
#ifndef ODESYS_BDF_EPOCH
#define ODESYS_BDF_EPOCH 6
#endif
#ifndef ODESYS_BDF_REFRESH_JAC
#define ODESYS_BDF_REFRESH_JAC 1
#endif
// Convergence tolerance of the simplified Newton iteration, in units of the local error tolerance (weighted rms
// norm).  scipy's BDF uses min(0.03, sqrt(rtol)) - 3e-5 at the calibrated rtol of 1e-9 -, which is BELOW the rounding
// floor of the corrections on kinetics networks: the residual c f(y) carries eps x (gross reaction rate) x c, and
// on a species that is controlled by atol that is 1e-3 .. 1e-2 tolerance units once c = h / alpha reaches ~10
// (radiolysis, t > 500: corrections stagnate at 7e-3, the rate test calls the iteration divergent, the step is
// halved seven times, grows back and fails again - hundreds of rejections on 1-4 % of the systems).  CVODE asks
// for 0.1; 0.03 keeps the algebraic error at 3 % of the tolerance and is above the floor.
#ifndef ODESYS_BDF_NEWTON_TOL
#define ODESYS_BDF_NEWTON_TOL 0.03
#endif

namespace AnyODE {
enum Status : int { success = 0, recoverable_error = 1, unrecoverable_error = -1 };
template <class T> ODESYS_FN void ignore(const T &) {}
}  // namespace AnyODE

template <int N> ODESYS_FN double odesys_ipow(double x) {
    if constexpr (N == 0) {
        return 1.0;
    } else if constexpr (N == 1) {
        return x;
    } else if constexpr (N % 2 == 0) {
        const double h = odesys_ipow<N / 2>(x);
        return h * h;
    } else {
        return x * odesys_ipow<N - 1>(x);
    }
}
