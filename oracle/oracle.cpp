// ORACLE - TEST INFRASTRUCTURE ONLY.  Never imported by the product path (pyodesys_b200/).
// Allowed users: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference.
//
// CPU restatement of the reference's batched native integration path, compiled per ODE system
// with g++ (-fopenmp) against the same generated ``odesys_generated.h`` the CUDA shim uses:
//
//   * batch driver: one integrator per system inside an OpenMP loop over systems, the role of
//     multi_predefined<OdeSys> / multi_adaptive<OdeSys> (call sites
//     pyodesys/native/sources/_odeint_wrapper.pyx:89-92,163-166, _cvode_wrapper.pyx:190-197,339-345);
//     thread count plays ANYODE_NUM_THREADS (README.rst:215-218);
//   * system callbacks: the generated OdeSys::rhs / dense_jac_rmaj
//     (pyodesys/native/sources/odesys_anyode_template.cpp:147-204,246-293);
//   * linear algebra: Doolittle LU with row pivoting + triangular solves following
//     AnyODE::getrf_callback / getrs_callback (external/anyode/include/anyode/anyode_blasless.hpp:10-95)
//     (the unmodified header itself is compiled into oracle/_ref and compared in tests);
//   * steppers: the reference delegates them to third-party code that is NOT in its tree:
//     pyodeint >= 0.10.3 -> Boost.Numeric.Odeint (Boost >= 1.72; setup.py:80, README.rst:84-86):
//     ``runge_kutta4``, ``runge_kutta_dopri5`` + ``controlled_runge_kutta`` + dense output,
//     ``rosenbrock4`` + ``rosenbrock4_controller`` + ``rosenbrock4_dense_output``.  Their published
//     algorithms are restated here TABLE-DRIVEN (Butcher / Rosenbrock coefficient arrays, generic
//     loops) - deliberately a different formulation from the hand-unrolled register code of the
//     CUDA kernels so that the two implementations check each other.
//
// Parity pinning: see tests/test_oracle_golden.py (reference known answers G1-G9, fixtures generated
// from the Python reference in this container by tests/golden/make_golden.py).
#include "odesys_generated.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using odesys_anyode::OdeSys;
constexpr int NY = ODESYS_NY;
constexpr int NP_TOT = ODESYS_NP_TOT;
typedef std::vector<double> vec;

enum { ST_OK = 0, ST_MXSTEPS = 1, ST_STEP_TOO_SMALL = 2, ST_RHS = 3, ST_ROWS = 4 };

struct Settings {
    double rtol;
    double atol[NY];
    double get_dx_max_factor;
    int error_outside_bounds;
    // set through oracle_set_extras (apply to the calls that follow)
    double max_invariant_violation;
    int return_on_root, max_roots;
    double *root_out;       // [n, max_roots, 1 + NY]
    int32_t *root_indices;  // [n, max_roots]
    int32_t *nroots_found;  // [n]
    int linear_solver;      // BDF: 0 dense LU, 1 matrix-free GMRES on OdeSys::jtimes (oracle_set_linear_solver)
};
Settings g_extras = {};

struct Work {
    OdeSys sys;
    const Settings *cfg;
    int nlu = 0, nsolve = 0;
    int rhs(double x, const double *y, double *f) {
        const int nfev0 = sys.nfev;
        int st = sys.rhs(x, y, f);
        sys.nfev = nfev0 + 1;  // every call counts (template:161 increments before the checks)
#if ODESYS_HAS_BOUNDS
        if (cfg->error_outside_bounds && sys.outside_bounds(y)) st = 1;
#endif
#if ODESYS_NINVAR > 0
        if (sys.invariant_violated(x, y)) st = 1;  // template:191-198
#endif
#if ODESYS_NONNEGATIVE
        for (int i = 0; i < NY; ++i)
            if (y[i] < 0) st = 1;
#endif
        return st;
    }
};

bool all_finite(const double *v) {
    for (int i = 0; i < NY; ++i)
        if (!std::isfinite(v[i])) return false;
    return true;
}

// ---- dense LU, column-major like anyode_blasless.hpp:21 (a[ci*lda + ri]) -----------------------
struct DenseLU {
    double a[NY * NY];
    int ipiv[NY];
    bool ok;
    double &A(int ri, int ci) { return a[ci * NY + ri]; }
    void factor() {
        ok = true;
        for (int i = 0; i < NY; ++i) {
            int pivrow = i;
            double absmax = std::fabs(A(i, i));
            for (int j = i + 1; j < NY; ++j) {
                const double cur = std::fabs(A(j, i));
                if (cur > absmax) {
                    absmax = cur;
                    pivrow = j;
                }
            }
            if (absmax == 0 || !std::isfinite(absmax)) ok = false;
            ipiv[i] = pivrow;
            if (pivrow != i)
                for (int ci = 0; ci < NY; ++ci) std::swap(A(i, ci), A(pivrow, ci));
            for (int ri = i + 1; ri < NY; ++ri) A(ri, i) /= A(i, i);
            for (int ci = i + 1; ci < NY; ++ci) {
                const double aici = A(i, ci);
                for (int ri = i + 1; ri < NY; ++ri) A(ri, ci) -= A(ri, i) * aici;
            }
        }
    }
    void solve(double *b) {
        for (int i = 0; i < NY; ++i)
            if (ipiv[i] != i) std::swap(b[i], b[ipiv[i]]);
        for (int i = 1; i < NY; ++i)
            for (int j = 0; j < i; ++j) b[i] -= A(i, j) * b[j];
        for (int i = NY - 1; i >= 0; --i) {
            for (int j = i + 1; j < NY; ++j) b[i] -= A(i, j) * b[j];
            b[i] /= A(i, i);
        }
    }
};

// ---- stepper interface ----------------------------------------------------------------------------
struct StepperBase {
    virtual ~StepperBase() {}
    virtual bool adaptive() const = 0;
    virtual int begin(Work &, double, const double *) { return 0; }
    virtual double auto_step(Work &, double t, const double *y, double span) = 0;
    virtual int attempt(Work &, double t, const double *y, double h, double *ynew, double *err) = 0;
    virtual bool control(double err, double h, double *hnext) = 0;
    virtual void accepted() {}
    virtual void dense(double theta, const double *yold, const double *ynew, double *out) = 0;
};

// Starting step size (Hairer, Norsett, Wanner, Solving ODEs I, II.4 "Starting Step Size"; max norm).
double starting_step(Work &w, double t0, const double *y, const double *f0, double span, int order) {
    const Settings &c = *w.cfg;
    double d0 = 0, d1 = 0;
    for (int i = 0; i < NY; ++i) {
        const double sk = c.atol[i] + c.rtol * std::fabs(y[i]);
        d0 = std::max(d0, std::fabs(y[i]) / sk);
        d1 = std::max(d1, std::fabs(f0[i]) / sk);
    }
    double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    h0 = std::min(h0, std::fabs(span));
    const double dir = span >= 0 ? 1.0 : -1.0;
    double y1[NY], f1[NY];
    for (int i = 0; i < NY; ++i) y1[i] = y[i] + dir * h0 * f0[i];
    const int st = w.rhs(t0 + dir * h0, y1, f1);
    double d2 = 0;
    for (int i = 0; i < NY; ++i) {
        const double sk = c.atol[i] + c.rtol * std::fabs(y[i]);
        d2 = std::max(d2, std::fabs(f1[i] - f0[i]) / sk);
    }
    d2 /= h0;
    if (st != 0 || !std::isfinite(d2)) d2 = 0;
    const double dmax = std::max(d1, d2);
    const double h1 = (dmax <= 1e-15) ? std::max(1e-6, h0 * 1e-3) : std::pow(0.01 / dmax, 1.0 / order);
    return std::min(std::min(100 * h0, h1), std::fabs(span));
}

// ---- explicit Runge-Kutta driven by a Butcher tableau ---------------------------------------------------
struct Tableau {
    int stages;
    const double *c;       // [stages]
    const double *a;       // [stages*stages] lower triangular
    const double *b;       // [stages]
    const double *e;       // [stages] error weights (b - bhat) or nullptr
};

const double rk4_c[4] = {0, 0.5, 0.5, 1};
const double rk4_a[16] = {0, 0, 0, 0, 0.5, 0, 0, 0, 0, 0.5, 0, 0, 0, 0, 1, 0};
const double rk4_b[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
const Tableau RK4 = {4, rk4_c, rk4_a, rk4_b, nullptr};

// Dormand & Prince (1980) RK5(4)7M
const double dp_c[7] = {0, 1.0 / 5, 3.0 / 10, 4.0 / 5, 8.0 / 9, 1, 1};
const double dp_a[49] = {
    0, 0, 0, 0, 0, 0, 0,
    1.0 / 5, 0, 0, 0, 0, 0, 0,
    3.0 / 40, 9.0 / 40, 0, 0, 0, 0, 0,
    44.0 / 45, -56.0 / 15, 32.0 / 9, 0, 0, 0, 0,
    19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729, 0, 0, 0,
    9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656, 0, 0,
    35.0 / 384, 0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84, 0};
const double dp_b[7] = {35.0 / 384, 0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84, 0};
const double dp_bhat[7] = {5179.0 / 57600, 0, 7571.0 / 16695, 393.0 / 640, -92097.0 / 339200, 187.0 / 2100, 1.0 / 40};
const Tableau DP5 = {7, dp_c, dp_a, dp_b, dp_bhat};
// dense output weights of dopri5.f (CONTD5)
const double dp_d[7] = {-12715105075.0 / 11282082432.0, 0, 87487479700.0 / 32700410799.0,
                        -10690763975.0 / 1880347072.0, 701980252875.0 / 199316789632.0,
                        -1453857185.0 / 822651844.0, 69997945.0 / 29380423.0};

struct ExplicitRK : StepperBase {
    const Tableau &tb;
    bool fsal;
    int order;
    double k[7][NY];
    double kfirst[NY];
    double hlast;
    double ylast_old[NY], ylast_new[NY];
    ExplicitRK(const Tableau &t, bool fsal_, int order_) : tb(t), fsal(fsal_), order(order_) {}
    bool adaptive() const override { return tb.e != nullptr || fsal; }
    int begin(Work &w, double t, const double *y) override {
        if (fsal) return w.rhs(t, y, kfirst);
        return 0;
    }
    double auto_step(Work &w, double t, const double *y, double span) override {
        return starting_step(w, t, y, kfirst, span, order) * (span >= 0 ? 1 : -1);
    }
    int attempt(Work &w, double t, const double *y, double h, double *ynew, double *err) override {
        int bad = 0;
        double tmp[NY];
        for (int s = 0; s < tb.stages; ++s) {
            if (s == 0 && fsal) {
                std::memcpy(k[0], kfirst, sizeof(kfirst));
                continue;
            }
            for (int i = 0; i < NY; ++i) {
                double acc = 0;
                for (int j = 0; j < s; ++j) acc += tb.a[s * tb.stages + j] * k[j][i];
                tmp[i] = y[i] + h * acc;
            }
            bad |= w.rhs(t + tb.c[s] * h, tmp, k[s]);
            if (fsal && s == tb.stages - 1) std::memcpy(ynew, tmp, sizeof(tmp));
        }
        if (!fsal) {
            for (int i = 0; i < NY; ++i) {
                double acc = 0;
                for (int s = 0; s < tb.stages; ++s) acc += tb.b[s] * k[s][i];
                ynew[i] = y[i] + h * acc;
            }
            *err = 0;
            return bad;
        }
        // odeint default_error_checker: max_i |err_i| / (eps_abs + eps_rel (|x_i| + |dt| |dxdt_i|))
        const Settings &c = *w.cfg;
        double e = 0;
        for (int i = 0; i < NY; ++i) {
            double acc = 0;
            for (int s = 0; s < tb.stages; ++s) acc += (dp_b[s] - dp_bhat[s]) * k[s][i];
            const double sc = c.atol[i] + c.rtol * (std::fabs(y[i]) + std::fabs(h) * std::fabs(k[0][i]));
            e = std::max(e, std::fabs(h * acc) / sc);
        }
        *err = e;
        hlast = h;
        std::memcpy(ylast_old, y, sizeof(ylast_old));
        std::memcpy(ylast_new, ynew, sizeof(ylast_new));
        if (!std::isfinite(e)) bad |= 1;
        return bad;
    }
    // odeint default_step_adjuster (stepper order 5, error order 4)
    bool control(double err, double h, double *hnext) override {
        if (!fsal) {
            *hnext = h;
            return true;
        }
        if (err > 1.0) {
            *hnext = h * std::max(0.9 * std::pow(err, -1.0 / 3.0), 0.2);
            return false;
        }
        *hnext = h;
        if (err < 0.5) {
            const double e = std::max(std::pow(5.0, -5.0), err);
            *hnext = h * std::min(0.9 * std::pow(e, -1.0 / 5.0), 5.0);
        }
        return true;
    }
    void accepted() override {
        if (fsal) std::memcpy(kfirst, k[tb.stages - 1], sizeof(kfirst));
    }
    void dense(double th, const double *yold, const double *ynew, double *out) override {
        if (!fsal) {
            for (int i = 0; i < NY; ++i) out[i] = yold[i] + th * (ynew[i] - yold[i]);
            return;
        }
        const double th1 = 1 - th;
        for (int i = 0; i < NY; ++i) {
            const double ydiff = ynew[i] - yold[i];
            const double bspl = hlast * k[0][i] - ydiff;
            const double r4 = ydiff - hlast * k[6][i] - bspl;
            double r5 = 0;
            for (int s = 0; s < 7; ++s) r5 += dp_d[s] * k[s][i];
            r5 *= hlast;
            out[i] = yold[i] + th * (ydiff + th1 * (bspl + th * (r4 + th1 * r5)));
        }
    }
};

#if ODESYS_HAS_JAC
// ---- Rosenbrock 4(3): (I/(gamma h) - J) u_i = f(y + sum a_ij u_j) + sum c_ij u_j / h + h d_i f_t -----------
// Coefficients: oracle/check_ros4_coeffs.py proves them against the order conditions.
const double ros_gamma = 0.25;
const double ros_a[6][6] = {
    {0, 0, 0, 0, 0, 0},
    {0.1544000000000000e+01, 0, 0, 0, 0, 0},
    {0.9466785280815826e+00, 0.2557011698983284e+00, 0, 0, 0, 0},
    {0.3314825187068521e+01, 0.2896124015972201e+01, 0.9986419139977817e+00, 0, 0, 0},
    {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950e+00, 0, 0},
    {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950e+00, 1, 0}};
const double ros_c[6][6] = {
    {0, 0, 0, 0, 0, 0},
    {-0.5668800000000000e+01, 0, 0, 0, 0, 0},
    {-0.2430093356833875e+01, -0.2063599157091915e+00, 0, 0, 0, 0},
    {-0.1073529058151375e+00, -0.9594562251023355e+01, -0.2047028614809616e+02, 0, 0, 0},
    {0.7496443313967647e+01, -0.1024680431464352e+02, -0.3399990352819905e+02, 0.1170890893206160e+02, 0, 0},
    {0.8083246795921522e+01, -0.7981132988064893e+01, -0.3152159432874371e+02, 0.1631930543123136e+02,
     -0.6058818238834054e+01, 0}};
const double ros_ct[6] = {0, 0.386, 0.21, 0.63, 1, 1};
const double ros_d[6] = {0.25, -0.1043, 0.1035, -0.3620000000000023e-01, 0, 0};
const double ros_dense[2][5] = {
    {0.1012623508344586e+02, -0.7487995877610167e+01, -0.3480091861555747e+02, -0.7992771707568823e+01,
     0.1025137723295662e+01},
    {-0.6762803392801253e+00, 0.6087714651680015e+01, 0.1643084320892478e+02, 0.2476722511418386e+02,
     -0.6594389125716872e+01}};

struct Rosenbrock4 : StepperBase {
    double u[6][NY];
    double cont3[NY], cont4[NY];
    bool first = true, last_rejected = false;
    double h_old = 0, err_old = 1;
    bool adaptive() const override { return true; }
    int begin(Work &, double, const double *) override {
        first = true;
        last_rejected = false;
        h_old = 0;
        err_old = 1;
        return 0;
    }
    double auto_step(Work &w, double t, const double *y, double span) override {
        double f0[NY];
        w.rhs(t, y, f0);
        return starting_step(w, t, y, f0, span, 4) * (span >= 0 ? 1 : -1);
    }
    int attempt(Work &w, double t, const double *y, double h, double *ynew, double *err) override {
        int bad = 0;
        double f[NY], J[NY * NY], dfdt[NY], ytmp[NY];
        bad |= w.rhs(t, y, f);
        bad |= w.sys.dense_jac_rmaj(t, y, f, J, NY, dfdt);
        DenseLU lu;
        for (int r = 0; r < NY; ++r)
            for (int cidx = 0; cidx < NY; ++cidx)
                lu.A(r, cidx) = (r == cidx ? 1.0 / (ros_gamma * h) : 0.0) - J[r * NY + cidx];
        lu.factor();
        w.nlu += 1;
        if (!lu.ok) bad |= 1;
        for (int s = 0; s < 6; ++s) {
            if (s > 0) {
                for (int i = 0; i < NY; ++i) {
                    double acc = y[i];
                    for (int j = 0; j < s; ++j) acc += ros_a[s][j] * u[j][i];
                    ytmp[i] = acc;
                }
                bad |= w.rhs(t + ros_ct[s] * h, ytmp, f);
            }
            for (int i = 0; i < NY; ++i) {
                double acc = 0;
                for (int j = 0; j < s; ++j) acc += ros_c[s][j] * u[j][i];
                u[s][i] = f[i] + acc / h + h * ros_d[s] * dfdt[i];
            }
            lu.solve(u[s]);
            w.nsolve += 1;
        }
        const Settings &c = *w.cfg;
        double e = 0;
        for (int i = 0; i < NY; ++i) {
            ynew[i] = ytmp[i] + u[5][i];  // ytmp = y + sum_{j<5} a_6j u_j  (a_65 = 1)
            const double sk = c.atol[i] + c.rtol * std::max(std::fabs(y[i]), std::fabs(ynew[i]));
            e = std::max(e, std::fabs(u[5][i]) / sk);
            cont3[i] = cont4[i] = 0;
            for (int j = 0; j < 5; ++j) {
                cont3[i] += ros_dense[0][j] * u[j][i];
                cont4[i] += ros_dense[1][j] * u[j][i];
            }
        }
        *err = e;
        if (!std::isfinite(e)) bad |= 1;
        return bad;
    }
    // rosenbrock4_controller::try_step
    bool control(double err, double h, double *hnext) override {
        const double safe = 0.9, fac1 = 5.0, fac2 = 1.0 / 6.0;
        double fac = std::max(fac2, std::min(fac1, std::pow(err, 0.25) / safe));
        double hn = h / fac;
        if (err <= 1.0) {
            if (!first) {
                double fp = (h_old / h) * std::pow(err * err / err_old, 0.25) / safe;
                fp = std::max(fac2, std::min(fac1, fp));
                fac = std::max(fac, fp);
                hn = h / fac;
            }
            first = false;
            h_old = h;
            err_old = std::max(0.01, err);
            if (last_rejected) hn = (h >= 0) ? std::min(hn, h) : std::max(hn, h);
            last_rejected = false;
            *hnext = hn;
            return true;
        }
        last_rejected = true;
        *hnext = hn;
        return false;
    }
    void dense(double s, const double *yold, const double *ynew, double *out) override {
        const double s1 = 1 - s;
        for (int i = 0; i < NY; ++i) out[i] = yold[i] * s1 + s * (ynew[i] + s1 * (cont3[i] + s * cont4[i]));
    }
};
// ---- BDF / NDF, variable step, variable order 1..5 ---------------------------------------------------------
// Role: cvode's default ``method='BDF'`` (pyodesys/native/sources/_cvode_wrapper.pyx:88,244).  SUNDIALS is not
// in the reference tree; this restates the published quasi-constant step size NDF method in backward
// difference form (Shampine & Reichelt, "The MATLAB ODE suite", SIAM J. Sci. Comput. 18 (1997), section 2;
// Byrne & Hindmarsh 1975 for the order/step selection; the same formulation scipy's ``BDF`` class uses):
//   D[j] = j-th backward difference of the solution on the current uniform grid of step h,
//   predictor  y_p = sum_{j<=k} D[j],   psi = sum_{j=1..k} gamma_j D[j] / alpha_k,
//   corrector  (I - c J) dy = c f(t+h, y) - psi - d,  c = h / alpha_k   (simplified Newton, stored J),
//   error      = errconst_k * d,  rms norm over atol + rtol |y_new|,
//   after k+1 equal steps: compare the estimates for orders k-1, k, k+1 and rescale D to the new step.
// Matrices (R, U of the rescaling) are built explicitly here; the CUDA kernel applies them column by column.
const int BDF_MAXORD = 5;
const int BDF_NEWTON_MAXITER = 4;
const double bdf_kappa[6] = {0, -0.1850, -1.0 / 9, -0.0823, -0.0415, 0};

#if ODESYS_HAS_JTIMES
// Full GMRES (x0 = 0, at most NY iterations, two-pass modified Gram-Schmidt, Givens rotations) on
// A = I - c J(t, yref) with A v = v - c jtimes(v); b: right-hand side in, solution out; returns the products used.
// This is this project's own matrix-free solve (role of cvode's SPGMR + user jtimes, test_cvode.py:198-217), stated
// here with plain sums (no fma) independently of the kernel's register code.
int gmres_solve(OdeSys &sys, double t, const double *yref, double c, double *b) {
    double V[NY + 1][NY], H[NY + 1][NY], cs_[NY], sn_[NY], g[NY + 1];
    double beta = 0;
    for (int i = 0; i < NY; ++i) beta += b[i] * b[i];
    beta = std::sqrt(beta);
    if (!(beta > 0)) return 0;
    for (int i = 0; i < NY; ++i) V[0][i] = b[i] / beta;
    for (int i = 0; i <= NY; ++i) g[i] = 0;
    g[0] = beta;
    int m = 0, nmv = 0;
    for (int j = 0; j < NY; ++j) {
        double w[NY], Jv[NY];
        sys.jtimes(V[j], Jv, t, yref, nullptr);
        nmv += 1;
        for (int i = 0; i < NY; ++i) w[i] = V[j][i] - c * Jv[i];
        for (int k = 0; k <= j; ++k) H[k][j] = 0;
        for (int pass = 0; pass < 2; ++pass)
            for (int k = 0; k <= j; ++k) {
                double dt = 0;
                for (int i = 0; i < NY; ++i) dt += w[i] * V[k][i];
                H[k][j] += dt;
                for (int i = 0; i < NY; ++i) w[i] -= dt * V[k][i];
            }
        double hn = 0;
        for (int i = 0; i < NY; ++i) hn += w[i] * w[i];
        hn = std::sqrt(hn);
        for (int k = 0; k < j; ++k) {
            const double hk = H[k][j], hk1 = H[k + 1][j];
            H[k][j] = cs_[k] * hk + sn_[k] * hk1;
            H[k + 1][j] = -sn_[k] * hk + cs_[k] * hk1;
        }
        const double den = std::sqrt(H[j][j] * H[j][j] + hn * hn);
        if (!(den > 0)) break;
        cs_[j] = H[j][j] / den;
        sn_[j] = hn / den;
        H[j][j] = den;
        g[j + 1] = -sn_[j] * g[j];
        g[j] = cs_[j] * g[j];
        m = j + 1;
        if (hn <= 1e-14 * den || std::fabs(g[j + 1]) <= 1e-16 * beta || j + 1 == NY) break;
        for (int i = 0; i < NY; ++i) V[j + 1][i] = w[i] / hn;
    }
    double yk[NY];
    for (int i = m - 1; i >= 0; --i) {
        double acc = g[i];
        for (int k = i + 1; k < m; ++k) acc -= H[i][k] * yk[k];
        yk[i] = acc / H[i][i];
    }
    for (int i = 0; i < NY; ++i) b[i] = 0;
    for (int k = 0; k < m; ++k)
        for (int i = 0; i < NY; ++i) b[i] += yk[k] * V[k][i];
    return nmv;
}
#endif

struct BDF : StepperBase {
    double D[BDF_MAXORD + 3][NY];
    double Jm[NY * NY];
    DenseLU lu;
    double gamma_[BDF_MAXORD + 2], alpha_[BDF_MAXORD + 2], errc_[BDF_MAXORD + 2];
    double f0[NY], d[NY], scale[NY];
    double h_cur = 0, c_lu = 0, newton_tol = 0, newton_floor = 0, safety = 0.9;
    // the tolerances arrive already calibrated (odesys_kappa, odesys_prelude.h)
    Settings cs;
    int order = 1, n_equal = 0, n_attempt = 0;
    bool have_lu = false, jac_current = false, fresh = true, newton_failed = false, refresh_jac = false;
    bool krylov = false;
    bool adaptive() const override { return true; }

    static double rms(const double *v, const double *sc) {
        double s = 0;
        for (int i = 0; i < NY; ++i) s += (v[i] / sc[i]) * (v[i] / sc[i]);
        return std::sqrt(s / NY);
    }
    int begin(Work &w, double t, const double *y) override {
        gamma_[0] = 0;
        for (int k = 1; k <= BDF_MAXORD + 1; ++k) gamma_[k] = gamma_[k - 1] + 1.0 / k;
        for (int k = 0; k <= BDF_MAXORD + 1; ++k) {
            const double kap = k <= BDF_MAXORD ? bdf_kappa[k] : 0.0;
            alpha_[k] = (1 - kap) * gamma_[k];
            errc_[k] = kap * gamma_[k] + 1.0 / (k + 1);
        }
        cs = *w.cfg;
        const double rtol = cs.rtol;
        newton_tol = std::max(10 * DBL_EPSILON / rtol, ODESYS_BDF_NEWTON_TOL);  // odesys_prelude.h
        // a correction this small is at the rounding level of the residual (eps / rtol in units of the tolerance):
        // the contraction-rate test would compare noise with noise and call a converged iteration divergent
        newton_floor = std::max(0.03 * newton_tol, 3 * DBL_EPSILON / rtol);
        order = 1;
        n_equal = 0;
        n_attempt = 0;
        refresh_jac = false;
        have_lu = false;
        fresh = true;
        newton_failed = false;
        std::memset(D, 0, sizeof(D));
        for (int i = 0; i < NY; ++i) D[0][i] = y[i];
        int bad = w.rhs(t, y, f0);
        double dfdt[NY];
        krylov = ODESYS_HAS_JTIMES && w.cfg->linear_solver == 1;
        if (!krylov) bad |= w.sys.dense_jac_rmaj(t, y, f0, Jm, NY, dfdt);
        jac_current = true;
        return bad;
    }
    double auto_step(Work &w, double t, const double *y, double span) override {
        const Settings *keep = w.cfg;
        w.cfg = &cs;
        const double h = starting_step(w, t, y, f0, span, 1) * (span >= 0 ? 1 : -1);
        w.cfg = keep;
        return h;
    }
    // D[0..order] <- (R(factor) U)^T D[0..order]
    void change_D(double factor) {
        const int n = order + 1;
        double R[6][6], U[6][6], RU[6][6], Dn[6][NY];
        for (int pass = 0; pass < 2; ++pass) {
            double(*M)[6] = pass == 0 ? R : U;
            const double fac = pass == 0 ? factor : 1.0;
            for (int j = 0; j < n; ++j) M[0][j] = 1.0;
            for (int i = 1; i < n; ++i)
                for (int j = 0; j < n; ++j) M[i][j] = (j == 0) ? 0.0 : M[i - 1][j] * (i - 1 - fac * j) / i;
        }
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                double s = 0;
                for (int k = 0; k < n; ++k) s += R[i][k] * U[k][j];
                RU[i][j] = s;
            }
        for (int j = 0; j < n; ++j)
            for (int v = 0; v < NY; ++v) {
                double s = 0;
                for (int i = 0; i < n; ++i) s += RU[i][j] * D[i][v];
                Dn[j][v] = s;
            }
        for (int j = 0; j < n; ++j) std::memcpy(D[j], Dn[j], sizeof(D[j]));
    }
    int attempt(Work &w, double t, const double *y, double h, double *ynew, double *err) override {
        (void)y;
        n_attempt += 1;
        if (fresh) {
            for (int i = 0; i < NY; ++i) D[1][i] = h * f0[i];
            h_cur = h;
            fresh = false;
        } else if (h != h_cur) {
            change_D(h / h_cur);
            h_cur = h;
            n_equal = 0;
            have_lu = false;
        }
        const double tnew = t + h;
        double ypred[NY], psi[NY];
        for (int i = 0; i < NY; ++i) {
            double s = 0, ps = 0;
            for (int j = 0; j <= order; ++j) s += D[j][i];
            for (int j = 1; j <= order; ++j) ps += gamma_[j] * D[j][i];
            ypred[i] = s;
            psi[i] = ps / alpha_[order];
            scale[i] = cs.atol[i] + cs.rtol * std::fabs(s);
        }
        const double c = h / alpha_[order];
        bool converged = false;
        int n_iter = 0;
        newton_failed = false;
        if (krylov) jac_current = true;
        if (refresh_jac) {  // a new order / step was selected: the iteration matrix is re-formed, with a new Jacobian
            refresh_jac = false;
            if (!krylov && !jac_current) {
                double dfdt[NY];
                if (w.sys.dense_jac_rmaj(tnew, ypred, nullptr, Jm, NY, dfdt)) return 1;
                jac_current = true;
                have_lu = false;
            }
        }
        while (!converged) {
            if (!krylov && (!have_lu || c != c_lu)) {
                for (int r = 0; r < NY; ++r)
                    for (int q = 0; q < NY; ++q) lu.A(r, q) = (r == q ? 1.0 : 0.0) - c * Jm[r * NY + q];
                lu.factor();
                w.nlu += 1;
                c_lu = c;
                have_lu = lu.ok;
                if (!lu.ok) return 1;
            }
            double yk[NY], f[NY], dy[NY], dy_norm_old = -1;
            for (int i = 0; i < NY; ++i) {
                d[i] = 0;
                yk[i] = ypred[i];
            }
            for (int k = 0; k < BDF_NEWTON_MAXITER; ++k) {
                n_iter = k + 1;
                const int st = w.rhs(tnew, yk, f);
                if (st != 0 || !all_finite(f)) break;
                for (int i = 0; i < NY; ++i) dy[i] = c * f[i] - psi[i] - d[i];
#if ODESYS_HAS_JTIMES
                if (krylov) w.nlu += gmres_solve(w.sys, tnew, ypred, c, dy);
                else
#endif
                lu.solve(dy);
                w.nsolve += 1;
                const double dy_norm = rms(dy, scale);
                double rate = -1;
                const bool tiny = dy_norm <= newton_floor;
                if (dy_norm_old >= 0) rate = dy_norm / dy_norm_old;
                if (!tiny && rate >= 0 && (rate >= 1 || std::pow(rate, BDF_NEWTON_MAXITER - k) / (1 - rate) * dy_norm > newton_tol)) break;
                for (int i = 0; i < NY; ++i) {
                    yk[i] += dy[i];
                    d[i] += dy[i];
                }
                if (tiny || (rate >= 0 && rate / (1 - rate) * dy_norm < newton_tol)) {
                    converged = true;
                    break;
                }
                dy_norm_old = dy_norm;
            }
            if (converged) {
                for (int i = 0; i < NY; ++i) ynew[i] = yk[i];
                break;
            }
            if (jac_current) break;
            double fp[NY], dfdt[NY];
            int bad = w.rhs(tnew, ypred, fp);
            bad |= w.sys.dense_jac_rmaj(tnew, ypred, fp, Jm, NY, dfdt);
            if (bad) return 1;
            jac_current = true;
            have_lu = false;
        }
        if (!converged) {
            newton_failed = true;
            for (int i = 0; i < NY; ++i) ynew[i] = D[0][i];
            *err = 2.0;
            return 0;
        }
        safety = 0.9 * (2 * BDF_NEWTON_MAXITER + 1) / (2.0 * BDF_NEWTON_MAXITER + n_iter);
        for (int i = 0; i < NY; ++i) scale[i] = cs.atol[i] + cs.rtol * std::fabs(ynew[i]);
        *err = errc_[order] * rms(d, scale);
        if (!std::isfinite(*err)) return 1;
        return 0;
    }
    bool control(double err, double h, double *hnext) override {
        if (newton_failed) {
            newton_failed = false;
            *hnext = 0.5 * h;
            return false;
        }
        if (err > 1.0) {
            *hnext = h * std::max(0.2, safety * std::pow(err, -1.0 / (order + 1)));
            return false;
        }
        // accepted: advance the difference table
        jac_current = false;
        n_equal += 1;
        for (int i = 0; i < NY; ++i) {
            D[order + 2][i] = d[i] - D[order + 1][i];
            D[order + 1][i] = d[i];
        }
        for (int j = order; j >= 0; --j)
            for (int i = 0; i < NY; ++i) D[j][i] += D[j + 1][i];
        *hnext = h;
        if (n_equal < order + 1 || n_attempt % (order <= 2 ? ODESYS_BDF_EPOCH / 2 : ODESYS_BDF_EPOCH) != 0) return true;  // odesys_prelude.h
        double tmp[NY];
        double em = INFINITY, ep = INFINITY;
        if (order > 1) {
            for (int i = 0; i < NY; ++i) tmp[i] = errc_[order - 1] * D[order][i];
            em = rms(tmp, scale);
        }
        if (order < BDF_MAXORD) {
            for (int i = 0; i < NY; ++i) tmp[i] = errc_[order + 1] * D[order + 2][i];
            ep = rms(tmp, scale);
        }
        const double norms[3] = {em, err, ep};
        double best = -1;
        int ibest = 1;
        for (int q = 0; q < 3; ++q) {
            const double fq = norms[q] == 0 ? INFINITY : std::pow(norms[q], -1.0 / (order + q));
            if (fq > best) {
                best = fq;
                ibest = q;
            }
        }
        order += ibest - 1;
        *hnext = h * std::min(10.0, safety * best);
        n_equal = 0;
        have_lu = false;
        refresh_jac = ODESYS_BDF_REFRESH_JAC != 0;
        return true;
    }
    // Interpolating polynomial through the last order+1 solution values (Newton backward form).
    void dense(double th, const double *, const double *, double *out) override {
        for (int i = 0; i < NY; ++i) out[i] = D[0][i];
        double p = 1;
        for (int j = 0; j < order; ++j) {
            p *= (th - 1 + j) / (1 + j);
            for (int i = 0; i < NY; ++i) out[i] += D[j + 1][i] * p;
        }
    }
};
#endif

StepperBase *make_stepper(int method) {
    switch (method) {
    case 0: return new ExplicitRK(RK4, false, 4);
    case 1: return new ExplicitRK(DP5, true, 5);
#if ODESYS_HAS_JAC
    case 2: return new Rosenbrock4();
    case 3: return new BDF();
#endif
    default: return nullptr;
    }
}

// ---- event functions (OdeSys::roots, template:377-396) ---------------------------------------------
// Sign changes of g over an accepted step are located on the stepper's dense output by a bracketing secant
// iteration with Illinois weights (the scheme documented for CVODE's rootfinding: tolerance
// 100 eps (|t| + |h|), earliest secant root drives, upper bracket end is reported).
#if ODESYS_NROOTS > 0
constexpr int NROOTS = ODESYS_NROOTS;
bool crossed(double from, double to) { return from != 0.0 && (to == 0.0 || (from < 0.0) != (to < 0.0)); }

struct RootEvent {
    double theta;  // position inside the step, 1.0 = its end
    double x;
    double y[NY];
    double g[NROOTS];
};

// first root in (lo, 1] given g at lo (g_lo) and at the end of the step (g_hi); false if there is none
bool first_root(Work &w, StepperBase *st, double t, double hh, double tnew, const double *y, const double *ynew,
                double lo, const double *g_lo, const double *g_hi, RootEvent *ev) {
    bool any = false;
    for (int r = 0; r < NROOTS; ++r) any = any || crossed(g_lo[r], g_hi[r]);
    if (!any) return false;
    const double ttol = 100.0 * DBL_EPSILON * (std::fabs(t) + std::fabs(hh)) / std::fabs(hh);
    double ta = lo, tb = 1.0;
    vec ga(g_lo, g_lo + NROOTS), gb(g_hi, g_hi + NROOTS), yb(ynew, ynew + NY), ym(NY), gm(NROOTS);
    double alpha = 1.0;
    int side = 0, sideprev = -1;
    for (int it = 0; it < 200 && (tb - ta) > ttol; ++it) {
        int imax = -1;
        double best = -1.0;
        for (int r = 0; r < NROOTS; ++r)
            if (crossed(ga[r], gb[r])) {
                const double frac = std::fabs(gb[r] / (gb[r] - ga[r]));
                if (frac > best) {
                    best = frac;
                    imax = r;
                }
            }
        alpha = (sideprev == side) ? ((side == 2) ? alpha * 2.0 : alpha * 0.5) : 1.0;
        double tm = tb - (tb - ta) * gb[imax] / (gb[imax] - alpha * ga[imax]);
        const double width = tb - ta, fracint = width / ttol, fracsub = fracint > 5.0 ? 0.1 : 0.5 / fracint;
        if (!(std::fabs(tm - ta) >= 0.5 * ttol)) tm = ta + fracsub * width;
        if (!(std::fabs(tb - tm) >= 0.5 * ttol)) tm = tb - fracsub * width;
        st->dense(tm, y, ynew, ym.data());
        w.sys.roots(std::fma(tm, hh, t), ym.data(), gm.data());
        bool chg = false;
        for (int r = 0; r < NROOTS; ++r) chg = chg || crossed(ga[r], gm[r]);
        sideprev = side;
        if (chg) {
            tb = tm;
            gb = gm;
            yb = ym;
            side = 1;
        } else {
            ta = tm;
            ga = gm;
            side = 2;
        }
    }
    ev->theta = tb;
    ev->x = (tb == 1.0) ? tnew : std::fma(tb, hh, t);
    std::copy(yb.begin(), yb.end(), ev->y);
    std::copy(gb.begin(), gb.end(), ev->g);
    return true;
}
#endif

struct Job {
    int64_t n;
    int nt;
    const double *y0, *params, *xout;
    int64_t xout_stride;
    const double *x0, *xend, *dx0, *dx_min, *dx_max;
    int64_t mxsteps;
    int autorestart;
    double *yout, *xyout;
    int64_t max_rows;
    int32_t *nrows, *nreached, *status, *counters;
};

// One system, start to finish: the body of the reference's OpenMP loop.
void run_one(const Job &J, const Settings &cfg, int method, int mode, int64_t s) {
    Work w;
    w.cfg = &cfg;
    for (int i = 0; i < NP_TOT; ++i) w.sys.m_p[i] = J.params[s * NP_TOT + i];
    w.sys.setup();
    StepperBase *st = make_stepper(method);
    double y[NY], ynew[NY];
    for (int i = 0; i < NY; ++i) y[i] = J.y0[s * NY + i];
    const double *xo = (mode == 0) ? J.xout + s * J.xout_stride : nullptr;
    double t = (mode == 0) ? xo[0] : J.x0[s];
    const double tend = (mode == 0) ? xo[J.nt - 1] : J.xend[s];
    const double dir = tend >= t ? 1.0 : -1.0;
    int k_out = 1;
    if (mode == 0) {
        for (int i = 0; i < NY; ++i) J.yout[(s * J.nt) * NY + i] = y[i];
        if (tend == t)
            for (; k_out < J.nt; ++k_out)
                for (int i = 0; i < NY; ++i) J.yout[(s * J.nt + k_out) * NY + i] = y[i];
    } else {
        double *row = J.xyout + (s * J.max_rows) * (1 + NY);
        row[0] = t;
        for (int i = 0; i < NY; ++i) row[1 + i] = y[i];
    }
    const double hmin = J.dx_min ? J.dx_min[s] : 0.0;
    double hmax = J.dx_max ? J.dx_max[s] : INFINITY;
    if (!(hmax > 0)) hmax = INFINITY;
    const int64_t budget = J.mxsteps > 0 ? J.mxsteps : std::numeric_limits<int64_t>::max();
    int64_t steps_left = budget;
    int restarts = J.autorestart, nfail = 0, nsteps = 0, nreject = 0;
#if ODESYS_NINVAR > 0
    w.sys.init_invariants(cfg.max_invariant_violation, t, y);
#endif
#if ODESYS_NROOTS > 0
    double g_old[NROOTS];
    int n_roots = 0;
    const bool find_roots = cfg.max_roots > 0;
    if (find_roots) w.sys.roots(t, y, g_old);
#endif
    bool stop_at_root = false;
    int bad0 = st->begin(w, t, y);
    double h = J.dx0 ? J.dx0[s] : 0.0;
#if ODESYS_HAS_FIRST_STEP
    if (h == 0.0) h = w.sys.get_dx0(t, y);
#endif
    if (h != 0.0) h = std::fabs(h) * dir;
    else if (method == 0) h = (mode == 0) ? dir * INFINITY : (tend - t) / 100.0;
    else h = st->auto_step(w, t, y, tend - t);
    if (bad0 || !all_finite(y)) nfail = 1 << 20;
    int status = -1;
    while (true) {
        if (nfail >= (1 << 20)) status = ST_RHS;
        else if (t == tend) status = ST_OK;
        else if (steps_left <= 0) {
            if (restarts > 0) {
                restarts -= 1;
                steps_left = J.mxsteps;
                st->begin(w, t, y);
                if (method != 0) h = st->auto_step(w, t, y, tend - t);
            } else {
                status = ST_MXSTEPS;
            }
        }
        if (status >= 0) break;
        double hh = h;
        if (std::fabs(hh) > hmax) hh = dir * hmax;
#if ODESYS_HAS_DX_MAX
        if (cfg.get_dx_max_factor > 0.0) {
            const double m = cfg.get_dx_max_factor * w.sys.get_dx_max(t, y);
            if (std::fabs(hh) > m) hh = dir * m;
        }
#endif
        if (st->adaptive() && std::fabs(hh) < hmin) hh = dir * hmin;
        double target = tend;
        if (method == 0 && mode == 0) target = xo[k_out];
        const double rem = target - t;
        const bool lands = std::fabs(hh) >= std::fabs(rem) * (1.0 - 1e-13);
        double tnew = t + hh;
        if (lands) {
            hh = rem;
            tnew = target;
        }
        const bool hits_end = lands && (target == tend);
        double err = 0, hnext = hh;
        const int bad = st->attempt(w, t, y, hh, ynew, &err);
        nsteps += 1;
        steps_left -= 1;
        bool ok;
        if (bad || !all_finite(ynew)) {
            ok = false;
            nfail += 1;
            if (nfail >= 12 || !st->adaptive()) nfail = 1 << 20;
            double ignored;
            st->control(2.0, hh, &ignored);
            hnext = 0.25 * hh;
        } else {
            ok = st->control(err, hh, &hnext);
            if (ok) nfail = 0;
        }
        if (!ok) {
            nreject += 1;
            if (st->adaptive() && (std::fabs(hnext) <= 16.0 * DBL_EPSILON * std::fabs(t) || hnext == 0.0 ||
                                   (hmin > 0.0 && std::fabs(hh) <= hmin)))
                status = ST_STEP_TOO_SMALL;
            h = hnext;
            if (status >= 0) break;
            continue;
        }
#if ODESYS_NROOTS > 0
        if (find_roots) {
            double g_hi[NROOTS], g_lo[NROOTS];
            w.sys.roots(tnew, ynew, g_hi);
            std::memcpy(g_lo, g_old, sizeof(g_lo));
            double lo = 0.0;
            RootEvent ev;
            for (int found = 0; found < 16 && !stop_at_root; ++found) {
                if (!first_root(w, st, t, hh, tnew, y, ynew, lo, g_lo, g_hi, &ev)) break;
                const bool at_end = ev.theta == 1.0;
                int idx = k_out;
                if (mode == 0)
                    while (idx < J.nt && (xo[idx] - ev.x) * dir < 0.0) idx += 1;
                const bool stored = n_roots < cfg.max_roots;
                if (stored) {
                    double *ro = cfg.root_out + (s * cfg.max_roots + n_roots) * (1 + NY);
                    ro[0] = ev.x;
                    for (int i = 0; i < NY; ++i) ro[1 + i] = ev.y[i];
                    cfg.root_indices[s * cfg.max_roots + n_roots] = idx;
                }
                n_roots += 1;
                if (mode == 1 && cfg.return_on_root) {
                    tnew = ev.x;
                    std::memcpy(ynew, ev.y, sizeof(ev.y));
                    stop_at_root = true;
                } else if (mode == 1 && !at_end && stored) {
                    if (k_out >= J.max_rows) {
                        status = ST_ROWS;
                        break;
                    }
                    double *row = J.xyout + (s * J.max_rows + k_out) * (1 + NY);
                    row[0] = ev.x;
                    for (int i = 0; i < NY; ++i) row[1 + i] = ev.y[i];
                    k_out += 1;
                }
                lo = ev.theta;
                std::memcpy(g_lo, ev.g, sizeof(g_lo));
                if (at_end) break;
            }
            std::memcpy(g_old, g_hi, sizeof(g_hi));
            if (status >= 0) break;
        }
#endif
        if (mode == 0) {
            while (k_out < J.nt) {
                const double xk = xo[k_out];
                if ((xk - tnew) * dir > 0.0) break;
                double yo[NY];
                if (xk == tnew) std::memcpy(yo, ynew, sizeof(yo));
                else st->dense((xk - t) / hh, y, ynew, yo);
                for (int i = 0; i < NY; ++i) J.yout[(s * J.nt + k_out) * NY + i] = yo[i];
                k_out += 1;
                steps_left = budget;
            }
        } else {
            if (k_out >= J.max_rows) {
                status = ST_ROWS;
                break;
            }
            double *row = J.xyout + (s * J.max_rows + k_out) * (1 + NY);
            row[0] = tnew;
            for (int i = 0; i < NY; ++i) row[1 + i] = ynew[i];
            k_out += 1;
        }
        st->accepted();
        std::memcpy(y, ynew, sizeof(y));
        t = (hits_end && !stop_at_root) ? tend : tnew;
        if (st->adaptive()) h = hnext;
        if (stop_at_root) {
            status = ST_OK;
            break;
        }
    }
    J.status[s] = status;
#if ODESYS_NROOTS > 0
    if (find_roots) cfg.nroots_found[s] = n_roots;
#endif
    if (mode == 0) {
        J.nreached[s] = k_out;
        if (status != ST_OK)
            for (int k = k_out; k < J.nt; ++k)
                for (int i = 0; i < NY; ++i) J.yout[(s * J.nt + k) * NY + i] = NAN;
    } else {
        J.nrows[s] = k_out;
    }
    if (J.counters) {
        J.counters[0 * J.n + s] = nsteps;
        J.counters[1 * J.n + s] = w.sys.nfev;
        J.counters[2 * J.n + s] = w.sys.njev;
        J.counters[3 * J.n + s] = w.nlu;
        J.counters[4 * J.n + s] = w.nsolve;
        J.counters[5 * J.n + s] = nreject;
    }
    delete st;
}

void run_batch(const Job &J, const Settings &cfg, int method, int mode, int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    // calibrated local-error tolerances (odesys_prelude.h, same table as the CUDA shim)
    Settings eff = cfg;
    eff.rtol = odesys_effective_rtol(method, cfg.rtol);
    for (int i = 0; i < NY; ++i) eff.atol[i] = odesys_kappa(method) * cfg.atol[i];
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t s = 0; s < J.n; ++s) run_one(J, eff, method, mode, s);
}

Settings make_settings(const double *atol, int natol, double rtol, double get_dx_max_factor, int eob) {
    Settings c;
    c.rtol = rtol;
    for (int i = 0; i < NY; ++i) c.atol[i] = atol[natol == 1 ? 0 : i];
    c.get_dx_max_factor = get_dx_max_factor;
    c.error_outside_bounds = eob;
    c.max_invariant_violation = g_extras.max_invariant_violation;
    c.return_on_root = g_extras.return_on_root;
    c.max_roots = (ODESYS_NROOTS > 0 && g_extras.root_out) ? g_extras.max_roots : 0;
    c.root_out = g_extras.root_out;
    c.root_indices = g_extras.root_indices;
    c.nroots_found = g_extras.nroots_found;
    c.linear_solver = g_extras.linear_solver;
    return c;
}

}  // namespace

extern "C" {

int oracle_info(int *ny, int *np_tot) {
    *ny = NY;
    *np_tot = NP_TOT;
    return 0;
}

// Options outside the positional signatures: invariant guard and event functions.  They stay in force for
// the integrate calls that follow (reset with zeros / NULL).
int oracle_set_extras(double max_invariant_violation, int return_on_root, int max_roots, double *root_out,
                      int32_t *root_indices, int32_t *nroots_found) {
    g_extras.max_invariant_violation = max_invariant_violation;
    g_extras.return_on_root = return_on_root;
    g_extras.max_roots = max_roots;
    g_extras.root_out = root_out;
    g_extras.root_indices = root_indices;
    g_extras.nroots_found = nroots_found;
    return 0;
}

// BDF only: 1 = the Newton corrections are solved by GMRES on OdeSys::jtimes (no Jacobian, no factorisation;
// counters[3] then counts the J v products) - the CPU restatement of csrc/bdf_integrator.cuh gmres_solve
int oracle_set_linear_solver(int which) {
    if (which != 0 && !(which == 1 && ODESYS_HAS_JTIMES)) return -1;
    g_extras.linear_solver = which;
    return 0;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int oracle_integrate_predefined(int64_t n, int nt, const double *y0, const double *xout, int xout_shared,
                                const double *params, const double *atol, int natol, double rtol,
                                const double *dx0, const double *dx_min, const double *dx_max, int64_t mxsteps,
                                int method, int autorestart, double get_dx_max_factor, int error_outside_bounds,
                                double *yout, int32_t *nreached, int32_t *status, int32_t *counters, int nthreads) {
    if (method < 0 || method > 3 || (method >= 2 && !ODESYS_HAS_JAC)) return -1;
    Job J;
    std::memset(&J, 0, sizeof(J));
    J.n = n; J.nt = nt; J.y0 = y0; J.params = params; J.xout = xout; J.xout_stride = xout_shared ? 0 : nt;
    J.dx0 = dx0; J.dx_min = dx_min; J.dx_max = dx_max; J.mxsteps = mxsteps; J.autorestart = autorestart;
    J.yout = yout; J.nreached = nreached; J.status = status; J.counters = counters;
    const Settings cfg = make_settings(atol, natol, rtol, get_dx_max_factor, error_outside_bounds);
    run_batch(J, cfg, method, 0, nthreads);
    return 0;
}

int oracle_integrate_adaptive(int64_t n, const double *y0, const double *x0, const double *xend,
                              const double *params, const double *atol, int natol, double rtol, const double *dx0,
                              const double *dx_min, const double *dx_max, int64_t mxsteps, int method,
                              int autorestart, double get_dx_max_factor, int error_outside_bounds, int64_t max_rows,
                              double *xyout, int32_t *nrows, int32_t *status, int32_t *counters, int nthreads) {
    if (method < 0 || method > 3 || (method >= 2 && !ODESYS_HAS_JAC)) return -1;
    Job J;
    std::memset(&J, 0, sizeof(J));
    J.n = n; J.y0 = y0; J.params = params; J.x0 = x0; J.xend = xend;
    J.dx0 = dx0; J.dx_min = dx_min; J.dx_max = dx_max; J.mxsteps = mxsteps; J.autorestart = autorestart;
    J.xyout = xyout; J.max_rows = max_rows; J.nrows = nrows; J.status = status; J.counters = counters;
    const Settings cfg = make_settings(atol, natol, rtol, get_dx_max_factor, error_outside_bounds);
    run_batch(J, cfg, method, 1, nthreads);
    return 0;
}

int oracle_eval(int64_t n, const double *x, const double *y, const double *params, double *f, double *jac,
                double *dfdt) {
    for (int64_t s = 0; s < n; ++s) {
        OdeSys sys;
        for (int i = 0; i < NP_TOT; ++i) sys.m_p[i] = params[s * NP_TOT + i];
        sys.setup();
        sys.rhs(x[s], y + s * NY, f + s * NY);
#if ODESYS_HAS_JAC
        if (jac) sys.dense_jac_rmaj(x[s], y + s * NY, f + s * NY, jac + s * NY * NY, NY, dfdt ? dfdt + s * NY : nullptr);
#endif
    }
    return 0;
}

// template:214-243 / :311-344 - the matrix-free product and the CSC Jacobian of the generated system
int oracle_eval_jtimes(int64_t n, const double *x, const double *y, const double *params, const double *v, double *Jv) {
#if ODESYS_HAS_JTIMES
    for (int64_t s = 0; s < n; ++s) {
        OdeSys sys;
        for (int i = 0; i < NP_TOT; ++i) sys.m_p[i] = params[s * NP_TOT + i];
        sys.setup();
        sys.jtimes(v + s * NY, Jv + s * NY, x[s], y + s * NY, nullptr);
    }
    return 0;
#else
    (void)n; (void)x; (void)y; (void)params; (void)v; (void)Jv;
    return -1;
#endif
}

int oracle_eval_sparse_jac(int64_t n, const double *x, const double *y, const double *params, double *data,
                           int32_t *colptrs, int32_t *rowvals) {
#if ODESYS_NNZ >= 0
    for (int64_t s = 0; s < n; ++s) {
        OdeSys sys;
        for (int i = 0; i < NP_TOT; ++i) sys.m_p[i] = params[s * NP_TOT + i];
        sys.setup();
        sys.sparse_jac_csc(x[s], y + s * NY, nullptr, data + s * ODESYS_NNZ, colptrs, rowvals);
    }
    return 0;
#else
    (void)n; (void)x; (void)y; (void)params; (void)data; (void)colptrs; (void)rowvals;
    return -1;
#endif
}

int oracle_lu_solve(int64_t n, const double *A, const double *b, double *x, int32_t *ok) {
    for (int64_t s = 0; s < n; ++s) {
        DenseLU lu;
        for (int r = 0; r < NY; ++r)
            for (int c = 0; c < NY; ++c) lu.A(r, c) = A[s * NY * NY + r * NY + c];
        lu.factor();
        for (int i = 0; i < NY; ++i) x[s * NY + i] = b[s * NY + i];
        lu.solve(x + s * NY);
        ok[s] = lu.ok ? 1 : 0;
    }
    return 0;
}
}
