"""
Verifies the Rosenbrock 4(3) coefficient set used by both the CUDA stepper and the CPU oracle.

The constants are the six-stage, stiffly accurate RODAS-type set that Boost.Numeric.Odeint's
``rosenbrock4`` stepper carries (``default_rosenbrock_coefficients``; Boost >= 1.72 per
README.rst:84-86 of the reference).  Boost is not vendored in the reference tree nor installed
here, so the numbers were restated from the published method and are *proven* here instead of
trusted: the transformed coefficients (a_ij, c_ij, gamma) are mapped back to the classical
(alpha_ij, gamma_ij, b_i) form and all order conditions up to order 4 (and order 3 for the
embedded solution) are evaluated.  Run: python oracle/check_ros4_coeffs.py
"""
import numpy as np

gam = 0.25
A = np.zeros((6, 6)); C = np.zeros((6, 6))
A[1, 0] = 0.1544000000000000e+01
A[2, 0] = 0.9466785280815826e+00; A[2, 1] = 0.2557011698983284e+00
A[3, 0] = 0.3314825187068521e+01; A[3, 1] = 0.2896124015972201e+01; A[3, 2] = 0.9986419139977817e+00
A[4, 0] = 0.1221224509226641e+01; A[4, 1] = 0.6019134481288629e+01; A[4, 2] = 0.1253708332932087e+02
A[4, 3] = -0.6878860361058950e+00
# stage 6 and the solution reuse the stage-5 row plus unit weights (stiffly accurate):
A[5, :] = A[4, :]; A[5, 4] = 1.0
C[1, 0] = -0.5668800000000000e+01
C[2, 0] = -0.2430093356833875e+01; C[2, 1] = -0.2063599157091915e+00
C[3, 0] = -0.1073529058151375e+00; C[3, 1] = -0.9594562251023355e+01; C[3, 2] = -0.2047028614809616e+02
C[4, 0] = 0.7496443313967647e+01; C[4, 1] = -0.1024680431464352e+02; C[4, 2] = -0.3399990352819905e+02
C[4, 3] = 0.1170890893206160e+02
C[5, 0] = 0.8083246795921522e+01; C[5, 1] = -0.7981132988064893e+01; C[5, 2] = -0.3152159432874371e+02
C[5, 3] = 0.1631930543123136e+02; C[5, 4] = -0.6058818238834054e+01
c = np.array([0, 0.386, 0.21, 0.63, 1.0, 1.0])
d = np.array([0.25, -0.1043, 0.1035, -0.3620000000000023e-01, 0.0, 0.0])
m = np.array(list(A[4, :4]) + [1.0, 1.0])      # new solution: y + sum(A5j g_j) + g5 + g6
mhat = np.array(list(A[4, :4]) + [1.0, 0.0])   # embedded: without the last stage

# u = Ginv k  with Ginv = diag(1/gam) - C   ->  Gamma = inv(Ginv), alpha = A Gamma, b = m Gamma
Ginv = np.eye(6)/gam - C
G = np.linalg.inv(Ginv)
alpha = A @ G
b = m @ G
bhat = mhat @ G
beta = alpha + G
ai = alpha.sum(1)
bi = beta.sum(1)


def conds(b, order):
    out = {'1': b.sum() - 1,
           '2': b @ bi - (0.5 - gam) - gam*0 - 0.0}
    # Rosenbrock order conditions (Hairer & Wanner, Solving ODEs II, IV.7, Table 7.1)
    out['2'] = b @ bi - 0.5
    if order >= 3:
        out['3a'] = b @ (ai**2) - 1/3
        out['3b'] = b @ (beta @ bi) - 1/6
    if order >= 4:
        out['4a'] = b @ (ai**3) - 1/4
        out['4b'] = b @ (ai*(alpha @ bi)) - 1/8
        out['4c'] = b @ (beta @ (ai**2)) - 1/12
        out['4d'] = b @ (beta @ (beta @ bi)) - 1/24
    return out


def main():
    worst = 0.0
    print('alpha_i vs c_i       :', np.abs(ai - c).max())
    print('gamma_i vs d_i       :', np.abs(G.sum(1) - d).max())
    worst = max(worst, np.abs(ai - c).max(), np.abs(G.sum(1) - d).max())
    for name, bb, order in (('solution (order 4)', b, 4), ('embedded (order 3)', bhat, 3)):
        for k, v in conds(bb, order).items():
            print('%-20s cond %-3s residual % .3e' % (name, k, v))
            worst = max(worst, abs(v))
    # stiffly accurate: last row of beta equals b
    print('stiff accuracy       :', np.abs(beta[5] - b).max())
    worst = max(worst, np.abs(beta[5] - b).max())
    print('worst residual', worst)
    return worst


if __name__ == '__main__':
    import sys
    sys.exit(0 if main() < 5e-13 else 1)
