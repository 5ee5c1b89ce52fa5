import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np
import systems_for_tests as tsys
s = tsys.decay_err()
mod = s._native.mod
y = np.array([[5.,4,2]]); p = np.array([[4.,3,5,4,2]])
print('eval', mod.eval_rhs_jac(0.1, y, p))
print('eval x>.5', mod.eval_rhs_jac(0.6, y, p))
for m in ['rk4','dopri5']:
    r = mod.integrate_predefined_arrays(y, np.array([0,.1,.2]), p, 1e-8, 1e-8, mxsteps=100, method=m, dx0=0.01)
    print(m, r)
r = mod.integrate_adaptive_arrays(y, [0.], [.2], p, 1e-6, 1e-6, mxsteps=30, method='dopri5')
print(r[0][0,:r[1][0]], r[1:])
s2 = tsys.decay3()
r = s2._native.mod.integrate_predefined_arrays(y, np.array([0,.1,.2]), np.array([[4.,3,0,5,4,2]]), 1e-8, 1e-8, mxsteps=100, method='dopri5')
print('decay3', r)
