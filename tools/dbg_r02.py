import sys; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from uno_b200 import KKTSolver, B200KKTError, generators as G
kkt = G.ocp_distributed(256, 2)
s = KKTSolver(device=0)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
for it in range(4):
    s.do_numerical_factorization(kkt.val)
    st = s.stats()
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    st2 = s.stats()
    print('cycle', it, 'inertia', s.get_inertia(), 'replans', st['num_replans_last'], 'delayed_total', st['num_delayed_total'], 'flops', st['flops_factor'], 't_factor', st['t_factor_last'], 'steps', st2['refine_steps_last'], 'berr', berr, 'omega', st2['backward_error_componentwise_last'], 't_solve', st2['t_solve_last'], flush=True)
print('--- NaN test', flush=True)
kkt = G.ocp_distributed(8, 2, seed=5)
s = KKTSolver(device=0)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
bad = kkt.val.copy(); bad[kkt.reg_size + 3] = np.nan
try:
    s.do_numerical_factorization(bad); print('no error! inertia', s.get_inertia(), s.stats()['num_null_pivots_last'])
except B200KKTError as e: print('factorize raised:', e)
s.do_numerical_factorization(kkt.val); print('inertia after', s.get_inertia())
rhs = kkt.rhs.copy(); rhs[0] = np.inf
try:
    x = s.solve_indefinite_system(rhs); print('no error on solve', x[:3])
except B200KKTError as e: print('solve raised:', e)
