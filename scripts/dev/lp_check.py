"""Phase-1 LP (scipy HiGHS) on the rows of a QP from Oracle.qp_debug: is the constraint set feasible?"""
import sys, os
import numpy as np
import scipy.sparse as sp
from scipy.optimize import linprog
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

def phase1(q, x0):
    """min t  s.t. dynamics, x_0 = x0, lb - t w <= C z + e <= ub + t w   (w = row width scale); returns t* (<=0: strictly feasible
    with that relative margin)."""
    N = q["A"].shape[0]; NZ = 10
    nv = (N + 1) * NZ + 1
    Aeq, beq, Aub, bub = [], [], [], []
    for k in range(N):
        for i in range(8):
            r = np.zeros(nv); r[k*NZ+2:k*NZ+10] = q["A"][k][i]; r[k*NZ:k*NZ+2] = q["B"][k][i]; r[(k+1)*NZ+2+i] = -1
            Aeq.append(r); beq.append(-q["c"][k][i])
    for i in range(8):
        r = np.zeros(nv); r[2+i] = 1; Aeq.append(r); beq.append(x0[i])
    for k in range(N + 1):
        for i in range(q["nrows"][k]):
            if q["soft"][k, i]: continue
            w = q["ub"][k, i] - q["lb"][k, i]
            r = np.zeros(nv); r[k*NZ:(k+1)*NZ] = q["C"][k, i]; r[-1] = -w
            Aub.append(r); bub.append(q["ub"][k, i] - q["e"][k, i])
            r = np.zeros(nv); r[k*NZ:(k+1)*NZ] = -q["C"][k, i]; r[-1] = -w
            Aub.append(r); bub.append(-(q["lb"][k, i] - q["e"][k, i]))
    c = np.zeros(nv); c[-1] = 1
    res = linprog(c, A_ub=sp.csr_matrix(np.array(Aub)), b_ub=bub, A_eq=sp.csr_matrix(np.array(Aeq)), b_eq=beq,
                  bounds=[(None, None)] * nv, method="highs")
    return res.x[-1] if res.status == 0 else np.nan

if __name__ == "__main__":
    from scripts.dev.ipm_experiment import _abi, mp, Oracle, x0, x, u, run
    kw = eval(sys.argv[1]) if len(sys.argv) > 1 else {}
    r, u0, st, xo = run(**kw)
    bad = np.nonzero(st != 0)[0]
    cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **kw)
    o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
    feas = []
    for i in bad:
        q = o.qp_debug(x0[i], x[i], u[i])
        t = phase1(q, x0[i])
        if t < 0: feas.append((int(i), float(t), q["info"][:2].tolist()))
    print(len(bad), "failed;", len(feas), "strictly feasible:", feas)
    ok = np.nonzero(st == 0)[0][:10]
    print("sanity on solved:", [round(float(phase1(o.qp_debug(x0[i], x[i], u[i]), x0[i])), 5) for i in ok])
