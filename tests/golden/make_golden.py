"""Writes tests/golden/*.npz from the CPU oracle (oracle/ihm2_oracle.cpp).

The reference ships no golden vectors and its numerical core (acados / HPIPM / CasADi) cannot
be run here, so these fixtures freeze the oracle's outputs ("parity unpinned").
They pin the oracle against regressions and give the GPU tests inputs/outputs that travel.
Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from ihm2_b200 import _abi  # noqa: E402
from ihm2_b200.track import load_packaged_track  # noqa: E402
from ihm2_b200.workloads import PERTURBATION  # noqa: E402
from oracle.lib import Oracle, initial_iterate  # noqa: E402


def main():
    rng = np.random.default_rng(1234)
    mp = load_packaged_track("fsds_default")
    ng = len(mp.s_ref)
    # ---- model + integrators at a handful of states
    cfg1 = _abi.make_config(batch=1, n_grid=ng)
    o1 = Oracle(cfg1, mp.s_ref, mp.kappa_ref)
    P = 6
    X = np.zeros((P, 8)); U = np.zeros((P, 2))
    X[:, 0] = rng.uniform(0, 300, P); X[:, 1] = rng.uniform(-1, 1, P); X[:, 2] = rng.uniform(-0.3, 0.3, P)
    X[:, 3] = rng.uniform(0.5, 20, P); X[:, 4] = rng.uniform(-1, 1, P); X[:, 5] = rng.uniform(-1, 1, P)
    X[:, 6] = rng.uniform(-400, 400, P); X[:, 7] = rng.uniform(-0.4, 0.4, P)
    U[:, 0] = X[:, 6] + rng.uniform(-100, 100, P); U[:, 1] = X[:, 7] + rng.uniform(-0.02, 0.02, P)
    F = np.zeros((P, 8)); J = np.zeros((P, 8, 10)); Fd = np.zeros((P, 8))
    Xe = np.zeros((P, 8)); Se = np.zeros((P, 8, 10)); Xi = np.zeros((P, 8)); Si = np.zeros((P, 8, 10))
    for i in range(P):
        F[i], J[i] = o1.fkin6(X[i], U[i])
        Fd[i] = o1.fdyn6(X[i], U[i])
        Xe[i], Se[i] = o1.integrate(0, X[i], U[i], cfg1.dt)
        Xi[i], Si[i] = o1.integrate(1, X[i], U[i], cfg1.dt)
    np.savez(os.path.join(HERE, "model.npz"), track="fsds_default", x=X, u=U, f=F, J=J, fdyn6=Fd,
             erk4_x=Xe, erk4_S=Se, irk_x=Xi, irk_S=Si)
    # ---- RTI: nominal after 30 perfect-model steps, then 8 perturbed instances
    x, u = initial_iterate(cfg1)
    xs = np.zeros((1, 8)); xs[0, 0] = -6.0
    for _ in range(30):
        _, st, _, x, u = o1.solve(xs, x, u, do_shift=True)
        assert st[0] == 0
        xs = x[:, 1, :].copy()
    B = 8
    cfg = _abi.make_config(batch=B, n_grid=ng)
    x0 = xs + rng.uniform(-1, 1, (B, 8)) * PERTURBATION * 0.5
    xw = np.repeat(x, B, 0); uw = np.repeat(u, B, 0)
    o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
    u0, st, stats, xo, uo = o.solve(x0, xw, uw, do_shift=True)
    A, Bm, phi, bad = o.linearize(xw, uw)
    assert bad == 0
    np.savez(os.path.join(HERE, "rti.npz"), track="fsds_default", x0=x0, x_in=xw, u_in=uw, u0=u0, status=st,
             qp_iters=stats[:, 0], x_out=xo, u_out=uo, A=A, B=Bm, phi=phi)
    # ---- plant steps (fkin6 / fdyn6 / switch) and a short closed loop
    Bs = 6
    cfgs = _abi.make_config(batch=Bs, n_grid=ng)
    os_ = Oracle(cfgs, mp.s_ref, mp.kappa_ref)
    Xs = X.copy(); Xs[:, 3] = np.array([1.0, 4.0, 8.0, 12.0, 15.0, 18.0]); Xs[:, 6] = rng.uniform(-100, 300, P)
    Us = np.stack([Xs[:, 6] + 20.0, Xs[:, 7] + 0.01], 1)
    out = {}
    for name, model in (("fkin6", _abi.SIM_FKIN6), ("fdyn6", _abi.SIM_FDYN6), ("switch", _abi.SIM_KIN6_DYN6)):
        cfgs.sim_model = model
        xn, md = os_.sim_step(Xs, Us)
        out[f"sim_{name}_x"] = xn; out[f"sim_{name}_model"] = md
    np.savez(os.path.join(HERE, "sim.npz"), track="fsds_default", x=Xs, u=Us, **out)
    Bc = 2
    cfgc = _abi.make_config(batch=Bc, n_grid=ng)
    oc = Oracle(cfgc, mp.s_ref, mp.kappa_ref)
    xc, uc = initial_iterate(cfgc)
    xs0 = np.zeros((Bc, 8)); xs0[:, 0] = -6.0; xs0[1, 1] = 0.3
    res = oc.closed_loop(xs0, xc, uc, [mp.lap_length], 40)
    np.savez(os.path.join(HERE, "closed_loop.npz"), track="fsds_default", x_start=xs0, n_steps=40,
             xlog=res["xlog"], ulog=res["ulog"], done=res["done"], fail=res["fail"])
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
