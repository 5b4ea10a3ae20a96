"""KKT residuals of the OCP-structured QP of python/mpc.py:29-101 (+ the soft rows of
old/generate_acaods_interface.py:312-449), written out in numpy from the problem DEFINITION -- no code shared with
the oracle or the CUDA kernels.  Used to certify what the GPU returns against mathematics: the QP is strictly convex, so
a point with vanishing KKT residuals is the solution HPIPM would return.

Variables per stage z_k = (u_k, x_k); x_0 fixed.  All arrays carry a leading batch dimension.
"""
import numpy as np

from ihm2_b200 import _abi


def cost_terms(cfg, s0, N):
    """Diagonal / cross entries of the Gauss-Newton Hessian and the linear term (controller.py:320-385)."""
    W, We = np.array(cfg.W[:]), np.array(cfg.W_e[:])
    cs, ct = cfg.cost_scale_stage, cfg.cost_scale_terminal
    Hx = cs * (W[:8] + np.array([0, 0, 0, 0, 0, 0, W[10], W[11]]))
    Hu = cs * np.array([W[8] + W[10], W[9] + W[11]])
    Hxu = -cs * np.array([W[10], W[11]])            # (T, u_T) and (delta, u_delta)
    k = np.arange(N)
    gs = -cs * W[0] * (s0[:, None] + cfg.s_target * k[None, :] / N)      # [B, N]
    gsN = -ct * We[0] * (s0 + cfg.s_target)
    return Hx, Hu, Hxu, gs, ct * We, gsN


def static_rows(cfg, N):
    """(C [N+1, 8 rows, 10], lb, ub [N+1, 8], active [N+1, 8]) of python/mpc.py:77-99 in (u, x) order."""
    C = np.zeros((N + 1, 8, 10)); lb = np.zeros((N + 1, 8)); ub = np.zeros((N + 1, 8)); act = np.zeros((N + 1, 8), bool)
    for k in range(N):
        C[k, 0, 0] = 1; C[k, 1, 1] = 1
        C[k, 2, 0] = 1; C[k, 2, 2 + 6] = -1
        C[k, 3, 1] = 1; C[k, 3, 2 + 7] = -1
        for r, ix in ((4, 1), (5, 3), (6, 6), (7, 7)):
            C[k, r, 2 + ix] = 1
        ub[k] = [cfg.T_max, cfg.delta_max, cfg.g_T_max, cfg.g_delta_max, cfg.n_max, cfg.v_x_max, cfg.T_max, cfg.delta_max]
        lb[k] = -ub[k]; lb[k, 5] = 0.0
        act[k, :4] = True
        act[k, 4:] = k > 0
    for r, ix in enumerate((1, 3, 4, 5)):       # idxbx_e = [1,3,4,5] (literal, mpc.py:81-83)
        C[N, r, 2 + ix] = 1
    ub[N, :4] = [cfg.n_max, cfg.v_x_max, cfg.T_max, cfg.delta_max]; lb[N, :4] = -ub[N, :4]
    act[N, :4] = True
    return C, lb, ub, act


def h_rows(cfg, hc, params, N):
    """Rows 8..12 from the linearised h (layout of Work::HC): C [B, N+1, 5, 10], e, lb, ub, active."""
    B = hc.shape[0]
    C = np.zeros((B, N + 1, 5, 10)); e = np.zeros((B, N + 1, 5))
    C[:, :, 0, 2:5] = hc[:, :, 0:3]; e[:, :, 0] = hc[:, :, 3]
    C[:, :, 1, 2:5] = hc[:, :, 4:7]; e[:, :, 1] = hc[:, :, 7]
    itT, itD = 1.0 / params[:, _abi.P_TT], 1.0 / params[:, _abi.P_TD]
    C[:, :N, 2, 0] = itT[:, None]; C[:, :N, 2, 2 + 6] = -itT[:, None]; e[:, :, 2] = hc[:, :, 13]
    C[:, :N, 3, 1] = itD[:, None]; C[:, :N, 3, 2 + 7] = -itD[:, None]; e[:, :, 3] = hc[:, :, 14]
    for j, ix in enumerate((3, 4, 6, 7)):
        C[:, :, 4, 2 + ix] = hc[:, :, 8 + j]
    e[:, :, 4] = hc[:, :, 12]
    ub = np.array([0.0, 0.0, cfg.T_dot_max, cfg.delta_dot_max, cfg.a_lat_max]); lb = np.array([-1e3, -1e3, -ub[2], -ub[3], -ub[4]])
    # rows 2, 3 (T_dot, delta_dot) are the linear rows 2, 3 of python/mpc.py:91-99, kept once in that form
    act = np.ones((N + 1, 5), bool); act[:, 2:4] = False; act[N, 4] = False
    return C, e, lb, ub, act


def kkt_residuals(cfg, x0, A, Bm, c, x, u, pi, lamL, lamU, params=None, hc=None, slacks=None):
    """Returns a dict of per-instance residuals [B]: stationarity (relative to the gradient scale), dynamics, primal
    feasibility (relative to 1 + |bound|), dual feasibility, complementarity; with soft rows also the stationarity in
    the slacks and the complementarity of s >= 0."""
    Bn, N = x.shape[0], u.shape[1]
    nl = hc is not None
    Hx, Hu, Hxu, gs, HxN, gsN = cost_terms(cfg, x0[:, 0], N)
    Cs, lbs, ubs, acts = static_rows(cfg, N)
    out = {}
    # ---- rows: values, multipliers
    z = np.concatenate([np.concatenate([u, np.zeros((Bn, 1, 2))], 1), x], 2)         # [B, N+1, 10]
    v = np.einsum("krj,bkj->bkr", Cs, z)                                          # [B, N+1, 8]
    L, U = lamL[:, :, :8] * acts, lamU[:, :, :8] * acts
    cl = np.einsum("krj,bkr->bkj", Cs, L - U)                                     # C'(lamL - lamU) [B, N+1, 10]
    soft = np.zeros((N + 1, 8), bool)
    if nl:
        soft[1:N, 6:8] = True; soft[N, 2:4] = True
        Ch, eh, lbh, ubh, acth = h_rows(cfg, hc, params, N)
        vh = np.einsum("bkrj,bkj->bkr", Ch, z) + eh
        Lh, Uh = lamL[:, :, 8:13] * acth, lamU[:, :, 8:13] * acth
        cl += np.einsum("bkrj,bkr->bkj", Ch, Lh - Uh)
    # ---- stationarity
    ru = Hu[None, None, :] * u + Hxu[None, None, :] * x[:, :N, 6:8] + np.einsum("bkij,bki->bkj", Bm, pi[:, 1:]) - cl[:, :N, :2]
    rx = Hx[None, None, :] * x[:, :N] + np.einsum("bkij,bki->bkj", A, pi[:, 1:]) - pi[:, :N] - cl[:, :N, 2:]
    rx[:, :, 6:8] += Hxu[None, None, :] * u
    rx[:, :, 0] += gs
    rxN = HxN[None, :] * x[:, N] - pi[:, N] - cl[:, N, 2:]
    rxN[:, 0] += gsN
    scale = 1.0 + np.maximum(np.abs(gs).max(1), np.abs(gsN))
    out["stat"] = np.maximum(np.maximum(np.abs(ru).reshape(Bn, -1).max(1), np.abs(rx[:, 1:]).reshape(Bn, -1).max(1)),
                             np.abs(rxN).max(1)) / scale
    # ---- dynamics, embedding
    pred = np.einsum("bkij,bkj->bki", A, x[:, :N]) + np.einsum("bkij,bkj->bki", Bm, u) + c
    out["dyn"] = np.abs(pred - x[:, 1:]).reshape(Bn, -1).max(1) / (1.0 + np.abs(c).reshape(Bn, -1).max(1))
    out["x0"] = np.abs(x[:, 0] - x0).max(1)

    def rows(v_, lb_, ub_, L_, U_, act_, soft_, sl):
        sL, sU, lsL, lsU = sl if sl is not None else (0.0, 0.0, 0.0, 0.0)
        sL = np.where(soft_, sL, 0.0); sU = np.where(soft_, sU, 0.0)
        w = 1.0 + np.minimum(np.abs(lb_), np.abs(ub_))
        prim = np.maximum(lb_ - sL - v_, v_ - ub_ - sU) / w
        comp = np.maximum(np.abs(L_ * (v_ + sL - lb_)), np.abs(U_ * (ub_ + sU - v_)))
        dual = -np.minimum(L_, U_)
        m = np.broadcast_to(act_, v_.shape)
        f = lambda a_: np.where(m, a_, 0.0).reshape(Bn, -1).max(1)
        return f(prim), f(comp), f(dual)

    sl_s = sl_h = None
    if nl:
        sl_s = tuple(t[:, :, :8] for t in slacks); sl_h = tuple(t[:, :, 8:13] for t in slacks)
    p1, c1, d1 = rows(v, lbs[None], ubs[None], L, U, acts[None], soft[None], sl_s)
    out["prim"], out["comp"], out["dual"] = p1, c1, d1
    if nl:
        p2, c2, d2 = rows(vh, lbh[None, None], ubh[None, None], Lh, Uh, acth[None], np.ones((1, N + 1, 5), bool), sl_h)
        out["prim"] = np.maximum(p1, p2); out["comp"] = np.maximum(c1, c2); out["dual"] = np.maximum(d1, d2)
        Z = np.where(np.arange(N + 1) < N, cfg.cost_scale_stage, cfg.cost_scale_terminal)[None, :, None] * cfg.slack_Z
        zz = np.where(np.arange(N + 1) < N, cfg.cost_scale_stage, cfg.cost_scale_terminal)[None, :, None] * cfg.slack_z
        sstat, scomp, sdual = [], [], []
        for (sL, sU, lsL, lsU), Lr, Ur, m in ((sl_s, L, U, (soft & acts)[None]), (sl_h, Lh, Uh, acth[None])):
            m = np.broadcast_to(m, sL.shape)
            f = lambda a_: np.where(m, a_, 0.0).reshape(Bn, -1).max(1)
            sstat.append(f(np.maximum(np.abs(Z * sL + zz - Lr - lsL), np.abs(Z * sU + zz - Ur - lsU))))
            scomp.append(f(np.maximum(np.abs(lsL * sL), np.abs(lsU * sU))))
            sdual.append(f(-np.minimum(np.minimum(sL, sU), np.minimum(lsL, lsU))))
        out["slack_stat"] = np.maximum(*sstat) / scale
        out["slack_comp"] = np.maximum(*scomp); out["slack_sign"] = np.maximum(*sdual)
    return out
