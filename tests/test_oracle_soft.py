"""CPU checks of the oracle's soft-constrained OCP and of its interior-point strategy, against
mathematics rather than against a second implementation of the same algorithm:

* h(x, u) as written in old/generate_acaods_interface.py:312-325 (numpy transcription) and its Jacobian (finite
  differences);
* the KKT conditions of the QP in (z, slacks) form -- stationarity in z and in every slack, primal / dual feasibility,
  complementarity of the rows AND of the slack bounds -- at the point the interior point returns (slacks eliminated
  inside it, so this certifies the elimination);
* a phase-1 LP (scipy HiGHS) on the hard rows of every QP the oracle reports as failed: none of them may be strictly
  feasible;
* a closed-loop lap that finishes, with the lap time of python/main.py:561-570.
"""
import numpy as np
import pytest
import scipy.sparse as sp
from scipy.optimize import linprog

from ihm2_b200 import _abi
from oracle.lib import Oracle, initial_iterate

N, NX, NU, NZ = 40, 8, 2, 10


def _widths(mp):
    return (mp.right_widths, mp.left_widths)


def h_reference(x, u, p, s_ref, wr, wl, half_L, half_W):
    """old/generate_acaods_interface.py:312-325 on the quantities of python/models.py:232-307 (F_Rx, F_Fx, beta)."""
    s, n, psi, vx, vy, r, T, dl = x
    right, left = np.interp(s, s_ref, wr), np.interp(s, s_ref, wl)
    m, lR, lF = p[_abi.P_M], p[_abi.P_LR], p[_abi.P_LF]
    C = lR / (lR + lF)
    Fm = p[_abi.P_CM0] * T
    Fd = -(p[_abi.P_CR0] + p[_abi.P_CR1] * vx + p[_abi.P_CR2] * vx * vx) * np.tanh(10 * vx)
    FRx, FFx = 0.5 * Fm + Fd, 0.5 * Fm
    beta = np.arctan(C * np.tan(dl))
    a_lat = (-FRx * np.sin(beta) + FFx * np.sin(dl - beta)) / m + (vx * vx + vy * vy) * np.sin(beta) / lR
    return np.array([
        n - half_L * np.sin(np.abs(psi)) + half_W * np.cos(psi) - right,
        -n + half_L * np.sin(np.abs(psi)) + half_W * np.cos(psi) - left,
        (u[0] - T) / p[_abi.P_TT],
        (u[1] - dl) / p[_abi.P_TD],
        a_lat,
    ])


def test_con_h_matches_reference_expression_and_finite_differences(built, fsds):
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref), nonlinear_constraints=True)
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref, widths=_widths(fsds))
    p = _abi.default_params()
    rng = np.random.default_rng(0)
    for _ in range(50):
        x = rng.uniform([0, -1.5, -0.5, 0.5, -1, -1, -400, -0.45], [380, 1.5, 0.5, 25, 1, 1, 400, 0.45])
        u = np.array([x[6] + rng.uniform(-100, 100), x[7] + rng.uniform(-0.02, 0.02)])
        h, J = o.con_h(x, u)
        ref = h_reference(x, u, p, fsds.s_ref, fsds.right_widths, fsds.left_widths, cfg.track_half_L, cfg.track_half_W)
        np.testing.assert_allclose(h, ref, rtol=1e-12, atol=1e-9)
        z = np.concatenate([x, u])
        for j in range(10):
            d = 1e-6 * max(1.0, abs(z[j]))
            zp, zm = z.copy(), z.copy(); zp[j] += d; zm[j] -= d
            fd = (h_reference(zp[:8], zp[8:], p, fsds.s_ref, fsds.right_widths, fsds.left_widths, cfg.track_half_L, cfg.track_half_W)
                  - h_reference(zm[:8], zm[8:], p, fsds.s_ref, fsds.right_widths, fsds.left_widths, cfg.track_half_L, cfg.track_half_W)) / (2 * d)
            if j == 0:
                continue   # the width tables are piecewise linear in s: the slope is the segment's (checked below)
            np.testing.assert_allclose(J[:, j], fd, rtol=2e-5, atol=2e-5 * max(1.0, np.abs(fd).max()))
        i = np.searchsorted(fsds.s_ref, x[0], side="right") - 1
        ds = fsds.s_ref[i + 1] - fsds.s_ref[i]
        assert abs(J[0, 0] + (fsds.right_widths[i + 1] - fsds.right_widths[i]) / ds) < 1e-12
        assert abs(J[1, 0] + (fsds.left_widths[i + 1] - fsds.left_widths[i]) / ds) < 1e-12


def _nl_problem(fsds, n_warm, dn):
    """One soft-constrained RTI problem down the track, the car pushed sideways by dn so that rows of h are active."""
    from oracle.workloads import nominal_iterate_cpu
    x_nom, xp, up = nominal_iterate_cpu("fsds_default", n_warm, nonlinear_constraints=True)
    from ihm2_b200.workloads import shift_np
    xs, us = shift_np(xp, up)
    x0 = x_nom.copy(); x0[1] += dn
    return x0, xs, us


@pytest.mark.parametrize("dn", [0.0, -0.5, 1.5])
def test_kkt_certificate_with_slacks(built, fsds, dn):
    """Stationarity in z and in the slacks, feasibility, complementarity of rows and of s >= 0, for the QP in its full
    (z, sL, sU) form: min 1/2 z'Hz + g'z + sum z_s (sL + sU) + Z_s/2 (sL^2 + sU^2)
    s.t. dynamics, lb - sL <= C z + e <= ub + sU, s >= 0."""
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref), nonlinear_constraints=True)
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref, widths=_widths(fsds))
    x0, xs, us = _nl_problem(fsds, 30, dn)
    q = o.qp_debug(x0, xs, us)
    assert q["status"] == 0
    # rows: python/mpc.py's (4 at stage 0, 8 on the path, 4 at the end) + h0, h1 (track) and h4 (a_lat); T_dot and
    # delta_dot of h are the two linear rows of mpc.py:91-99, kept once
    assert list(q["nrows"]) == [7] + [11] * (N - 1) + [6]
    assert q["soft"][0, :7].tolist() == [0, 0, 0, 0, 1, 1, 1]
    assert q["soft"][5, :11].tolist() == [0] * 6 + [1] * 5 and q["soft"][N, :6].tolist() == [0, 0, 1, 1, 1, 1]
    np.testing.assert_allclose(q["zZ"][3], [cfg.dt * 100.0, cfg.dt * 100.0]); np.testing.assert_allclose(q["zZ"][N], [100.0, 100.0])
    z, pi = q["z"], q["pi"]
    scale_g = 1 + np.abs(q["g"]).max()
    used = 0
    for k in range(N + 1):
        r = q["H"][k] @ z[k] + q["g"][k]
        if k < N:
            r[:2] += q["B"][k].T @ pi[k + 1]; r[2:] += q["A"][k].T @ pi[k + 1]
            np.testing.assert_allclose(q["A"][k] @ z[k, 2:] + q["B"][k] @ z[k, :2] + q["c"][k], z[k + 1, 2:],
                                       atol=1e-9 * (1 + np.abs(q["c"]).max()))
        r[2:] -= pi[k]
        zs, Zs = q["zZ"][k]
        for i in range(q["nrows"][k]):
            lamL, lamU, sL, sU, lsL, lsU = (q[n_][k, i] for n_ in ("lamL", "lamU", "sL", "sU", "lamSL", "lamSU"))
            r -= q["C"][k, i] * (lamL - lamU)
            v = q["C"][k, i] @ z[k] + q["e"][k, i]
            w = 1.0 + min(abs(q["lb"][k, i]), abs(q["ub"][k, i]))
            assert min(lamL, lamU) >= 0
            if q["soft"][k, i]:
                assert min(sL, sU, lsL, lsU) >= 0
                assert q["lb"][k, i] - sL - 1e-8 * w <= v <= q["ub"][k, i] + sU + 1e-8 * w
                assert abs(Zs * sL + zs - lamL - lsL) < 1e-7 * scale_g and abs(Zs * sU + zs - lamU - lsU) < 1e-7 * scale_g
                assert lamL * (v + sL - q["lb"][k, i]) < 1e-6 and lamU * (q["ub"][k, i] + sU - v) < 1e-6
                assert lsL * sL < 1e-6 and lsU * sU < 1e-6
                used += (sL > 1e-6) + (sU > 1e-6)
            else:
                assert sL == 0 and sU == 0
                assert q["lb"][k, i] - 1e-8 * w <= v <= q["ub"][k, i] + 1e-8 * w
                assert lamL * (v - q["lb"][k, i]) < 1e-6 and lamU * (q["ub"][k, i] - v) < 1e-6
        sl = slice(0, 2) if k == 0 else (slice(0, NZ) if k < N else slice(2, NZ))
        assert np.abs(r[sl]).max() < 1e-6 * scale_g
    assert used > 0, "the case must exercise the slacks"


def _phase1(q, x0):
    """min t  s.t. dynamics, x_0 = x0, lb - t w <= C z + e <= ub + t w over the HARD rows (w: row width); t* <= 0 means
    the constraint set has an interior of that relative margin."""
    n = q["A"].shape[0]
    nv = (n + 1) * NZ + 1
    Aeq, beq, Aub, bub = [], [], [], []
    for k in range(n):
        for i in range(8):
            r = np.zeros(nv); r[k * NZ + 2:k * NZ + 10] = q["A"][k][i]; r[k * NZ:k * NZ + 2] = q["B"][k][i]; r[(k + 1) * NZ + 2 + i] = -1
            Aeq.append(r); beq.append(-q["c"][k][i])
    for i in range(8):
        r = np.zeros(nv); r[2 + i] = 1; Aeq.append(r); beq.append(x0[i])
    for k in range(n + 1):
        for i in range(q["nrows"][k]):
            if q["soft"][k, i]:
                continue
            w = q["ub"][k, i] - q["lb"][k, i]
            r = np.zeros(nv); r[k * NZ:(k + 1) * NZ] = q["C"][k, i]; r[-1] = -w
            Aub.append(r); bub.append(q["ub"][k, i] - q["e"][k, i])
            r = np.zeros(nv); r[k * NZ:(k + 1) * NZ] = -q["C"][k, i]; r[-1] = -w
            Aub.append(r); bub.append(-(q["lb"][k, i] - q["e"][k, i]))
    c = np.zeros(nv); c[-1] = 1
    res = linprog(c, A_ub=sp.csr_matrix(np.array(Aub)), b_ub=bub, A_eq=sp.csr_matrix(np.array(Aeq)), b_eq=beq,
                  bounds=[(None, None)] * nv, method="highs")
    return res.x[-1] if res.status == 0 else np.nan


def test_no_strictly_feasible_qp_is_reported_failed(built, skidpad):
    """Config 2 (first 192 problems): every QP the interior point gives up on -- early exit or cap -- is checked with an
    LP; none may have an interior.  (Round 1: 6 of 48 failures among the first 256 problems were strictly feasible.)"""
    from ihm2_b200.workloads import perturbed_batch
    from oracle.workloads import nominal_iterate_cpu
    B = 192
    x_nom, xp, up = nominal_iterate_cpu("skidpad", 20)
    cfg = _abi.make_config(batch=B, n_grid=len(skidpad.s_ref))
    x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
    _, st, stats, _, _ = Oracle(cfg, skidpad.s_ref, skidpad.kappa_ref).solve(x0, x.copy(), u.copy())
    ok = st == 0
    assert 0.6 < ok.mean() < 0.95                         # the workload has its infeasible share
    assert stats[ok, 0].mean() <= 15.0                    # 8-15 interior-point iterations
    assert stats[~ok, 0].mean() <= 20.0                   # and the infeasible ones stop early
    o1 = Oracle(_abi.make_config(batch=1, n_grid=len(skidpad.s_ref)), skidpad.s_ref, skidpad.kappa_ref)
    margins = np.array([_phase1(o1.qp_debug(x0[i], x[i], u[i]), x0[i]) for i in np.nonzero(~ok)[0]])
    assert np.isfinite(margins).all()
    assert (margins > 0).all(), f"strictly feasible QPs reported as failed: {np.nonzero(margins <= 0)[0]}"
    # sanity of the LP itself: solved problems have an interior
    m_ok = np.array([_phase1(o1.qp_debug(x0[i], x[i], u[i]), x0[i]) for i in np.nonzero(ok)[0][:8]])
    assert (m_ok < 0).all()


def test_hpipm_mode_presets_agree_on_the_solution(built, skidpad):
    """The QP is strictly convex: every preset that converges returns the same point; they differ in iteration counts."""
    from ihm2_b200.workloads import perturbed_batch
    from oracle.workloads import nominal_iterate_cpu
    B = 48
    x_nom, xp, up = nominal_iterate_cpu("skidpad", 20)
    out = {}
    for mode in (None, "SPEED_ABS", "BALANCE", "ROBUST"):
        cfg = _abi.make_config(batch=B, n_grid=len(skidpad.s_ref), hpipm_mode=mode)
        x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
        out[mode] = Oracle(cfg, skidpad.s_ref, skidpad.kappa_ref).solve(x0, x.copy(), u.copy())
    ref = out[None]
    for mode in ("SPEED_ABS", "BALANCE", "ROBUST"):
        both = (ref[1] == 0) & (out[mode][1] == 0)
        assert both.sum() > 0.6 * B
        assert np.abs(out[mode][0][both] - ref[0][both]).max() / max(1.0, np.abs(ref[0][both]).max()) < 1e-6
    assert out[None][2][ref[1] == 0, 0].mean() < out["SPEED_ABS"][2][out["SPEED_ABS"][1] == 0, 0].mean()


def test_oracle_completes_a_lap_and_times_it(built, fsds):
    """BASELINE configs[0] on the oracle: with the soft constraints, acados' tolerance of a QP at its iteration limit and
    the straight slip-angle assignment in the plant the lap finishes; the lap time follows python/main.py:561-570."""
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref), nonlinear_constraints=True, qp_tolerate_maxiter=True, s_target=24.0)
    params = _abi.default_params()[None, :].copy(); params[0, _abi.P_CROSS] = 0.0
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref, params=params, widths=_widths(fsds))
    x, u = initial_iterate(cfg)
    xs = np.zeros((1, 8)); xs[0, 0] = -6.0
    res = o.closed_loop(xs, x, u, [fsds.lap_length], 1000, log=True)
    assert res["fail"][0] < 0 and res["done"][0] > 0
    s = res["xlog"][: res["done"][0] + 2, 0, 0]
    assert s[-1] > fsds.lap_length + 1.0                                        # main.py:515
    idx = np.argwhere(np.abs(np.diff(np.mod(s, fsds.lap_length))) > 1.0).ravel()   # main.py:561-566
    assert idx.size >= 2
    assert abs(res["lap_time"][0] - cfg.dt * (idx[1] - idx[0])) < 1e-12
    assert 30.0 < res["lap_time"][0] < 60.0
    assert np.abs(res["xlog"][: res["done"][0], 0, 1]).max() <= cfg.n_max + 0.05     # on the track all the way


def _slots_from_qp_debug(q, nl):
    """Oracle rows (static rows first, then h rows) -> the 8 / 16 row slots per stage of the C ABI, batch of one."""
    RS = 16 if nl else 8
    out = {k: np.zeros((1, N + 1, RS)) for k in ("lamL", "lamU", "sL", "sU", "lamSL", "lamSU")}
    hc = np.zeros((1, N + 1, 16))
    for k in range(N + 1):
        nst = 4 if k in (0, N) else 8
        for r in range(q["nrows"][k]):
            slot = r if r < nst else 8 + (0, 1, 4)[r - nst]
            for f in out:
                out[f][0, k, slot] = q[f][k, r]
            if r >= nst:
                i, Cr = (0, 1, 4)[r - nst], q["C"][k, r]
                if i == 0: hc[0, k, 0:3] = Cr[2:5]; hc[0, k, 3] = q["e"][k, r]
                if i == 1: hc[0, k, 4:7] = Cr[2:5]; hc[0, k, 7] = q["e"][k, r]
                if i == 4: hc[0, k, 8:12] = Cr[[2 + 3, 2 + 4, 2 + 6, 2 + 7]]; hc[0, k, 12] = q["e"][k, r]
    return out, hc


@pytest.mark.parametrize("nl", [False, True])
def test_vectorised_kkt_check_agrees_on_the_oracle_solution(built, fsds, nl):
    """tests/kkt_check.py (the certificate the GPU tests apply to whole batches) on the oracle's own solution."""
    from kkt_check import kkt_residuals
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref), nonlinear_constraints=nl)
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref, widths=_widths(fsds))
    x0, xs, us = _nl_problem(fsds, 30, 0.3)
    q = o.qp_debug(x0, xs, us)
    assert q["status"] == 0
    sl, hc = _slots_from_qp_debug(q, nl)
    z = q["z"]
    r = kkt_residuals(cfg, x0[None], q["A"][None], q["B"][None], q["c"][None], z[None, :, 2:], z[None, :N, :2], q["pi"][None],
                      sl["lamL"], sl["lamU"], params=_abi.default_params()[None], hc=hc if nl else None,
                      slacks=(sl["sL"], sl["sU"], sl["lamSL"], sl["lamSU"]) if nl else None)
    print({k: float(v[0]) for k, v in r.items()})
    assert r["stat"][0] < 1e-6 and r["dyn"][0] < 1e-9 and r["x0"][0] == 0 and r["prim"][0] < 1e-8
    assert r["comp"][0] < 1e-6 and r["dual"][0] <= 0
    if nl:
        assert r["slack_stat"][0] < 1e-6 and r["slack_comp"][0] < 1e-6 and r["slack_sign"][0] <= 0
    # and it is not vacuous: a perturbed solution fails it
    zb = z.copy(); zb[5, 3] += 1e-3
    rb = kkt_residuals(cfg, x0[None], q["A"][None], q["B"][None], q["c"][None], zb[None, :, 2:], zb[None, :N, :2], q["pi"][None],
                       sl["lamL"], sl["lamU"], params=_abi.default_params()[None], hc=hc if nl else None,
                       slacks=(sl["sL"], sl["sU"], sl["lamSL"], sl["lamSU"]) if nl else None)
    assert rb["dyn"][0] > 1e-6
