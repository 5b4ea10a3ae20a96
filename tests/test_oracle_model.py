"""Independent checks that pin the CPU oracle's model and integrators (the reference has no
tests of its own): finite differences, a restatement of fkin6 in numpy written directly from
python/models.py:232-307, scipy ODE solutions, and the committed golden vectors."""
import os

import numpy as np
import pytest
from scipy.integrate import solve_ivp

from ihm2_b200 import _abi
from oracle.lib import Oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def fkin6_numpy(x, u, s_ref, kappa_ref, p):
    """python/models.py:232-307 transcribed line by line (numpy instead of CasADi)."""
    s, n, psi, v_x, v_y, r, T, delta = x
    u_T, u_delta = u
    m, l_R, l_F = p[_abi.P_M], p[_abi.P_LR], p[_abi.P_LF]
    C = l_R / (l_R + l_F)
    delta_dot = (u_delta - delta) / p[_abi.P_TD]
    T_dot = (u_T - T) / p[_abi.P_TT]
    F_motor = p[_abi.P_CM0] * T
    F_drag = -(p[_abi.P_CR0] + p[_abi.P_CR1] * v_x + p[_abi.P_CR2] * v_x**2) * np.tanh(10 * v_x)
    F_Rx = 0.5 * F_motor + F_drag
    F_Fx = 0.5 * F_motor
    beta = np.arctan(C * np.tan(delta))
    v_dot = (F_Rx * np.cos(beta) + F_Fx * np.cos(delta - beta)) / m
    beta_dot = C * (1 + np.tan(delta) ** 2) / (1 + C**2 * np.tan(delta) ** 2) * delta_dot
    i = int(np.clip(np.searchsorted(s_ref, s, side="right") - 1, 0, len(s_ref) - 2))
    kappa = kappa_ref[i] + (kappa_ref[i + 1] - kappa_ref[i]) / (s_ref[i + 1] - s_ref[i]) * (s - s_ref[i])
    s_dot = (v_x * np.cos(psi) - v_y * np.sin(psi)) / (1 + kappa * n)
    v_y_dot = v_dot * np.sin(beta) + beta_dot * v_x
    return np.array([s_dot, v_x * np.sin(psi) + v_y * np.cos(psi), r - kappa * s_dot,
                     v_dot * np.cos(beta) - beta_dot * v_y, v_y_dot, l_R * v_y_dot - beta_dot, T_dot, delta_dot])


@pytest.fixture(scope="module")
def orc(built, fsds):
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref))
    return Oracle(cfg, fsds.s_ref, fsds.kappa_ref)


def _points(n, seed=0):
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(0, 300, n), rng.uniform(-1, 1, n), rng.uniform(-0.3, 0.3, n), rng.uniform(0.5, 20, n),
                  rng.uniform(-1, 1, n), rng.uniform(-1, 1, n), rng.uniform(-400, 400, n), rng.uniform(-0.4, 0.4, n)], 1)
    U = np.stack([X[:, 6] + rng.uniform(-100, 100, n), X[:, 7] + rng.uniform(-0.02, 0.02, n)], 1)
    return X, U


def test_fkin6_matches_reference_transcription(orc, fsds):
    X, U = _points(50)
    p = _abi.default_params()
    for x, u in zip(X, U):
        f, _ = orc.fkin6(x, u)
        np.testing.assert_allclose(f, fkin6_numpy(x, u, fsds.s_ref, fsds.kappa_ref, p), rtol=1e-12, atol=1e-10)


def test_fkin6_jacobian_finite_differences(orc, fsds):
    X, U = _points(20, 1)
    for x, u in zip(X, U):
        # stay inside one table segment so the piecewise-linear kappa is smooth
        i = np.searchsorted(fsds.s_ref, x[0], side="right") - 1
        x = x.copy(); x[0] = 0.5 * (fsds.s_ref[i] + fsds.s_ref[i + 1])
        _, J = orc.fkin6(x, u)
        z = np.concatenate([x, u])
        for j in range(10):
            e = 1e-6 * max(1.0, abs(z[j]))
            if j == 0:
                e = min(e, 0.2 * (fsds.s_ref[i + 1] - fsds.s_ref[i]))
            zp, zm = z.copy(), z.copy(); zp[j] += e; zm[j] -= e
            fd = (orc.fkin6(zp[:8], zp[8:], jac=False) - orc.fkin6(zm[:8], zm[8:], jac=False)) / (2 * e)
            np.testing.assert_allclose(J[:, j], fd, rtol=2e-6, atol=2e-6 * max(1.0, np.abs(fd).max()))


def test_kappa_lookup_extrapolates_linearly(orc, fsds):
    s_ref, k_ref = fsds.s_ref, fsds.kappa_ref
    x = np.array([s_ref[-1] + 3.0, 0.5, 0.0, 5.0, 0.0, 0.0, 0.0, 0.0]); u = np.zeros(2)
    f = orc.fkin6(x, u, jac=False)
    kap = k_ref[-2] + (k_ref[-1] - k_ref[-2]) / (s_ref[-1] - s_ref[-2]) * (x[0] - s_ref[-2])
    assert abs(f[0] - 5.0 / (1 + kap * 0.5)) < 1e-12
    x[0] = s_ref[0] - 2.0
    f = orc.fkin6(x, u, jac=False)
    kap = k_ref[0] + (k_ref[1] - k_ref[0]) / (s_ref[1] - s_ref[0]) * (x[0] - s_ref[0])
    assert abs(f[0] - 5.0 / (1 + kap * 0.5)) < 1e-12


def _slow_lag_oracle(fsds):
    """Actuator lags slowed to 0.5 s so that explicit RK4 and a scipy reference are meaningful."""
    p = _abi.default_params(); p[_abi.P_TT] = 0.5; p[_abi.P_TD] = 0.5
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref))
    return Oracle(cfg, fsds.s_ref, fsds.kappa_ref, params=p[None]), p


def test_integrators_against_scipy(built, fsds):
    o, p = _slow_lag_oracle(fsds)
    X, U = _points(5, 2)
    for x, u in zip(X, U):
        sol = solve_ivp(lambda t, y: fkin6_numpy(y, u, fsds.s_ref, fsds.kappa_ref, p), (0, 0.05), x, rtol=1e-12,
                        atol=1e-12, method="DOP853")
        ref = sol.y[:, -1]
        xe, _ = o.integrate(0, x, u, 0.05)
        xi, _ = o.integrate(1, x, u, 0.05)
        # kappa' jumps at the table knots limit the observable order; both must still be accurate
        assert np.max(np.abs(xe - ref) / np.maximum(1, np.abs(ref))) < 2e-5
        assert np.max(np.abs(xi - ref) / np.maximum(1, np.abs(ref))) < 2e-6


def test_erk4_is_fourth_order_and_irk_is_more_accurate(built, fsds):
    o, p = _slow_lag_oracle(fsds)
    # constant-curvature stretch: put the state where kappa is locally linear (inside a segment)
    x = np.array([100.3, 0.2, 0.05, 8.0, 0.1, 0.2, 100.0, 0.05]); u = np.array([150.0, 0.06])
    h = 0.02
    def rollout(kind, h, n):
        y = x.copy()
        for _ in range(n):
            y = o.integrate(kind, y, u, h, sens=False) if kind == 2 else o.integrate(kind, y, u, h)[0]
        return y
    ref = rollout(1, h / 8, 8)
    e1 = np.abs(rollout(0, h, 1) - ref).max(); e2 = np.abs(rollout(0, h / 2, 2) - ref).max()
    assert 8 < e1 / e2 < 40, (e1, e2)  # ~16 for order 4 (local 5)
    assert np.abs(rollout(1, h, 1) - ref).max() < 1e-3 * e1 + 1e-12


@pytest.mark.parametrize("kind", [0, 1])
def test_integrator_sensitivities_finite_differences(built, fsds, kind):
    o, _ = _slow_lag_oracle(fsds)
    x = np.array([100.3, 0.2, 0.05, 8.0, 0.1, 0.2, 100.0, 0.05]); u = np.array([150.0, 0.06])
    _, S = o.integrate(kind, x, u, 0.05)
    z = np.concatenate([x, u])
    for j in range(10):
        e = 1e-6 * max(1.0, abs(z[j])) if j else 1e-4
        zp, zm = z.copy(), z.copy(); zp[j] += e; zm[j] -= e
        fd = (o.integrate(kind, zp[:8], zp[8:], 0.05)[0] - o.integrate(kind, zm[:8], zm[8:], 0.05)[0]) / (2 * e)
        np.testing.assert_allclose(S[:, j], fd, rtol=1e-6, atol=1e-7)


def test_irk_stiff_lag_matches_pade(orc):
    """T' = (u_T - T)/t_T with h/t_T = 50: Gauss-Legendre-4 gives the (4,4) Pade value of exp(-50)."""
    z = -50.0
    R = (1 + z / 2 + 3 * z**2 / 28 + z**3 / 84 + z**4 / 1680) / (1 - z / 2 + 3 * z**2 / 28 - z**3 / 84 + z**4 / 1680)
    x = np.array([10.0, 0, 0, 5.0, 0, 0, 100.0, 0.0]); u = np.array([300.0, 0.0])
    xn, S = orc.integrate(1, x, u, 0.05)
    assert abs(xn[6] - (R * 100.0 + (1 - R) * 300.0)) < 1e-9
    assert abs(S[6, 6] - R) < 1e-12 and abs(S[6, 8] - (1 - R)) < 1e-12


def test_golden_model_vectors(orc):
    g = np.load(os.path.join(GOLD, "model.npz"))
    for i in range(len(g["x"])):
        f, J = orc.fkin6(g["x"][i], g["u"][i])
        np.testing.assert_allclose(f, g["f"][i], rtol=1e-13, atol=1e-13)
        np.testing.assert_allclose(J, g["J"][i], rtol=1e-13, atol=1e-13)
        np.testing.assert_allclose(orc.fdyn6(g["x"][i], g["u"][i]), g["fdyn6"][i], rtol=1e-12, atol=1e-12)
        xe, Se = orc.integrate(0, g["x"][i], g["u"][i], 0.05)
        np.testing.assert_allclose(xe, g["erk4_x"][i], rtol=1e-12, atol=1e-9)
        xi, Si = orc.integrate(1, g["x"][i], g["u"][i], 0.05)
        np.testing.assert_allclose(xi, g["irk_x"][i], rtol=1e-12, atol=1e-11)
        np.testing.assert_allclose(Si, g["irk_S"][i], rtol=1e-10, atol=1e-11)


def test_cartesian_dyn6_satisfies_the_reference_implicit_form():
    """Independent pin of the oracle's Cartesian dyn6: a numpy transcription of the implicit residual of
    python/models.py:310-452, evaluated at the oracle's state derivative, must vanish; kin6 (:168-229) explicitly."""
    import ctypes as C
    from ihm2_b200 import _abi
    from ihm2_b200.track import load_packaged_track
    from oracle.lib import Oracle, _p, _f
    mp = load_packaged_track("fsds_default")
    B = 64
    rng = np.random.default_rng(8)
    x = np.stack([rng.uniform(-50, 50, B), rng.uniform(-50, 50, B), rng.uniform(-3, 3, B), rng.uniform(0.5, 25, B),
                  rng.uniform(-1, 1, B), rng.uniform(-1, 1, B), rng.uniform(-400, 400, B), rng.uniform(-0.4, 0.4, B)], 1)
    u = np.stack([rng.uniform(-500, 500, B), rng.uniform(-0.5, 0.5, B)], 1)
    P = _abi.default_params()
    m, g, lR, lF, tw, zcg = [P[i] for i in (_abi.P_M, _abi.P_G, _abi.P_LR, _abi.P_LF, _abi.P_TRACKW, _abi.P_ZCG)]
    wb, C_ = lR + lF, lR / (lR + lF)
    h = 1e-7
    for kind, model in ((3, _abi.SIM_KIN6C), (4, _abi.SIM_DYN6C)):
        cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), sim_model=model, sim_steps=1)
        orc = Oracle(cfg, mp.s_ref, mp.kappa_ref)
        for i in range(B):
            # state derivative from one tiny explicit-Euler-sized plant step: (x+ - x) / h with h -> 0 is not exact,
            # so ask the integrator for a very short step and compare to first order
            cfg.dt = h
            xn, _ = Oracle(cfg, mp.s_ref, mp.kappa_ref).sim_step(x[i:i + 1], u[i:i + 1])
            xd = (xn[0] - x[i]) / h
            X, Y, phi, vx, vy, r, T, dl = x[i]
            if model == _abi.SIM_KIN6C:
                td = np.tan(dl)
                beta = np.arctan(C_ * td)
                ddl = (u[i, 1] - dl) / P[_abi.P_TD]
                bd = C_ * (1 + td * td) / (1 + C_ * C_ * td * td) * ddl
                Fm = P[_abi.P_CM0] * T
                Fd = -(P[_abi.P_CR0] + P[_abi.P_CR1] * vx + P[_abi.P_CR2] * vx * vx) * np.tanh(10 * vx)
                vd = ((0.5 * Fm + Fd) * np.cos(beta) + 0.5 * Fm * np.cos(dl - beta)) / m
                ref = np.array([vx * np.cos(phi) - vy * np.sin(phi), vx * np.sin(phi) + vy * np.cos(phi), r,
                                vd * np.cos(beta) - bd * vy, vd * np.sin(beta) + bd * vx, (vd * np.sin(beta) + bd * vx) / lR])
                assert np.abs(xd[:6] - ref).max() < 2e-4 * max(1.0, np.abs(ref).max()), (i, xd[:6], ref)
                continue
            ax, ay = xd[3] - vy * r, xd[4] + vx * r
            Fdown = 0.5 * P[_abi.P_CDOWN] * vx * vx
            stat = 0.5 * m * g * lF / wb
            lon, lat = 0.5 * m * ax * zcg / wb, 0.5 * m * ay * zcg / tw
            Fz = dict(FL=-(stat - lon + lat + 0.25 * Fdown), FR=-(stat - lon - lat + 0.25 * Fdown),
                      RL=-(stat + lon + lat + 0.25 * Fdown), RR=-(stat + lon - lat + 0.25 * Fdown))
            vxw = dict(FL=vx - 0.5 * tw * r, FR=vx + 0.5 * tw * r, RL=vx - 0.5 * tw * r, RR=vx + 0.5 * tw * r)
            vyw = dict(FL=vy + lF * r, FR=vy + lF * r, RL=vy - lR * r, RR=vy - lR * r)
            al = {k: np.arctan2(vyw[k], vxw[k]) - (dl if k[0] == "F" else 0.0) for k in Fz}
            Ba, Ca, Da, Ea = [P[j] for j in (_abi.P_BA, _abi.P_CA, _abi.P_DA, _abi.P_EA)]
            Fy = {k: Fz[k] * Da * np.sin(Ca * np.arctan(Ba * al[k] - Ea * (Ba * al[k] - np.arctan(Ba * al[k])))) for k in Fz}
            Fdrag = -(P[_abi.P_CR0] + P[_abi.P_CR1] * vx + P[_abi.P_CR2] * vx * vx) * np.tanh(1000 * vx)
            td = np.tan(dl)
            sb = C_ * td / np.sqrt(1 + C_ * C_ * td * td)
            dtau = P[_abi.P_KTV] * (vx * sb / lR - r)
            dn = -m * g - 0.25 * Fdown
            Fx = {k: P[_abi.P_CM0] * (T + (dtau if k[1] == "R" else -dtau)) * Fz[k] / dn for k in Fz}
            cd, sd = np.cos(dl), np.sin(dl)
            res = np.array([
                xd[0] - (vx * np.cos(phi) - vy * np.sin(phi)), xd[1] - (vx * np.sin(phi) + vy * np.cos(phi)), xd[2] - r,
                m * ax - ((Fx["FR"] + Fx["FL"]) * cd - (Fy["FR"] + Fy["FL"]) * sd + Fx["RR"] + Fx["RL"] + Fdrag),
                m * ay - ((Fx["FR"] + Fx["FL"]) * sd + (Fy["FR"] + Fy["FL"]) * cd + Fy["RR"] + Fy["RL"]),
                P[_abi.P_IZ] * xd[5] - ((Fx["FR"] * cd - Fy["FR"] * sd) * tw / 2 + (Fx["FR"] * sd + Fy["FR"] * cd) * lF
                                        - (Fx["FL"] * cd - Fy["FL"] * sd) * tw / 2 + (Fx["FL"] * sd + Fy["FL"] * cd) * lF
                                        + Fx["RR"] * tw / 2 - Fy["RR"] * lR - Fx["RL"] * tw / 2 - Fy["RL"] * lR)])
            scale = np.array([1, 1, 1, m * 10, m * 10, P[_abi.P_IZ] * 10])
            assert np.abs(res / scale).max() < 5e-4, (i, res)


def test_frenet_fdyn6_satisfies_the_reference_implicit_form():
    """Independent pin of the plant of BASELINE configs[0] and [2]: a numpy transcription of the implicit residual of
    `fdyn6_model` (python/models.py:455-606: wheel loads with load transfer and downforce :486-515, slip angles with
    smooth_abs_nonzero :518-540, the CROSSWISE slip-angle assignment :543-546, torque vectoring :548-560, Frenet
    kinematics :562-575), evaluated at the state derivative the oracle returns, must vanish.  The oracle solves the
    implicit form for xdot as an affine system in (a_x, a_y), so this is a different route to the same equations."""
    from ihm2_b200 import _abi
    from ihm2_b200.track import load_packaged_track
    from oracle.lib import Oracle
    mp = load_packaged_track("fsds_default")
    rng = np.random.default_rng(11)
    B = 200
    x = np.stack([rng.uniform(0, 380, B), rng.uniform(-1.5, 1.5, B), rng.uniform(-0.5, 0.5, B), rng.uniform(0.3, 25, B),
                  rng.uniform(-1.5, 1.5, B), rng.uniform(-1.2, 1.2, B), rng.uniform(-400, 400, B), rng.uniform(-0.45, 0.45, B)], 1)
    u = np.stack([rng.uniform(-500, 500, B), rng.uniform(-0.5, 0.5, B)], 1)
    sabs = lambda v: np.tanh(10 * v) * v + 1e-6 * np.exp(-v * v)      # utils.py:23-32 (smooth_sgn = tanh(10 x))
    for cross in (1.0, 0.0):
        P = _abi.default_params(); P[_abi.P_CROSS] = cross
        cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref))
        orc = Oracle(cfg, mp.s_ref, mp.kappa_ref, params=P[None, :].copy())
        m, g, lR, lF, tw, zcg, Iz = [P[i] for i in (_abi.P_M, _abi.P_G, _abi.P_LR, _abi.P_LF, _abi.P_TRACKW, _abi.P_ZCG, _abi.P_IZ)]
        wb = lR + lF
        Ba, Ca, Da, Ea = [P[j] for j in (_abi.P_BA, _abi.P_CA, _abi.P_DA, _abi.P_EA)]
        pac = lambda al: Da * np.sin(Ca * np.arctan(Ba * al - Ea * (Ba * al - np.arctan(Ba * al))))
        worst = 0.0
        for i in range(B):
            xd = orc.fdyn6(x[i], u[i])
            s, n, psi, vx, vy, r, T, dl = x[i]
            ax, ay = xd[3] - vy * r, xd[4] + vx * r
            Fdown = 0.5 * P[_abi.P_CDOWN] * vx * vx
            stat = 0.5 * m * g * lF / wb
            lon, lat = 0.5 * m * ax * zcg / wb, 0.5 * m * ay * zcg / tw
            Fz = dict(FL=-(stat - lon + lat + 0.25 * Fdown), FR=-(stat - lon - lat + 0.25 * Fdown),
                      RL=-(stat + lon + lat + 0.25 * Fdown), RR=-(stat + lon - lat + 0.25 * Fdown))
            cd, sd = np.cos(dl), np.sin(dl)
            vxF = dict(FL=vx - 0.5 * tw * r, FR=vx + 0.5 * tw * r); vyF = vy + lF * r
            vlon = dict(FL=cd * vxF["FL"] + sd * vyF, FR=cd * vxF["FR"] + sd * vyF, RL=vx - 0.5 * tw * r, RR=vx + 0.5 * tw * r)
            vlat = dict(FL=-sd * vxF["FL"] + cd * vyF, FR=-sd * vxF["FR"] + cd * vyF, RL=vy - lR * r, RR=vy - lR * r)
            al = {k: np.arctan2(vlat[k], sabs(vlon[k])) for k in Fz}
            src = dict(FL="RR", FR="RL", RL="FR", RR="FL") if cross else dict(FL="FL", FR="FR", RL="RL", RR="RR")
            Fy = {k: Fz[k] * pac(al[src[k]]) for k in Fz}
            Fdrag = -(P[_abi.P_CR0] + P[_abi.P_CR1] * vx + P[_abi.P_CR2] * vx * vx) * np.tanh(10 * vx)
            beta = np.arctan(lR / wb * np.tan(dl))
            dtau = P[_abi.P_KTV] * (np.hypot(vx, vy) * np.sin(beta) / lR - r)
            dn = -m * g - 0.25 * Fdown
            Fx = {k: P[_abi.P_CM0] * (T + (dtau if k[1] == "R" else -dtau)) * Fz[k] / dn for k in Fz}
            i0 = np.searchsorted(mp.s_ref, s, side="right") - 1
            kap = mp.kappa_ref[i0] + (mp.kappa_ref[i0 + 1] - mp.kappa_ref[i0]) / (mp.s_ref[i0 + 1] - mp.s_ref[i0]) * (s - mp.s_ref[i0])
            sdot = (vx * np.cos(psi) - vy * np.sin(psi)) / (1 + kap * n)
            res = np.array([
                xd[0] - sdot, xd[1] - (vx * np.sin(psi) + vy * np.cos(psi)), xd[2] - (r - kap * sdot),
                m * ax - ((Fx["FR"] + Fx["FL"]) * cd - (Fy["FR"] + Fy["FL"]) * sd + Fx["RR"] + Fx["RL"] + Fdrag),
                m * ay - ((Fx["FR"] + Fx["FL"]) * sd + (Fy["FR"] + Fy["FL"]) * cd + Fy["RR"] + Fy["RL"]),
                Iz * xd[5] - ((Fx["FR"] * cd - Fy["FR"] * sd) * tw / 2 + (Fx["FR"] * sd + Fy["FR"] * cd) * lF
                              - (Fx["FL"] * cd - Fy["FL"] * sd) * tw / 2 + (Fx["FL"] * sd + Fy["FL"] * cd) * lF
                              + Fx["RR"] * tw / 2 - Fy["RR"] * lR - Fx["RL"] * tw / 2 - Fy["RL"] * lR),
                xd[6] - (u[i, 0] - T) / P[_abi.P_TT], xd[7] - (u[i, 1] - dl) / P[_abi.P_TD]])
            scale = np.array([1, 1, 1, m * 10, m * 10, Iz * 10, 1 + abs(xd[6]), 1 + abs(xd[7])])
            worst = max(worst, np.abs(res / scale).max())
        assert worst < 1e-10, (cross, worst)


def test_fdyn10_satisfies_the_reference_implicit_form():
    """Independent pin of the dyn10 plant: numpy transcription of the implicit residual of `fdyn10_model`
    (python/models.py:609-801: slip ratios and longitudinal Pacejka :722-732, straight slip angles :717-720, wheel-speed
    rows :781-784, four torque lags) at the oracle's state derivative."""
    from ihm2_b200 import _abi
    from ihm2_b200.track import load_packaged_track
    from oracle.lib import Oracle
    mp = load_packaged_track("fsds_default")
    P = _abi.default_params()
    cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref))
    orc = Oracle(cfg, mp.s_ref, mp.kappa_ref)
    m, g, lR, lF, tw, zcg, Iz = [P[i] for i in (_abi.P_M, _abi.P_G, _abi.P_LR, _abi.P_LF, _abi.P_TRACKW, _abi.P_ZCG, _abi.P_IZ)]
    Rw, Iw, kd, ks = [P[i] for i in (_abi.P_RW, _abi.P_IW, _abi.P_KD, _abi.P_KS)]
    wb = lR + lF
    Ba, Ca, Da, Ea = [P[j] for j in (_abi.P_BA, _abi.P_CA, _abi.P_DA, _abi.P_EA)]
    Bs, Cs, Ds, Es = [P[j] for j in (_abi.P_BS, _abi.P_CS, _abi.P_DS, _abi.P_ES)]
    # constants.py:85-95 by value (the parameter vector must carry the reference's numbers)
    assert abs(Ds - (1.12e-4 * 0.5 * 230 * 9.81 * 0.7853 / 1.5706 + 1.57)) < 1e-12 and Cs == 1.86 and (Rw, Iw, kd, ks) == (0.20809, 0.3, 0.17, 15.0)
    latp = lambda al: Da * np.sin(Ca * np.arctan(Ba * al - Ea * (Ba * al - np.arctan(Ba * al))))
    lonp = lambda sl: Ds * np.sin(Cs * np.arctan(Bs * sl - Es * (Bs * sl - np.arctan(Bs * sl))))
    sabs = lambda v: np.tanh(10 * v) * v + 1e-6 * np.exp(-v * v)
    rng = np.random.default_rng(12)
    worst = 0.0
    for _ in range(200):
        v = rng.uniform(0.5, 25)
        x = np.concatenate([[rng.uniform(0, 380), rng.uniform(-1.5, 1.5), rng.uniform(-0.5, 0.5), v, rng.uniform(-1.5, 1.5), rng.uniform(-1.2, 1.2)],
                            v / Rw * rng.uniform(0.9, 1.15, 4), rng.uniform(-100, 100, 4), [rng.uniform(-0.45, 0.45)]])
        u = np.concatenate([rng.uniform(-125, 125, 4), [rng.uniform(-0.5, 0.5)]])
        xd = orc.fdyn10(x, u)
        s, n, psi, vx, vy, r = x[:6]; om, tau, dl = x[6:10], x[10:14], x[14]
        ax, ay = xd[3] - vy * r, xd[4] + vx * r
        Fdrag = -(P[_abi.P_CR0] + P[_abi.P_CR1] * vx + P[_abi.P_CR2] * vx * vx) * np.tanh(1000 * vx)
        Fdown = 0.5 * P[_abi.P_CDOWN] * vx * vx
        stat = 0.5 * m * g * lF / wb
        lon, lat = 0.5 * m * ax * zcg / wb, 0.5 * m * ay * zcg / tw
        Fz = np.array([-(stat - lon + lat + 0.25 * Fdown), -(stat - lon - lat + 0.25 * Fdown),
                       -(stat + lon + lat + 0.25 * Fdown), -(stat + lon - lat + 0.25 * Fdown)])
        cd, sd = np.cos(dl), np.sin(dl)
        vxFL, vxFR, vyF = vx - 0.5 * tw * r, vx + 0.5 * tw * r, vy + lF * r
        vlon = np.array([cd * vxFL + sd * vyF, cd * vxFR + sd * vyF, vx - 0.5 * tw * r, vx + 0.5 * tw * r])
        vlat = np.array([-sd * vxFL + cd * vyF, -sd * vxFR + cd * vyF, vy - lR * r, vy - lR * r])
        va = sabs(vlon)
        Flat = Fz * latp(np.arctan2(vlat, va))
        Flon = -Fz * lonp(om * Rw / va - 1.0)
        i0 = np.searchsorted(mp.s_ref, s, side="right") - 1
        kap = mp.kappa_ref[i0] + (mp.kappa_ref[i0 + 1] - mp.kappa_ref[i0]) / (mp.s_ref[i0 + 1] - mp.s_ref[i0]) * (s - mp.s_ref[i0])
        sdot = (vx * np.cos(psi) - vy * np.sin(psi)) / (1 + kap * n)
        FL, FR, RL, RR = 0, 1, 2, 3
        res = np.concatenate([[
            xd[0] - sdot, xd[1] - (vx * np.sin(psi) + vy * np.cos(psi)), xd[2] - (r - kap * sdot),
            m * ax - (Fdrag + cd * (Flon[FL] + Flon[FR]) - sd * (Flat[FL] + Flat[FR]) + Flon[RL] + Flon[RR]),
            m * ay - (sd * (Flon[FL] + Flon[FR]) + cd * (Flat[FL] + Flat[FR]) + Flat[RL] + Flat[RR]),
            Iz * xd[5] - ((Flon[FR] * cd - Flat[FR] * sd) * tw / 2 + (Flon[FR] * sd + Flat[FR] * cd) * lF
                          - (Flon[FL] * cd - Flat[FL] * sd) * tw / 2 + (Flon[FL] * sd + Flat[FL] * cd) * lF
                          + Flon[RR] * tw / 2 - Flat[RR] * lR - Flon[RL] * tw / 2 - Flat[RL] * lR)],
            Iw * xd[6:10] - (tau - (kd * om + ks + Rw * Flon)),
            xd[10:14] - (u[:4] - tau) / P[_abi.P_TT], [xd[14] - (u[4] - dl) / P[_abi.P_TD]]])
        scale = np.concatenate([[1, 1, 1, m * 10, m * 10, Iz * 10], np.full(4, 100.0), 1 + np.abs(xd[10:15])])
        worst = max(worst, np.abs(res / scale).max())
    assert worst < 1e-10, worst
