"""Optimality certificates for the oracle's QP / RTI step: the QP is strictly convex, so a point
that satisfies the KKT conditions IS the solution the reference's HPIPM would return."""
import os

import numpy as np
import pytest

from ihm2_b200 import _abi
from oracle.lib import Oracle, initial_iterate

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
N, NX, NU, NZ = 40, 8, 2, 10


@pytest.fixture(scope="module")
def qp(built, fsds):
    g = np.load(os.path.join(GOLD, "rti.npz"))
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref))
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref)
    from ihm2_b200.workloads import shift_np
    xs, us = shift_np(g["x_in"][0], g["u_in"][0])
    return o.qp_debug(g["x0"][0], xs, us), g, cfg


def test_cost_matrices_match_reference_definition(qp):
    q, g, cfg = qp
    # python/mpc.py:52-58 selectors and controller.py:320-366 weights, scaled by dt on the path
    Vx = np.zeros((12, 8)); Vx[:8] = np.eye(8); Vx[10:, 6:] = np.eye(2)
    Vu = np.zeros((12, 2)); Vu[8:10] = np.eye(2); Vu[10:] = -np.eye(2)
    V = np.hstack([Vu, Vx])
    W = np.diag([1, 1, 1, 1, 1, 1, 1, 100, 1, 100, 0, 500.0])
    np.testing.assert_allclose(q["H"][3], cfg.dt * V.T @ W @ V, atol=1e-14)
    np.testing.assert_allclose(q["H"][N][2:, 2:], np.diag([1000, 100, 100, 1, 1, 1, 1, 100.0]), atol=1e-14)
    s0 = g["x0"][0][0]
    assert abs(q["g"][5][2] + cfg.dt * 1.0 * (s0 + 40.0 * 5 / N)) < 1e-12
    assert abs(q["g"][N][2] + 1000.0 * (s0 + 40.0)) < 1e-9


def test_constraint_rows_match_reference_definition(qp):
    q, _, _ = qp
    assert list(q["nrows"]) == [4] + [8] * (N - 1) + [4]
    C = q["C"][7]
    assert C[2, 0] == 1 and C[2, 2 + 6] == -1 and C[3, 1] == 1 and C[3, 2 + 7] == -1   # mpc.py:91-97
    np.testing.assert_allclose(q["ub"][7, :8], [500, 0.5, 1000, 0.02, 2, 31, 500, 0.5])
    np.testing.assert_allclose(q["lb"][7, :8], [-500, -0.5, -1000, -0.02, -2, 0, -500, -0.5])
    # terminal idxbx_e = [1,3,4,5] with the n / v_x / T / delta limits (literal quirk, mpc.py:81-83)
    idx = [int(np.argmax(q["C"][N, r])) - 2 for r in range(4)]
    assert idx == [1, 3, 4, 5]
    np.testing.assert_allclose(q["ub"][N, :4], [2, 31, 500, 0.5])
    np.testing.assert_allclose(q["lb"][N, :4], [-2, -31, -500, -0.5])


def test_kkt_certificate(qp):
    q, g, _ = qp
    assert q["status"] == 0
    z, pi = q["z"], q["pi"]
    np.testing.assert_allclose(z[0, 2:], g["x0"][0], atol=0)
    scale_g = 1 + np.abs(q["g"]).max()
    for k in range(N + 1):
        r = q["H"][k] @ z[k] + q["g"][k]
        if k < N:
            r[:2] += q["B"][k].T @ pi[k + 1]
            r[2:] += q["A"][k].T @ pi[k + 1]
            np.testing.assert_allclose(q["A"][k] @ z[k, 2:] + q["B"][k] @ z[k, :2] + q["c"][k], z[k + 1, 2:],
                                       atol=1e-9 * (1 + np.abs(q["c"]).max()))
        r[2:] -= pi[k]
        for i in range(q["nrows"][k]):
            r -= q["C"][k, i] * (q["lamL"][k, i] - q["lamU"][k, i])
            v = q["C"][k, i] @ z[k]
            assert q["lb"][k, i] - 1e-9 <= v <= q["ub"][k, i] + 1e-9
            assert q["lamL"][k, i] >= 0 and q["lamU"][k, i] >= 0
            assert q["lamL"][k, i] * (v - q["lb"][k, i]) < 1e-7 and q["lamU"][k, i] * (q["ub"][k, i] - v) < 1e-7
        sl = slice(0, 2) if k == 0 else (slice(0, NZ) if k < N else slice(2, NZ))
        assert np.abs(r[sl]).max() < 1e-8 * scale_g


def test_solution_matches_dense_active_set_solve(qp):
    """Re-solve the QP as one dense equality-constrained problem on the active set the IPM found."""
    q, g, _ = qp
    nv = (N + 1) * NZ
    H = np.zeros((nv, nv)); gv = np.zeros(nv)
    rows, rhs = [], []
    for k in range(N + 1):
        sl = slice(k * NZ, (k + 1) * NZ)
        H[sl, sl] = q["H"][k]; gv[sl] = q["g"][k]
        if k < N:
            for i in range(NX):
                r = np.zeros(nv); r[k * NZ + 2: k * NZ + NZ] = q["A"][k][i]; r[k * NZ: k * NZ + 2] = q["B"][k][i]
                r[(k + 1) * NZ + 2 + i] = -1.0
                rows.append(r); rhs.append(-q["c"][k][i])
        for i in range(q["nrows"][k]):
            v = q["C"][k, i] @ q["z"][k]
            for bound, lam in ((q["lb"][k, i], q["lamL"][k, i]), (q["ub"][k, i], q["lamU"][k, i])):
                if abs(v - bound) < 1e-6 and lam > 1e-6:
                    r = np.zeros(nv); r[sl] = q["C"][k, i]; rows.append(r); rhs.append(bound)
    for i in range(NX):
        r = np.zeros(nv); r[2 + i] = 1.0; rows.append(r); rhs.append(g["x0"][0][i])
    r = np.zeros(nv); r[N * NZ] = 1; rows.append(r); rhs.append(0.0)      # unused u_N
    r = np.zeros(nv); r[N * NZ + 1] = 1; rows.append(r); rhs.append(0.0)
    E = np.array(rows); m = len(rows)
    KKT = np.block([[H, E.T], [E, np.zeros((m, m))]])
    sol = np.linalg.lstsq(KKT, np.concatenate([-gv, rhs]), rcond=None)[0][:nv].reshape(N + 1, NZ)
    err = np.abs(sol - q["z"]).max(axis=0) / np.maximum(1, np.abs(q["z"]).max(axis=0))
    assert err.max() < 1e-6, err


def test_golden_rti_vectors(built, fsds):
    g = np.load(os.path.join(GOLD, "rti.npz"))
    cfg = _abi.make_config(batch=len(g["x0"]), n_grid=len(fsds.s_ref))
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref)
    u0, st, stats, xo, uo = o.solve(g["x0"], g["x_in"], g["u_in"], do_shift=True)
    assert (st == g["status"]).all() and (st == 0).all()
    np.testing.assert_allclose(u0, g["u0"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(xo, g["x_out"], rtol=1e-9, atol=1e-8)
    np.testing.assert_allclose(uo, g["u_out"], rtol=1e-9, atol=1e-7)


def test_shift_is_the_literal_warm_start(built, fsds):
    from ihm2_b200.workloads import shift_np
    cfg = _abi.make_config(batch=1, n_grid=len(fsds.s_ref))
    x, u = initial_iterate(cfg)
    x += np.arange(41)[None, :, None] * 0.01; u += np.arange(40)[None, :, None] * 0.1
    xs, us = shift_np(x, u)
    # controller.py:388-393: x_j <- x_{j+1} (j < N-1), x_{N-1} <- x_N, x_N kept, u_{N-1} <- 0
    np.testing.assert_array_equal(xs[0, :39], x[0, 1:40]); np.testing.assert_array_equal(xs[0, 39], x[0, 40])
    np.testing.assert_array_equal(xs[0, 40], x[0, 40]); np.testing.assert_array_equal(us[0, :39], u[0, 1:])
    assert (us[0, 39] == 0).all()


def test_closed_loop_golden_and_lap_progress(built, fsds):
    g = np.load(os.path.join(GOLD, "closed_loop.npz"))
    cfg = _abi.make_config(batch=2, n_grid=len(fsds.s_ref))
    o = Oracle(cfg, fsds.s_ref, fsds.kappa_ref)
    x, u = initial_iterate(cfg)
    res = o.closed_loop(g["x_start"], x, u, [fsds.lap_length], int(g["n_steps"]))
    np.testing.assert_allclose(res["xlog"], g["xlog"], rtol=1e-8, atol=1e-7)
    # with the reference's hard state bounds (no slacks) the QP becomes infeasible once the fdyn6
    # plant departs from the kinematic prediction: the instance freezes at that step (main.py:503-512)
    np.testing.assert_array_equal(res["fail"], g["fail"]); np.testing.assert_array_equal(res["done"], g["done"])
    assert (res["fail"] >= 30).all()
    assert (np.diff(res["xlog"][:, :, 0], axis=0) > -1e-9).all()          # s never decreases
    assert res["xlog"][30, 0, 3] > 5.0                                   # the car is up to speed


def test_oracle_sqp_merit_backtracking_steps():
    """SQP with MERIT_BACKTRACKING (new_python/controller.py:301-308): accepted step lengths come from
    {0.7^j} down to the floor 0.05, the multiplier-weighted L1 merit makes some of these inputs backtrack, and a
    converged instance (status 0) satisfies the KKT tolerances."""
    import os
    from ihm2_b200 import _abi
    from ihm2_b200.track import load_packaged_track
    from oracle.lib import Oracle
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "rti.npz"))
    mp = load_packaged_track(str(d["track"]))
    allowed = np.array([0.7**j for j in range(9)] + [0.05])
    seen = []
    for iters in (1, 2, 10):
        cfg = _abi.make_config(batch=8, n_grid=len(mp.s_ref), nlp_solver=_abi.SQP, nlp_max_iter=iters,
                               globalization=_abi.MERIT_BACKTRACKING)
        _, status, stats, x, u = Oracle(cfg, mp.s_ref, mp.kappa_ref).solve(d["x0"], d["x_in"].copy(), d["u_in"].copy())
        assert set(np.unique(status)) <= {0, 2}
        run = status == 2
        assert np.all(np.abs(stats[run, 2][:, None] - allowed[None, :]).min(axis=1) < 1e-12)
        assert np.all(stats[status == 0, 3:7] <= cfg.nlp_tol)
        assert np.all(np.isfinite(x)) and np.all(np.isfinite(u))
        seen += list(stats[run, 2])
    assert min(seen) < 1.0 and max(seen) == 1.0
