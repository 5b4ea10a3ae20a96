"""Pins the oracle's DPC rollout / gradient with independent checks: a torch fp64 transcription of
dpc/main.py:56-97 and :374-409 differentiated by autograd, and central finite differences."""
import numpy as np
import torch

from oracle import lib as olib

M, WHEELBASE, CM0, CR0, CR1, CR2 = 230.0, 1.5706, 4.950, 297.030, 16.665, 0.6784  # dpc/main.py:23-29
DT, NF = 1 / 20, 40  # dpc/main.py:31-34


def _f(x, u):  # dpc/main.py:56-84
    phi, v, T, delta = x[..., 2], x[..., 3], u[..., 0], u[..., 1]
    beta = 0.5 * delta
    v_x = v * torch.cos(beta)
    l_R = 0.5 * WHEELBASE
    return torch.stack([
        v * torch.cos(phi + beta),
        v * torch.sin(phi + beta),
        v * torch.sin(beta) / l_R,
        (CM0 * T - (CR0 + CR1 * v_x + CR2 * v_x**2) * torch.tanh(10 * v_x)) / M,
    ], -1)


def _step(x, u):  # dpc/main.py:87-97
    k1 = _f(x, u)
    k2 = _f(x + DT / 2 * k1, u)
    k3 = _f(x + DT / 2 * k2, u)
    k4 = _f(x + DT * k3, u)
    return x + DT / 6 * (k1 + 2 * k2 + 2 * k3 + k4)


def _cost(x0, u, ref, w):  # dpc/main.py:374-409
    x, J, xs = x0, 0.0, [x0]
    for i in range(NF + 1):
        if i > 0:
            q = w[6:10] if i == NF else w[0:4]
            cp, sp = torch.cos(ref[:, i, 2]), torch.sin(ref[:, i, 2])
            dX, dY = x[:, 0] - ref[:, i, 0], x[:, 1] - ref[:, i, 1]
            e_lon, e_lat = cp * dX + sp * dY, -sp * dX + cp * dY
            J = J + q[0] * e_lon**2 + q[1] * e_lat**2 + q[2] * (x[:, 2] - ref[:, i, 2])**2 \
                + q[3] * (x[:, 3] - ref[:, i, 3])**2
        if i < NF:
            J = J + w[4] * u[:, i, 0]**2 + w[5] * u[:, i, 1]**2
            x = _step(x, u[:, i])
            xs.append(x)
    return J, torch.stack(xs, 1)


def dpc_inputs(B, seed=3):
    """Rollouts starting near a reference that moves at ~5 m/s along a curving path."""
    g = np.random.default_rng(seed)
    t = DT * np.arange(NF + 1)
    v_ref = g.uniform(3.0, 12.0, (B, 1))
    kappa = g.uniform(-0.15, 0.15, (B, 1))
    phi0 = g.uniform(-np.pi, np.pi, (B, 1))
    phi_ref = phi0 + kappa * v_ref * t
    safe = np.where(np.abs(kappa) < 1e-3, 1e-3, kappa)
    X_ref = (np.sin(phi_ref) - np.sin(phi0)) / safe
    Y_ref = -(np.cos(phi_ref) - np.cos(phi0)) / safe
    ref = np.stack([X_ref, Y_ref, phi_ref, np.broadcast_to(v_ref, phi_ref.shape)], -1)
    x0 = ref[:, 0] + g.uniform(-1, 1, (B, 4)) * np.array([0.5, 0.5, 0.2, 2.0])
    x0[:, 3] = np.abs(x0[:, 3])
    x0[: B // 8, 3] = g.uniform(-0.05, 0.05, B // 8)  # around the tanh(10 v_x) switch
    u = np.stack([g.uniform(-500, 500, (B, NF)), g.uniform(-0.5, 0.5, (B, NF))], -1)  # dpc/main.py:36-37
    w = np.array([1.0, 2.0, 3.0, 0.5, 1e-5, 2.0, 4.0, 5.0, 6.0, 0.7])
    return x0, u, ref, w


def test_dpc_oracle_matches_autograd():
    x0, u, ref, w = dpc_inputs(64)
    cost, grad, xtraj = olib.dpc_rollout_grad(x0, u, ref, w)
    ut = torch.tensor(u, requires_grad=True)
    J, xs = _cost(torch.tensor(x0), ut, torch.tensor(ref), torch.tensor(w))
    J.sum().backward()
    assert np.abs(xs.detach().numpy() - xtraj).max() < 1e-11
    assert np.abs(J.detach().numpy() - cost).max() / np.abs(cost).max() < 1e-13
    g = ut.grad.numpy()
    assert np.abs(g - grad).max() / np.abs(g).max() < 1e-12


def test_dpc_oracle_gradient_finite_differences():
    x0, u, ref, w = dpc_inputs(4, seed=5)
    _, grad, _ = olib.dpc_rollout_grad(x0, u, ref, w)
    for (k, j, h) in [(0, 0, 1e-2), (0, 1, 1e-5), (17, 0, 1e-2), (17, 1, 1e-5), (39, 0, 1e-2), (39, 1, 1e-5)]:
        up, um = u.copy(), u.copy()
        up[:, k, j] += h
        um[:, k, j] -= h
        fd = (olib.dpc_rollout_grad(x0, up, ref, w)[0] - olib.dpc_rollout_grad(x0, um, ref, w)[0]) / (2 * h)
        assert np.abs(fd - grad[:, k, j]).max() <= 1e-5 * max(1.0, np.abs(grad[:, k, j]).max())


def test_dpc_oracle_zero_horizon_terms():
    """Stage 0 carries no tracking term (dpc/main.py:375) and the terminal stage uses the _f weights (:391-409)."""
    x0, u, ref, w = dpc_inputs(8, seed=7)
    c0 = olib.dpc_rollout_grad(x0, u, ref, w)[0]
    ref2 = ref.copy()
    ref2[:, 0] += 100.0
    assert np.array_equal(c0, olib.dpc_rollout_grad(x0, u, ref2, w)[0])
    w2 = w.copy()
    w2[6:] = 0.0
    ref3 = ref.copy()
    ref3[:, NF] += 100.0
    a = olib.dpc_rollout_grad(x0, u, ref, w2)[0]
    assert np.allclose(a, olib.dpc_rollout_grad(x0, u, ref3, w2)[0], rtol=1e-15, atol=0)


def test_dpc_motion_planner_samples_in_time():
    """MotionPlanner.plan (dpc/main.py:996-1011): references v_ref * dt apart along the centre line, heading within
    pi of the car's."""
    from ihm2_b200 import dpc

    mp = dpc.MotionPlanner("fsds_competition_1", v_ref=5.0)
    s0 = np.linspace(0.0, mp.lap_length, 50, endpoint=False)
    phi = np.interp(s0, mp.s_ref, mp.phi_ref) + 0.3
    ref = mp.plan(s0, phi)
    assert ref.shape == (50, NF + 1, 4)
    step = np.hypot(np.diff(ref[..., 0], axis=1), np.diff(ref[..., 1], axis=1))
    assert np.all(np.abs(step / (5.0 * dpc.dt) - 1.0) < 0.03)  # chords of an arc-length table, so only nearly equal
    assert np.all(np.abs(ref[..., 2] - phi[:, None]) <= np.pi)
