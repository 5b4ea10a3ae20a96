"""Closed loop with the Cartesian plants of the ROS stack, for a batch of cars.

What the two ROS nodes of the reference do per control period, with every per-car operation a batched kernel:
  sim node      src/ihm2/src/sim_node.cpp:145-151 (controls clipped, steering rate-limited per message),
                :197-260 (kin6 / dyn6 by speed, no-reverse clamp)              -> `ihm2b_sim_step_dev`, Cartesian models
  control node  src/ihm2/src/mpc_control_node.cpp:142-157 (projection, Frenet state), :109-124 shift, :177-206 solve
                                                                                -> `ihm2b_project_dev`, `ihm2b_solve_dev`
The host loop below only sequences launches and keeps the per-car bookkeeping (lap count, failure) on the device.
The sim node times laps with a start-line intersection (:262-273); here a lap is counted when the projected arc
length wraps around the lap length, which is the same event up to the position of the line.
"""
from __future__ import annotations

import numpy as np
import torch

from .engine import Engine
from .projection import Track


def closed_loop_cartesian(eng: Engine, track: Track, state: torch.Tensor, n_steps: int, T_max: float = 500.0,
                          delta_max: float = 0.5, s_tol: float = 2.0, track_index: torch.Tensor | None = None,
                          log: bool = False, controller=None):
    """state [B,8] = (X, Y, phi, v_x, v_y, r, T, delta) on the device, updated in place.

    `controller(frenet [B,3], state [B,8]) -> u [B,2]`, when given, replaces the NMPC (e.g. a batched
    `StanleyController`, the reference's fallback of new_python/controller.py:162-234); `eng` then only steps the plant.

    Returns dict(state, laps [B] int, lap_step [B] step index of the first completed lap or -1, fail [B] step of the
    first controller failure / plant NaN / lost projection or -1, s [B] last projected arc length, and with log=True
    poses [n_steps+1, B, 8] and controls [n_steps, B, 2])."""
    B, dev = eng.B, state.device
    assert state.shape == (B, 8) and state.dtype == torch.float64
    dt = eng.cfg.dt
    L = torch.as_tensor(track.lap_length, dtype=torch.float64, device=dev)
    L = L[track_index.long()] if track_index is not None else L[0].expand(B)
    s_guess = torch.zeros(B, dtype=torch.float64, device=dev)
    u_prev = torch.zeros(B, 2, dtype=torch.float64, device=dev)
    laps = torch.zeros(B, dtype=torch.int32, device=dev)
    lap_step = torch.full((B,), -1, dtype=torch.int32, device=dev)
    fail = torch.full((B,), -1, dtype=torch.int32, device=dev)
    s_last = torch.zeros(B, dtype=torch.float64, device=dev)
    # sim_node.cpp:147 allows 0.01 * 2 delta_max / 1.0 per message, i.e. a full steering sweep per second at the node's
    # 100 Hz; one message per control period dt here
    ddelta_max = dt * 2 * delta_max / 1.0
    poses = [state.clone()] if log else None
    controls = [] if log else None
    for step in range(n_steps):
        alive = fail < 0
        frenet, flag = track.project(state[:, :3].contiguous(), s_guess, s_tol, track_index)
        s = frenet[:, 0]
        wrapped = alive & (s < s_last - 0.5 * L)                              # crossed the start of the table
        if controller is not None:
            laps += wrapped.int()
            lap_step = torch.where(wrapped & (lap_step < 0), torch.full_like(lap_step, step), lap_step)
        elif bool(wrapped.any()):
            # the warm start lives one lap ahead of the new measurement: bring it back
            x_it, u_it = eng.get_iterate()
            x_it[wrapped, :, 0] -= L[wrapped, None]
            eng.set_iterate(x_it, None)
            laps += wrapped.int()
            lap_step = torch.where(wrapped & (lap_step < 0), torch.full_like(lap_step, step), lap_step)
        s_last = torch.where(alive, s, s_last)
        s_guess = torch.where(alive, torch.fmod(s + state[:, 3] * dt, L), s_guess)   # mpc_control_node.cpp:145
        if controller is not None:
            u0 = controller(frenet, state)
            ok = torch.isfinite(u0).all(1)
        else:
            x0 = torch.cat([frenet, state[:, 3:]], 1).contiguous()
            eng.shift_iterate()
            u0, status = eng.solve(x0)
            ok = (status == 0) | (status == 2)                                # mpc_control_node.cpp:190-197
        u = u_prev.clone()
        u[:, 0] = torch.clamp(u0[:, 0], -T_max, T_max)                        # sim_node.cpp:148-149
        u[:, 1] = torch.clamp(torch.maximum(torch.minimum(u0[:, 1], u_prev[:, 1] + ddelta_max), u_prev[:, 1] - ddelta_max),
                              -delta_max, delta_max)
        u = torch.where((alive & ok)[:, None], u, u_prev)
        xn, _ = eng.sim_step(state, u)
        lost = (flag != 0) if controller is None else torch.zeros_like(ok)   # the exception is the MPC node's (:154-156)
        bad = alive & (~ok | lost | torch.isnan(xn).any(1))
        fail = torch.where(bad, torch.full_like(fail, step), fail)
        move = (alive & ~bad)[:, None]
        state.copy_(torch.where(move, xn, state))
        u_prev = torch.where(move, u, u_prev)
        if log:
            poses.append(state.clone())
            controls.append(u_prev.clone())
    out = dict(state=state, laps=laps, lap_step=lap_step, fail=fail, s=s_last)
    if log:
        out["poses"], out["controls"] = torch.stack(poses), torch.stack(controls)
    return out
