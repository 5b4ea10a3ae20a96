"""DPC rollout of dpc/main.py on the device (BASELINE.json configs[4]).

The reference fixes the model (kin4, dpc/main.py:56-84), its discretisation (one RK4 step of 1/20 s per stage,
:87-97), the loss (IHM1.5 tracking cost, :374-409) and how references are sampled in time along the centre line
(`MotionPlanner`, :903-1011); the policy itself is an empty class (:617-635).  This module mirrors those pieces:
`MotionPlanner.plan` for whole batches on the host, and `rollout_grad` / `DPCLoss` which call
`ihm2b_dpc_rollout_grad_dev` for the loss and its gradient in the controls, so that any torch policy
`u = policy(x0, ref)` trains through the rollout without autograd tracing 40 RK4 steps.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import astuple, dataclass

import numpy as np
import torch

from . import engine
from .track import MotionPlan, load_packaged_track

Nf, nx, nu, dt = 40, 4, 2, 1 / 20  # dpc/main.py:31-34
T_max, delta_max = 500.0, 0.5  # dpc/main.py:36-37
_L_R = 0.7853  # python/constants.py:52, used to undo the slip-angle shift of the packaged heading tables


@dataclass
class DPCWeights:  # keyword names of NMPCControllerIpopt.__init__, dpc/main.py:332-345
    q_lon: float = 1.0
    q_lat: float = 1.0
    q_phi: float = 1.0
    q_v: float = 1.0
    r_T: float = 1.0
    r_delta: float = 1.0
    q_lon_f: float = 1.0
    q_lat_f: float = 1.0
    q_phi_f: float = 1.0
    q_v_f: float = 1.0

    def as_array(self) -> np.ndarray:
        return np.array(astuple(self), dtype=np.float64)


def unwrap_to_pi(x: np.ndarray) -> np.ndarray:  # dpc/main.py:48-53
    diffs = np.diff(x)
    diffs[diffs > 1.5 * np.pi] -= 2 * np.pi
    diffs[diffs < -1.5 * np.pi] += 2 * np.pi
    return np.insert(x[0] + np.cumsum(diffs), 0, x[0])


def teds_projection(x: np.ndarray, a) -> np.ndarray:  # dpc/main.py:43-45
    return np.mod(x - a, 2 * np.pi) + a


class MotionPlanner:
    """dpc/main.py:903-953 from a single-lap table (s, X, Y, tangent heading), `plan` (:996-1011) for a batch.

    `project` (:955-994) belongs to the Cartesian closed loop; here the arc length s0 is an input."""

    def __init__(self, track: MotionPlan | str, v_ref: float = 5.0):
        if isinstance(track, str):
            track = load_packaged_track(track, tripled=False)
            # the packaged tables hold phi - asin(l_R kappa) (python/motion_planning.py:380); dpc uses the tangent
            heading = track.phi_ref + np.arcsin(_L_R * track.kappa_ref)
        else:
            heading = track.phi_ref
        s, X, Y = track.s_ref, track.X_ref, track.Y_ref
        chord = float(np.hypot(X[-1] - X[0], Y[-1] - Y[0]))
        lap_length = float(s[-1] + chord)
        t_diff = np.append(np.diff(s), chord) / v_ref
        t_extra = np.insert(np.cumsum(t_diff), 0, 0.0)
        lap_time = float(t_extra[-1])
        t = t_extra[:-1]
        self.lap_length, self.lap_time, self.v_ref = lap_length, lap_time, v_ref
        self.s_ref = np.concatenate((s - lap_length, s, s + lap_length))
        self.t_ref = np.concatenate((t - lap_time, t, t + lap_time))
        self.X_ref = np.concatenate((X, X, X))
        self.Y_ref = np.concatenate((Y, Y, Y))
        self.phi_ref = unwrap_to_pi(np.concatenate((heading, heading, heading)))

    def plan(self, s0: np.ndarray, phi: np.ndarray) -> np.ndarray:
        """s0 [B], phi [B] -> ref [B, Nf+1, 4] = (X_ref, Y_ref, phi_ref, v_ref)."""
        s0 = np.atleast_1d(np.asarray(s0, dtype=np.float64))
        phi = np.atleast_1d(np.asarray(phi, dtype=np.float64))
        t0 = np.interp(s0, self.s_ref, self.t_ref)
        t = dt * np.arange(Nf + 1)[None, :] + t0[:, None]
        s = np.interp(t, self.t_ref, self.s_ref)
        ref = np.empty(s.shape + (4,))
        ref[..., 0] = np.interp(s, self.s_ref, self.X_ref)
        ref[..., 1] = np.interp(s, self.s_ref, self.Y_ref)
        ref[..., 2] = teds_projection(np.interp(s, self.s_ref, self.phi_ref), phi[:, None] - np.pi)
        ref[..., 3] = self.v_ref
        return ref


def rollout_grad(x0: torch.Tensor, u: torch.Tensor, ref: torch.Tensor, weights: DPCWeights | np.ndarray | None = None,
                 step: float = dt, want_trajectory: bool = False):
    """x0 [B,4], u [B,Nf,2], ref [B,Nf+1,4] (contiguous CUDA tensors, all fp64 like the reference or all fp32) ->
    cost [B], grad [B,Nf,2] (, x [B,Nf+1,4]) in the same dtype."""
    if not torch.cuda.is_available():
        raise engine.Ihm2bError("ihm2_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    lib = engine.load_library()
    B, n = u.shape[0], u.shape[1]
    dtype = u.dtype
    if dtype not in (torch.float64, torch.float32):
        raise ValueError(f"fp64 or fp32 tensors expected, got {dtype}")
    for t, shape in ((x0, (B, nx)), (u, (B, n, nu)), (ref, (B, n + 1, nx))):
        if tuple(t.shape) != shape or t.dtype != dtype:
            raise ValueError(f"expected {dtype} tensor of shape {shape}, got {t.dtype} {tuple(t.shape)}")
    w = weights.as_array() if isinstance(weights, DPCWeights) else np.ascontiguousarray(
        DPCWeights().as_array() if weights is None else weights, dtype=np.float64)
    assert w.shape == (10,)
    cost = torch.empty(B, dtype=dtype, device=u.device)
    grad = torch.empty_like(u)
    xtraj = torch.empty((B, n + 1, nx), dtype=dtype, device=u.device) if want_trajectory else None
    entry = lib.ihm2b_dpc_rollout_grad_dev if dtype == torch.float64 else lib.ihm2b_dpc_rollout_grad_f32_dev
    rc = entry(
        C.c_int64(B), C.c_int32(n), C.c_double(step), engine._dptr(x0), engine._dptr(u), engine._dptr(ref),
        w.ctypes.data_as(C.c_void_p), engine._dptr(cost), engine._dptr(grad), engine._dptr(xtraj), engine._stream())
    if rc != 0:
        raise engine.Ihm2bError(f"libihm2b error {rc}: {lib.ihm2b_last_error().decode()}")
    return (cost, grad, xtraj) if want_trajectory else (cost, grad)


class DPCLoss(torch.autograd.Function):
    """loss[B] = rollout cost of the controls `u`; the backward pass hands the kernel's d loss / d u to autograd."""

    @staticmethod
    def forward(ctx, x0, u, ref, weights=None):
        cost, grad = rollout_grad(x0.contiguous(), u.detach().contiguous(), ref.contiguous(), weights)
        ctx.save_for_backward(grad)
        return cost

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return None, grad * g[:, None, None], None, None
