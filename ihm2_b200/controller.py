"""Batched mirror of the reference's controller surface (new_python/controller.py).

Same class, method and keyword names; arguments gain a leading batch dimension.  The NMPC
(`IHM2Controller`) runs entirely in libihm2b.so through `engine.Engine`; the Stanley baseline is
elementwise torch on the device.  PODS types are plain dataclasses (strongpods is not a
dependency here); `ModelBounds` follows python/mpc.py:18-26.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass, field

import numpy as np
import torch

from . import _abi
from .engine import Engine

__all__ = ["CarGeometry", "DrivetrainParams", "AerodynamicParams", "ActuatorParams", "CarParams", "ModelBounds",
           "default_car_params", "Controller", "StanleyController", "IHM2Controller"]


@dataclass
class CarGeometry:  # controller.py:40-48
    length: float = 3.19
    width: float = 1.55
    wheelbase: float = 1.5706
    track: float = 1.24
    cog_to_rear_axle: float = 0.7853
    cog_to_front_axle: float = 0.7853
    cog_height: float = 0.295


@dataclass
class DrivetrainParams:  # controller.py:51-56
    C_m0: float = 4.950
    C_r0: float = 297.030
    C_r1: float = 16.665
    C_r2: float = 0.6784


@dataclass
class AerodynamicParams:  # controller.py:59-61
    downforce_coeff: float = 3.96864


@dataclass
class ActuatorParams:  # controller.py:64-70
    wheel_torque_max: float = 125.0
    total_torque_max: float = 500.0
    steering_max: float = 0.5
    steering_rate_max: float = 1.0
    steering_time_constant: float = 0.02
    torque_time_constant: float = 1e-3


@dataclass
class CarParams:  # controller.py:73-81 (tyre model kept as the Pacejka scalars the plant uses)
    mass: float = 230.0
    yaw_inertia: float = 137.583
    geometry: CarGeometry = field(default_factory=CarGeometry)
    drivetrain_params: DrivetrainParams = field(default_factory=DrivetrainParams)
    aerodynamic_params: AerodynamicParams = field(default_factory=AerodynamicParams)
    actuator_params: ActuatorParams = field(default_factory=ActuatorParams)

    def to_vector(self) -> np.ndarray:
        """The per-instance parameter vector of include/ihm2b.h (enum ihm2b_param_index)."""
        p = _abi.default_params()
        p[_abi.P_M], p[_abi.P_IZ] = self.mass, self.yaw_inertia
        p[_abi.P_LR], p[_abi.P_LF] = self.geometry.cog_to_rear_axle, self.geometry.cog_to_front_axle
        p[_abi.P_ZCG], p[_abi.P_TRACKW] = self.geometry.cog_height, self.geometry.track
        d = self.drivetrain_params
        p[_abi.P_CM0], p[_abi.P_CR0], p[_abi.P_CR1], p[_abi.P_CR2] = d.C_m0, d.C_r0, d.C_r1, d.C_r2
        p[_abi.P_CDOWN] = self.aerodynamic_params.downforce_coeff
        p[_abi.P_TT], p[_abi.P_TD] = self.actuator_params.torque_time_constant, self.actuator_params.steering_time_constant
        return p


def default_car_params() -> CarParams:
    return CarParams()


@dataclass
class ModelBounds:  # python/mpc.py:18-26
    n_max: float = 2.0
    v_x_min: float = 0.0
    v_x_max: float = 31.0
    T_max: float = 500.0
    delta_max: float = 0.5
    T_dot_max: float = 1e6
    delta_dot_max: float = 1.0
    a_lat_max: float = 5.0


class Controller(ABC):  # controller.py:134-159
    id: str = "controller"

    @dataclass
    class Config:
        horizon_size: int = 40
        sampling_time: float = 1 / 20

    def __init__(self, car_params: CarParams | None = None) -> None:
        self.car_params = default_car_params() if car_params is None else car_params
        self.config = self.Config()

    @abstractmethod
    def compute_control(self, *args, **kwargs):
        ...

    @property
    def sampling_time(self) -> float:
        return self.config.sampling_time


class StanleyController(Controller):
    """controller.py:162-234, batched.  The reference keeps the integrator history in a class-level
    deque shared by all instances (controller.py:184); here it is a per-object ring buffer [400, B]."""
    id = "stanley"

    @dataclass
    class Config(Controller.Config):
        k_P: float = 90.0
        k_I: float = 20.0
        k_offset: float = 1.0
        k_n: float = 5.5
        k_psi: float = 1.8
        k_kappa: float = 1.0

    def __init__(self, car_params: CarParams | None = None, config: "StanleyController.Config | None" = None,
                 T_max: float = 500.0, delta_max: float = 0.5) -> None:
        super().__init__(car_params)
        self.config = self.Config() if config is None else config
        self.T_max = min(T_max, self.car_params.actuator_params.total_torque_max)
        self.delta_max = min(delta_max, self.car_params.actuator_params.steering_max)
        self._eps: list[torch.Tensor] = []

    def compute_control(self, n, psi, v_x, v_x_ref, kappa_ref) -> torch.Tensor:
        n, psi, v_x, v_x_ref, kappa_ref = [torch.as_tensor(v, dtype=torch.float64) for v in (n, psi, v_x, v_x_ref, kappa_ref)]
        epsilon = v_x_ref - v_x
        u_T = self.config.k_P * epsilon
        if len(self._eps) > 1:  # trapezoid-like sum without the end points (controller.py:211-221)
            u_T = u_T + self.config.k_I * (torch.stack(self._eps).sum(0) - self._eps[0] - self._eps[-1]) * self.sampling_time
        self._eps.append(epsilon)
        self._eps = self._eps[-400:]
        l_R = self.car_params.geometry.cog_to_rear_axle
        u_delta = self.config.k_kappa * torch.atan(2 * torch.tan(torch.asin(kappa_ref * l_R)))
        u_delta = u_delta - self.config.k_psi * psi - torch.atan(self.config.k_n * n / (2.0 + v_x))
        return torch.stack([u_T.clamp(-self.T_max, self.T_max), u_delta.clamp(-self.delta_max, self.delta_max)], -1)


class IHM2Controller(Controller):
    """controller.py:237-405 with a leading batch dimension.

    `compute_control(x)` takes `x [B, 8]` (CUDA tensor or numpy) and returns `u [B, 2]` plus a
    per-instance status (the scalar call of the reference returns None on failure; with B = 1 and
    a numpy input this convention is kept).  `x_pred [B, Nf+1, 8]` / `u_pred [B, Nf, 2]` stay
    readable attributes.  `integrator`, `params`, `track_index` are additions (defaults reproduce
    the reference: IRK 4 stages, python/constants.py, one track).

    `nonlinear_constraints=True` adds the soft rows h(x, u) of the gen-1 OCP (track boundaries with the car footprint
    from `right_width` / `left_width` tables shaped like `s_ref`, T_dot, delta_dot, a_lat <= `a_lat_max`:
    old/generate_acaods_interface.py:312-325, 351-449) and makes the state bounds on T and delta soft.
    `hpipm_mode` picks an interior-point preset (controller.py:304 selects "SPEED_ABS"; None: this library's own);
    `qp_tolerate_maxiter=True` is acados' rule of carrying on with a QP that stopped at its iteration limit
    (mpc_control_node.cpp:190-197).
    """
    id = "ihm2"
    nx = 8
    nu = 2

    def __init__(self, s_ref, kappa_ref, Nf: int = 40, dt: float = 1 / 20, s_target: float = 40.0, n_max: float = 2.0,
                 v_x_max: float = 31.0, T_max: float = 500.0, delta_max: float = 0.5, T_dot_max: float = 1e6,
                 delta_dot_max: float = 1.0, a_lat_max: float = 5.0, q_s: float = 1.0, q_n: float = 1.0,
                 q_psi: float = 1.0, q_v_x: float = 1.0, q_v_y: float = 1.0, q_r: float = 1.0, q_T: float = 1.0,
                 q_delta: float = 100.0, q_s_f: float = 1000.0, q_n_f: float = 100.0, q_psi_f: float = 100.0,
                 q_v_x_f: float = 1.0, q_v_y_f: float = 1.0, q_r_f: float = 1.0, q_T_f: float = 1.0,
                 q_delta_f: float = 100.0, q_T_dot: float = 0.0, q_delta_dot: float = 500.0, *, batch: int = 1,
                 car_params: CarParams | None = None, params=None, track_index=None, lap_length=None,
                 integrator: str = "IRK", sim_model: int = _abi.SIM_KIN6_DYN6, nlp_solver_type: str = "SQP_RTI",
                 nlp_solver_max_iter: int = 2, globalization: str = "MERIT_BACKTRACKING",
                 nonlinear_constraints: bool = False, right_width=None, left_width=None, hpipm_mode: str | None = None,
                 qp_tolerate_maxiter: bool = False) -> None:
        super().__init__(car_params)
        self.Nf, self.dt, self.s_target = Nf, dt, s_target
        self.config = Controller.Config(horizon_size=Nf, sampling_time=dt)
        s_ref = np.atleast_2d(np.asarray(s_ref, dtype=np.float64))
        kappa_ref = np.atleast_2d(np.asarray(kappa_ref, dtype=np.float64))
        # controller.py:326-343 (stage W) and :350-363 (terminal W); a_lat_max is accepted and unused (:263)
        W = [q_s, q_n, q_psi, q_v_x, q_v_y, q_r, q_T, q_delta, q_T, q_delta, q_T_dot, q_delta_dot]
        W_e = [q_s_f, q_n_f, q_psi_f, q_v_x_f, q_v_y_f, q_r_f, q_T_f, q_delta_f]
        cfg = _abi.make_config(batch=batch, n_grid=s_ref.shape[1], n_tracks=s_ref.shape[0], Nf=Nf, dt=dt,
                               s_target=s_target, n_max=n_max, v_x_max=v_x_max, T_max=T_max, delta_max=delta_max,
                               T_dot_max=T_dot_max, delta_dot_max=delta_dot_max, a_lat_max=a_lat_max, W=W, W_e=W_e,
                               integrator=_abi.IRK_GL4 if integrator.upper().startswith("IRK") else _abi.ERK4,
                               sim_model=sim_model,
                               # controller.py:301-302, 308 hard-codes "SQP" with 2 iterations and MERIT_BACKTRACKING; the
                               # default here is the RTI step BASELINE.json's metric is quoted on.  "SQP" runs the loop
                               # with the KKT test and, as in the reference, the merit-function line search.
                               nlp_solver=_abi.SQP if nlp_solver_type.upper() == "SQP" else _abi.SQP_RTI,
                               nlp_max_iter=nlp_solver_max_iter,
                               globalization=_abi.MERIT_BACKTRACKING if globalization.upper() == "MERIT_BACKTRACKING"
                               else _abi.FIXED_STEP,
                               nonlinear_constraints=nonlinear_constraints, hpipm_mode=hpipm_mode,
                               qp_tolerate_maxiter=qp_tolerate_maxiter)
        widths = None
        if nonlinear_constraints:
            if right_width is None or left_width is None:
                raise ValueError("nonlinear_constraints needs the right_width / left_width tables of the track")
            widths = (np.atleast_2d(np.asarray(right_width, dtype=np.float64)), np.atleast_2d(np.asarray(left_width, dtype=np.float64)))
        if params is None:
            params = np.tile(self.car_params.to_vector(), (batch, 1))
        self.solver = Engine(cfg, s_ref, kappa_ref, track_index=track_index, params=params, lap_length=lap_length,
                             widths=widths)
        self.batch = batch

    @property
    def x_pred(self) -> torch.Tensor:
        return self.solver.get_iterate()[0]

    @property
    def u_pred(self) -> torch.Tensor:
        return self.solver.get_iterate()[1]

    def compute_control(self, x, return_status: bool = False):
        """controller.py:368-405: x0 embedding + yref, shift warm start, solve, read back u_0."""
        scalar = isinstance(x, np.ndarray) and x.ndim == 1
        if isinstance(x, np.ndarray):
            u0, status = self.solver.compute_control_host(np.atleast_2d(x), do_shift=True)
            if scalar:
                return u0[0] if status[0] in (_abi.STATUS_SUCCESS, _abi.STATUS_MAXITER) else None
            return (u0, status) if return_status else u0
        self.solver.shift_iterate()
        u0, status = self.solver.solve(x.contiguous())
        return (u0, status) if return_status else u0


class BatchedOcpSolver:
    """The `AcadosOcpSolver` verbs the reference uses on its solver object (new_python/controller.py:320-405,
    dpc/main.py:226-302) -- `set`, `cost_set`, `constraints_set`, `solve`, `get` -- on whole-batch tensors, for code that
    pokes `controller.solver` directly.  What acados lets the caller change per call and this engine fixes at creation
    (weights W, parameters p, yref -- it follows s_0 and s_target) is checked against the engine's configuration instead
    of being silently ignored.  Stage values are staged on the device and pushed with the next `solve()`."""

    def __init__(self, engine: Engine):
        self.engine = engine
        self.B, self.N = engine.B, engine.N
        self._x, self._u = engine.get_iterate()
        self._x0 = self._x[:, 0, :].clone()
        self._dirty = False
        self.status = None

    def _t(self, v, n):
        v = torch.as_tensor(v, dtype=torch.float64, device=self._x.device)
        return v.expand(self.B, n) if v.dim() == 1 else v

    def set(self, stage: int, field: str, value):
        if field == "x":
            self._x[:, stage, :] = self._t(value, _abi.NX); self._dirty = True
        elif field == "u":
            self._u[:, stage, :] = self._t(value, _abi.NU); self._dirty = True
        elif field in ("lbx", "ubx"):
            self.constraints_set(stage, field, value)
        elif field == "yref":
            # controller.py:373-385: yref_k = (s_0 + s_target k / N, 0, ...): the engine forms it from x_0 itself
            v = torch.as_tensor(value, dtype=torch.float64, device=self._x.device)
            v = v.expand(self.B, v.shape[-1]) if v.dim() == 1 else v
            want = self._x0[:, 0] + self.engine.cfg.s_target * stage / self.N
            if not torch.allclose(v[:, 0], want, atol=1e-9) or bool((v[:, 1:] != 0).any()):
                raise ValueError("yref differs from (s_0 + s_target k / N, 0, ...), the reference trajectory this engine is built for")
        elif field == "p":
            pass   # the track tables are given at construction (ihm2b_set_tracks); per-stage copies are redundant
        else:
            raise KeyError(f"BatchedOcpSolver.set: unsupported field {field!r}")

    def cost_set(self, stage: int, field: str, value):
        if field != "W":
            raise KeyError(f"BatchedOcpSolver.cost_set: unsupported field {field!r}")
        W = np.asarray(value.cpu() if torch.is_tensor(value) else value, dtype=np.float64)
        want = np.array(self.engine.cfg.W_e[:] if stage == self.N else self.engine.cfg.W[:])
        if W.shape != (len(want), len(want)) or not np.allclose(W, np.diag(want)):
            raise ValueError("W differs from the weights the engine was created with (ihm2b_config.W / W_e)")

    def constraints_set(self, stage: int, field: str, value):
        if stage != 0 or field not in ("lbx", "ubx"):
            raise KeyError("BatchedOcpSolver.constraints_set: only the x_0 embedding (stage 0, lbx / ubx) changes per call")
        self._x0 = self._t(value, _abi.NX).contiguous().clone()   # controller.py:370-371: lbx_0 = ubx_0 = measured state

    def solve(self):
        if self._dirty:
            self.engine.set_iterate(self._x, self._u); self._dirty = False
        _, self.status = self.engine.solve(self._x0)
        self._x, self._u = self.engine.get_iterate()
        return self.status

    def get(self, stage: int, field: str):
        if field == "x":
            return self._x[:, stage, :]
        if field == "u":
            return self._u[:, stage, :]
        raise KeyError(f"BatchedOcpSolver.get: unsupported field {field!r}")
