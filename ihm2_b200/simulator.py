"""Batched mirror of the reference's simulator surface (new_python/simulator.py): `Simulator` ABC,
`SingleModelSimulator`, `MultiModelSimulator`, `SimulatorConfig`, `SimulatorModel`.

The reference's classes wrap `AcadosSimSolver`s (IRK, GAUSS_RADAU_IIA, 4 stages, 100 steps; the
intended options are in the comments at simulator.py:65-70 and live in python/main.py:395-400).
Here the step runs in libihm2b.so (`ihm2b_sim_step_dev`) on Frenet states `x [B, 8] =
(s, n, psi, v_x, v_y, r, T, delta)` -- the coordinates python/main.py simulates in.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from enum import Enum, auto

import numpy as np
import torch

from . import _abi
from .engine import Engine

__all__ = ["SimulatorConfig", "SimulatorModel", "Simulator", "SingleModelSimulator", "MultiModelSimulator",
           "dyn10_state_from_controller_state", "controller_state_from_dyn10_state"]


@dataclass
class SimulatorConfig:  # simulator.py:25-29 (the attribute name keeps the reference's spelling)
    sampling_time: float = 1 / 20
    integrator_type: str = "IRK"
    colloaction_type: str = "GAUSS_RADAU_IIA"
    num_steps: int = 100


class SimulatorModel(Enum):  # simulator.py:32-36
    KIN6 = auto()
    DYN6 = auto()
    KIN6_DYN6 = auto()
    DYN10 = auto()


_MODEL = {SimulatorModel.KIN6: _abi.SIM_FKIN6, SimulatorModel.DYN6: _abi.SIM_FDYN6, SimulatorModel.KIN6_DYN6: _abi.SIM_KIN6_DYN6,
          SimulatorModel.DYN10: _abi.SIM_FDYN6}   # the engine's sim_model is unused by the dyn10 step (its own entry point)


def dyn10_state_from_controller_state(x8, wheel_radius: float = 0.20809):
    """(s, n, psi, v_x, v_y, r, T, delta) -> the 15 states of `fdyn10_model` (python/models.py:615-629): wheels rolling
    without slip (omega = v_x / R_w), the torque spread evenly (tau_ij = T / 4, as python/main.py:490-500 spreads u_T)."""
    x8 = torch.as_tensor(x8, dtype=torch.float64)
    om = (x8[:, 3] / wheel_radius)[:, None].expand(-1, 4)
    return torch.cat([x8[:, :6], om, (0.25 * x8[:, 6])[:, None].expand(-1, 4), x8[:, 7:8]], 1).contiguous()


def controller_state_from_dyn10_state(x10):
    """The controller's 8 states of a dyn10 plant state: T = sum of the wheel torques."""
    return torch.cat([x10[:, :6], x10[:, 10:14].sum(1, keepdim=True), x10[:, 14:15]], 1).contiguous()


class Simulator(ABC):  # simulator.py:39-45
    config: SimulatorConfig

    @abstractmethod
    def simulate(self, x, u):
        ...


class _EngineSimulator(Simulator):
    def __init__(self, model: SimulatorModel, s_ref, kappa_ref, batch: int = 1, config: SimulatorConfig | None = None,
                 params=None, track_index=None, switch_threshold: float = 3.0) -> None:
        self.config = SimulatorConfig() if config is None else config
        s_ref = np.atleast_2d(np.asarray(s_ref, dtype=np.float64))
        cfg = _abi.make_config(batch=batch, n_grid=s_ref.shape[1], n_tracks=s_ref.shape[0], dt=self.config.sampling_time,
                               sim_model=_MODEL[model], sim_steps=self.config.num_steps)
        cfg.sim_switch = switch_threshold
        self.model = model
        self.engine = Engine(cfg, s_ref, np.atleast_2d(np.asarray(kappa_ref, dtype=np.float64)), track_index=track_index,
                             params=params)

    def simulate(self, x, u):
        """x [B, 8], u [B, 2] -> x_next [B, 8] (NaN rows where the integrator failed, main.py:503-512).
        DYN10: x [B, 15]; u [B, 5] = (u_tau_FL .. u_tau_RR, u_delta), or the controller's [B, 2] = (u_T, u_delta), which is
        spread as python/main.py:490-500 does (0.25 u_T to every wheel)."""
        as_np = isinstance(x, np.ndarray)
        xt = torch.as_tensor(np.atleast_2d(x) if as_np else x, dtype=torch.float64, device=self.engine.device).contiguous()
        ut = torch.as_tensor(np.atleast_2d(u) if as_np else u, dtype=torch.float64, device=self.engine.device).contiguous()
        if self.model == SimulatorModel.DYN10:
            if ut.shape[1] == 2:
                ut = torch.cat([(0.25 * ut[:, 0])[:, None].expand(-1, 4), ut[:, 1:2]], 1).contiguous()
            xn = self.engine.sim_step_dyn10(xt, ut)
        else:
            xn, self.last_model = self.engine.sim_step(xt, ut)
        return xn.cpu().numpy() if as_np else xn


class SingleModelSimulator(_EngineSimulator):  # simulator.py:48-50
    def __init__(self, model: SimulatorModel, *args, **kwargs) -> None:
        assert model in (SimulatorModel.KIN6, SimulatorModel.DYN6, SimulatorModel.DYN10)
        super().__init__(model, *args, **kwargs)


class MultiModelSimulator(_EngineSimulator):
    """simulator.py:53-62.  The reference switches with a user callable on the speed; the live rule
    (python/main.py:482-489) is v^2 sin(beta) / l_R <= 3 -> kinematic, else dynamic, evaluated per
    instance inside the kernel; `switch_threshold` is that constant."""

    def __init__(self, *args, **kwargs) -> None:
        super().__init__(SimulatorModel.KIN6_DYN6, *args, **kwargs)
