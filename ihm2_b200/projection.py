"""Cartesian pose -> Frenet state for batches of cars.

Mirrors `Track` of src/ihm2/src/common/tracks.cpp (constructor :132-181 from the track CSV columns, `project`
:200-288, `length`) and the conversion the control node applies to each state message
(src/ihm2/src/mpc_control_node.cpp:142-157): `s` from the projection, `n` = distance to the projected point signed by
tangent x offset, `psi` = wrap_to_pi(phi - phi_ref).  The work is done by `ihm2b_project_dev`."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import engine
from .track import MotionPlan, load_packaged_track


class Track:
    def __init__(self, tracks: MotionPlan | str | list, device=None):
        if not torch.cuda.is_available():
            raise engine.Ihm2bError("ihm2_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = engine.load_library()
        plans = tracks if isinstance(tracks, (list, tuple)) else [tracks]
        plans = [load_packaged_track(t) if isinstance(t, str) else t for t in plans]
        n = {len(p.s_ref) for p in plans}
        assert len(n) == 1, "all tracks of one Track object share the table length"
        self.n_grid = n.pop()
        dev = torch.device("cuda" if device is None else device)
        tab = lambda f: torch.tensor(np.stack([getattr(p, f) for p in plans]), dtype=torch.float64, device=dev)  # noqa: E731
        self.s_ref, self.X_ref, self.Y_ref, self.phi_ref = tab("s_ref"), tab("X_ref"), tab("Y_ref"), tab("phi_ref")
        self.lap_length = np.array([p.lap_length for p in plans])

    def length(self, track: int = 0) -> float:
        return float(self.lap_length[track])

    def project(self, pose: torch.Tensor, s_guess: torch.Tensor, s_tol: float = 2.0, track_index: torch.Tensor | None = None,
                want_ref: bool = False):
        """pose [B,3] = (X, Y, phi), s_guess [B] (fp64, CUDA) -> frenet [B,3] = (s, n, psi), flag [B] (int32: 1 where
        the control node would throw "tangent x normal is too small") and, on request, (X_ref, Y_ref, phi_ref) [B,3]."""
        B = pose.shape[0]
        assert pose.shape == (B, 3) and s_guess.shape == (B,) and pose.dtype == s_guess.dtype == torch.float64
        pose, s_guess = pose.contiguous(), s_guess.contiguous()
        if track_index is not None:
            assert track_index.shape == (B,) and track_index.dtype == torch.int32
        frenet = torch.empty(B, 3, dtype=torch.float64, device=pose.device)
        flag = torch.empty(B, dtype=torch.int32, device=pose.device)
        ref = torch.empty(B, 3, dtype=torch.float64, device=pose.device) if want_ref else None
        rc = self.lib.ihm2b_project_dev(C.c_int64(B), C.c_int32(self.n_grid), engine._dptr(self.s_ref), engine._dptr(self.X_ref),
                                        engine._dptr(self.Y_ref), engine._dptr(self.phi_ref), engine._dptr(track_index),
                                        engine._dptr(pose), engine._dptr(s_guess), C.c_double(s_tol), engine._dptr(frenet),
                                        engine._dptr(ref), engine._dptr(flag), engine._stream())
        if rc != 0:
            raise engine.Ihm2bError(f"libihm2b error {rc}: {self.lib.ihm2b_last_error().decode()}")
        return (frenet, flag, ref) if want_ref else (frenet, flag)

    def frenet_to_cartesian(self, s: torch.Tensor, n: torch.Tensor, track_index: torch.Tensor | None = None):
        """`Track::frenet_to_cartesian` (tracks.cpp:290-306) for a batch: (s, n) [B] -> (X, Y) [B]."""
        t = torch.zeros_like(s, dtype=torch.long) if track_index is None else track_index.long()
        return frenet_to_cartesian(self.s_ref[t], self.X_ref[t], self.Y_ref[t], self.phi_ref[t], s, n)


def _interp_rows(s_ref: torch.Tensor, tab: torch.Tensor, s: torch.Tensor, ind: torch.Tensor) -> torch.Tensor:
    """`Track::interp` (tracks.cpp:192-199): value of the linear spline of segment `ind` at s, per row."""
    s0, s1 = s_ref.gather(1, ind[:, None])[:, 0], s_ref.gather(1, ind[:, None] + 1)[:, 0]
    v0, v1 = tab.gather(1, ind[:, None])[:, 0], tab.gather(1, ind[:, None] + 1)[:, 0]
    return v0 + (v1 - v0) / (s1 - s0) * (s - s0)


def frenet_to_cartesian(s_ref: torch.Tensor, X_ref: torch.Tensor, Y_ref: torch.Tensor, phi_ref: torch.Tensor,
                        s: torch.Tensor, n: torch.Tensor):
    """`Track::frenet_to_cartesian` (src/ihm2/src/common/tracks.cpp:290-306): the reference point at arc length s on the
    linear splines of the track CSV (`locate_index`: upper_bound - 1, tracks.cpp:183-185), displaced by n along the
    normal (-sin phi_ref, cos phi_ref).  Tables [B, n_grid] (one row per car, or [n_grid] shared), s, n [B]; any device."""
    if s_ref.dim() == 1:
        s_ref, X_ref, Y_ref, phi_ref = (t[None, :].expand(s.shape[0], -1) for t in (s_ref, X_ref, Y_ref, phi_ref))
    ind = (torch.searchsorted(s_ref.contiguous(), s[:, None].contiguous(), right=True)[:, 0] - 1).clamp(0, s_ref.shape[1] - 2)
    Xr, Yr, pr = (_interp_rows(s_ref, t, s, ind) for t in (X_ref, Y_ref, phi_ref))
    return Xr - n * torch.sin(pr), Yr + n * torch.cos(pr)


class StartLineTimer:
    """Lap timing of the ROS sim node (src/ihm2/src/sim_node.cpp:46-51, 261-274) for a batch of cars: a lap ends when
    the segment last position -> position crosses the start line (two cones) while v_x > 0; the first crossing starts
    the clock, later ones record last / best lap time.  `update` is called once per plant step with the step's time."""

    def __init__(self, start_line_pos_1, start_line_pos_2, batch: int, device=None):
        dev = torch.device("cpu" if device is None else device)
        p1 = torch.as_tensor(start_line_pos_1, dtype=torch.float64, device=dev)
        p2 = torch.as_tensor(start_line_pos_2, dtype=torch.float64, device=dev)
        self.p1 = p1.expand(batch, 2) if p1.dim() == 1 else p1
        self.p2 = p2.expand(batch, 2) if p2.dim() == 1 else p2
        self.last_pos = None
        self.last_stamp = torch.full((batch,), float("nan"), dtype=torch.float64, device=dev)   # NaN: no crossing yet
        self.last_lap_time = torch.full((batch,), float("nan"), dtype=torch.float64, device=dev)
        self.best_lap_time = torch.full((batch,), float("nan"), dtype=torch.float64, device=dev)
        self.laps = torch.zeros(batch, dtype=torch.int32, device=dev)

    @staticmethod
    def _ccw(a, b, c):      # sim_node.cpp:46-48
        return (c[:, 1] - a[:, 1]) * (b[:, 0] - a[:, 0]) > (b[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0])

    @classmethod
    def _intersect(cls, a, b, c, d):   # sim_node.cpp:49-51
        return (cls._ccw(a, c, d) != cls._ccw(b, c, d)) & (cls._ccw(a, b, c) != cls._ccw(a, b, d))

    def update(self, pos: torch.Tensor, v_x: torch.Tensor, t: float):
        """pos [B,2], v_x [B], t: time stamp of this plant step (s).  Returns the mask of cars that completed a lap."""
        if self.last_pos is None:
            self.last_pos = pos.clone()
            return torch.zeros_like(v_x, dtype=torch.bool)
        crossed = (v_x > 0.0) & self._intersect(self.p1, self.p2, self.last_pos, pos)
        timed = crossed & ~torch.isnan(self.last_stamp)
        lap = t - self.last_stamp
        self.last_lap_time = torch.where(timed, lap, self.last_lap_time)
        better = timed & (torch.isnan(self.best_lap_time) | (lap < self.best_lap_time))
        self.best_lap_time = torch.where(better, lap, self.best_lap_time)
        self.laps += timed.int()
        self.last_stamp = torch.where(crossed, torch.full_like(self.last_stamp, t), self.last_stamp)
        self.last_pos = pos.clone()
        return timed
