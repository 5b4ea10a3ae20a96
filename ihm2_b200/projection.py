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
