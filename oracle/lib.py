"""ctypes driver of oracle/_build/libihm2_oracle.so (test infrastructure only)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libihm2_oracle.so")
NX, NU, NZ, NP = 8, 2, 10, 28


def build_oracle(force: bool = False) -> str:
    src = os.path.join(_HERE, "ihm2_oracle.cpp")
    hdr = os.path.join(_HERE, "..", "include", "ihm2b.h")
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(f) and os.path.getmtime(f) > os.path.getmtime(_SO) for f in (src, hdr)
    )
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B", "_build/libihm2_oracle.so"], check=True,
                       capture_output=True)
    return _SO


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class _Tracks(C.Structure):   # struct ihm2o_tracks
    _fields_ = [("s_ref", C.c_void_p), ("kappa_ref", C.c_void_p), ("right_width", C.c_void_p), ("left_width", C.c_void_p)]


class Oracle:
    """Holds one problem set (config + tracks + per-instance params) like an ihm2b handle."""

    def __init__(self, cfg, s_ref, kappa_ref, track_index=None, params=None, widths=None):
        """widths = (right_width, left_width) tables shaped like s_ref (needed with nonlinear_constraints)."""
        from ihm2_b200._abi import default_params

        self.lib = C.CDLL(build_oracle())
        assert self.lib.ihm2o_np() == NP
        self.cfg = cfg
        self.s_ref = _f(np.atleast_2d(s_ref))
        self.kappa_ref = _f(np.atleast_2d(kappa_ref))
        assert self.s_ref.shape == (cfg.n_tracks, cfg.n_grid)
        B = cfg.batch
        self.track_index = np.ascontiguousarray(
            np.zeros(B, np.int32) if track_index is None else track_index, dtype=np.int32)
        self.params = _f(np.tile(default_params(), (B, 1)) if params is None else params)
        assert self.params.shape == (B, NP)
        self.right_width = self.left_width = None
        if widths is not None:
            self.right_width, self.left_width = (_f(np.atleast_2d(w)) for w in widths)
            assert self.right_width.shape == self.s_ref.shape and self.left_width.shape == self.s_ref.shape
        elif cfg.nonlinear_constraints:
            raise ValueError("nonlinear_constraints needs the track widths")
        self._tk = _Tracks(_p(self.s_ref), _p(self.kappa_ref), _p(self.right_width), _p(self.left_width))

    # ---- single-point model / integrator evaluations (track 0, instance 0 params) ----
    def fkin6(self, x, u, jac=True):
        f = np.zeros(NX)
        J = np.zeros((NX, NZ)) if jac else None
        self.lib.ihm2o_fkin6(_p(_f(x)), _p(_f(u)), _p(self.params[0]), _p(self.s_ref[0]),
                             _p(self.kappa_ref[0]), C.c_int(self.cfg.n_grid), _p(f), _p(J))
        return (f, J) if jac else f

    def fdyn6(self, x, u):
        f = np.zeros(NX)
        self.lib.ihm2o_fdyn6(_p(_f(x)), _p(_f(u)), _p(self.params[0]), _p(self.s_ref[0]),
                             _p(self.kappa_ref[0]), C.c_int(self.cfg.n_grid), _p(f))
        return f

    def integrate(self, kind, x, u, h, tableau=None, sens=True):
        """kind: 0 ERK4, 1 IRK (fkin6, analytic Newton), 2 IRK on fdyn6."""
        A = _f(tableau[0] if tableau else np.array(self.cfg.irk_A[:]))
        b = _f(tableau[1] if tableau else np.array(self.cfg.irk_b[:]))
        xn = np.zeros(NX)
        S = np.zeros((NX, NZ)) if (sens and kind != 2) else None
        rc = self.lib.ihm2o_integrate(C.c_int(kind), _p(A), _p(b), C.c_double(h), _p(_f(x)),
                                      _p(_f(u)), _p(self.params[0]), _p(self.s_ref[0]),
                                      _p(self.kappa_ref[0]), C.c_int(self.cfg.n_grid), _p(xn), _p(S))
        assert rc >= 0, f"integrator failed rc={rc}"
        return (xn, S) if S is not None else xn

    # ---- batched path ----
    def _common(self):
        return (C.byref(self.cfg), C.byref(self._tk), _p(self.track_index), _p(self.params))

    def linearize(self, x, u):
        B, N = self.cfg.batch, self.cfg.N
        x, u = _f(x), _f(u)
        A = np.zeros((B, N, NX, NX)); Bm = np.zeros((B, N, NX, NU)); phi = np.zeros((B, N, NX))
        bad = self.lib.ihm2o_linearize(*self._common(), _p(x), _p(u), _p(A), _p(Bm), _p(phi))
        return A, Bm, phi, bad

    def solve(self, x0, x, u, do_shift=False):
        """Returns u0 [B,2], status [B], stats [B,8]; x, u are updated copies."""
        B = self.cfg.batch
        x0 = _f(x0); x = _f(x).copy(); u = _f(u).copy()
        u0 = np.zeros((B, NU)); status = np.zeros(B, np.int32); stats = np.zeros((B, 8))
        self.lib.ihm2o_solve(*self._common(), _p(x0), _p(x), _p(u), _p(u0), _p(status), _p(stats),
                             C.c_int(int(do_shift)))
        return u0, status, stats, x, u

    def qp_debug(self, x0, x, u):
        """First QP of one problem and its IPM solution; rows padded to 16 per stage."""
        N, PR = self.cfg.N, 16
        out = dict(
            A=np.zeros((N, NX, NX)), B=np.zeros((N, NX, NU)), c=np.zeros((N, NX)),
            H=np.zeros((N + 1, NZ, NZ)), g=np.zeros((N + 1, NZ)), C=np.zeros((N + 1, PR, NZ)),
            e=np.zeros((N + 1, PR)), lb=np.zeros((N + 1, PR)), ub=np.zeros((N + 1, PR)),
            soft=np.zeros((N + 1, PR), np.int32), nrows=np.zeros(N + 1, np.int32), zZ=np.zeros((N + 1, 2)),
            z=np.zeros((N + 1, NZ)), pi=np.zeros((N + 1, NX)), lamL=np.zeros((N + 1, PR)),
            lamU=np.zeros((N + 1, PR)), tL=np.zeros((N + 1, PR)), tU=np.zeros((N + 1, PR)),
            sL=np.zeros((N + 1, PR)), sU=np.zeros((N + 1, PR)), lamSL=np.zeros((N + 1, PR)),
            lamSU=np.zeros((N + 1, PR)), info=np.zeros(7))
        order = ["A", "B", "c", "H", "g", "C", "e", "lb", "ub", "soft", "nrows", "zZ", "z", "pi", "lamL", "lamU", "tL",
                 "tU", "sL", "sU", "lamSL", "lamSU", "info"]
        st = self.lib.ihm2o_qp_debug(C.byref(self.cfg), C.byref(self._tk), _p(self.params), _p(_f(x0)), _p(_f(x)),
                                     _p(_f(u)), *[_p(out[k]) for k in order])
        out["status"] = st
        return out

    def con_h(self, x, u):
        h = np.zeros(5); J = np.zeros((5, NZ))
        self.lib.ihm2o_con_h(C.byref(self.cfg), C.byref(self._tk), _p(self.params), _p(_f(x)), _p(_f(u)), _p(h), _p(J))
        return h, J

    def fdyn10(self, x, u):
        f = np.zeros(15)
        self.lib.ihm2o_fdyn10(C.byref(self.cfg), C.byref(self._tk), _p(self.params), _p(_f(x)), _p(_f(u)), _p(f))
        return f

    def sim_step_dyn10(self, x, u):
        B = self.cfg.batch
        xn = np.zeros((B, 15))
        self.lib.ihm2o_sim_step_dyn10(*self._common(), _p(_f(x)), _p(_f(u)), _p(xn))
        return xn

    def sim_step(self, x, u):
        B = self.cfg.batch
        xn = np.zeros((B, NX)); model = np.zeros(B, np.int32)
        self.lib.ihm2o_sim_step(*self._common(), _p(_f(x)), _p(_f(u)), _p(xn), _p(model))
        return xn, model

    def closed_loop(self, xs, x, u, lap_length, n_steps, log=True):
        B = self.cfg.batch
        xs = _f(xs).copy(); x = _f(x).copy(); u = _f(u).copy()
        lap = _f(np.atleast_1d(lap_length))
        xlog = np.zeros((n_steps + 1, B, NX)) if log else None
        ulog = np.zeros((n_steps, B, NU)) if log else None
        done = np.zeros(B, np.int32); fail = np.zeros(B, np.int32)
        lap_time = np.zeros(B); cost_sum = np.zeros(B)
        self.lib.ihm2o_closed_loop(*self._common(), _p(lap), C.c_int(n_steps), _p(xs), _p(x), _p(u),
                                   _p(xlog), _p(ulog), _p(done), _p(fail), _p(lap_time), _p(cost_sum))
        return dict(x=xs, x_pred=x, u_pred=u, xlog=xlog, ulog=ulog, done=done, fail=fail, lap_time=lap_time,
                    cost_sum=cost_sum)


def initial_iterate(cfg, batch=None):
    """controller.py:313-317."""
    B = cfg.batch if batch is None else batch
    N = cfg.N
    x = np.zeros((B, N + 1, NX)); u = np.zeros((B, N, NU))
    x[:, :, 0] = -6.0 + np.arange(N + 1) * cfg.dt
    u[:, :, 0] = cfg.T_max
    return x, u


def dpc_rollout_grad(x0, u, ref, weights, dt=1.0 / 20):
    """DPC rollout (dpc/main.py:56-97) + IHM1.5 cost (dpc/main.py:374-409) and its gradient in the controls.

    x0 [B,4], u [B,Nf,2], ref [B,Nf+1,4], weights [10] -> cost [B], grad [B,Nf,2], xtraj [B,Nf+1,4]."""
    lib = C.CDLL(build_oracle())
    x0, u, ref, weights = _f(x0), _f(u), _f(ref), _f(weights)
    B, Nf = u.shape[0], u.shape[1]
    assert x0.shape == (B, 4) and ref.shape == (B, Nf + 1, 4) and weights.shape == (10,)
    cost, grad, xtraj = np.zeros(B), np.zeros((B, Nf, 2)), np.zeros((B, Nf + 1, 4))
    rc = lib.ihm2o_dpc_rollout_grad(C.c_int(B), C.c_int(Nf), C.c_double(dt), _p(x0), _p(u), _p(ref), _p(weights),
                                    _p(cost), _p(grad), _p(xtraj))
    assert rc == 0
    return cost, grad, xtraj


def project(s_ref, X_ref, Y_ref, phi_ref, pose, s_guess, s_tol=2.0, track=None):
    """Track::project + the control node's Frenet conversion (tracks.cpp:200-288, mpc_control_node.cpp:142-157).

    tables [n_tracks, n_grid] (or [n_grid]); pose [B,3] = (X, Y, phi); s_guess [B] -> frenet [B,3] = (s, n, psi),
    ref [B,3] = (X_ref, Y_ref, phi_ref) at the projection, flag [B]."""
    lib = C.CDLL(build_oracle())
    s_ref, X_ref, Y_ref, phi_ref = [_f(np.atleast_2d(a)) for a in (s_ref, X_ref, Y_ref, phi_ref)]
    pose, s_guess = _f(pose), _f(s_guess)
    B = pose.shape[0]
    tr = None if track is None else np.ascontiguousarray(track, dtype=np.int32)
    frenet, ref, flag = np.zeros((B, 3)), np.zeros((B, 3)), np.zeros(B, np.int32)
    rc = lib.ihm2o_project(C.c_int(B), C.c_int(s_ref.shape[1]), _p(s_ref), _p(X_ref), _p(Y_ref), _p(phi_ref), _p(tr),
                           _p(pose), _p(s_guess), C.c_double(s_tol), _p(frenet), _p(ref), _p(flag))
    assert rc == 0
    return frenet, ref, flag
