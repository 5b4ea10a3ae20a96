"""ctypes binding of ihm2_b200/lib/libihm2b.so (the C ABI of include/ihm2b.h).

PyTorch supplies device memory and streams only; every computation happens in the library's
hand-written sm_100a kernels.  There is no CPU fallback: a missing library or a missing GPU
raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _abi
from ._abi import NP, NU, NX, Config

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libihm2b.so")
_lib = None

# every symbol include/ihm2b.h declares
EXPORTS = (
    "ihm2b_create", "ihm2b_destroy", "ihm2b_last_error", "ihm2b_version", "ihm2b_set_tracks",
    "ihm2b_set_track_index", "ihm2b_set_params", "ihm2b_set_iterate_dev", "ihm2b_get_iterate_dev", "ihm2b_get_iterate_host",
    "ihm2b_reset_iterate", "ihm2b_shift_iterate", "ihm2b_solve_dev", "ihm2b_compute_control_host",
    "ihm2b_linearize_dev", "ihm2b_get_stats_dev", "ihm2b_sim_step_dev", "ihm2b_sim_step_static_dev", "ihm2b_sim_step_dyn10_dev", "ihm2b_closed_loop_dev",
    "ihm2b_get_multipliers_dev", "ihm2b_get_slacks_dev", "ihm2b_get_h_rows_dev",
    "ihm2b_dpc_rollout_grad_dev", "ihm2b_dpc_rollout_grad_f32_dev", "ihm2b_project_dev", "ihm2b_launch_count",
    "ihm2b_profile_enable", "ihm2b_profile_read", "ihm2b_debug_trace", "ihm2b_measure_fp64_tflops",
)


class Ihm2bError(RuntimeError):
    pass


def load_library() -> C.CDLL:
    """dlopen libihm2b.so; raises if it has not been built (`python -c 'import
    __graft_entry__ as g; g.build()'`)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise Ihm2bError(f"{_LIB_PATH} not found: build it with __graft_entry__.build(); "
                             "there is no CPU fallback")
        lib = C.CDLL(_LIB_PATH)
        lib.ihm2b_last_error.restype = C.c_char_p
        lib.ihm2b_launch_count.restype = C.c_int64
        lib.ihm2b_launch_count.argtypes = [C.c_void_p]
        for name in EXPORTS:
            getattr(lib, name)  # AttributeError if the header and the library disagree
        _lib = lib
    return _lib


def _dptr(t: torch.Tensor | None):
    if t is None:
        return C.c_void_p(0)
    assert t.is_cuda and t.is_contiguous(), "device buffers must be contiguous CUDA tensors"
    return C.c_void_p(t.data_ptr())


def _hptr(a: np.ndarray | None):
    return a.ctypes.data_as(C.c_void_p) if a is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def measure_fp64_tflops() -> float:
    """DFMA- and DMMA-chain microbenchmarks on the current device (FP64 roofline denominator: the larger)."""
    lib = load_library()
    out = C.c_double(0.0)
    rc = lib.ihm2b_measure_fp64_tflops(C.byref(out))
    if rc != 0:
        raise Ihm2bError(f"libihm2b error {rc}: {lib.ihm2b_last_error().decode()}")
    return float(out.value)


class Engine:
    """One ihm2b handle: B independent NMPC problems on the current CUDA device."""

    def __init__(self, cfg: Config, s_ref, kappa_ref, track_index=None, params=None,
                 lap_length=None, device=None, widths=None):
        """widths = (right_width, left_width) tables shaped like s_ref: needed with cfg.nonlinear_constraints (the
        track rows of h, old/generate_acaods_interface.py:312-318)."""
        if not torch.cuda.is_available():
            raise Ihm2bError("ihm2_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = load_library()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        torch.cuda.set_device(self.device)
        torch.cuda.init()
        torch.zeros(1, device=self.device)  # make sure the primary context exists
        self.cfg = cfg
        self.B, self.N = cfg.batch, cfg.N
        self._h = C.c_void_p(0)
        self._check(self.lib.ihm2b_create(C.byref(cfg), C.byref(self._h)))
        s_ref = np.ascontiguousarray(np.atleast_2d(s_ref), dtype=np.float64)
        kappa_ref = np.ascontiguousarray(np.atleast_2d(kappa_ref), dtype=np.float64)
        if s_ref.shape != (cfg.n_tracks, cfg.n_grid) or kappa_ref.shape != s_ref.shape:
            raise ValueError(f"track tables must be [{cfg.n_tracks}, {cfg.n_grid}]")
        rw = lw = None
        if widths is not None:
            rw, lw = (np.ascontiguousarray(np.atleast_2d(t), dtype=np.float64) for t in widths)
            if rw.shape != s_ref.shape or lw.shape != s_ref.shape:
                raise ValueError(f"width tables must be [{cfg.n_tracks}, {cfg.n_grid}]")
        self._check(self.lib.ihm2b_set_tracks(self._h, _hptr(s_ref), _hptr(kappa_ref), _hptr(rw), _hptr(lw)))
        ti = np.zeros(self.B, np.int32) if track_index is None else np.ascontiguousarray(track_index, dtype=np.int32)
        self._check(self.lib.ihm2b_set_track_index(self._h, _hptr(ti)))
        p = np.tile(_abi.default_params(), (self.B, 1)) if params is None else np.ascontiguousarray(params, dtype=np.float64)
        if p.shape != (self.B, NP):
            raise ValueError(f"params must be [{self.B}, {NP}]")
        self._check(self.lib.ihm2b_set_params(self._h, _hptr(p)))
        self.lap_length = None if lap_length is None else np.ascontiguousarray(np.atleast_1d(lap_length), dtype=np.float64)

    def _check(self, rc: int):
        if rc != 0:
            raise Ihm2bError(f"libihm2b error {rc}: {self.lib.ihm2b_last_error().decode()}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            torch.cuda.synchronize(self.device)
            self.lib.ihm2b_destroy(self._h)
            self._h = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- iterate ----
    def _new(self, *shape, dtype=torch.float64):
        return torch.empty(*shape, dtype=dtype, device=self.device)

    def set_iterate(self, x=None, u=None):
        x = None if x is None else torch.as_tensor(x, dtype=torch.float64, device=self.device).contiguous()
        u = None if u is None else torch.as_tensor(u, dtype=torch.float64, device=self.device).contiguous()
        if x is not None:
            assert x.shape == (self.B, self.N + 1, NX)
        if u is not None:
            assert u.shape == (self.B, self.N, NU)
        self._check(self.lib.ihm2b_set_iterate_dev(self._h, _dptr(x), _dptr(u), _stream()))

    def get_iterate(self):
        x, u = self._new(self.B, self.N + 1, NX), self._new(self.B, self.N, NU)
        self._check(self.lib.ihm2b_get_iterate_dev(self._h, _dptr(x), _dptr(u), _stream()))
        return x, u

    def get_iterate_host(self, x: np.ndarray | None = None, u: np.ndarray | None = None):
        """x_pred / u_pred into host arrays (pass views of pinned tensors for the full PCIe rate); synchronous."""
        x = np.empty((self.B, self.N + 1, NX)) if x is None else x
        u = np.empty((self.B, self.N, NU)) if u is None else u
        assert x.flags.c_contiguous and u.flags.c_contiguous and x.dtype == np.float64 and u.dtype == np.float64
        self._check(self.lib.ihm2b_get_iterate_host(self._h, _hptr(x), _hptr(u)))
        return x, u

    def reset_iterate(self):
        self._check(self.lib.ihm2b_reset_iterate(self._h, _stream()))

    def shift_iterate(self):
        self._check(self.lib.ihm2b_shift_iterate(self._h, _stream()))

    # ---- hot path ----
    def solve(self, x0: torch.Tensor, u0: torch.Tensor | None = None, status: torch.Tensor | None = None):
        """Device in, device out: x0 [B,8] -> (u0 [B,2], status [B]); asynchronous."""
        assert x0.shape == (self.B, NX) and x0.dtype == torch.float64
        u0 = self._new(self.B, NU) if u0 is None else u0
        status = self._new(self.B, dtype=torch.int32) if status is None else status
        self._check(self.lib.ihm2b_solve_dev(self._h, _dptr(x0), _dptr(u0), _dptr(status), _stream()))
        return u0, status

    def compute_control_host(self, x0: np.ndarray, do_shift: bool = True):
        """Host in, host out (the reference's numpy call shape); synchronous."""
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        assert x0.shape == (self.B, NX)
        u0 = np.empty((self.B, NU))
        status = np.empty(self.B, np.int32)
        self._check(self.lib.ihm2b_compute_control_host(self._h, _hptr(x0), _hptr(u0), _hptr(status),
                                                        C.c_int(int(do_shift))))
        return u0, status

    def linearize(self):
        A = self._new(self.B, self.N, NX, NX)
        Bm = self._new(self.B, self.N, NX, NU)
        phi = self._new(self.B, self.N, NX)
        self._check(self.lib.ihm2b_linearize_dev(self._h, _dptr(A), _dptr(Bm), _dptr(phi), _stream()))
        return A, Bm, phi

    def stats(self):
        st = self._new(self.B, 8)
        self._check(self.lib.ihm2b_get_stats_dev(self._h, _dptr(st), _stream()))
        return st

    def multipliers(self):
        """(pi [B,N+1,8], lamL [B,N+1,RS], lamU [B,N+1,RS]) of the last QP; RS = 8 row slots per stage, 16 with h rows."""
        RS = 16 if self.cfg.nonlinear_constraints else 8
        pi, lamL, lamU = self._new(self.B, self.N + 1, NX), self._new(self.B, self.N + 1, RS), self._new(self.B, self.N + 1, RS)
        self._check(self.lib.ihm2b_get_multipliers_dev(self._h, _dptr(pi), _dptr(lamL), _dptr(lamU), _stream()))
        return pi, lamL, lamU

    def slacks(self):
        """(sL, sU, lamSL, lamSU) of the soft rows (nonlinear_constraints only), expanded to the 16 row slots per stage
        of `multipliers()`: [B,N+1,16], zeros where a slot has no slack.  (The ABI returns 8 packed slots per stage.)"""
        packed = [self._new(self.B, self.N + 1, 8) for _ in range(4)]
        self._check(self.lib.ihm2b_get_slacks_dev(self._h, *[_dptr(t) for t in packed], _stream()))
        out = []
        N = self.N
        for pk in packed:
            t = torch.zeros(self.B, N + 1, 16, dtype=torch.float64, device=self.device)
            t[:, :, 8] = pk[:, :, 0]; t[:, :, 9] = pk[:, :, 1]; t[:, :N, 12] = pk[:, :N, 4]      # h0, h1, a_lat
            t[:, 1:N, 6:8] = pk[:, 1:N, 6:8]                                                   # soft bounds on T, delta
            t[:, N, 2:4] = pk[:, N, 2:4]                                                       # terminal: v_y, r
            out.append(t)
        return out

    def h_rows(self):
        """Linearised rows of h at the last linearisation point, [B,N+1,16] (layout of Work::HC)."""
        hc = self._new(self.B, self.N + 1, 16)
        self._check(self.lib.ihm2b_get_h_rows_dev(self._h, _dptr(hc), _stream()))
        return hc

    # ---- plant / closed loop ----
    def sim_step(self, x: torch.Tensor, u: torch.Tensor):
        xn = self._new(self.B, NX)
        model = self._new(self.B, dtype=torch.int32)
        self._check(self.lib.ihm2b_sim_step_dev(self._h, _dptr(x.contiguous()), _dptr(u.contiguous()),
                                                _dptr(xn), _dptr(model), _stream()))
        return xn, model

    def sim_step_static(self, x: torch.Tensor, u: torch.Tensor, xn: torch.Tensor, model: "torch.Tensor | None" = None):
        """ihm2b_sim_step_static_dev: the plant step with its scratch in static device memory.  Allocates nothing:
        the caller passes the outputs xn [B,8] (float64) and, optionally, model [B] (int32); B <= 65536."""
        assert x.shape == (self.B, NX) and u.shape == (self.B, NU) and xn.shape == (self.B, NX)
        assert x.is_contiguous() and u.is_contiguous() and xn.is_contiguous() and xn.dtype == torch.float64
        assert model is None or (model.shape == (self.B,) and model.dtype == torch.int32 and model.is_contiguous())
        self._check(self.lib.ihm2b_sim_step_static_dev(self._h, _dptr(x), _dptr(u), _dptr(xn), _dptr(model), _stream()))
        return xn

    def sim_step_dyn10(self, x10: torch.Tensor, u5: torch.Tensor):
        """dyn10 plant: x10 [B,15], u5 [B,5] -> x_next [B,15] (NaN rows where the integrator failed)."""
        assert x10.shape == (self.B, _abi.NX10) and u5.shape == (self.B, _abi.NU10)
        xn = self._new(self.B, _abi.NX10)
        self._check(self.lib.ihm2b_sim_step_dyn10_dev(self._h, _dptr(x10.contiguous()), _dptr(u5.contiguous()), _dptr(xn), _stream()))
        return xn

    def closed_loop(self, x: torch.Tensor, n_steps: int, log: bool = False):
        """n_steps of shift -> solve -> plant step on the device.  x [B,8] is updated in place."""
        assert self.lap_length is not None, "lap_length is required for the closed loop"
        assert x.shape == (self.B, NX) and x.is_contiguous()
        xlog = self._new(n_steps + 1, self.B, NX) if log else None
        ulog = self._new(n_steps, self.B, NU) if log else None
        done = self._new(self.B, dtype=torch.int32)
        fail = self._new(self.B, dtype=torch.int32)
        lap_time, cost_sum = self._new(self.B), self._new(self.B)
        self._check(self.lib.ihm2b_closed_loop_dev(self._h, C.c_int32(n_steps), _dptr(x), _hptr(self.lap_length),
                                                   _dptr(xlog), _dptr(ulog), _dptr(done), _dptr(fail), _dptr(lap_time),
                                                   _dptr(cost_sum), _stream()))
        return dict(x=x, xlog=xlog, ulog=ulog, done=done, fail=fail, lap_time=lap_time, cost_sum=cost_sum)

    # ---- measurement aids ----
    def profile(self, on: bool = True):
        self._check(self.lib.ihm2b_profile_enable(self._h, C.c_int(int(on))))

    def profile_read(self):
        """Mean device time (ms) of the linearisation and QP kernels since profile(True)."""
        ms = (C.c_double * 2)()
        self._check(self.lib.ihm2b_profile_read(self._h, ms))
        return float(ms[0]), float(ms[1])

    @property
    def launch_count(self) -> int:
        return int(self.lib.ihm2b_launch_count(self._h))
