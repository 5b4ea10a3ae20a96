"""Track pre-processing: centre line -> (s_ref, kappa_ref, ...) tables.

Follows the reference's offline motion plan (python/motion_planning.py:28-124 spline fit,
:139-232 arc-length resampling, :235-289 heading / curvature, :345-399 offline_motion_plan,
:402-428 triple_motion_plan_ref).  Runs once per track on the CPU (numpy / scipy); it feeds
the hot path but is not on it.

The reference solves the spline fit with a generic QP solver (qpsolvers / proxqp); the QP has
equality constraints only, so the exact solution is one sparse KKT solve, which is what is
done here.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

__all__ = [
    "NUMBER_SPLINE_INTERVALS",
    "MotionPlan",
    "load_center_line",
    "fit_spline",
    "offline_motion_plan",
    "triple_motion_plan_ref",
    "load_packaged_track",
    "write_track_data_file",
    "PACKAGED_TRACKS",
]

NUMBER_SPLINE_INTERVALS = 500  # python/motion_planning.py:25
_L_R = 0.7853  # python/constants.py:52 (used in the heading offset, motion_planning.py:381)
_DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "tracks")
PACKAGED_TRACKS = (
    "fsds_default",
    "fsds_competition_1",
    "fsds_competition_2",
    "fsds_competition_3",
    "skidpad",
    "short_skidpad",
)


@dataclass
class MotionPlan:  # python/motion_planning.py:292-301
    s_ref: np.ndarray
    X_ref: np.ndarray
    Y_ref: np.ndarray
    phi_ref: np.ndarray
    kappa_ref: np.ndarray
    right_widths: np.ndarray
    left_widths: np.ndarray
    lap_length: float


def load_center_line(path: str) -> tuple[np.ndarray, np.ndarray]:
    """Read `center_line.csv` (x, y, right_width, left_width; dpc/main.py:1052-1089).

    One header line, which may be commented with '#' (data/short_skidpad).  Returns
    (center_line [n,2], widths [n,2]).
    """
    rows = []
    with open(path) as f:
        for i, line in enumerate(f):
            line = line.strip()
            if not line:
                continue
            if i == 0 and (line.startswith("#") or line[0].isalpha()):
                continue
            rows.append([float(v) for v in line.split(",")[:4]])
    arr = np.asarray(rows, dtype=np.float64)
    return arr[:, :2].copy(), arr[:, 2:4].copy()


def fit_spline(path: np.ndarray, curv_weight: float = 1.0) -> tuple[np.ndarray, np.ndarray]:
    """Closed cubic spline through `path` minimising interpolation error + curvature.

    Same QP as python/motion_planning.py:28-124: unknowns (a_i, b_i, c_i, d_i) per segment,
    C0/C1/C2 continuity as equality constraints (with the chord-length ratios rho_i),
    cost |B p - path|^2 + curv_weight |C p|^2 + 1e-10 |p|^2.
    """
    assert path.ndim == 2 and path.shape[1] == 2
    N = path.shape[0]
    delta_s = np.linalg.norm(path[1:] - path[:-1], axis=1)
    delta_s = np.append(delta_s, np.linalg.norm(path[0] - path[-1]))
    rho = np.empty(N)
    rho[:-1] = delta_s[:-1] / delta_s[1:]
    rho[-1] = delta_s[-1] / delta_s[0]
    idx = np.arange(N)
    nxt = np.roll(idx, -1)
    blk = np.array([[1.0, 1.0, 1.0, 1.0], [0.0, 1.0, 2.0, 3.0], [0.0, 0.0, 2.0, 6.0]])
    A = sp.kron(sp.eye(N), blk, format="lil")
    for i in range(N):
        A[3 * i, 4 * nxt[i]] += -1.0
        A[3 * i + 1, 4 * nxt[i] + 1] += -rho[i]
        A[3 * i + 2, 4 * nxt[i] + 2] += -2.0 * rho[i] ** 2
    A = A.tocsc()
    Bm = sp.kron(sp.eye(N), np.array([[1.0, 0.0, 0.0, 0.0]]), format="csc")
    C = sp.csc_matrix(
        (
            np.concatenate((2 / np.square(delta_s), 6 / np.square(delta_s))),
            (np.concatenate((idx, idx)), np.concatenate((2 + 4 * idx, 3 + 4 * idx))),
        ),
        shape=(N, 4 * N),
    )
    P = Bm.T @ Bm + curv_weight * (C.T @ C) + 1e-10 * sp.eye(4 * N, format="csc")
    q = -(Bm.T @ path)
    K = sp.bmat([[P, A.T], [A, None]], format="csc")
    rhs = np.zeros((7 * N, 2))
    rhs[: 4 * N] = -q
    sol = spla.splu(K).solve(rhs)
    return sol[: 4 * N, 0].reshape(N, 4), sol[: 4 * N, 1].reshape(N, 4)


def _poly(c: np.ndarray, idx: np.ndarray, t: np.ndarray, order: int) -> np.ndarray:
    if order == 0:
        return c[idx, 0] + c[idx, 1] * t + c[idx, 2] * t**2 + c[idx, 3] * t**3
    if order == 1:
        return c[idx, 1] + 2 * c[idx, 2] * t + 3 * c[idx, 3] * t**2
    return 2 * c[idx, 2] + 6 * c[idx, 3] * t


def spline_interval_lengths(cX: np.ndarray, cY: np.ndarray, n_interp: int = 100) -> np.ndarray:
    """Polyline length of each segment from n_interp samples (motion_planning.py:139-177)."""
    t = np.linspace(0.0, 1.0, n_interp)
    T = np.stack([np.ones_like(t), t, t**2, t**3])  # [4, n_interp]
    X = cX @ T
    Y = cY @ T
    return np.sum(np.hypot(np.diff(X, axis=1), np.diff(Y, axis=1)), axis=1)


def offline_motion_plan(
    center_line: np.ndarray, track_widths: np.ndarray, n_samples: int = NUMBER_SPLINE_INTERVALS
) -> MotionPlan:
    """python/motion_planning.py:345-399 (curv_weight = 2.0, arc-length resampling to
    n_samples points with endpoint excluded, analytic curvature, phi - asin(l_R kappa),
    constant min widths, lap_length = s[-1] + closing chord)."""
    cX, cY = fit_spline(center_line, curv_weight=2.0)
    delta_s = spline_interval_lengths(cX, cY)
    s = np.cumsum(delta_s)
    s_interp = np.linspace(0.0, s[-1], n_samples, endpoint=False)
    idx = np.argmax(s_interp[:, None] < s[None, :], axis=1)
    s_prev = np.where(idx > 0, s[idx - 1], 0.0)
    t = (s_interp - s_prev) / delta_s[idx]
    X = _poly(cX, idx, t, 0)
    Y = _poly(cY, idx, t, 0)
    xd, yd = _poly(cX, idx, t, 1), _poly(cY, idx, t, 1)
    xdd, ydd = _poly(cX, idx, t, 2), _poly(cY, idx, t, 2)
    kappa = (xd * ydd - yd * xdd) / np.power(xd**2 + yd**2, 1.5)
    phi = np.arctan2(yd, xd) - np.arcsin(_L_R * kappa)
    lap_length = float(s_interp[-1] + np.hypot(X[-1] - X[0], Y[-1] - Y[0]))
    return MotionPlan(
        s_ref=s_interp,
        X_ref=X,
        Y_ref=Y,
        phi_ref=phi,
        kappa_ref=kappa,
        right_widths=np.full(n_samples, np.min(track_widths[:, 0])),
        left_widths=np.full(n_samples, np.min(track_widths[:, 1])),
        lap_length=lap_length,
    )


def triple_motion_plan_ref(mp: MotionPlan) -> MotionPlan:
    """Three laps back to back, s shifted by -L, 0, +L (motion_planning.py:402-428)."""
    L = mp.lap_length
    t3 = lambda a: np.hstack((a, a, a))  # noqa: E731
    return MotionPlan(
        s_ref=np.hstack((mp.s_ref - L, mp.s_ref, mp.s_ref + L)),
        X_ref=t3(mp.X_ref),
        Y_ref=t3(mp.Y_ref),
        phi_ref=t3(mp.phi_ref),
        kappa_ref=t3(mp.kappa_ref),
        right_widths=t3(mp.right_widths),
        left_widths=t3(mp.left_widths),
        lap_length=L,
    )


def load_packaged_track(name: str, tripled: bool = True) -> MotionPlan:
    """Single-lap tables precomputed by scripts/make_tracks.py from the reference's
    data/<name>/center_line.csv; tripled on load."""
    d = np.load(os.path.join(_DATA_DIR, f"{name}.npz"))
    mp = MotionPlan(
        s_ref=d["s_ref"],
        X_ref=d["X_ref"],
        Y_ref=d["Y_ref"],
        phi_ref=d["phi_ref"],
        kappa_ref=d["kappa_ref"],
        right_widths=d["right_widths"],
        left_widths=d["left_widths"],
        lap_length=float(d["lap_length"]),
    )
    return triple_motion_plan_ref(mp) if tripled else mp


def write_track_data_file(motion_plan: MotionPlan, outfile: str) -> None:
    """python/motion_planning.py:431-456 (`generate_track_data_file`): the single-lap tables as the CSV the C++ nodes
    and dpc/main.py read -- header `s_ref,X_ref,Y_ref,phi_ref,kappa_ref,right_width,left_width`, rows in `%.6f`."""
    parent = os.path.dirname(outfile)
    if parent and not os.path.exists(parent):
        os.makedirs(parent)
    with open(outfile, "w") as f:
        f.write("s_ref,X_ref,Y_ref,phi_ref,kappa_ref,right_width,left_width\n")
        np.savetxt(f, np.column_stack((motion_plan.s_ref, motion_plan.X_ref, motion_plan.Y_ref, motion_plan.phi_ref,
                                       motion_plan.kappa_ref, motion_plan.right_widths, motion_plan.left_widths)),
                   delimiter=",", fmt="%.6f")
