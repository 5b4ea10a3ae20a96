"""ctypes mirror of include/ihm2b.h plus the constants that fill it.

Vehicle constants: python/constants.py:43-111 of the reference.  Solver options:
new_python/controller.py:250-309, python/mpc.py:29-101.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

NX, NU, NY, NYN, NP = 8, 2, 12, 8, 28
NX10, NU10 = 15, 5   # dyn10 plant

# enum ihm2b_integrator / nlp_solver / globalization / sim_model / status
ERK4, IRK_GL4 = 0, 1
SQP_RTI, SQP = 0, 1
FIXED_STEP, MERIT_BACKTRACKING = 0, 1
SIM_FKIN6, SIM_FDYN6, SIM_KIN6_DYN6 = 0, 1, 2
SIM_KIN6C, SIM_DYN6C, SIM_KIN6C_DYN6C = 3, 4, 5  # Cartesian plants of the ROS sim node
STATUS_SUCCESS, STATUS_FAILURE, STATUS_MAXITER, STATUS_MINSTEP, STATUS_QP_FAILURE = 0, 1, 2, 3, 4

(P_M, P_IZ, P_LR, P_LF, P_CM0, P_CR0, P_CR1, P_CR2, P_TT, P_TD, P_ZCG, P_TRACKW, P_CDOWN,
 P_KTV, P_BA, P_CA, P_DA, P_EA, P_G, P_CROSS, P_RW, P_IW, P_KD, P_KS, P_BS, P_CS, P_DS, P_ES) = range(NP)


class Config(C.Structure):
    """struct ihm2b_config (include/ihm2b.h)."""

    _fields_ = [
        ("batch", C.c_int32),
        ("N", C.c_int32),
        ("n_tracks", C.c_int32),
        ("n_grid", C.c_int32),
        ("integrator", C.c_int32),
        ("nlp_solver", C.c_int32),
        ("nlp_max_iter", C.c_int32),
        ("globalization", C.c_int32),
        ("qp_max_iter", C.c_int32),
        ("nonlinear_constraints", C.c_int32),
        ("sim_model", C.c_int32),
        ("sim_steps", C.c_int32),
        ("schedule_by_history", C.c_int32),
        ("qp_tolerate_maxiter", C.c_int32),
        ("dt", C.c_double),
        ("s_target", C.c_double),
        ("n_max", C.c_double),
        ("v_x_max", C.c_double),
        ("T_max", C.c_double),
        ("delta_max", C.c_double),
        ("g_T_max", C.c_double),
        ("g_delta_max", C.c_double),
        ("a_lat_max", C.c_double),
        ("W", C.c_double * NY),
        ("W_e", C.c_double * NYN),
        ("cost_scale_stage", C.c_double),
        ("cost_scale_terminal", C.c_double),
        ("qp_mu0", C.c_double),
        ("qp_tol", C.c_double),
        ("qp_tol_comp", C.c_double),
        ("nlp_tol", C.c_double),
        ("irk_A", C.c_double * 16),
        ("irk_b", C.c_double * 4),
        ("sim_A", C.c_double * 16),
        ("sim_b", C.c_double * 4),
        ("sim_switch", C.c_double),
        ("track_half_L", C.c_double),
        ("track_half_W", C.c_double),
        ("slack_z", C.c_double),
        ("slack_Z", C.c_double),
        ("T_dot_max", C.c_double),
        ("delta_dot_max", C.c_double),
        ("qp_split_step", C.c_int32),
        ("qp_infeas_exit", C.c_int32),
        ("qp_thr0", C.c_double),
        ("linearize_fp32", C.c_int32),
        ("reserved_", C.c_int32),
    ]


def default_params() -> np.ndarray:
    """Nominal vehicle parameter vector (python/constants.py:43-111)."""
    g, m = 9.81, 230.0
    l_R = l_F = 0.7853
    wheelbase = 1.5706
    static_weight = 0.5 * m * g * l_F / wheelbase
    b1a, b2a, c1a, d1a, d2a, e1a, e2a = 3.79e1, 5.28e2, 1.57, -2.03e-4, 1.77, -2.24e-3, 1.81
    BCDa = b1a * np.sin(2 * np.arctan(static_weight / b2a))
    Ca = c1a
    Da = d1a * static_weight + d2a
    Ea = e1a * static_weight + e2a
    Ba = BCDa / (Ca * Da)
    p = np.zeros(NP)
    p[P_M], p[P_IZ], p[P_LR], p[P_LF] = m, 137.583, l_R, l_F
    p[P_CM0], p[P_CR0], p[P_CR1], p[P_CR2] = 4.950, 297.030, 16.665, 0.6784
    p[P_TT], p[P_TD] = 1e-3, 0.02
    p[P_ZCG], p[P_TRACKW], p[P_CDOWN], p[P_KTV] = 0.295, 1.24, 3.96864, 300.0
    p[P_BA], p[P_CA], p[P_DA], p[P_EA] = Ba, Ca, Da, Ea
    p[P_G], p[P_CROSS] = g, 1.0
    # dyn10 plant only: wheel parameters and the longitudinal Pacejka coefficients (constants.py:69-74, 85-101)
    b1s, b2s, b3s, c1s, d1s, d2s, e1s, e2s, e3s = -6.75e-6, 1.35e-1, 1.2e-3, 1.86, 1.12e-4, 1.57, -5.38e-6, 1.11e-2, -4.26
    BCDs = (b1s * static_weight**2 + b2s * static_weight) * np.exp(-b3s * static_weight)
    Cs, Ds = c1s, d1s * static_weight + d2s
    Es = e1s * static_weight**2 + e2s * static_weight + e3s
    p[P_RW], p[P_IW], p[P_KD], p[P_KS] = 0.20809, 0.3, 0.17, 15.0
    p[P_BS], p[P_CS], p[P_DS], p[P_ES] = BCDs / (Cs * Ds), Cs, Ds, Es
    return p


def gauss_legendre_tableau(ns: int = 4) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Collocation tableau on the Gauss-Legendre nodes (acados IRK default collocation)."""
    x, _ = np.polynomial.legendre.leggauss(ns)
    return _collocation_tableau(0.5 * (x + 1.0))


def radau_iia_tableau(ns: int = 4) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Collocation tableau on the Radau-IIA nodes (collocation_type GAUSS_RADAU_IIA,
    python/main.py:400): roots of P_{s-1}^{(1,0)} mapped to (0,1), plus c_s = 1."""
    from scipy.special import roots_jacobi

    x, _ = roots_jacobi(ns - 1, 1.0, 0.0)
    return _collocation_tableau(np.append(np.sort(0.5 * (x + 1.0)), 1.0))


def _collocation_tableau(c: np.ndarray):
    ns = len(c)
    # a_ij = int_0^{c_i} l_j, b_j = int_0^1 l_j  via the moment conditions
    V = np.vander(c, ns, increasing=True).T  # V[k, j] = c_j^k
    A = np.zeros((ns, ns))
    for i in range(ns):
        rhs = np.array([c[i] ** (k + 1) / (k + 1) for k in range(ns)])
        A[i] = np.linalg.solve(V, rhs)
    b = np.linalg.solve(V, np.array([1.0 / (k + 1) for k in range(ns)]))
    return A, b, c


# Interior-point presets.  The HPIPM rows restate `d_ocp_qp_ipm_arg_set_default` (mu0, residual tolerances, split
# step) with acados' own iteration cap of 50 (`qp_solver_iter_max`) in place of HPIPM's 15 / 30 / 100, as remembered:
# HPIPM / acados are not in this image, so the numbers are unpinned.  The QP is strictly convex, so every
# preset converges to the same point; they differ in iteration count and in what happens to infeasible QPs.
# "B200" (default): start barrier parameter matched to the size of the cost gradient of this OCP (the terminal
# weight times the progress reference, ~ 3e4), and an early exit for primal-infeasible QPs.
HPIPM_MODES = {
    "B200":      dict(mu0=3e3, tol=1e-9, tol_comp=1e-8, max_iter=50, split_step=False, infeas_exit=True),
    "SPEED_ABS": dict(mu0=1e1, tol=1e-9, tol_comp=1e-8, max_iter=50, split_step=True, infeas_exit=False),
    "SPEED":     dict(mu0=1e1, tol=1e-6, tol_comp=1e-6, max_iter=50, split_step=True, infeas_exit=False),
    "BALANCE":   dict(mu0=1e1, tol=1e-8, tol_comp=1e-8, max_iter=50, split_step=False, infeas_exit=False),
    "ROBUST":    dict(mu0=1e2, tol=1e-8, tol_comp=1e-8, max_iter=100, split_step=False, infeas_exit=False),
}


def make_config(
    batch: int,
    n_grid: int,
    n_tracks: int = 1,
    Nf: int = 40,
    dt: float = 1 / 20,
    s_target: float = 40.0,
    n_max: float = 2.0,
    v_x_max: float = 31.0,
    T_max: float = 500.0,
    delta_max: float = 0.5,
    T_dot_max: float = 1e6,
    delta_dot_max: float = 1.0,
    a_lat_max: float = 5.0,
    W=None,
    W_e=None,
    integrator: int = IRK_GL4,
    nlp_solver: int = SQP_RTI,
    nlp_max_iter: int = 2,
    globalization: int = FIXED_STEP,
    qp_max_iter: int | None = None,
    qp_mu0: float | None = None,
    qp_tol: float | None = None,
    qp_tol_comp: float | None = None,
    hpipm_mode: str | None = None,
    nlp_tol: float = 1e-6,
    cost_scaling: bool = True,
    sim_model: int = SIM_KIN6_DYN6,
    sim_steps: int = 100,
    nonlinear_constraints: bool = False,
    schedule_by_history: bool = True,
    qp_tolerate_maxiter: bool = False,
    qp_split_step: bool | None = None,
    qp_infeas_exit: bool | None = None,
    qp_thr0: float = 0.1,
    linearize_fp32: bool = False,
) -> Config:
    """Fill ihm2b_config from `IHM2Controller.__init__` style arguments
    (new_python/controller.py:250-309; python/mpc.py:77-99 for the bounds).

    `hpipm_mode` (new_python/controller.py:304 selects "SPEED_ABS"; python/generate_all.py:107-116 "ROBUST") picks
    one of the interior-point presets of HPIPM_MODES; None is this library's own default.  Explicit qp_* arguments
    override the preset."""
    mode = dict(HPIPM_MODES["B200" if hpipm_mode is None else hpipm_mode])
    qp_max_iter = mode["max_iter"] if qp_max_iter is None else qp_max_iter
    qp_mu0 = mode["mu0"] if qp_mu0 is None else qp_mu0
    qp_tol = mode["tol"] if qp_tol is None else qp_tol
    qp_tol_comp = mode["tol_comp"] if qp_tol_comp is None else qp_tol_comp
    qp_split_step = mode["split_step"] if qp_split_step is None else qp_split_step
    qp_infeas_exit = mode["infeas_exit"] if qp_infeas_exit is None else qp_infeas_exit
    p = default_params()
    cf = Config()
    cf.batch, cf.N, cf.n_tracks, cf.n_grid = batch, Nf, n_tracks, n_grid
    cf.integrator, cf.nlp_solver, cf.nlp_max_iter = integrator, nlp_solver, nlp_max_iter
    cf.globalization, cf.qp_max_iter = globalization, qp_max_iter
    cf.nonlinear_constraints = int(nonlinear_constraints)
    cf.schedule_by_history = int(schedule_by_history)
    cf.qp_tolerate_maxiter = int(qp_tolerate_maxiter)
    cf.sim_model, cf.sim_steps = sim_model, sim_steps
    cf.dt, cf.s_target = dt, s_target
    cf.n_max, cf.v_x_max, cf.T_max, cf.delta_max = n_max, v_x_max, T_max, delta_max
    cf.g_T_max = p[P_TT] * T_dot_max  # python/mpc.py:98-99 uses the module constants
    cf.g_delta_max = p[P_TD] * delta_dot_max
    cf.a_lat_max = a_lat_max
    W = np.array([1, 1, 1, 1, 1, 1, 1, 100, 1, 100, 0, 500], float) if W is None else np.asarray(W, float)
    W_e = np.array([1000, 100, 100, 1, 1, 1, 1, 100], float) if W_e is None else np.asarray(W_e, float)
    cf.W[:] = W.tolist()
    cf.W_e[:] = W_e.tolist()
    cf.cost_scale_stage = dt if cost_scaling else 1.0
    cf.cost_scale_terminal = 1.0
    cf.qp_mu0, cf.qp_tol, cf.qp_tol_comp, cf.nlp_tol = qp_mu0, qp_tol, qp_tol_comp, nlp_tol
    A, b, _ = gauss_legendre_tableau(4)
    cf.irk_A[:] = A.ravel().tolist()
    cf.irk_b[:] = b.tolist()
    A, b, _ = radau_iia_tableau(4)
    cf.sim_A[:] = A.ravel().tolist()
    cf.sim_b[:] = b.tolist()
    cf.sim_switch = 3.0
    cf.track_half_L, cf.track_half_W = 0.5 * 3.19, 0.5 * 1.55  # python/constants.py:56-57
    cf.slack_z = cf.slack_Z = 100.0  # zl = zu = Zl = Zu = 100 (old/generate_acaods_interface.py:383-390)
    cf.T_dot_max, cf.delta_dot_max = T_dot_max, delta_dot_max
    cf.qp_split_step, cf.qp_infeas_exit, cf.qp_thr0 = int(qp_split_step), int(qp_infeas_exit), qp_thr0
    cf.linearize_fp32 = int(linearize_fp32)
    return cf
