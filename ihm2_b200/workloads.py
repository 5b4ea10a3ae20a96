"""Synthetic inputs of the BASELINE.json configurations, generated with the
product path itself (no oracle): seeds fixed, independent of the number of GPUs.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _abi
from ._abi import NU, NX
from .engine import Engine
from .track import load_packaged_track

# perturbation amplitudes of config 2: (s, n, psi, v_x, v_y, r, T, delta)
PERTURBATION = np.array([0.5, 0.3, 0.1, 1.0, 0.2, 0.2, 50.0, 0.05])


def nominal_iterate(track: str = "skidpad", n_steps: int = 20, **cfg_kwargs):
    """Controller's initial guess (controller.py:313-317) advanced `n_steps` RTI steps in a
    perfect-model loop (next state = first predicted state), so the nominal is on track and
    moving.  Returns (x_nom [8], x_pred [N+1,8], u_pred [N,2]) as numpy arrays."""
    mp = load_packaged_track(track)
    cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **cfg_kwargs)
    widths = (mp.right_widths, mp.left_widths) if cfg.nonlinear_constraints else None
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=widths)
    x = torch.zeros(1, NX, dtype=torch.float64, device=eng.device)
    x[0, 0] = -6.0
    for _ in range(n_steps):
        eng.shift_iterate()
        _, status = eng.solve(x)
        xp, _ = eng.get_iterate()
        if int(status[0]) not in (0, 2):
            raise RuntimeError(f"nominal RTI step failed with status {int(status[0])}")
        x = xp[:, 1, :].contiguous()
    xp, up = eng.get_iterate()
    out = x[0].cpu().numpy(), xp[0].cpu().numpy(), up[0].cpu().numpy()
    eng.close()
    return out


def shift_np(x: np.ndarray, u: np.ndarray):
    """controller.py:388-393 (literal) on [..., N+1, 8] / [..., N, 2] arrays."""
    xs, us = x.copy(), u.copy()
    xs[..., :-1, :] = x[..., 1:, :]
    us[..., :-1, :] = u[..., 1:, :]
    us[..., -1, :] = 0.0
    return xs, us


def perturbed_batch(x_nom: np.ndarray, x_pred: np.ndarray, u_pred: np.ndarray, batch: int, seed: int = 0,
                    first: int = 0, cfg=None):
    """Config 2: x0_i = x_nom + eps_i, eps ~ U(-1,1) * PERTURBATION, clipped to the bounds;
    warm start = shifted nominal iterate for every instance.  Instance i depends only on
    (seed, first + i), so shards of a larger batch are reproducible on any rank."""
    g = torch.Generator().manual_seed(seed)
    eps = torch.rand(first + batch, NX, generator=g, dtype=torch.float64)[first:].numpy() * 2.0 - 1.0
    x0 = x_nom[None, :] + eps * PERTURBATION[None, :]
    n_max, vmax, Tmax, dmax = (2.0, 31.0, 500.0, 0.5) if cfg is None else (cfg.n_max, cfg.v_x_max, cfg.T_max, cfg.delta_max)
    x0[:, 1] = np.clip(x0[:, 1], -n_max, n_max)
    x0[:, 3] = np.clip(x0[:, 3], 0.0, vmax)
    x0[:, 6] = np.clip(x0[:, 6], -Tmax, Tmax)
    x0[:, 7] = np.clip(x0[:, 7], -dmax, dmax)
    xs, us = shift_np(x_pred, u_pred)
    x = np.broadcast_to(xs, (batch,) + xs.shape).copy()
    u = np.broadcast_to(us, (batch,) + us.shape).copy()
    return x0, x, u
