"""CPU twins of ihm2_b200/workloads.py (test infrastructure only): the same synthetic inputs, generated with
the oracle instead of the GPU engine, so that CPU-only tests and experiments can build the BASELINE.json
configurations without a device."""
from __future__ import annotations

import numpy as np

from ihm2_b200 import _abi
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import perturbed_batch  # noqa: F401  (pure numpy / torch-CPU)

from .lib import Oracle, initial_iterate


def nominal_iterate_cpu(track: str = "skidpad", n_steps: int = 20, **cfg_kwargs):
    """ihm2_b200.workloads.nominal_iterate on the oracle: the controller's initial guess advanced n_steps RTI
    steps in a perfect-model loop.  Returns (x_nom [8], x_pred [N+1,8], u_pred [N,2])."""
    mp = load_packaged_track(track)
    cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **cfg_kwargs)
    o = Oracle(cfg, mp.s_ref, mp.kappa_ref, widths=(mp.right_widths, mp.left_widths))
    x, u = initial_iterate(cfg)
    xs = np.zeros((1, 8))
    xs[0, 0] = -6.0
    for _ in range(n_steps):
        _, st, _, x, u = o.solve(xs, x, u, do_shift=True)
        if int(st[0]) not in (0, 2):
            raise RuntimeError(f"nominal RTI step failed with status {int(st[0])}")
        xs = x[:, 1, :].copy()
    return xs[0], x[0], u[0]
