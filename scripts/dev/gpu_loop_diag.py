"""Closed loop on the GPU, the oracle re-solving every step from the GPU's own state and iterate (development aid):
finds the first step at which the two solves differ and shows what the failing GPU solves look like."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from oracle.lib import Oracle

mp = load_packaged_track("fsds_default")
B, n_steps = 8, int(os.environ.get("STEPS", 540))
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nonlinear_constraints=True, qp_tolerate_maxiter=True, s_target=24.0)
params = np.tile(_abi.default_params(), (B, 1)); params[:, _abi.P_CROSS] = 0.0
params[:, _abi.P_M] *= np.linspace(0.97, 1.03, B)
W = (mp.right_widths, mp.left_widths)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, params=params, lap_length=[mp.lap_length], widths=W)
orc = Oracle(cfg, mp.s_ref, mp.kappa_ref, params=params, widths=W)
xs = torch.zeros(B, 8, dtype=torch.float64, device="cuda"); xs[:, 0] = -6.0
alive = np.ones(B, bool)
for step in range(n_steps):
    xi, ui = [t.cpu().numpy() for t in eng.get_iterate()]
    eng.shift_iterate()
    u0, status = eng.solve(xs)
    sg, stg, u0g = status.cpu().numpy(), eng.stats().cpu().numpy(), u0.cpu().numpy()
    u0o, so, sto, xo, uo = orc.solve(xs.cpu().numpy(), xi.copy(), ui.copy(), do_shift=True)
    du = np.abs(u0g - u0o).max(1) / np.maximum(1.0, np.abs(u0o).max(1))
    bad = alive & ((sg != so) | (du > 1e-6) | (stg[:, 0] != sto[:, 0]))
    for i in np.nonzero(bad)[0]:
        print(f"step {step} inst {i}: status gpu {sg[i]} cpu {so[i]}  ipm {stg[i,0]:.0f}/{sto[i,0]:.0f}  du0 {du[i]:.2e}  "
              f"x = {np.array2string(xs[i].cpu().numpy(), precision=3)}  gpu res {np.array2string(stg[i,2:7], precision=2)}")
    alive &= np.isin(sg, (0, 2))
    xn, _ = eng.sim_step(xs, u0)
    ok = torch.isfinite(xn).all(1) & torch.as_tensor(alive, device="cuda")
    xs = torch.where(ok[:, None], xn, xs)
    if not alive.any():
        break
print("alive", alive, "s", xs[:, 0].cpu().numpy())
