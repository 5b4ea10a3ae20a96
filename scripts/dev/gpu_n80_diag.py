"""N = 80 soft-constrained nominal loop: the GPU solve fed with the oracle's iterate at every step, side by side."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import shift_np
from oracle.lib import Oracle, initial_iterate
track = sys.argv[1] if len(sys.argv) > 1 else "fsds_competition_1"
Nf = int(sys.argv[2]) if len(sys.argv) > 2 else 80
st_ = float(sys.argv[3]) if len(sys.argv) > 3 else 32.0
mp = load_packaged_track(track)
cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), Nf=Nf, s_target=st_, nonlinear_constraints=True)
W = (mp.right_widths, mp.left_widths)
o = Oracle(cfg, mp.s_ref, mp.kappa_ref, widths=W)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=W)
x, u = initial_iterate(cfg)
xs = np.zeros((1, 8)); xs[0, 0] = -6.0
for step in range(25):
    xsh, ush = shift_np(x, u)
    eng.set_iterate(xsh, ush)
    u0g, stg = eng.solve(torch.as_tensor(xs, device="cuda"))
    sg = eng.stats().cpu().numpy()[0]
    xg = eng.get_iterate()[0].cpu().numpy()
    u0o, sto, so, x, u = o.solve(xs, x, u, do_shift=True)
    print(f"step {step:2d} oracle st {int(sto[0])} its {so[0,0]:.0f} res {so[0,3]:.1e} {so[0,4]:.1e} {so[0,5]:.1e} {so[0,6]:.1e} mu {so[0,2]:.1e} | gpu st {int(stg[0])} its {sg[0]:.0f} res {sg[3]:.1e} {sg[4]:.1e} {sg[5]:.1e} {sg[6]:.1e} mu {sg[2]:.1e} | dx {np.abs(xg - x).max():.1e}", flush=True)
    if int(sto[0]) not in (0, 2): break
    xs = x[:, 1, :].copy()
