"""Closed loop on the CPU oracle, stepped from Python (diagnostics)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.track import load_packaged_track
from oracle.lib import Oracle, initial_iterate

track = sys.argv[1] if len(sys.argv) > 1 else "fsds_default"
kw = eval(sys.argv[2]) if len(sys.argv) > 2 else {}
pk = eval(sys.argv[3]) if len(sys.argv) > 3 else {}
n_steps = int(os.environ.get("STEPS", 600))
every = int(os.environ.get("EVERY", 10))
mp = load_packaged_track(track)
cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **kw)
params = _abi.default_params()[None, :].copy()
for k, v in pk.items(): params[0, getattr(_abi, k)] = v
o = Oracle(cfg, mp.s_ref, mp.kappa_ref, params=params, widths=(mp.right_widths, mp.left_widths))
x, u = initial_iterate(cfg)
xs = np.zeros((1, 8)); xs[0, 0] = -6.0
slog = []
for i in range(n_steps):
    u0, st, stats, x, u = o.solve(xs, x, u, do_shift=True)
    if st[0] not in (0, 2):
        print("step", i, "status", st, "stats", stats[0]); break
    xn, model = o.sim_step(xs, u0)
    if not np.isfinite(xn).all():
        print("step", i, "plant NaN"); break
    xs = xn; slog.append(xs[0, 0])
    if i % every == 0:
        print(i, "m%d it%2d" % (model[0], stats[0, 0]), np.array2string(xs[0], precision=3, suppress_small=True, max_line_width=200), "u", np.round(u0[0], 3))
    if xs[0, 0] > mp.lap_length + 1.0:
        print("lap done at step", i); break
s = np.array([-6.0] + slog)
idx = np.argwhere(np.abs(np.diff(np.mod(s, mp.lap_length))) > 1.0).ravel()
if idx.size > 1: print("lap time %.3f s" % (cfg.dt * (idx[1] - idx[0])))
