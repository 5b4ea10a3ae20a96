import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
os.environ["IHM2O_TRACE"] = "1"
from ihm2_b200 import _abi
from ihm2_b200.track import load_packaged_track
from oracle.lib import Oracle, initial_iterate
kw = eval(sys.argv[1]) if len(sys.argv) > 1 else {}
mp = load_packaged_track("fsds_default")
cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), nonlinear_constraints=True, **kw)
o = Oracle(cfg, mp.s_ref, mp.kappa_ref, widths=(mp.right_widths, mp.left_widths))
x, u = initial_iterate(cfg)
xs = np.zeros((1, 8)); xs[0, 0] = -6.0
u0, st, stats, xo, uo = o.solve(xs, x, u, do_shift=True)
print(st, stats)
