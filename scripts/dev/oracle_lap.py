"""Closed loop on the CPU oracle (config 1): does a lap finish?"""
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
mp = load_packaged_track(track)
cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **kw)
params = _abi.default_params()[None, :].copy()
for k, v in pk.items(): params[0, getattr(_abi, k)] = v
o = Oracle(cfg, mp.s_ref, mp.kappa_ref, params=params, widths=(mp.right_widths, mp.left_widths))
x, u = initial_iterate(cfg)
xs = np.zeros((1, 8)); xs[0, 0] = -6.0
t = time.time()
res = o.closed_loop(xs, x, u, [mp.lap_length], n_steps)
print("time %.1fs" % (time.time() - t), "done", res["done"], "fail", res["fail"], "lap_length", mp.lap_length)
xl = res["xlog"][:, 0]
last = max(int(res["done"][0]), int(res["fail"][0]))
last = last if last >= 0 else n_steps - 1
print("s_final %.2f  max|n| %.2f  max v_x %.2f" % (xl[last, 0], np.abs(xl[:last + 1, 1]).max(), xl[:last + 1, 3].max()))
for i in range(0, last + 1, max(1, last // 25)):
    print(i, np.array2string(xl[i], precision=3, suppress_small=True, max_line_width=200))
