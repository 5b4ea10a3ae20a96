import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
os.environ["IHM2O_TRACE"] = "1"
from scripts.dev.ipm_experiment import _abi, mp, Oracle, x0, x, u
i = int(sys.argv[1]); kw = eval(sys.argv[2]) if len(sys.argv) > 2 else {}
cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **kw)
o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
u0, st, stats, xo, uo = o.solve(x0[i:i+1], x[i:i+1], u[i:i+1])
print("status", st, "iters", stats[0, 0])
