import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from scripts.dev.ipm_experiment import run
base, u0b, stb, xob = run()
for kw in eval(sys.argv[1]):
    r, u0, st, xo = run(**kw)
    both = (st == 0) & (stb == 0)
    err = np.abs(u0[both] - u0b[both]).max() / max(1, np.abs(u0b[both]).max())
    print(kw, {k: round(v, 2) for k, v in r.items()}, "du0 %.2e" % err, "newly ok", int(((st == 0) & (stb != 0)).sum()), "newly failed", int(((st != 0) & (stb == 0)).sum()))
