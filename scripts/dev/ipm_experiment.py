"""IPM strategy experiments on the CPU oracle with the config-2 workload (first B problems)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import perturbed_batch
from oracle.lib import Oracle
from oracle.workloads import nominal_iterate_cpu

B = int(os.environ.get("B", 512))
mp = load_packaged_track("skidpad")
cache = "/tmp/nominal_skidpad.npz"
if os.path.exists(cache):
    d = np.load(cache); x_nom, xp, up = d["x_nom"], d["xp"], d["up"]
else:
    x_nom, xp, up = nominal_iterate_cpu("skidpad", 20)
    np.savez(cache, x_nom=x_nom, xp=xp, up=up)
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0)

def run(**kw):
    cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), **kw)
    o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
    t = time.time()
    u0, st, stats, xo, uo = o.solve(x0, x, u)
    dt = time.time() - t
    ok = st == 0
    return dict(ok=int(ok.sum()), fail=int((~ok).sum()), it_ok=float(stats[ok, 0].mean()), it_fail=float(stats[~ok, 0].mean()) if (~ok).any() else 0.0,
                it_all=float(stats[:, 0].mean()), max_ok=float(stats[ok, 0].max()), t=dt), u0, st, xo

if __name__ == "__main__":
    base, u0b, stb, xob = run()
    print("base", base)
    for kw in (dict(qp_mu0=100.0), dict(qp_mu0=1000.0), dict(qp_split_step=True), dict(qp_split_step=True, qp_mu0=100.0),
               dict(qp_split_step=True, qp_mu0=1000.0), dict(qp_infeas_exit=True), dict(qp_infeas_exit=True, qp_split_step=True, qp_mu0=1000.0)):
        r, u0, st, xo = run(**kw)
        both = (st == 0) & (stb == 0)
        err = np.abs(u0[both] - u0b[both]).max() / max(1, np.abs(u0b[both]).max())
        print(kw, r, "du0 %.2e" % err, "newly ok", int(((st == 0) & (stb != 0)).sum()), "newly failed", int(((st != 0) & (stb == 0)).sum()))
