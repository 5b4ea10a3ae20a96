"""GPU vs oracle on the soft-constrained RTI step: per-instance diagnostics (development aid)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import perturbed_batch
from oracle.lib import Oracle
from oracle.workloads import nominal_iterate_cpu

mp = load_packaged_track("fsds_default")
B = 256
kw = eval(sys.argv[1]) if len(sys.argv) > 1 else {}
x_nom, xp, up = nominal_iterate_cpu("fsds_default", 30, nonlinear_constraints=True)
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nonlinear_constraints=True, **kw)
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=5, cfg=cfg)
x0[:, 1] = np.clip(x0[:, 1] + np.linspace(-0.9, 0.9, B), -1.8, 1.8)
W = (mp.right_widths, mp.left_widths)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=W)
eng.set_iterate(x, u)
u0, status = eng.solve(torch.as_tensor(x0, device="cuda"))
xg, ug = [t.cpu().numpy() for t in eng.get_iterate()]
sg = status.cpu().numpy(); stg = eng.stats().cpu().numpy()
u0o, so, sto, xo, uo = Oracle(cfg, mp.s_ref, mp.kappa_ref, widths=W).solve(x0, x, u)
ok = (sg == 0) & (so == 0)
print("status agree", np.mean(sg == so), "ok", ok.sum())
ex = np.abs(xg - xo).reshape(B, -1).max(1); eu = np.abs(ug - uo).reshape(B, -1).max(1)
print("same iters", np.mean(stg[ok, 0] == sto[ok, 0]))
worst = np.argsort(-(ex * ok))[:8]
for i in worst:
    d = np.abs(xg[i] - xo[i]).max(0)
    print(i, "it", stg[i, 0], sto[i, 0], "ex %.2e eu %.2e" % (ex[i], eu[i]), "per comp", np.array2string(d, precision=1),
          "res g/b/d/m gpu", np.array2string(stg[i, 3:7], precision=1), "cpu", np.array2string(sto[i, 3:7], precision=1))
print("median ex %.2e, 90%% %.2e" % (np.median(ex[ok]), np.quantile(ex[ok], 0.9)))
