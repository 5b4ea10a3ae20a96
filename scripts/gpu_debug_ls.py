import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch
from oracle.lib import Oracle
mp = load_packaged_track("skidpad")
x_nom, xp, up = nominal_iterate("skidpad", 20)
B = 384
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=2)
for iters in (2, 8):
    cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nlp_solver=_abi.SQP, nlp_max_iter=iters, globalization=_abi.MERIT_BACKTRACKING)
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref); eng.set_iterate(x, u)
    u0, status = eng.solve(torch.as_tensor(x0, device="cuda"))
    sg = status.cpu().numpy(); stg = eng.stats().cpu().numpy(); xg = eng.get_iterate()[0].cpu().numpy()
    eng.close()
    u0o, so, sto, xo, uo = Oracle(cfg, mp.s_ref, mp.kappa_ref).solve(x0, x, u)
    print("iters", iters, "status agree", np.mean(sg == so))
    for a in np.unique(so):
        print("  oracle", a, "-> gpu", np.unique(sg[so == a], return_counts=True))
    both = (sg == 2) & (so == 2)
    print("  alpha agree", np.mean(stg[both, 2] == sto[both, 2]), "oracle alphas", np.unique(np.round(sto[both, 2], 3), return_counts=True))
    bad = np.where(both & (stg[:, 2] != sto[:, 2]))[0][:8]
    print("  mismatch", [(i, stg[i, 2], sto[i, 2]) for i in bad])
    dif = np.where(sg != so)[0][:8]
    print("  status diff", [(i, sg[i], so[i], stg[i, 1], sto[i, 1], stg[i,2], sto[i,2]) for i in dif])
    for i in list(bad[:5]) + list(np.where(both & (stg[:, 2] == sto[:, 2]))[0][:3]):
        print("   inst", i, "gpu m0,m1", stg[i, 3], stg[i, 4], "oracle m0,m1(accepted)", sto[i, 3], sto[i, 4], "alpha", stg[i, 2], sto[i, 2])
    ok = both & (stg[:, 2] == sto[:, 2])
    print("  x rel on agree", np.abs(xg[ok] - xo[ok]).max())
