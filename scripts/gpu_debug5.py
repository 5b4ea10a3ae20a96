import os, sys, numpy as np, torch, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from oracle.lib import Oracle, initial_iterate
np.set_printoptions(precision=4, linewidth=220)
mp = load_packaged_track("skidpad"); ng = len(mp.s_ref)
cfg = _abi.make_config(batch=1, n_grid=ng)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
x = torch.zeros(1, 8, dtype=torch.float64, device="cuda"); x[0, 0] = -6
for t in range(20):
    eng.shift_iterate()
    xi, ui = [v.cpu().numpy() for v in eng.get_iterate()]
    u0, st = eng.solve(x); torch.cuda.synchronize()
    u0o, so, sto, xo, uo = o.solve(x.cpu().numpy(), xi, ui)
    xg, ug = [v.cpu().numpy() for v in eng.get_iterate()]
    stats = eng.stats().cpu().numpy()[0]
    print(t, "gpu st", int(st[0]), "it", stats[0], "| ora st", so[0], "it", sto[0, 0], "| dx", np.abs(xg - xo).max(), "du", np.abs(ug - uo).max())
    if int(st[0]) != 0:
        tr = np.zeros((64, 8)); eng.lib.ihm2b_debug_trace(eng._h, tr.ctypes.data_as(C.c_void_p))
        print(tr[:52, :7])
        break
    x = xg_t = torch.as_tensor(xg[:, 1, :].copy(), device="cuda")
