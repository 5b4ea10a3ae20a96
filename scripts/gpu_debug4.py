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
x, u = initial_iterate(cfg)
x0 = np.zeros((1, 8)); x0[0, 0] = -6
from ihm2_b200.workloads import shift_np
xs, us = shift_np(x, u)
eng.set_iterate(xs, us)
u0, st = eng.solve(torch.as_tensor(x0, device="cuda")); torch.cuda.synchronize()
tr = np.zeros((64, 8)); eng.lib.ihm2b_debug_trace(eng._h, tr.ctypes.data_as(C.c_void_p))
print("gpu status", st.cpu().numpy(), "stats", eng.stats().cpu().numpy())
print("trace: res_g res_b res_d res_m mu alpha mu_t")
print(tr[:30, :7])
o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
q = o.qp_debug(x0[0], xs[0], us[0])
print("oracle info", q["info"], "status", q["status"])
xg, ug = [t.cpu().numpy() for t in eng.get_iterate()]
print("u gpu", ug[0, :4].ravel(), "\nu ora", q["z"][:4, :2].ravel())
