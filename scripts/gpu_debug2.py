import os, sys, numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from oracle.lib import Oracle
from scripts.gpu_debug import random_iterate
np.set_printoptions(precision=5, linewidth=200, suppress=False)
mp = load_packaged_track("skidpad"); ng = len(mp.s_ref)
cfg = _abi.make_config(batch=4, n_grid=ng, integrator=_abi.IRK_GL4)
x, u = random_iterate(4, cfg.N, 5)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref); eng.set_iterate(x, u)
A, Bm, phi = [t.cpu().numpy() for t in eng.linearize()]
o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
Ao, Bo, po, bad = o.linearize(x, u)
print("bad", bad)
d = np.abs(phi - po).max(axis=2); print("phi diff per (b,k):\n", d)
b, k = np.unravel_index(np.argmin(d), d.shape)
for (b, k) in [(0, 0), tuple(np.unravel_index(np.argmax(d), d.shape))]:
    print("point", b, k, "x", x[b, k], "u", u[b, k])
    print("phi gpu", phi[b, k]); print("phi ora", po[b, k])
    print("A gpu\n", A[b, k]); print("A ora\n", Ao[b, k]); print("B gpu\n", Bm[b, k].T); print("B ora\n", Bo[b, k].T)
