"""Small driver for ncu: a few RTI steps of config 2 at a chosen batch size."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 444
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
mp = load_packaged_track("skidpad")
x_nom, xp, up = nominal_iterate("skidpad", 20)
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref))
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
xd, ud, x0d = [torch.as_tensor(v, device="cuda") for v in (x, u, x0)]
for _ in range(reps):
    eng.set_iterate(xd, ud)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.solve(x0d); e1.record(); torch.cuda.synchronize()
    print(f"B={B} solve {e0.elapsed_time(e1):.3f} ms", flush=True)
print("ipm iters mean", eng.stats()[:, 0].mean().item())
import ctypes as C
tr = np.zeros((64, 8)); eng.lib.ihm2b_debug_trace(eng._h, tr.ctypes.data_as(C.c_void_p))
names = ["A rows", "B residuals", "C riccati", "D/G fwd", "E/H rows(+chi)", "F bwd2", "I costate", "J update"]
print("phase cycles per iteration (problem 0):", {n: int(v) for n, v in zip(names, tr[63])}, "sum", int(tr[63].sum()))
