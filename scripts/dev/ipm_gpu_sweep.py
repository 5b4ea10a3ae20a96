"""Step time of config 2 on the GPU under different interior-point start / step settings (make_config keywords)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch

mp = load_packaged_track("skidpad")
x_nom, xp, up = nominal_iterate("skidpad", 20)
B = 4096
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=_abi.make_config(batch=B, n_grid=len(mp.s_ref)))
xd, ud, x0d = [torch.as_tensor(v, device="cuda") for v in (x, u, x0)]
ref = None
for kw in eval(sys.argv[1]):
    eng = Engine(_abi.make_config(batch=B, n_grid=len(mp.s_ref), **kw), mp.s_ref, mp.kappa_ref)
    ts = []
    for _ in range(5):
        eng.set_iterate(xd, ud)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); u0, st = eng.solve(x0d); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    its = eng.stats()[:, 0].cpu().numpy(); st = st.cpu().numpy(); u0 = u0.cpu().numpy()
    if ref is None: ref = (st, u0)
    both = (st == 0) & (ref[0] == 0)
    print(kw, f"{min(ts[1:]):.3f} ms  ok {(st == 0).sum()} mean {its.mean():.2f} max {its.max():.0f}  du0 {np.abs(u0[both] - ref[1][both]).max():.1e}", flush=True)
    eng.close()
