"""Small driver for ncu: the soft-constrained path at B problems -- RTI solves (con_h_kernel, qp_rti_kernel<40, true>) and
one SQP x 2 + merit solve (nlp_residual_kernel, merit_base_kernel, merit_trial_kernel, solve_finish_kernel)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
mp = load_packaged_track("fsds_default")
W = (mp.right_widths, mp.left_widths)
x_nom, xp, up = nominal_iterate("fsds_default", 30, nonlinear_constraints=True)
for mode in ("rti", "sqp"):
    kw = dict(nlp_solver=_abi.SQP, nlp_max_iter=2, globalization=_abi.MERIT_BACKTRACKING) if mode == "sqp" else {}
    cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nonlinear_constraints=True, **kw)
    x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=5, cfg=cfg)
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=W)
    xd, ud, x0d = [torch.as_tensor(v, device="cuda") for v in (x, u, x0)]
    for _ in range(2):
        eng.set_iterate(xd, ud)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); _, st = eng.solve(x0d); e1.record(); torch.cuda.synchronize()
        print(f"{mode} B={B} solve {e0.elapsed_time(e1):.3f} ms  ipm its/solve {eng.stats()[:, 0].mean().item():.2f}  status {np.bincount(st.cpu().numpy(), minlength=5).tolist()}", flush=True)
    eng.close()
