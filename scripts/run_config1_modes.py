"""BASELINE.json configs[0] (one closed-loop lap on fsds_default, python/main.py:438-517) under the three solver
set-ups: RTI, SQP x2 fixed step, SQP x2 with merit backtracking (the reference's own, controller.py:301-308).
Prints where each run ends.   python scripts/run_config1_modes.py [track=fsds_default] [n_steps=600]"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track

name = sys.argv[1] if len(sys.argv) > 1 else "fsds_default"
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 600
mp = load_packaged_track(name)
res = {}
tol = dict(qp_tolerate_maxiter=True)
for label, kw in (("rti", dict()), ("rti_tolerant", dict(**tol)),
                  ("sqp2_merit_tolerant", dict(nlp_solver=_abi.SQP, nlp_max_iter=2, globalization=_abi.MERIT_BACKTRACKING, **tol)),
                  ("sqp2_fixed", dict(nlp_solver=_abi.SQP, nlp_max_iter=2, globalization=_abi.FIXED_STEP)),
                  ("sqp2_merit", dict(nlp_solver=_abi.SQP, nlp_max_iter=2, globalization=_abi.MERIT_BACKTRACKING)),
                  ("sqp10_merit", dict(nlp_solver=_abi.SQP, nlp_max_iter=10, globalization=_abi.MERIT_BACKTRACKING))):
    cfg = _abi.make_config(batch=1, n_grid=len(mp.s_ref), **kw)
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref, lap_length=np.array([mp.lap_length]))
    eng.reset_iterate()
    x = torch.zeros(1, 8, dtype=torch.float64, device="cuda"); x[:, 0] = -6.0
    out = eng.closed_loop(x, n_steps, log=True)
    done, fail = int(out["done"][0]), int(out["fail"][0])
    xl = out["xlog"][:, 0].cpu().numpy()
    last = (fail if fail >= 0 else done if done >= 0 else n_steps)
    res[label] = {"done_step": done, "fail_step": fail, "s_end": float(xl[last, 0]), "lap_length": mp.lap_length,
                  "lap_time_s": (done + 1) * cfg.dt if done >= 0 else None, "v_x_max": float(np.nanmax(xl[:last + 1, 3])),
                  "n_abs_max": float(np.nanmax(np.abs(xl[:last + 1, 1])))}
    eng.close()
print(json.dumps({"track": name, **res}))
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"track": name, **res}, open(f"gpurun_out/config1_modes_{name}.json", "w"), indent=1)
