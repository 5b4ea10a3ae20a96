"""BASELINE.json configs[3] as far as this build goes: full SQP (fixed step) to the KKT tolerance, N = 80, batch 16384,
on fsds_competition_1 -- WITHOUT the nonlinear h(x, u) rows and slacks of old/generate_acaods_interface.py (not built),
i.e. the new_python OCP (python/mpc.py:29-101) on the longer horizon.

  python scripts/run_config4.py [B=16384] [max_sqp_iter=50] [fixed|merit] [Nf=80]  -> one JSON line (and gpurun_out/config4[_merit].json)

Inputs: slow straight warm start (v = 5 m/s, T = 30), measured states perturbed around its first point (seed 2).
"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 50
glob = sys.argv[3] if len(sys.argv) > 3 else "fixed"
Nf = int(sys.argv[4]) if len(sys.argv) > 4 else 80
mp = load_packaged_track("fsds_competition_1")
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), Nf=Nf, s_target=40.0, nlp_solver=_abi.SQP, nlp_max_iter=iters,
                       globalization=_abi.MERIT_BACKTRACKING if glob == "merit" else _abi.FIXED_STEP, nlp_tol=1e-6)
rng = np.random.default_rng(2)
v0 = 5.0
x = np.zeros((B, Nf + 1, 8)); u = np.zeros((B, Nf, 2))
x[:, :, 0] = -6 + v0 * cfg.dt * np.arange(Nf + 1)[None, :]; x[:, :, 3] = v0; x[:, :, 6] = 30.0; u[:, :, 0] = 30.0
x0 = x[:, 0, :] + rng.uniform(-1, 1, (B, 8)) * np.array([0.5, 0.3, 0.05, 1.0, 0.1, 0.1, 20, 0.02])
eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
xd, ud, x0d = [torch.as_tensor(v, device="cuda") for v in (x, u, x0)]
times = []
for rep in range(2):
    eng.set_iterate(xd, ud)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = eng.launch_count
    e0.record(); u0, status = eng.solve(x0d); e1.record(); torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
ms = min(times)
st = status.cpu().numpy(); stats = eng.stats().cpu().numpy()
qps = float(stats[:, 1].sum())
res = {"config": f"configs[3] without h(x,u) / slacks: full SQP ({glob} step), N = {Nf}, fsds_competition_1, seed 2",
       "instances": B, "max_sqp_iter": iters, "device_ms": ms, "gpu_launches": int(eng.launch_count - l0),
       "status_counts": {str(k): int(v) for k, v in enumerate(np.bincount(st, minlength=5)) if v},
       "sqp_iterations_mean": float(stats[:, 1].mean()), "sqp_iterations_converged_mean": float(stats[st == 0, 1].mean()) if (st == 0).any() else None,
       "qp_solves": qps, "qp_solves_per_s": qps / (ms * 1e-3), "ipm_iterations_per_qp": float(stats[:, 0].sum() / max(qps, 1)),
       "kkt_residual_max_converged": float(stats[st == 0, 3:7].max()) if (st == 0).any() else None}
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
if glob == "merit":
    res["accepted_alpha_last_iteration_hist"] = {f"{a:.3f}": int(c) for a, c in zip(*np.unique(np.round(stats[st == 2, 2], 3), return_counts=True))}
json.dump(res, open("gpurun_out/config4%s%s.json" % ("_merit" if glob == "merit" else "", "" if Nf == 80 else f"_n{Nf}"), "w"))
eng.close()
