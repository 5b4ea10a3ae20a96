"""BASELINE.json configs[3]: full SQP to the KKT tolerance on the soft-constrained OCP (fkin6 + h(x, u) of
old/generate_acaods_interface.py:312-325 with slacks), N = 80, batch 16384, fsds_competition_1.

  python scripts/run_config4.py [B=16384] [max_sqp_iter=50] [fixed|merit] [Nf=80] [s_target=32] [warm_steps=25]
      -> one JSON line (and gpurun_out/config4_<step>_n<Nf>.json)

Inputs as config 2: the controller's own closed loop advanced warm_steps RTI steps gives the nominal iterate, the
measured states are perturbed around its first point (seed 2), warm start = shifted nominal.  SQP iterations and QPs / s
are reported separately, with the distribution of the KKT residuals the loop stops on.
"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 50
glob = sys.argv[3] if len(sys.argv) > 3 else "fixed"
Nf = int(sys.argv[4]) if len(sys.argv) > 4 else 80
s_target = float(sys.argv[5]) if len(sys.argv) > 5 else 32.0
warm = int(sys.argv[6]) if len(sys.argv) > 6 else 25
track = "fsds_competition_1"
mp = load_packaged_track(track)
kw = dict(Nf=Nf, s_target=s_target, nonlinear_constraints=True)
x_nom, xp, up = nominal_iterate(track, warm, **kw)
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nlp_solver=_abi.SQP, nlp_max_iter=iters,
                       globalization=_abi.MERIT_BACKTRACKING if glob == "merit" else _abi.FIXED_STEP, nlp_tol=1e-6, **kw)
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=2, cfg=cfg)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=(mp.right_widths, mp.left_widths))
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
run = np.isin(st, (0, 2))
q = lambda v: [float(t) for t in np.quantile(v, [0.1, 0.5, 0.9])] if len(v) else None
res = {"config": f"configs[3]: full SQP ({glob} step) on the soft-constrained OCP, N = {Nf}, s_target = {s_target}, {track}, seed 2",
       "nominal_state": [float(t) for t in x_nom],
       "instances": B, "max_sqp_iter": iters, "device_ms": ms, "gpu_launches": int(eng.launch_count - l0),
       "status_counts": {str(k): int(v) for k, v in enumerate(np.bincount(st, minlength=5)) if v},
       "reached_1e-6_kkt_tolerance": int((st == 0).sum()),
       "sqp_iterations_mean": float(stats[:, 1].mean()),
       "sqp_iterations_converged_mean": float(stats[st == 0, 1].mean()) if (st == 0).any() else None,
       "qp_solves": qps, "qp_solves_per_s": qps / (ms * 1e-3), "ipm_iterations_per_qp": float(stats[:, 0].sum() / max(qps, 1)),
       "kkt_residuals_q10_q50_q90": {n: q(stats[run, 3 + i]) for i, n in enumerate(("res_stat", "res_eq", "res_ineq", "res_comp"))},
       "cost_q10_q50_q90": q(stats[run, 7])}
if glob == "merit":
    res["accepted_alpha_last_iteration_hist"] = {f"{a:.3f}": int(c) for a, c in zip(*np.unique(np.round(stats[st == 2, 2], 3), return_counts=True))}
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(f"gpurun_out/config4_{glob}_n{Nf}_s{int(s_target)}.json", "w"), indent=1)
eng.close()
