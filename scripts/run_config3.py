"""BASELINE.json configs[2]: closed-loop laps over fsds_competition_1/2/3 with randomised tyre / mass
parameters, all on the device: shift -> RTI solve -> plant step per 50 ms.

  python scripts/run_config3.py [B=65536] [n_steps=1200] [setup=lapping|literal|hard] [s_target=24]
      -> one JSON line (and gpurun_out/config3_<setup>.json)

Instance i: track i % 3; m, I_z * U(0.9, 1.1) (same factor); Pacejka Da, Ba * U(0.85, 1.15); C_r0..2 *
U(0.9, 1.1); C_m0 * U(0.95, 1.05); seed 1.  The controller model sees the instance's mass / drivetrain
parameters too (one parameter vector per instance).  An instance is done when s > L + 1 (python/main.py:515) and
stops at its first solver failure (status not in {0, 2}), as the reference loop does; the lap time is the
reference's (python/main.py:561-570).

Set-ups:
  literal  soft constraints h(x, u) + slacks, acados' tolerance of a QP at its iteration limit
           (mpc_control_node.cpp:190-197), literal fdyn6 plant (crosswise slip angles of python/models.py:543-546)
  lapping  the same controller, straight slip-angle assignment in the plant (P_CROSS = 0)
  hard     the new_python OCP (hard bounds, no h rows) on the literal plant: the round-1 set-up
"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1200
setup = sys.argv[3] if len(sys.argv) > 3 else "lapping"
s_target = float(sys.argv[4]) if len(sys.argv) > 4 else 24.0
names = ["fsds_competition_1", "fsds_competition_2", "fsds_competition_3"]
mps = [load_packaged_track(n) for n in names]
s_ref = np.stack([m.s_ref for m in mps]); kappa_ref = np.stack([m.kappa_ref for m in mps])
widths = (np.stack([m.right_widths for m in mps]), np.stack([m.left_widths for m in mps]))
track = (np.arange(B) % 3).astype(np.int32)
lap = np.array([m.lap_length for m in mps])      # per track
g = torch.Generator().manual_seed(1)
r = torch.rand(B, 8, generator=g, dtype=torch.float64).numpy()
p = np.tile(_abi.default_params(), (B, 1))
f_m = 0.9 + 0.2 * r[:, 0]
p[:, _abi.P_M] *= f_m; p[:, _abi.P_IZ] *= f_m
p[:, _abi.P_DA] *= 0.85 + 0.3 * r[:, 1]; p[:, _abi.P_BA] *= 0.85 + 0.3 * r[:, 2]
for j, c in enumerate((_abi.P_CR0, _abi.P_CR1, _abi.P_CR2)):
    p[:, c] *= 0.9 + 0.2 * r[:, 3 + j]
p[:, _abi.P_CM0] *= 0.95 + 0.1 * r[:, 6]
nl = setup != "hard"
if setup == "lapping":
    p[:, _abi.P_CROSS] = 0.0
cfg = _abi.make_config(batch=B, n_grid=s_ref.shape[1], n_tracks=3, s_target=s_target, nonlinear_constraints=nl,
                       qp_tolerate_maxiter=nl)
eng = Engine(cfg, s_ref, kappa_ref, track_index=track, params=p, lap_length=lap, widths=widths if nl else None)
eng.reset_iterate()
x = torch.zeros(B, 8, dtype=torch.float64, device="cuda"); x[:, 0] = -6.0     # python/main.py:438-439
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
l0 = eng.launch_count
e0.record(); out = eng.closed_loop(x, n_steps); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
done, fail = out["done"].cpu().numpy(), out["fail"].cpu().numpy()
lap_t, cost = out["lap_time"].cpu().numpy(), out["cost_sum"].cpu().numpy()
steps_run = np.where(fail >= 0, fail + 1, np.where(done >= 0, done + 1, n_steps))   # controller calls per instance
fin = done >= 0
q = lambda v: [float(t) for t in np.quantile(v, [0.0, 0.1, 0.5, 0.9, 1.0])] if len(v) else None
res = {"config": f"configs[2]: closed-loop laps, fsds_competition_1/2/3, randomised parameters, seed 1; set-up '{setup}', s_target = {s_target}",
       "instances": B, "n_steps": n_steps, "device_ms": ms, "gpu_launches": int(eng.launch_count - l0),
       "laps_finished": int(fin.sum()), "dnf": int((~fin).sum()), "stopped_on_solver_failure": int((fail >= 0).sum()),
       "still_running": int(((done < 0) & (fail < 0)).sum()),
       "median_failure_step": float(np.median(fail[fail >= 0])) if (fail >= 0).any() else None,
       "lap_time_s_min_q10_q50_q90_max": {n: q(lap_t[fin & (track == i) & np.isfinite(lap_t)]) for i, n in enumerate(names)},
       "finished_per_track": {n: int((fin & (track == i)).sum()) for i, n in enumerate(names)},
       "mean_cost_per_step_finished": float((cost[fin] / steps_run[fin]).mean()) if fin.any() else None,
       "rti_solves": int(steps_run.sum()), "rti_solves_per_s": float(steps_run.sum() / (ms * 1e-3)),
       "s_final_median": float(np.median(out["x"][:, 0].cpu().numpy()))}
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(f"gpurun_out/config3_{setup}.json", "w"), indent=1)
eng.close()
