"""BASELINE.json configs[2]: closed-loop laps over fsds_competition_1/2/3 with randomised tyre / mass
parameters, all on the device: shift -> RTI solve -> plant step per 50 ms.

  python scripts/run_config3.py [B=65536] [n_steps=600]     -> one JSON line (and gpurun_out/config3.json)

Instance i: track i % 3; m, I_z * U(0.9, 1.1) (same factor); Pacejka Da, Ba * U(0.85, 1.15); C_r0..2 *
U(0.9, 1.1); C_m0 * U(0.95, 1.05); seed 1.  The controller model sees the instance's mass / drivetrain
parameters too (one parameter vector per instance).  A lap counts when s > L + 1 (python/main.py:561-570);
an instance stops at its first solver failure (status not in {0, 2}), as the reference loop does.
"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 600
names = ["fsds_competition_1", "fsds_competition_2", "fsds_competition_3"]
mps = [load_packaged_track(n) for n in names]
s_ref = np.stack([m.s_ref for m in mps]); kappa_ref = np.stack([m.kappa_ref for m in mps])
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
cfg = _abi.make_config(batch=B, n_grid=s_ref.shape[1], n_tracks=3)
eng = Engine(cfg, s_ref, kappa_ref, track_index=track, params=p, lap_length=lap)
eng.reset_iterate()
x = torch.zeros(B, 8, dtype=torch.float64, device="cuda"); x[:, 0] = -6.0     # python/main.py:438-439
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
l0 = eng.launch_count
e0.record(); out = eng.closed_loop(x, n_steps); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
done, fail = out["done"].cpu().numpy(), out["fail"].cpu().numpy()
steps_run = np.where(fail >= 0, fail + 1, np.where(done >= 0, done + 1, n_steps))   # controller calls per instance
res = {"config": "configs[2]: closed-loop laps, fsds_competition_1/2/3, randomised parameters, seed 1",
       "instances": B, "n_steps": n_steps, "device_ms": ms, "gpu_launches": int(eng.launch_count - l0),
       "laps_finished": int((done >= 0).sum()), "stopped_on_solver_failure": int((fail >= 0).sum()),
       "still_running": int(((done < 0) & (fail < 0)).sum()),
       "median_failure_step": float(np.median(fail[fail >= 0])) if (fail >= 0).any() else None,
       "lap_time_s": {n: (float((np.median(done[(done >= 0) & (track == i)]) + 1) * cfg.dt) if ((done >= 0) & (track == i)).any() else None)
                      for i, n in enumerate(names)},
       "rti_solves": int(steps_run.sum()), "rti_solves_per_s": float(steps_run.sum() / (ms * 1e-3)),
       "s_final_median": float(np.median(out["x"][:, 0].cpu().numpy()))}
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/config3.json", "w"))
eng.close()
