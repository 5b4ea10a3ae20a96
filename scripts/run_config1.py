"""BASELINE.json configs[0]: one closed-loop lap on data/fsds_default, N = 40, on the device and
on the CPU oracle, with the lap time of python/main.py:561-570 on both sides.

  python scripts/run_config1.py [n_steps=1000] [s_target=24] [P_CROSS=1]  -> one JSON line (and gpurun_out/config1_lap_cross<P_CROSS>.json)

Set-up that laps: soft constraints h(x, u) + slacks, acados' tolerance of a QP at its iteration limit, on the literal
fdyn6 plant (P_CROSS = 1: crosswise slip angles, python/models.py:543-546) or with the straight assignment (0).  B = 1 like the reference's loop; 8 more cars with masses +-3 % show the spread
between two closed loops that differ in the last bits."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from oracle.lib import Oracle, initial_iterate

n_steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
s_target = float(sys.argv[2]) if len(sys.argv) > 2 else 24.0
cross = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
mp = load_packaged_track("fsds_default")
B = 9
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nonlinear_constraints=True, qp_tolerate_maxiter=True, s_target=s_target)
params = np.tile(_abi.default_params(), (B, 1)); params[:, _abi.P_CROSS] = cross
params[1:, _abi.P_M] *= np.linspace(0.97, 1.03, B - 1)
W = (mp.right_widths, mp.left_widths)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, params=params, lap_length=[mp.lap_length], widths=W)
x = torch.zeros(B, 8, dtype=torch.float64, device="cuda"); x[:, 0] = -6.0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); out = eng.closed_loop(x, n_steps); e1.record(); torch.cuda.synchronize()
gpu_ms = e0.elapsed_time(e1)
xi, ui = initial_iterate(cfg)
xs = np.zeros((B, 8)); xs[:, 0] = -6.0
t = time.time()
o = Oracle(cfg, mp.s_ref, mp.kappa_ref, params=params, widths=W).closed_loop(xs, xi, ui, [mp.lap_length], n_steps)
cpu_s = time.time() - t
res = {"config": f"configs[0]: one closed-loop lap, fsds_default, N = 40, s_target = {s_target}, soft constraints, qp_tolerate_maxiter, P_CROSS = {cross:g}",
       "lap_length_m": float(mp.lap_length),
       "gpu": {"done_step": out["done"].cpu().tolist(), "fail_step": out["fail"].cpu().tolist(), "lap_time_s": out["lap_time"].cpu().tolist(),
               "device_ms_9_cars": gpu_ms},
       "oracle": {"done_step": o["done"].tolist(), "fail_step": o["fail"].tolist(), "lap_time_s": o["lap_time"].tolist(), "wall_s_9_cars": cpu_s},
       "nominal_car_lap_time_s": {"gpu": float(out["lap_time"][0]), "oracle": float(o["lap_time"][0])}}
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(f"gpurun_out/config1_lap_cross{int(cross)}.json", "w"), indent=1)
eng.close()
