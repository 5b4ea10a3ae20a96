"""Closed loop with the Cartesian kin6 / dyn6 plants (ROS sim node semantics) + projection + NMPC, one car per
track (or B copies).  python scripts/run_cartesian_loop.py [track=fsds_competition_1] [n_steps=1200] [B=1] [mode=rti] [s_target=40]"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.cartesian_loop import closed_loop_cartesian
from ihm2_b200.engine import Engine
from ihm2_b200.projection import Track
from ihm2_b200.track import load_packaged_track

name = sys.argv[1] if len(sys.argv) > 1 else "fsds_competition_1"
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1200
B = int(sys.argv[3]) if len(sys.argv) > 3 else 1
mode = sys.argv[4] if len(sys.argv) > 4 else "rti"
s_target = float(sys.argv[5]) if len(sys.argv) > 5 else 40.0
mp = load_packaged_track(name)
kw = dict(sim_model=_abi.SIM_KIN6C_DYN6C, s_target=s_target)
if mode.startswith("sqp"):
    kw.update(nlp_solver=_abi.SQP, nlp_max_iter=2, globalization=_abi.MERIT_BACKTRACKING)
if mode.endswith("tolerant"):
    kw.update(qp_tolerate_maxiter=True)
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), **kw)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, lap_length=np.array([mp.lap_length]))
eng.reset_iterate()
trk = Track(mp)
# start on the centre line a little after the start of the table, standing still, heading along the reference
i0 = len(mp.s_ref) // 3        # first sample of the middle copy: s = 0
state = torch.zeros(B, 8, dtype=torch.float64, device="cuda")
state[:, 0], state[:, 1], state[:, 2] = float(mp.X_ref[i0]), float(mp.Y_ref[i0]), float(mp.phi_ref[i0])
state[:, 1] += 0.05            # off the centre line: the control node refuses an exact zero offset
if B > 1:
    g = torch.Generator().manual_seed(0)
    state[:, 1] += 0.3 * (torch.rand(B, generator=g, dtype=torch.float64).cuda() - 0.5)
x_it, u_it = eng.get_iterate()
x_it[:, :, 0] += 6.0           # the reference's initial guess starts at s = -6 (controller.py:313-317)
eng.set_iterate(x_it, None)
torch.cuda.synchronize(); t0 = time.perf_counter()
out = closed_loop_cartesian(eng, trk, state, n_steps, log=(B == 1))
torch.cuda.synchronize(); wall = time.perf_counter() - t0
laps, lap_step, fail = [out[k].cpu().numpy() for k in ("laps", "lap_step", "fail")]
res = {"track": name, "mode": mode, "s_target": s_target, "instances": B, "n_steps": n_steps, "lap_length": mp.lap_length, "wall_s": wall,
       "laps_completed": int((laps > 0).sum()), "failed": int((fail >= 0).sum()),
       "first_lap_time_s": float(np.median((lap_step[lap_step >= 0] + 1) * cfg.dt)) if (lap_step >= 0).any() else None,
       "median_fail_step": float(np.median(fail[fail >= 0])) if (fail >= 0).any() else None,
       "s_end_median": float(np.median(out["s"].cpu().numpy())), "laps_max": int(laps.max())}
if B == 1:
    P = out["poses"][:, 0].cpu().numpy()
    res["v_x_max"] = float(np.nanmax(P[:, 3])); res["v_x_mean"] = float(np.nanmean(P[:, 3]))
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(f"gpurun_out/cartesian_loop_{name}_{mode}_{B}_st{int(s_target)}.json", "w"), indent=1)
eng.close()
