"""Batched Stanley baseline (new_python/controller.py:162-234) around the Cartesian kin6 / dyn6 plant and the projection
kernel: laps of B cars with randomised mass / tyre parameters.   python scripts/run_stanley_loop.py [track] [B] [v_ref]"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.cartesian_loop import closed_loop_cartesian
from ihm2_b200.controller import StanleyController
from ihm2_b200.engine import Engine
from ihm2_b200.projection import Track
from ihm2_b200.track import load_packaged_track

name = sys.argv[1] if len(sys.argv) > 1 else "fsds_competition_1"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
v_ref = float(sys.argv[3]) if len(sys.argv) > 3 else 5.0
mp = load_packaged_track(name)
n_steps = int(1.25 * mp.lap_length / v_ref / 0.05) + 100
g = torch.Generator().manual_seed(1)
r = torch.rand(B, 4, generator=g, dtype=torch.float64).numpy()
p = np.tile(_abi.default_params(), (B, 1))
f_m = 0.9 + 0.2 * r[:, 0]
p[:, _abi.P_M] *= f_m; p[:, _abi.P_IZ] *= f_m
p[:, _abi.P_DA] *= 0.85 + 0.3 * r[:, 1]; p[:, _abi.P_BA] *= 0.85 + 0.3 * r[:, 2]
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), sim_model=_abi.SIM_KIN6C_DYN6C)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, params=p, lap_length=np.array([mp.lap_length]))
trk = Track(mp)
i0 = len(mp.s_ref) // 3
state = torch.zeros(B, 8, dtype=torch.float64, device="cuda")
state[:, 0], state[:, 1], state[:, 2] = float(mp.X_ref[i0]), float(mp.Y_ref[i0]), float(mp.phi_ref[i0])
state[:, 1] += 0.3 * (torch.rand(B, generator=g, dtype=torch.float64).cuda() - 0.5) + 0.01
stanley = StanleyController()
s_tab = torch.tensor(mp.s_ref, device="cuda"); k_tab = torch.tensor(mp.kappa_ref, device="cuda")


def controller(frenet, st):
    idx = torch.searchsorted(s_tab, frenet[:, 0].contiguous()).clamp(1, len(s_tab) - 1)
    w = (frenet[:, 0] - s_tab[idx - 1]) / (s_tab[idx] - s_tab[idx - 1])
    kappa = k_tab[idx - 1] + w * (k_tab[idx] - k_tab[idx - 1])
    return stanley.compute_control(frenet[:, 1], frenet[:, 2], st[:, 3], torch.full_like(kappa, v_ref), kappa).to(st.device)


torch.cuda.synchronize(); t0 = time.perf_counter()
out = closed_loop_cartesian(eng, trk, state, n_steps, controller=controller)
torch.cuda.synchronize(); wall = time.perf_counter() - t0
laps, lap_step, fail = [out[k].cpu().numpy() for k in ("laps", "lap_step", "fail")]
lt = (lap_step[lap_step >= 0] + 1) * cfg.dt
res = {"what": "batched Stanley baseline, Cartesian kin6/dyn6 plant (speed switch at 3 m/s), projection kernel; randomised mass and tyre parameters (seed 1)",
       "track": name, "instances": B, "v_ref": v_ref, "n_steps": n_steps, "lap_length": mp.lap_length, "wall_s": wall,
       "laps_completed": int((laps > 0).sum()), "failed": int((fail >= 0).sum()),
       "lap_time_s": {"median": float(np.median(lt)), "min": float(lt.min()), "max": float(lt.max())} if len(lt) else None,
       "ideal_lap_time_at_v_ref_s": mp.lap_length / v_ref}
print(json.dumps(res))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(f"gpurun_out/stanley_loop_{name}.json", "w"), indent=1)
eng.close()
