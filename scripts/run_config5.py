"""BASELINE.json configs[4]: DPC rollouts + gradients per second at B = 2^20.

x0 sampled along fsds_competition_1 with noise (seed 3), references from MotionPlanner.plan, controls random
within the bounds of dpc/main.py:36-37.  Writes gpurun_out/config5_dpc.json."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ihm2_b200 import dpc  # noqa: E402
from oracle import lib as olib  # noqa: E402  (CPU baseline leg only)

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
DT = torch.float32 if (len(sys.argv) > 2 and sys.argv[2] == "f32") else torch.float64
ESZ = 4 if DT == torch.float32 else 8
g = np.random.default_rng(3)
mp = dpc.MotionPlanner("fsds_competition_1", v_ref=5.0)
s0 = g.uniform(0.0, mp.lap_length, B)
phi_c = np.interp(s0, mp.s_ref, mp.phi_ref)
x0 = np.stack([np.interp(s0, mp.s_ref, mp.X_ref) + g.normal(0, 0.3, B), np.interp(s0, mp.s_ref, mp.Y_ref) + g.normal(0, 0.3, B),
               phi_c + g.normal(0, 0.1, B), np.abs(5.0 + g.normal(0, 1.0, B))], 1)
ref = mp.plan(s0, x0[:, 2])
u = np.stack([g.uniform(-dpc.T_max, dpc.T_max, (B, dpc.Nf)), g.uniform(-dpc.delta_max, dpc.delta_max, (B, dpc.Nf))], -1)
w = dpc.DPCWeights(q_lon=1.0, q_lat=10.0, q_phi=5.0, q_v=1.0, r_T=1e-4, r_delta=2.0, q_lon_f=2.0, q_lat_f=20.0, q_phi_f=10.0, q_v_f=2.0)

dx0, du, dref = [torch.tensor(a, dtype=DT, device="cuda") for a in (x0, u, ref)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    cost, grad = dpc.rollout_grad(dx0, du, dref, w)
ms = []
for _ in range(10):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cost, grad = dpc.rollout_grad(dx0, du, dref, w)
    e1.record()
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
ms_med = float(np.median(ms))
# end to end from pinned host buffers
hx0, hu, href = [torch.tensor(a, dtype=DT).pin_memory() for a in (x0, u, ref)]
hc, hg = torch.empty(B, dtype=DT).pin_memory(), torch.empty(B, dpc.Nf, 2, dtype=DT).pin_memory()
e2e = []
for _ in range(5):
    t0 = time.perf_counter()
    a, b, c = hx0.cuda(non_blocking=True), hu.cuda(non_blocking=True), href.cuda(non_blocking=True)
    cc, gg = dpc.rollout_grad(a, b, c, w)
    hc.copy_(cc, non_blocking=True)
    hg.copy_(gg, non_blocking=True)
    torch.cuda.synchronize()
    e2e.append(time.perf_counter() - t0)
# CPU oracle on a bounded sample, all host cores
n_cpu = min(B, 1 << 16)
olib.dpc_rollout_grad(x0[:1024], u[:1024], ref[:1024], w.as_array())
t0 = time.perf_counter()
oc, og, _ = olib.dpc_rollout_grad(x0[:n_cpu], u[:n_cpu], ref[:n_cpu], w.as_array())
t_cpu = time.perf_counter() - t0
rel = lambda a, b: float(np.abs(a - b).max() / max(1.0, np.abs(b).max()))  # noqa: E731
bytes_per = ESZ * (4 + 2 * dpc.Nf + 4 * (dpc.Nf + 1) + 1 + 2 * dpc.Nf)
out = {
    "workload": f"configs[4]: DPC kin4 rollout N=40 + IHM1.5 cost + gradient in the controls, B = {B}, {'fp32' if ESZ == 4 else 'fp64'}, fsds_competition_1, seed 3",
    "metric": "rollouts+grads/s", "value": B / (ms_med * 1e-3), "ms_per_batch": ms_med, "ms_all": ms,
    "algorithmic_bytes_per_rollout": bytes_per, "algorithmic_GBps": B * bytes_per / (ms_med * 1e-3) / 1e9,
    "e2e": {"value": B / float(np.median(e2e)), "unit": "rollouts+grads/s", "h2d_bytes": int(ESZ * (x0.size + u.size + ref.size)),
            "d2h_bytes": int(ESZ * (B + u.size))},
    "cpu_baseline": {"value": n_cpu / t_cpu, "unit": "rollouts+grads/s", "cores": len(os.sched_getaffinity(0)), "kind": "port",
                     "sample": f"first {n_cpu} rollouts, OpenMP over rollouts"},
    "parity_on_sample": {"cost_rel": rel(cost[:n_cpu].double().cpu().numpy(), oc), "grad_rel": rel(grad[:n_cpu].double().cpu().numpy(), og)},
    "mean_cost": float(cost.mean()),
}
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/config5_dpc%s.json" % ("_f32" if ESZ == 4 else ""), "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k != "ms_all"}))
