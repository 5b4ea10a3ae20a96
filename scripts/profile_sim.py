"""Times the plant-step kernel (Radau-IIA, 100 sub-steps, fkin6 <-> fdyn6 switch) at B = 4096."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
mp = load_packaged_track("fsds_default")
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref))
eng = Engine(cfg, mp.s_ref, mp.kappa_ref, lap_length=[mp.lap_length])
rng = np.random.default_rng(0)
x = np.zeros((B, 8)); x[:, 0] = rng.uniform(0, 100, B); x[:, 3] = rng.uniform(2, 12, B); x[:, 6] = rng.uniform(-100, 300, B)
x[:, 7] = rng.uniform(-0.1, 0.1, B)
u = np.stack([rng.uniform(-300, 400, B), rng.uniform(-0.2, 0.2, B)], axis=1)
xd, ud = torch.as_tensor(x, device="cuda"), torch.as_tensor(u, device="cuda")
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); xn, model = eng.sim_step(xd, ud); e1.record(); torch.cuda.synchronize()
    print(f"B={B} sim_step {e0.elapsed_time(e1):.3f} ms  (fdyn6 share {float((model == 1).float().mean()):.2f})", flush=True)
