"""Times the linearisation kernel alone (config 2 inputs, B = 4096)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
mp = load_packaged_track("skidpad")
x_nom, xp, up = nominal_iterate("skidpad", 20)
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref))
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
eng.set_iterate(torch.as_tensor(x, device="cuda"), torch.as_tensor(u, device="cuda"))
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.lib.ihm2b_linearize_dev(eng._h, None, None, None, None); e1.record(); torch.cuda.synchronize()
    print(f"B={B} linearize {e0.elapsed_time(e1):.3f} ms", flush=True)
