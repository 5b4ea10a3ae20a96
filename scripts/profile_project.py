"""Device time of the batched projection kernel.  python scripts/profile_project.py [B=1048576]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from ihm2_b200.projection import Track
from ihm2_b200.track import load_packaged_track
from test_oracle_project import sample
from oracle import lib as olib
import time
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
mp = load_packaged_track("fsds_competition_2")
s, n, psi, pose, s_guess = sample(mp, B, seed=4)
trk = Track(mp)
dp, dg = torch.tensor(pose, device="cuda"), torch.tensor(s_guess, device="cuda")
for _ in range(3): trk.project(dp, dg)
ms = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); trk.project(dp, dg); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
nc = min(B, 1 << 18)
t0 = time.perf_counter(); olib.project(mp.s_ref, mp.X_ref, mp.Y_ref, mp.phi_ref, pose[:nc], s_guess[:nc]); tc = time.perf_counter() - t0
print(f"B={B} project {np.median(ms):.4f} ms = {B / np.median(ms) / 1e3:.1f} M poses/s; algorithmic {B * (32 + 28) / np.median(ms) / 1e6:.1f} GB/s; CPU oracle {nc / tc / 1e6:.2f} M poses/s on {len(os.sched_getaffinity(0))} cores")
