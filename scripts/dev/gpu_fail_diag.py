"""Config-3 style closed loop stepped from Python: which statuses stop the instances, and where (development aid)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
B = int(os.environ.get("B", 3072)); n_steps = int(os.environ.get("STEPS", 700))
kw = eval(sys.argv[1]) if len(sys.argv) > 1 else {}
names = ["fsds_competition_1", "fsds_competition_2", "fsds_competition_3"]
mps = [load_packaged_track(n) for n in names]
s_ref = np.stack([m.s_ref for m in mps]); kappa_ref = np.stack([m.kappa_ref for m in mps])
widths = (np.stack([m.right_widths for m in mps]), np.stack([m.left_widths for m in mps]))
track = (np.arange(B) % 3).astype(np.int32)
g = torch.Generator().manual_seed(1)
r = torch.rand(B, 8, generator=g, dtype=torch.float64).numpy()
p = np.tile(_abi.default_params(), (B, 1))
f_m = 0.9 + 0.2 * r[:, 0]
p[:, _abi.P_M] *= f_m; p[:, _abi.P_IZ] *= f_m
p[:, _abi.P_DA] *= 0.85 + 0.3 * r[:, 1]; p[:, _abi.P_BA] *= 0.85 + 0.3 * r[:, 2]
for j, c in enumerate((_abi.P_CR0, _abi.P_CR1, _abi.P_CR2)):
    p[:, c] *= 0.9 + 0.2 * r[:, 3 + j]
p[:, _abi.P_CM0] *= 0.95 + 0.1 * r[:, 6]
p[:, _abi.P_CROSS] = 0.0
cfg = _abi.make_config(batch=B, n_grid=s_ref.shape[1], n_tracks=3, s_target=24.0, nonlinear_constraints=True, qp_tolerate_maxiter=True, **kw)
eng = Engine(cfg, s_ref, kappa_ref, track_index=track, params=p, widths=widths)
x = torch.zeros(B, 8, dtype=torch.float64, device="cuda"); x[:, 0] = -6.0
alive = torch.ones(B, dtype=torch.bool, device="cuda")
fail_status = np.full(B, -1); fail_step = np.full(B, -1); fail_x = np.zeros((B, 8)); fail_stats = np.zeros((B, 8))
for step in range(n_steps):
    eng.shift_iterate()
    u0, status = eng.solve(x)
    bad = alive & ~((status == 0) | (status == 2))
    if bad.any():
        idx = bad.nonzero().flatten().cpu().numpy()
        fail_status[idx] = status[bad].cpu().numpy(); fail_step[idx] = step
        fail_x[idx] = x[bad].cpu().numpy(); fail_stats[idx] = eng.stats()[bad].cpu().numpy()
    alive &= ~bad
    xn, _ = eng.sim_step(x, u0)
    nan = alive & ~torch.isfinite(xn).all(1)
    if nan.any():
        idx = nan.nonzero().flatten().cpu().numpy(); fail_status[idx] = 99; fail_step[idx] = step
    alive &= ~nan
    x = torch.where(alive[:, None], xn, x)
print("alive", int(alive.sum()), "of", B, "after", n_steps, "steps; failure statuses", dict(zip(*np.unique(fail_status[fail_status >= 0], return_counts=True))))
f = fail_status >= 0
for t in range(3):
    m = f & (track == t)
    print(names[t], "failed", m.sum(), "of", (track == t).sum(), "median step", np.median(fail_step[m]) if m.any() else None,
          "median s", np.median(fail_x[m, 0]) if m.any() else None)
i = np.nonzero(f)[0][:12]
for j in i:
    print(j, "status", fail_status[j], "step", fail_step[j], "x", np.array2string(fail_x[j], precision=2), "stats", np.array2string(fail_stats[j, :7], precision=2))
