"""Parity statistics of the library in ihm2_b200/lib against the oracle on three workloads: config 2 (RTI), the SQP
N = 40 test case (25 iterations), the soft-constrained RTI case -- status agreement, equal iteration counts, max field error."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch
from oracle.lib import Oracle
from oracle.workloads import nominal_iterate_cpu

def run(mp, cfg, x0, x, u, widths=None):
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=widths)
    eng.set_iterate(x, u)
    u0, st = eng.solve(torch.as_tensor(x0, device="cuda"))
    g = dict(st=st.cpu().numpy(), stats=eng.stats().cpu().numpy(), x=eng.get_iterate()[0].cpu().numpy(), u0=u0.cpu().numpy())
    eng.close()
    key = "/tmp/oracle_%s.npz" % abs(hash((cfg.batch, cfg.N, cfg.nlp_solver, cfg.nonlinear_constraints, float(x0.sum())))) 
    if os.path.exists(key):
        d = np.load(key); o = dict(st=d["st"], stats=d["stats"], x=d["x"], u0=d["u0"])
    else:
        u0o, so, sto, xo, uo = Oracle(cfg, mp.s_ref, mp.kappa_ref, widths=widths).solve(x0, x, u)
        o = dict(st=so, stats=sto, x=xo, u0=u0o); np.savez(key, **o)
    return g, o

def report(name, g, o, col):
    ok = (g["st"] == 0) & (o["st"] == 0)
    d = np.abs(g["stats"][ok, col] - o["stats"][ok, col])
    ex = np.abs(g["x"][ok] - o["x"][ok]).reshape(ok.sum(), -1).max(1) / max(1.0, np.abs(o["x"][ok]).max())
    print(f"{name}: status equal {np.mean(g['st'] == o['st']):.4f} (gpu ok {np.sum(g['st'] == 0)}, oracle ok {np.sum(o['st'] == 0)}); "
          f"same count {np.mean(d == 0):.4f}, within 1 {np.mean(d <= 1):.4f}, max diff {d.max():.0f}; x err q50 {np.median(ex):.1e} q99 {np.quantile(ex, 0.99):.1e} max {ex.max():.1e}", flush=True)

sk = load_packaged_track("skidpad"); fs = load_packaged_track("fsds_default")
x_nom, xp, up = nominal_iterate_cpu("skidpad", 20)
B = 4096
cfg = _abi.make_config(batch=B, n_grid=len(sk.s_ref))
g, o = run(sk, cfg, *perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg))
report("config 2 (IPM iterations)", g, o, 0)
B = 512
cfg = _abi.make_config(batch=B, n_grid=len(sk.s_ref), nlp_solver=_abi.SQP, nlp_max_iter=25, globalization=_abi.FIXED_STEP)
g, o = run(sk, cfg, *perturbed_batch(x_nom, xp, up, B, seed=2))
report("SQP N=40 x 25 (SQP iterations)", g, o, 1)
B = 1024
x_nom, xp, up = nominal_iterate_cpu("fsds_default", 30, nonlinear_constraints=True)
cfg = _abi.make_config(batch=B, n_grid=len(fs.s_ref), nonlinear_constraints=True)
g, o = run(fs, cfg, *perturbed_batch(x_nom, xp, up, B, seed=5, cfg=cfg), widths=(fs.right_widths, fs.left_widths))
report("soft RTI (IPM iterations)", g, o, 0)
