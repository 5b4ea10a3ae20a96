"""Development check on a GPU box: product kernels vs the CPU oracle, with detailed diffs.
Not a test and not a bench; run as `python scripts/gpu_debug.py [B]`."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi  # noqa: E402
from ihm2_b200.engine import Engine  # noqa: E402
from ihm2_b200.track import load_packaged_track  # noqa: E402
from ihm2_b200.workloads import nominal_iterate, perturbed_batch  # noqa: E402
from oracle.lib import Oracle  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
mp = load_packaged_track("skidpad")
ng = len(mp.s_ref)
dev = torch.device("cuda")


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(1.0, np.max(np.abs(b))))


def random_iterate(B, N, seed):
    rng = np.random.default_rng(seed)
    x = np.zeros((B, N + 1, 8)); u = np.zeros((B, N, 2))
    x[..., 0] = rng.uniform(-5, 60, (B, N + 1)); x[..., 1] = rng.uniform(-1.5, 1.5, (B, N + 1))
    x[..., 2] = rng.uniform(-0.4, 0.4, (B, N + 1)); x[..., 3] = rng.uniform(0.0, 15, (B, N + 1))
    x[..., 4] = rng.uniform(-1, 1, (B, N + 1)); x[..., 5] = rng.uniform(-1, 1, (B, N + 1))
    x[..., 6] = rng.uniform(-300, 300, (B, N + 1)); x[..., 7] = rng.uniform(-0.4, 0.4, (B, N + 1))
    u[..., 0] = x[:, :-1, 6] + rng.uniform(-0.5, 0.5, (B, N)); u[..., 1] = x[:, :-1, 7] + rng.uniform(-0.02, 0.02, (B, N))
    return x, u



if __name__ == "__main__":
    for name, integ in (("ERK4", _abi.ERK4), ("IRK_GL4", _abi.IRK_GL4)):
        Bl = 64
        cfg = _abi.make_config(batch=Bl, n_grid=ng, integrator=integ)
        x, u = random_iterate(Bl, cfg.N, 5)
        eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
        eng.set_iterate(x, u)
        A, Bm, phi = [t.cpu().numpy() for t in eng.linearize()]
        o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
        Ao, Bo, po, bad = o.linearize(x, u)
        print(f"[linearize {name}] bad={bad} relA={rel(A, Ao):.3e} relB={rel(Bm, Bo):.3e} relphi={rel(phi, po):.3e} "
              f"absA(no T row)={np.max(np.abs(A - Ao)[:, :, :6]):.3e} nan={np.isnan(A).sum()}", flush=True)
        eng.close()

    # ---- RTI parity (IRK)
    t0 = time.time()
    x_nom, xp, up = nominal_iterate("skidpad", 20)
    print("nominal x:", x_nom, f"({time.time() - t0:.1f}s)", flush=True)
    cfg = _abi.make_config(batch=B, n_grid=ng)
    x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
    eng.set_iterate(x, u)
    x0d = torch.as_tensor(x0, device=dev)
    u0, status = eng.solve(x0d)
    torch.cuda.synchronize()
    xg, ug = [t.cpu().numpy() for t in eng.get_iterate()]
    st = eng.stats().cpu().numpy()
    o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
    t0 = time.time()
    u0o, so, sto, xo, uo = o.solve(x0, x, u)
    tcpu = time.time() - t0
    print(f"[rti] gpu status counts {np.bincount(status.cpu().numpy(), minlength=5)} oracle {np.bincount(so, minlength=5)}")
    print(f"[rti] ipm iters gpu mean {st[:, 0].mean():.2f} max {st[:, 0].max():.0f} | oracle mean {sto[:, 0].mean():.2f} max {sto[:, 0].max():.0f}")
    sg = status.cpu().numpy()
    ok = (sg == 0) & (so == 0)
    print(f"[rti] status agree {np.mean(sg == so):.4f}; both ok {ok.sum()}")
    u0g = u0.cpu().numpy()
    u0g, u0o, xg, xo, ug, uo = u0g[ok], u0o[ok], xg[ok], xo[ok], ug[ok], uo[ok]
    print(f"[rti] rel u0 {rel(u0g, u0o):.3e} x {rel(xg, xo):.3e} u {rel(ug, uo):.3e}")
    err = np.abs(xg - xo).reshape(len(xg), -1).max(axis=1)
    print("[rti] per-problem max abs x err quantiles", np.quantile(err, [0.5, 0.9, 0.99, 1.0]))
    for c in range(8):
        print(f"   x[{c}] rel {rel(xg[..., c], xo[..., c]):.3e}", end="")
    print()
    for c in range(2):
        print(f"   u[{c}] rel {rel(ug[..., c], uo[..., c]):.3e}", end="")
    print(f"\n[rti] oracle cpu time {tcpu:.2f}s for {B} problems", flush=True)

    # ---- timing
    for rep in range(3):
        eng.set_iterate(x, u)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.solve(x0d)
        e1.record()
        torch.cuda.synchronize()
        e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eng.set_iterate(x, u)
        e2.record(); eng.linearize(); e3.record(); torch.cuda.synchronize()
        print(f"[time] linearize(+split) {e2.elapsed_time(e3):.3f} ms")
        print(f"[time] B={B} solve {e0.elapsed_time(e1):.3f} ms -> {B / e0.elapsed_time(e1) * 1e3:.0f} solves/s")
