"""How much of the QP launch is its longest problem?  Config 2: the 4096 batch, its 64 longest problems alone, and the
batch with the problems above a given iteration count replaced by typical ones."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch

mp = load_packaged_track("skidpad")
x_nom, xp, up = nominal_iterate("skidpad", 20)
B = 4096
cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref))
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)

def timed(x0, x, u, reps=4):
    b = len(x0)
    eng = Engine(_abi.make_config(batch=b, n_grid=len(mp.s_ref)), mp.s_ref, mp.kappa_ref)
    xd, ud, x0d = [torch.as_tensor(np.ascontiguousarray(v), device="cuda") for v in (x, u, x0)]
    eng.profile_enable(True) if hasattr(eng, "profile_enable") else None
    ts = []
    for _ in range(reps):
        eng.set_iterate(xd, ud)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng.solve(x0d); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    its = eng.stats()[:, 0].cpu().numpy()
    st = eng.status().cpu().numpy() if hasattr(eng, "status") else None
    eng.close()
    return min(ts[1:]), its

t_all, its = timed(x0, x, u)
print(f"all 4096: {t_all:.3f} ms, iterations mean {its.mean():.2f} max {its.max():.0f}; hist >=20: {(its >= 20).sum()}, >=25: {(its >= 25).sum()}, >=30: {(its >= 30).sum()}")
top = np.argsort(-its)[:64]
t_top, it_top = timed(x0[top], x[top], u[top])
print(f"64 longest alone: {t_top:.3f} ms (iterations {it_top.min():.0f}..{it_top.max():.0f})")
one = top[:1]
t_one, it_one = timed(x0[one], x[one], u[one])
print(f"the longest alone: {t_one:.3f} ms ({it_one[0]:.0f} iterations) -> {t_one / it_one[0] * 1e3:.1f} us per iteration, lineariser included")
typ = np.argsort(its)[B // 2]
for cap in (30, 25, 20, 16):
    idx = np.where(its > cap, typ, np.arange(B))
    t_c, it_c = timed(x0[idx], x[idx], u[idx])
    print(f"problems above {cap} iterations replaced by a median one ({(its > cap).sum()}): {t_c:.3f} ms, mean {it_c.mean():.2f}")
