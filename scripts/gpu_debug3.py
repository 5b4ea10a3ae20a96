import os, sys, numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ihm2_b200 import _abi
from ihm2_b200.engine import Engine
from ihm2_b200.track import load_packaged_track
from ihm2_b200.workloads import nominal_iterate, perturbed_batch
from oracle.lib import Oracle
np.set_printoptions(precision=4, linewidth=220)
mp = load_packaged_track("skidpad"); ng = len(mp.s_ref)
x_nom, xp, up = nominal_iterate("skidpad", 20)
B = 64
cfg = _abi.make_config(batch=B, n_grid=ng)
x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
eng = Engine(cfg, mp.s_ref, mp.kappa_ref); eng.set_iterate(x, u)
u0, status = eng.solve(torch.as_tensor(x0, device="cuda")); torch.cuda.synchronize()
xg, ug = [t.cpu().numpy() for t in eng.get_iterate()]
st = eng.stats().cpu().numpy(); sg = status.cpu().numpy()
cfg1 = _abi.make_config(batch=1, n_grid=ng)
o1 = Oracle(cfg1, mp.s_ref, mp.kappa_ref)
N = cfg.N
def objective(q, z):
    return sum(0.5 * z[k] @ q["H"][k] @ z[k] + q["g"][k] @ z[k] for k in range(N + 1))
for b in range(12):
    q = o1.qp_debug(x0[b], x[b], u[b])
    zo = q["z"]
    zg = np.zeros_like(zo); zg[:, 2:] = xg[b]; zg[:N, :2] = ug[b]
    # feasibility of gpu solution wrt oracle QP data
    rb = max(np.abs(q["A"][k] @ zg[k, 2:] + q["B"][k] @ zg[k, :2] + q["c"][k] - zg[k + 1, 2:]).max() for k in range(N))
    viol = 0.0
    for k in range(N + 1):
        for r in range(q["nrows"][k]):
            v = q["C"][k, r] @ zg[k]
            viol = max(viol, q["lb"][k, r] - v, v - q["ub"][k, r])
    d = np.abs(zg - zo)
    k, c = np.unravel_index(np.argmax(d), d.shape)
    print(f"b={b} gpu st {sg[b]} it {st[b,0]:.0f} mu {st[b,2]:.2e} | ora st {q['status']} it {q['info'][0]:.0f} mu {q['info'][2]:.2e} resg {q['info'][3]:.1e} | "
          f"max|dz| {d.max():.3e} at k={k} c={c} | gpu rb {rb:.1e} viol {viol:.1e} | obj gpu {objective(q, zg):.10e} ora {objective(q, zo):.10e}")
    if b == 0:
        print("dz by stage (max over comps):", d.max(axis=1))
        print("dz by comp (max over stages):", d.max(axis=0))
