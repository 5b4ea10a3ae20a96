"""KKT certificates of what the CUDA path returns, on whole batches: the primal point, the equality multipliers and the
inequality multipliers (ihm2b_get_multipliers_dev / ihm2b_get_slacks_dev) are put into the optimality conditions of the
QP as python/mpc.py defines it, in numpy (tests/kkt_check.py) -- the CUDA output against mathematics, not against its
algorithmic twin.  The QP is strictly convex: a KKT point is THE solution, whoever computes it."""
import numpy as np
import pytest

from kkt_check import kkt_residuals

pytestmark = pytest.mark.gpu


def _solve_with_multipliers(torch_cuda, cfg, mp, x0, x, u, widths=None, params=None):
    from ihm2_b200.engine import Engine
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref, widths=widths, params=params)
    eng.set_iterate(x, u)
    A, Bm, phi = [t.cpu().numpy() for t in eng.linearize()]
    c = phi - np.einsum("bkij,bkj->bki", A, x[:, :-1]) - np.einsum("bkij,bkj->bki", Bm, u)
    u0, status = eng.solve(torch_cuda.as_tensor(x0, device="cuda"))
    xs, us = [t.cpu().numpy() for t in eng.get_iterate()]
    pi, lamL, lamU = [t.cpu().numpy() for t in eng.multipliers()]
    out = dict(A=A, B=Bm, c=c, x=xs, u=us, pi=pi, lamL=lamL, lamU=lamU, status=status.cpu().numpy(), stats=eng.stats().cpu().numpy())
    if cfg.nonlinear_constraints:
        out["slacks"] = tuple(t.cpu().numpy() for t in eng.slacks())
        out["hc"] = eng.h_rows().cpu().numpy()
    eng.close()
    return out


def test_kkt_certificate_on_the_whole_config2_batch(torch_cuda, skidpad):
    """All 4096 problems of BASELINE configs[1]: every solve reported as success is a KKT point of its QP."""
    from ihm2_b200 import _abi
    from ihm2_b200.workloads import nominal_iterate, perturbed_batch
    B = 4096
    x_nom, xp, up = nominal_iterate("skidpad", 20)
    cfg = _abi.make_config(batch=B, n_grid=len(skidpad.s_ref))
    x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=0, cfg=cfg)
    s = _solve_with_multipliers(torch_cuda, cfg, skidpad, x0, x, u)
    ok = s["status"] == 0
    assert ok.sum() > 0.7 * B
    r = kkt_residuals(cfg, x0[ok], s["A"][ok], s["B"][ok], s["c"][ok], s["x"][ok], s["u"][ok], s["pi"][ok], s["lamL"][ok], s["lamU"][ok])
    print({k: (float(np.median(v)), float(v.max())) for k, v in r.items()})
    assert r["x0"].max() == 0.0
    assert r["dyn"].max() < 1e-9                # relative to 1 + |c|
    assert r["prim"].max() < 1e-8 and r["dual"].max() <= 0.0
    assert r["comp"].max() < 1e-6               # lam * t, the interior point's own tolerance is 1e-8
    # relative to the gradient scale; the few at 1e-7 .. 1e-6 are the thin interiors that stop at the 1e-6 exit
    assert r["stat"].max() < 1e-6 and np.quantile(r["stat"], 0.99) < 1e-7 and np.median(r["stat"]) < 1e-11


def test_kkt_certificate_with_soft_rows(torch_cuda, fsds):
    """The soft-constrained QP (h rows + soft bounds): stationarity in z and in every slack, feasibility with the slacks,
    complementarity of the rows and of s >= 0."""
    from ihm2_b200 import _abi
    from ihm2_b200.workloads import nominal_iterate, perturbed_batch
    B = 512
    x_nom, xp, up = nominal_iterate("fsds_default", 30, nonlinear_constraints=True)
    cfg = _abi.make_config(batch=B, n_grid=len(fsds.s_ref), nonlinear_constraints=True)
    x0, x, u = perturbed_batch(x_nom, xp, up, B, seed=7, cfg=cfg)
    x0[:, 1] = np.clip(x0[:, 1] + np.linspace(-0.6, 0.9, B), -1.8, 1.8)
    params = np.tile(_abi.default_params(), (B, 1))
    s = _solve_with_multipliers(torch_cuda, cfg, fsds, x0, x, u, widths=(fsds.right_widths, fsds.left_widths), params=params)
    ok = s["status"] == 0
    assert ok.sum() > 0.7 * B
    sl = tuple(t[ok] for t in s["slacks"])
    r = kkt_residuals(cfg, x0[ok], s["A"][ok], s["B"][ok], s["c"][ok], s["x"][ok], s["u"][ok], s["pi"][ok], s["lamL"][ok], s["lamU"][ok],
                      params=params[ok], hc=s["hc"][ok], slacks=sl)
    print({k: (float(np.median(v)), float(v.max())) for k, v in r.items()})
    used = (np.maximum(sl[0], sl[1]) > 1e-6).reshape(ok.sum(), -1).sum(1)
    assert (used > 0).mean() > 0.5, "the batch must exercise the slacks"
    assert r["x0"].max() == 0.0 and r["dyn"].max() < 1e-9
    assert r["prim"].max() < 1e-8 and r["dual"].max() <= 0.0 and r["slack_sign"].max() <= 0.0
    assert r["comp"].max() < 1e-6 and r["slack_comp"].max() < 1e-6
    assert r["slack_stat"].max() < 1e-6
    # thin interiors (barrier weights 1e9+ next to an O(1) Hessian) stop at the 1e-6 relative stationarity exit
    assert r["stat"].max() < 2e-6 and np.quantile(r["stat"], 0.9) < 1e-8
