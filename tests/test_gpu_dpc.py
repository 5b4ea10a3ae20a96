"""DPC rollout kernel against the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import lib as olib
from test_oracle_dpc import NF, dpc_inputs

pytestmark = pytest.mark.gpu

# fp64 on both sides; the two implementations differ in method (hand adjoint through the RK4 stages here, forward
# sensitivities per step in the oracle) and in libm, so the bar is a few hundred ulp of the largest entry
RTOL = 1e-12


def _dev(*arrays):
    return [torch.tensor(a, dtype=torch.float64, device="cuda") for a in arrays]


def _rel(a, b):
    return np.abs(a - b).max() / max(1.0, np.abs(b).max())


@pytest.mark.parametrize("B", [1, 127, 4096])
def test_dpc_rollout_matches_oracle(B):
    from ihm2_b200 import dpc

    x0, u, ref, w = dpc_inputs(B, seed=3 + B)
    cost, grad, xtraj = dpc.rollout_grad(*_dev(x0, u, ref), w, want_trajectory=True)
    oc, og, ox = olib.dpc_rollout_grad(x0, u, ref, w)
    assert _rel(xtraj.cpu().numpy(), ox) < RTOL
    assert _rel(cost.cpu().numpy(), oc) < RTOL
    assert _rel(grad.cpu().numpy(), og) < RTOL


def test_dpc_rollout_fp32_within_stated_tolerance():
    """The optional fp32 path: north_star's bar for fp32 is 1e-4 relative (per field, against the fp64 oracle)."""
    from ihm2_b200 import dpc

    B = 4096
    x0, u, ref, w = dpc_inputs(B, seed=21)
    f32 = [torch.tensor(a, dtype=torch.float32, device="cuda") for a in (x0, u, ref)]
    cost, grad, xtraj = dpc.rollout_grad(*f32, w, want_trajectory=True)
    assert cost.dtype == grad.dtype == torch.float32
    # the oracle sees the same (rounded) inputs
    oc, og, ox = olib.dpc_rollout_grad(*[t.double().cpu().numpy() for t in f32], w)
    assert _rel(xtraj.double().cpu().numpy(), ox) < 1e-4
    assert _rel(cost.double().cpu().numpy(), oc) < 1e-4
    g = grad.double().cpu().numpy()
    assert _rel(g[..., 0], og[..., 0]) < 1e-4 and _rel(g[..., 1], og[..., 1]) < 1e-4


def test_dpc_rollout_n80_and_bad_horizon():
    from ihm2_b200 import dpc, engine

    g = np.random.default_rng(0)
    B = 64
    x0, u, ref, w = dpc_inputs(B, seed=9)
    u80 = np.concatenate([u, u[:, ::-1]], 1)
    ref80 = np.concatenate([ref, ref[:, 1:] + g.normal(0, 0.1, (B, NF, 4))], 1)
    cost, grad = dpc.rollout_grad(*_dev(x0, u80, ref80), w)
    oc, og, _ = olib.dpc_rollout_grad(x0, u80, ref80, w)
    assert _rel(cost.cpu().numpy(), oc) < RTOL and _rel(grad.cpu().numpy(), og) < RTOL
    with pytest.raises(engine.Ihm2bError, match="Nf = 30"):
        dpc.rollout_grad(*_dev(x0, u[:, :30], ref[:, :31]), w)


def test_dpc_full_size_properties():
    """BASELINE configs[4] size (2^20 rollouts): permutation equivariance (bit-exact), a sampled oracle check and the
    directional derivative of the whole-batch loss against central differences of the kernel's own costs."""
    from ihm2_b200 import dpc

    B = 1 << 20
    x0, u, ref, w = dpc_inputs(B, seed=3)
    dx0, du, dref = _dev(x0, u, ref)
    cost, grad = dpc.rollout_grad(dx0, du, dref, w)
    perm = torch.randperm(B, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
    cp, gp = dpc.rollout_grad(dx0[perm].contiguous(), du[perm].contiguous(), dref[perm].contiguous(), w)
    assert torch.equal(cp, cost[perm]) and torch.equal(gp, grad[perm])
    idx = np.random.default_rng(1).choice(B, 2048, replace=False)
    oc, og, _ = olib.dpc_rollout_grad(x0[idx], u[idx], ref[idx], w)
    assert _rel(cost[idx].cpu().numpy(), oc) < RTOL and _rel(grad[idx].cpu().numpy(), og) < RTOL
    d = torch.randn(du.shape, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    d[..., 0] *= 100.0
    d[..., 1] *= 0.01
    h = 1e-4
    fd = (dpc.rollout_grad(dx0, du + h * d, dref, w)[0] - dpc.rollout_grad(dx0, du - h * d, dref, w)[0]) / (2 * h)
    an = (grad * d).sum((1, 2))
    assert ((fd - an).abs() / (1.0 + an.abs())).max().item() < 1e-4


def test_dpc_loss_trains_a_policy_through_the_rollout():
    """DPCLoss hands the kernel gradient to autograd: a linear policy's parameter gradient equals the chain rule."""
    from ihm2_b200 import dpc

    B = 256
    x0, u, ref, w = dpc_inputs(B, seed=11)
    dx0, _, dref = _dev(x0, u, ref)
    feat = torch.cat([dx0, dref[:, 1, :] - dx0], 1)
    Wp = torch.zeros(8, NF * 2, dtype=torch.float64, device="cuda", requires_grad=True)
    bp = torch.tensor(u.mean(0).reshape(-1), device="cuda", requires_grad=True)
    uu = (feat @ Wp + bp).reshape(B, NF, 2)
    loss = dpc.DPCLoss.apply(dx0, uu, dref, w).mean()
    loss.backward()
    _, grad = dpc.rollout_grad(dx0, uu.detach().contiguous(), dref, w)
    assert torch.allclose(bp.grad, grad.reshape(B, -1).mean(0), rtol=1e-12, atol=0)
    assert torch.allclose(Wp.grad, feat.T @ grad.reshape(B, -1) / B, rtol=1e-10, atol=1e-12)
