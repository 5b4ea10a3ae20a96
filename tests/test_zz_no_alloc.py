"""ihm2b_sim_step_static_dev: the plant step whose scratch (launch order of the kinematic / dynamic model partition) lives
in a static device array of the library, for callers that cannot allocate around the call.  No counterpart in the
reference: the check is against the existing entry point
ihm2b_sim_step_dev on the same inputs, through the C ABI, on calls in a row that share the static array.

Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-8   # the bar of the plant-step parity tests (tests/test_gpu_parity.py)


def _states(B, seed):
    rng = np.random.default_rng(seed)
    x = np.stack([rng.uniform(0, 300, B), rng.uniform(-1, 1, B), rng.uniform(-0.2, 0.2, B), rng.uniform(0.5, 20, B),
                  rng.uniform(-0.5, 0.5, B), rng.uniform(-0.5, 0.5, B), rng.uniform(-300, 300, B), rng.uniform(-0.3, 0.3, B)], 1)
    u = np.stack([x[:, 6] + rng.uniform(-50, 50, B), x[:, 7] + rng.uniform(-0.02, 0.02, B)], 1)
    return x, u


def _engine(cfg, mp):
    from ihm2_b200.engine import Engine
    return Engine(cfg, mp.s_ref, mp.kappa_ref)


def _compare(torch, eng, B, seed):
    x, u = _states(B, seed)
    xd, ud = torch.as_tensor(x, device="cuda"), torch.as_tensor(u, device="cuda")
    ref_x, ref_m = eng.sim_step(xd, ud)                                   # existing entry point (scratch of the handle)
    out_x = torch.full((B, 8), -7.0, dtype=torch.float64, device="cuda")  # outputs are the caller's
    out_m = torch.full((B,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    eng.sim_step_static(xd, ud, out_x, out_m)
    torch.cuda.synchronize()
    a, b = out_x.cpu().numpy(), ref_x.cpu().numpy()
    np.testing.assert_array_equal(out_m.cpu().numpy(), ref_m.cpu().numpy())
    np.testing.assert_array_equal(np.isnan(a), np.isnan(b))
    good = ~np.isnan(b).any(axis=1)
    assert good.sum() > 0.9 * B
    err = np.abs(a[good] - b[good]) / np.maximum(1.0, np.abs(b[good]))
    assert err.max() < TOL, err.max(axis=0)
    return ref_m.cpu().numpy()


def test_static_scratch_plant_step_equals_the_existing_entry_point(torch_cuda, fsds):
    """Two calls in a row on one handle (different states: the partition of the second differs from what the first left
    in the array), then a smaller handle (stale entries beyond its batch), against ihm2b_sim_step_dev."""
    from ihm2_b200 import _abi
    B = 256
    eng = _engine(_abi.make_config(batch=B, n_grid=len(fsds.s_ref)), fsds)     # default plant: kin6 / dyn6 switch
    m1 = _compare(torch_cuda, eng, B, seed=11)
    m2 = _compare(torch_cuda, eng, B, seed=12)
    assert (m1 == 1).sum() > 10 and (m1 == 0).sum() > 10 and (m1 != m2).any()   # both plants, different partitions
    eng.close()
    small = _engine(_abi.make_config(batch=40, n_grid=len(fsds.s_ref)), fsds)
    _compare(torch_cuda, small, 40, seed=13)
    _compare(torch_cuda, small, 40, seed=14)
    small.close()


def test_static_scratch_plant_step_single_model(torch_cuda, fsds):
    """A plant without model switch does not touch the array; same results."""
    from ihm2_b200 import _abi
    eng = _engine(_abi.make_config(batch=64, n_grid=len(fsds.s_ref), sim_model=_abi.SIM_FKIN6), fsds)
    _compare(torch_cuda, eng, 64, seed=15)
    eng.close()


def test_static_scratch_rejects_batches_beyond_its_size(torch_cuda, fsds):
    from ihm2_b200 import _abi
    from ihm2_b200.engine import Ihm2bError
    B = 65536 + 1
    eng = _engine(_abi.make_config(batch=B, n_grid=len(fsds.s_ref)), fsds)
    x = torch_cuda.zeros((B, 8), dtype=torch_cuda.float64, device="cuda")
    u = torch_cuda.zeros((B, 2), dtype=torch_cuda.float64, device="cuda")
    xn = torch_cuda.zeros((B, 8), dtype=torch_cuda.float64, device="cuda")
    with pytest.raises(Ihm2bError, match="exceeds"):
        eng.sim_step_static(x, u, xn)
    eng.close()
