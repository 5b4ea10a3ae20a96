"""Batched Track::project kernel against the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from ihm2_b200.track import load_packaged_track
from oracle import lib as olib
from test_oracle_project import sample

pytestmark = pytest.mark.gpu


def test_project_matches_oracle_on_three_tracks():
    from ihm2_b200.projection import Track

    names = ["fsds_competition_1", "fsds_competition_2", "fsds_competition_3"]
    mps = [load_packaged_track(n) for n in names]
    B = 30000
    idx = (np.arange(B) % 3).astype(np.int32)
    pose, s_guess = np.zeros((B, 3)), np.zeros(B)
    for t, mp in enumerate(mps):
        _, _, _, p, sg = sample(mp, B // 3, seed=t)
        pose[idx == t], s_guess[idx == t] = p, sg
    trk = Track(mps)
    frenet, flag, ref = trk.project(torch.tensor(pose, device="cuda"), torch.tensor(s_guess, device="cuda"),
                                    track_index=torch.tensor(idx, device="cuda"), want_ref=True)
    tabs = [np.stack([getattr(m, f) for m in mps]) for f in ("s_ref", "X_ref", "Y_ref", "phi_ref")]
    of, oref, oflag = olib.project(*tabs, pose, s_guess, track=idx)
    # same branches (closest point, neighbour) on both sides, then a few flops: only libm's last bits differ
    assert np.abs(frenet.cpu().numpy() - of).max() < 1e-11
    assert np.abs(ref.cpu().numpy() - oref).max() < 1e-11
    assert np.array_equal(flag.cpu().numpy(), oflag)


def test_project_edge_cases():
    """Car exactly on a table point, on the centre line (flagged: tangent x normal ~ 0), guesses at both ends of the
    table, a window that reaches past the table."""
    from ihm2_b200.projection import Track

    mp = load_packaged_track("fsds_default")
    n = len(mp.s_ref)
    k = np.array([0, 1, n // 2, n - 2, n - 1, 700, 701])
    pose = np.stack([mp.X_ref[k], mp.Y_ref[k], mp.phi_ref[k]], 1)
    pose[5:, 1] += 0.7
    s_guess = mp.s_ref[k] + np.array([0.0, -1.0, 0.3, 1.5, 5.0, 0.0, -3.0])
    trk = Track(mp)
    frenet, flag = trk.project(torch.tensor(pose, device="cuda"), torch.tensor(s_guess, device="cuda"), s_tol=2.0)
    of, _, oflag = olib.project(mp.s_ref, mp.X_ref, mp.Y_ref, mp.phi_ref, pose, s_guess)
    assert np.allclose(frenet.cpu().numpy(), of, rtol=0, atol=1e-11, equal_nan=True)
    assert np.array_equal(flag.cpu().numpy(), oflag) and oflag[:5].all() and not oflag[5:].any()
    assert trk.length() == mp.lap_length


def test_project_full_size_roundtrip():
    """2^20 cars: poses built from (s, n, psi) come back (size-independent property), and the time per batch."""
    from ihm2_b200.projection import Track

    mp = load_packaged_track("fsds_competition_2")
    B = 1 << 20
    s, n, psi, pose, s_guess = sample(mp, B, seed=4)
    trk = Track(mp)
    dp, dg = torch.tensor(pose, device="cuda"), torch.tensor(s_guess, device="cuda")
    frenet, flag = trk.project(dp, dg)
    f = frenet.cpu().numpy()
    assert np.abs(f[:, 0] - s).max() < 0.2 and np.abs(f[:, 1] - n).max() < 0.03 and np.abs(f[:, 2] - psi).max() < 0.03
    assert int(flag.sum()) == 0


def test_stanley_baseline_follows_the_centre_line_on_the_cartesian_plant():
    """The two-node loop (projection -> controller -> Cartesian kin6/dyn6 plant) with the reference's Stanley fallback
    (new_python/controller.py:162-234) as the controller: the cars settle at v_ref on the centre line."""
    from ihm2_b200 import _abi
    from ihm2_b200.cartesian_loop import closed_loop_cartesian
    from ihm2_b200.controller import StanleyController
    from ihm2_b200.engine import Engine
    from ihm2_b200.projection import Track

    mp = load_packaged_track("fsds_competition_1")
    B, v_ref, n_steps = 8, 5.0, 400
    cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref), sim_model=_abi.SIM_KIN6C_DYN6C)
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref, lap_length=np.array([mp.lap_length]))
    trk = Track(mp)
    i0 = len(mp.s_ref) // 3
    state = torch.zeros(B, 8, dtype=torch.float64, device="cuda")
    state[:, 0], state[:, 1], state[:, 2] = float(mp.X_ref[i0]), float(mp.Y_ref[i0]), float(mp.phi_ref[i0])
    state[:, 1] += torch.linspace(-0.3, 0.3, B, dtype=torch.float64, device="cuda")
    stanley = StanleyController()
    s_tab, k_tab = torch.tensor(mp.s_ref, device="cuda"), torch.tensor(mp.kappa_ref, device="cuda")

    def controller(frenet, st):
        idx = torch.searchsorted(s_tab, frenet[:, 0].contiguous()).clamp(1, len(s_tab) - 1)
        w = (frenet[:, 0] - s_tab[idx - 1]) / (s_tab[idx] - s_tab[idx - 1])
        kappa = k_tab[idx - 1] + w * (k_tab[idx] - k_tab[idx - 1])
        return stanley.compute_control(frenet[:, 1], frenet[:, 2], st[:, 3], torch.full_like(kappa, v_ref), kappa)

    out = closed_loop_cartesian(eng, trk, state, n_steps, controller=controller, log=True)
    eng.close()
    assert (out["fail"] < 0).all()
    s_end = out["s"].cpu().numpy()
    assert np.all(np.abs(s_end - (v_ref * n_steps * cfg.dt - 6.0)) < 6.0)      # ~ 5 m/s minus the start-up
    frenet, _ = trk.project(out["state"][:, :3].contiguous(), out["s"])
    assert frenet[:, 1].abs().max().item() < 0.1 and (out["state"][:, 3] - v_ref).abs().max().item() < 0.2
