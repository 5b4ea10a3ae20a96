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
