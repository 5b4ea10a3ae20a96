"""Pins the oracle's restatement of Track::project + the control node's Frenet conversion by
geometry: poses built from known (s, n, psi) on a packaged track must come back."""
import numpy as np

from ihm2_b200.track import load_packaged_track
from oracle import lib as olib

L_R = 0.7853  # python/constants.py:52


def poses_from_frenet(mp, s, n, psi):
    """Cartesian pose at arc length s, lateral offset n (left positive) and heading error psi w.r.t. phi_ref."""
    X = np.interp(s, mp.s_ref, mp.X_ref)
    Y = np.interp(s, mp.s_ref, mp.Y_ref)
    phi_ref = np.interp(s, mp.s_ref, np.unwrap(mp.phi_ref))   # the table wraps at +-pi
    tangent = phi_ref + np.arcsin(L_R * np.interp(s, mp.s_ref, mp.kappa_ref))  # the table holds phi - asin(l_R kappa)
    return np.stack([X - n * np.sin(tangent), Y + n * np.cos(tangent), phi_ref + psi], 1)


def sample(mp, B, seed=0):
    g = np.random.default_rng(seed)
    s = g.uniform(0.0, mp.lap_length, B)
    n = g.uniform(-1.5, 1.5, B)
    n[np.abs(n) < 1e-3] = 0.5
    psi = g.uniform(-0.5, 0.5, B)
    # the reference interpolates phi_ref linearly across its +-pi wrap (tracks.cpp:169-170): keep the samples out of
    # those table cells (and their neighbours)
    jump = np.where(np.abs(np.diff(mp.phi_ref)) > np.pi)[0]
    for j in jump:
        near = (s > mp.s_ref[max(j - 2, 0)]) & (s < mp.s_ref[min(j + 3, len(mp.s_ref) - 1)])
        s[near] = mp.s_ref[min(j + 8, len(mp.s_ref) - 1)]
    return s, n, psi, poses_from_frenet(mp, s, n, psi), s + g.uniform(-1.0, 1.0, B)


def test_projection_recovers_frenet_coordinates():
    mp = load_packaged_track("fsds_competition_1")
    s, n, psi, pose, s_guess = sample(mp, 2000)
    frenet, ref, flag = olib.project(mp.s_ref, mp.X_ref, mp.Y_ref, mp.phi_ref, pose, s_guess)
    assert np.abs(frenet[:, 0] - s).max() < 0.08          # polyline with ~0.7 m spacing under 1.5 m of offset
    assert np.abs(frenet[:, 1] - n).max() < 0.02
    assert np.abs(frenet[:, 2] - psi).max() < 0.02        # phi_ref taken at the projected s
    assert flag.sum() == 0
    assert np.abs(np.hypot(ref[:, 0] - pose[:, 0], ref[:, 1] - pose[:, 1]) - np.abs(frenet[:, 1])).max() < 1e-12


def test_projection_wraps_heading_and_keeps_window():
    mp = load_packaged_track("fsds_default")
    s, n, psi, pose, s_guess = sample(mp, 500, seed=1)
    pose2 = pose.copy()
    pose2[:, 2] += 4 * np.pi
    a = olib.project(mp.s_ref, mp.X_ref, mp.Y_ref, mp.phi_ref, pose, s_guess)[0]
    b = olib.project(mp.s_ref, mp.X_ref, mp.Y_ref, mp.phi_ref, pose2, s_guess)[0]
    assert np.array_equal(a[:, :2], b[:, :2]) and np.abs(a[:, 2] - b[:, 2]).max() < 1e-12
    # a guess a whole lap behind finds the same point one lap earlier on the tripled table (tracks are stored 3x)
    c = olib.project(mp.s_ref, mp.X_ref, mp.Y_ref, mp.phi_ref, pose, s_guess - mp.lap_length)[0]
    inner = (s > 5) & (s < mp.lap_length - 5)             # away from the lap seam (closing chord)
    assert np.abs(c[inner, 0] + mp.lap_length - a[inner, 0]).max() < 1e-6
    assert np.abs(c[inner, 1] - a[inner, 1]).max() < 1e-6
