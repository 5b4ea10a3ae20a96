"""Host-side logic: track pre-processing (python/motion_planning.py restated) and the synthetic
workload generators."""
import numpy as np

from ihm2_b200 import track as T
from ihm2_b200.workloads import perturbed_batch, shift_np


def test_closed_spline_fit_is_c2_and_interpolates():
    th = np.linspace(0, 2 * np.pi, 40, endpoint=False)
    path = np.stack([30 * np.cos(th), 20 * np.sin(th)], 1)
    cX, cY = T.fit_spline(path, curv_weight=2.0)
    # value continuity at the knots, and the curve stays close to the points (smoothing spline)
    end = cX.sum(axis=1)
    np.testing.assert_allclose(end, np.roll(cX[:, 0], -1), atol=1e-8)
    assert np.abs(cX[:, 0] - path[:, 0]).max() < 0.5 and np.abs(cY[:, 0] - path[:, 1]).max() < 0.5


def test_offline_motion_plan_on_an_ellipse():
    th = np.linspace(0, 2 * np.pi, 60, endpoint=False)
    path = np.stack([30 * np.cos(th), 30 * np.sin(th)], 1)
    mp = T.offline_motion_plan(path, np.full((60, 2), 1.5))
    assert len(mp.s_ref) == T.NUMBER_SPLINE_INTERVALS and np.all(np.diff(mp.s_ref) > 0) and mp.s_ref[0] == 0
    assert abs(mp.lap_length - 2 * np.pi * 30) < 0.5
    np.testing.assert_allclose(mp.kappa_ref, 1 / 30, atol=2e-3)       # counter-clockwise circle
    m3 = T.triple_motion_plan_ref(mp)
    assert len(m3.s_ref) == 1500 and np.all(np.diff(m3.s_ref) > 0)
    np.testing.assert_allclose(m3.s_ref[500:1000], mp.s_ref)
    np.testing.assert_allclose(m3.s_ref[:500], mp.s_ref - mp.lap_length)


def test_packaged_tracks_are_consistent():
    for name in T.PACKAGED_TRACKS:
        mp = T.load_packaged_track(name)
        assert len(mp.s_ref) == 1500 and np.all(np.diff(mp.s_ref) > 0), name
        assert np.isfinite(mp.kappa_ref).all() and mp.lap_length > 50


def test_perturbed_batch_is_shard_independent():
    rng = np.random.default_rng(0)
    xn = rng.normal(size=8); xp = rng.normal(size=(41, 8)); up = rng.normal(size=(40, 2))
    x0, x, u = perturbed_batch(xn, xp, up, 64, seed=0)
    a0, _, _ = perturbed_batch(xn, xp, up, 32, seed=0, first=0)
    b0, _, _ = perturbed_batch(xn, xp, up, 32, seed=0, first=32)
    np.testing.assert_array_equal(np.vstack([a0, b0]), x0)
    xs, us = shift_np(xp, up)
    np.testing.assert_array_equal(x[5], xs); np.testing.assert_array_equal(u[7], us)
    assert (np.abs(x0[:, 1]) <= 2).all() and (x0[:, 3] >= 0).all()
