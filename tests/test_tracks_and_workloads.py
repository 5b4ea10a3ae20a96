"""Host-side logic: track pre-processing (python/motion_planning.py restated) and the synthetic
workload generators."""
import os

import numpy as np
import pytest

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


GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_FIELDS = ("s_ref", "X_ref", "Y_ref", "phi_ref", "kappa_ref", "right_widths", "left_widths")


@pytest.mark.parametrize("name", ["fsds_default", "fsds_competition_1", "fsds_competition_2", "fsds_competition_3", "skidpad", "short_skidpad"])
def test_packaged_tables_match_the_reference_run(name):
    """The tables every config is fed with (ihm2_b200/data/tracks/*.npz) against what the REFERENCE's own
    fit_spline -> uniformly_sample_spline -> get_curvature / get_heading produce on data/<name>/center_line.csv
    (tests/golden/tracks_reference.npz, written by tests/golden/make_track_golden.py from /root/reference)."""
    from ihm2_b200.track import load_packaged_track
    g = np.load(os.path.join(GOLD, "tracks_reference.npz"))
    mp = load_packaged_track(name, tripled=False)
    for f in _FIELDS:
        ref = g[f"{name}/{f}"]
        assert np.abs(getattr(mp, f) - ref).max() <= 1e-8 * max(1.0, np.abs(ref).max()), f
    assert abs(mp.lap_length - float(g[f"{name}/lap_length"])) < 1e-9
    # triple_motion_plan_ref (motion_planning.py:402-428): the non-uniform seam included
    np.testing.assert_allclose(load_packaged_track(name).s_ref, g[f"{name}/s_ref_tripled"], rtol=0, atol=1e-9)


def test_offline_motion_plan_from_the_centre_line_matches_the_reference_run(tmp_path):
    """Our own pre-processing (one sparse KKT solve instead of proxqp) from the raw centre line kept in the package,
    and the CSV writer, against the reference's outputs."""
    from ihm2_b200.track import offline_motion_plan, write_track_data_file
    g = np.load(os.path.join(GOLD, "tracks_reference.npz"))
    d = np.load(os.path.join(os.path.dirname(GOLD), "..", "ihm2_b200", "data", "tracks", "fsds_default.npz"))
    mp = offline_motion_plan(d["center_line"], d["track_widths"])
    for f in _FIELDS:
        ref = g[f"fsds_default/{f}"]
        assert np.abs(getattr(mp, f) - ref).max() <= 1e-8 * max(1.0, np.abs(ref).max()), f
    out = tmp_path / "track.csv"
    write_track_data_file(mp, str(out))
    ours = open(out).read().splitlines()
    theirs = open(os.path.join(GOLD, "fsds_default_track_data_reference.csv")).read().splitlines()
    assert ours[0] == theirs[0] and len(ours) == len(theirs) == 501
    a = np.array([[float(v) for v in l.split(",")] for l in ours[1:]])
    b = np.array([[float(v) for v in l.split(",")] for l in theirs[1:]])
    assert np.abs(a - b).max() <= 1.0000001e-6      # "%.6f": at most one unit in the last printed digit


@pytest.mark.skipif(not os.path.isdir("/root/reference/python"), reason="the reference tree is only mounted in the build container")
def test_golden_track_tables_are_what_the_reference_produces_here():
    """Re-runs the reference's motion_planning.py (unmodified, third-party imports replaced by stand-ins) and checks the
    committed fixture against it."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_track_golden", os.path.join(GOLD, "make_track_golden.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    live = mod.reference_tables("/root/reference")
    g = np.load(os.path.join(GOLD, "tracks_reference.npz"))
    assert set(live) == set(g.files)
    for k in g.files:
        np.testing.assert_allclose(live[k], g[k], rtol=0, atol=1e-12)
