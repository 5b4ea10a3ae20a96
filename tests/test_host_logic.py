"""Host-side logic that needs no GPU: the boundary classes' surface, the Stanley baseline against a
scalar transcription of the reference, sharding and the gloo world_size-2 gather."""
import inspect
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ihm2_b200 import controller as C
from ihm2_b200 import simulator as S
from ihm2_b200.dist import gather_results, shard_range
from ihm2_b200.workloads import perturbed_batch


def test_ihm2controller_keeps_the_reference_keywords():
    # new_python/controller.py:250-282
    ref = ["s_ref", "kappa_ref", "Nf", "dt", "s_target", "n_max", "v_x_max", "T_max", "delta_max", "T_dot_max",
           "delta_dot_max", "a_lat_max", "q_s", "q_n", "q_psi", "q_v_x", "q_v_y", "q_r", "q_T", "q_delta", "q_s_f", "q_n_f",
           "q_psi_f", "q_v_x_f", "q_v_y_f", "q_r_f", "q_T_f", "q_delta_f", "q_T_dot", "q_delta_dot"]
    sig = inspect.signature(C.IHM2Controller.__init__)
    names = [n for n in sig.parameters if n != "self"]
    assert names[:len(ref)] == ref
    d = {n: sig.parameters[n].default for n in ref[2:]}
    assert d["Nf"] == 40 and d["dt"] == 1 / 20 and d["q_delta_dot"] == 500.0 and d["q_s_f"] == 1000.0 and d["q_T_dot"] == 0.0
    assert C.IHM2Controller.nx == 8 and C.IHM2Controller.nu == 2
    assert issubclass(C.IHM2Controller, C.Controller) and issubclass(S.MultiModelSimulator, S.Simulator)
    assert [m.name for m in S.SimulatorModel] == ["KIN6", "DYN6", "KIN6_DYN6", "DYN10"]
    assert inspect.isabstract(C.Controller) and inspect.isabstract(S.Simulator)


def test_car_params_vector_matches_reference_constants():
    from ihm2_b200 import _abi
    np.testing.assert_array_equal(C.default_car_params().to_vector(), _abi.default_params())
    mb = C.ModelBounds()
    assert (mb.n_max, mb.v_x_max, mb.T_max, mb.delta_max, mb.T_dot_max, mb.delta_dot_max, mb.a_lat_max) == (2, 31, 500, 0.5, 1e6, 1, 5)


def test_stanley_matches_scalar_transcription():
    """new_python/controller.py:201-234 transcribed for one car, against the batched version."""
    from collections import deque
    rng = np.random.default_rng(0)
    T, B = 30, 3
    n, psi, vx = rng.normal(0, 0.3, (T, B)), rng.normal(0, 0.1, (T, B)), rng.uniform(1, 10, (T, B))
    vref, kap = np.full((T, B), 6.0), rng.normal(0, 0.05, (T, B))
    ctl = C.StanleyController()
    out = np.stack([ctl.compute_control(*[torch.tensor(a[t]) for a in (n, psi, vx, vref, kap)]).numpy() for t in range(T)])
    cfg, dt, lR = C.StanleyController.Config(), 1 / 20, 0.7853
    for b in range(B):
        hist = deque(maxlen=400)
        for t in range(T):
            eps = vref[t, b] - vx[t, b]
            uT = cfg.k_P * eps
            if len(hist) > 1:
                uT += cfg.k_I * (np.sum(hist) - hist[0] - hist[-1]) * dt
            hist.append(eps)
            ud = cfg.k_kappa * np.arctan(2 * np.tan(np.arcsin(kap[t, b] * lR))) - cfg.k_psi * psi[t, b] - np.arctan(cfg.k_n * n[t, b] / (2.0 + vx[t, b]))
            np.testing.assert_allclose(out[t, b], np.clip([uT, ud], [-500, -0.5], [500, 0.5]), rtol=1e-12, atol=1e-12)


def test_shard_range_partitions_everything():
    for total, world in ((4096, 8), (10, 3), (5, 8), (65536, 7)):
        spans = [shard_range(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == total
        for (f0, c0), (f1, _) in zip(spans, spans[1:]):
            assert f0 + c0 == f1


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count = shard_range(total, rank, world)
    xn = np.arange(8.0); xp = np.zeros((41, 8)); up = np.zeros((40, 2))
    x0, _, _ = perturbed_batch(xn, xp, up, count, seed=0, first=first)
    local_u = torch.as_tensor(x0[:, :2] * 2.0)           # stands for the per-shard solve
    local_s = torch.full((count,), rank, dtype=torch.int32)
    gu, gs = gather_results(local_u), gather_results(local_s)
    if rank == 0:
        q.put((gu.numpy(), gs.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_shard_and_gather_equals_single_process():
    total, world, port = 11, 2, 29533
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    [p.start() for p in procs]
    gu, gs = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    x0, _, _ = perturbed_batch(np.arange(8.0), np.zeros((41, 8)), np.zeros((40, 2)), total, seed=0)
    np.testing.assert_array_equal(gu, x0[:, :2] * 2.0)     # independent of the number of ranks
    np.testing.assert_array_equal(gs, [0] * 6 + [1] * 5)


def test_frenet_to_cartesian_follows_the_track_class():
    """tracks.cpp:290-306: reference point on the linear splines + n along the normal; inverse of the projection's (s, n)."""
    import torch
    from ihm2_b200.projection import frenet_to_cartesian
    from ihm2_b200.track import load_packaged_track
    mp = load_packaged_track("fsds_default")
    tabs = [torch.tensor(getattr(mp, f)) for f in ("s_ref", "X_ref", "Y_ref", "phi_ref")]
    g = torch.Generator().manual_seed(0)
    s = torch.rand(200, generator=g, dtype=torch.float64) * mp.lap_length
    n = (torch.rand(200, generator=g, dtype=torch.float64) - 0.5) * 3
    X, Y = frenet_to_cartesian(*tabs, s, n)
    for i in range(0, 200, 17):      # scalar restatement of Track::interp + the normal
        k = int(np.searchsorted(mp.s_ref, float(s[i]), side="right")) - 1
        f = lambda t: t[k] + (t[k + 1] - t[k]) / (mp.s_ref[k + 1] - mp.s_ref[k]) * (float(s[i]) - mp.s_ref[k])   # noqa: E731
        Xr, Yr, pr = f(mp.X_ref), f(mp.Y_ref), f(mp.phi_ref)
        assert abs(float(X[i]) - (Xr - float(n[i]) * np.sin(pr))) < 1e-12 and abs(float(Y[i]) - (Yr + float(n[i]) * np.cos(pr))) < 1e-12
    # on a table point with n = 0 it is the table point
    X0, Y0 = frenet_to_cartesian(*tabs, torch.tensor(mp.s_ref[600:603]), torch.zeros(3, dtype=torch.float64))
    np.testing.assert_allclose(X0.numpy(), mp.X_ref[600:603], atol=1e-12); np.testing.assert_allclose(Y0.numpy(), mp.Y_ref[600:603], atol=1e-12)


def test_start_line_lap_timer_like_the_sim_node():
    """sim_node.cpp:46-51, 261-274: the first crossing starts the clock, later crossings with v_x > 0 give lap times."""
    import torch
    from ihm2_b200.projection import StartLineTimer
    tm = StartLineTimer([0.0, -2.0], [0.0, 2.0], batch=3)
    v = torch.tensor([5.0, 5.0, -1.0], dtype=torch.float64)
    circle = lambda th: torch.stack([10 * torch.sin(th), 10 - 10 * torch.cos(th)], 1)    # noqa: E731  passes (0, 0) heading +x
    done = []
    for step in range(0, 260):
        t = 0.05 * step
        th = torch.tensor([0.05 * step - 0.3, 0.1 * step - 0.3, 0.05 * step - 0.3], dtype=torch.float64)
        done.append(tm.update(circle(th), v, t))
    laps = tm.laps.tolist()
    assert laps[2] == 0 and np.isnan(float(tm.last_lap_time[2]))          # v_x <= 0 never counts
    assert laps[0] == 2 and abs(float(tm.last_lap_time[0]) - 2 * np.pi / 0.05 * 0.05) <= 0.05 + 1e-9
    assert laps[1] == 4 and abs(float(tm.best_lap_time[1]) - 2 * np.pi / 0.1 * 0.05) <= 0.05 + 1e-9
