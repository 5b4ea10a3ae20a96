"""bench.py -- BASELINE.json metric: NMPC SQP-RTI solves/s (N = 40) on configs[1]
(4096 perturbed initial states, one SQP-RTI step, data/skidpad, fp64) per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  torchrun ... bench.py --gpus N ...       (one rank per GPU, weak scaling: 4096 problems per GPU)

One step = one pass of the hot path over one batch: restore the warm start, linearise
(IRK-GL4 + sensitivities), solve the QP (IPM / Riccati), RTI update.  `value` is timed with
inputs resident in HBM; `e2e` goes through the host entry point (ihm2b_compute_control_host:
pinned H2D of x0, solve, D2H of u0 / status inside the timed region).
The roofline is the FP64 CUDA-core (DFMA) one; its peak is measured live.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU = 4096
N_HORIZON = 40
TRACK = "skidpad"
METRIC = "NMPC SQP-RTI solves/sec (N=40, batch)"
WORKLOAD = "configs[1]: 4096 perturbed initial states, single SQP-RTI step, data/skidpad, N=40"


def flops_per_solve(n_ipm_mean: float, N: int = N_HORIZON, irk: bool = True) -> dict:
    """Algorithmic flops of one RTI solve, SURVEY.md section 8(d): dense small-block formulas,
    FMA = 2, no credit for sparsity; the linearisation term is the ERK4 figure (the IRK cascade
    does more work per interval, so using it here understates the achieved rate)."""
    nx, nu, nz = 8, 2, 10
    f_lin = N * (4 * (65 + 160 + 2 * nx * nx * nz + 2 * nx * nz) + 9 * nx * nz)
    f_fact = N * (7 / 3 * nx**3 + 4 * nx * nx * nu + 2 * nx * nu * nu + nu**3 / 3)
    f_solve = N * (8 * nx * nx + 8 * nx * nu + 2 * nu * nu)
    f_res = N * 2 * (nz * nz + 2 * nx * nz)
    f_ineq = 30 * (8 + 16 * (N - 1) + 8)
    f_ipm = f_fact + 2 * f_solve + f_res + f_ineq
    return dict(lin=f_lin, ipm=f_ipm, total=f_lin + n_ipm_mean * f_ipm + 20 * N * nz, qp=n_ipm_mean * f_ipm + 20 * N * nz)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([v.strip() for v in out.split(",")])
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i] == "Active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


CPU_THREADS = 1


def host_cores() -> int:
    return len(os.sched_getaffinity(0))


def cpu_oracle_rate(x0, x, u, mp, n_sample: int, repeats: int = 1):
    """Times the CPU oracle (OpenMP over problems, all host cores) on the first n_sample problems."""
    from ihm2_b200 import _abi
    from oracle.lib import Oracle

    cfg = _abi.make_config(batch=n_sample, n_grid=len(mp.s_ref))
    o = Oracle(cfg, mp.s_ref, mp.kappa_ref)
    global CPU_THREADS
    CPU_THREADS = int(o.lib.ihm2o_set_threads(host_cores()))   # torchrun exports OMP_NUM_THREADS=1
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        o.solve(x0[:n_sample], x[:n_sample], u[:n_sample])
        best = min(best, time.perf_counter() - t0)
    return n_sample / best, best


def build_oracle_only():
    """The reference arm must not load libihm2b.so: build / load the CPU oracle alone."""
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout[-2000:] + r.stderr[-2000:])


def reference_arm(args, rank: int):
    """--impl reference: the reference's own CPU implementation of the path cannot be built here
    (acados / HPIPM / CasADi absent), so this times the CPU oracle port, all host threads, on the same
    4096 problems of one step as the GPU arm."""
    if rank != 0:
        return
    build_oracle_only()
    from ihm2_b200 import _abi                      # ctypes mirror of the config struct: no library is loaded
    from ihm2_b200.track import load_packaged_track
    from oracle.lib import Oracle
    from oracle.workloads import nominal_iterate_cpu, perturbed_batch

    mp = load_packaged_track(TRACK)
    x_nom, xp, up = nominal_iterate_cpu(TRACK, 20)   # same nominal as the GPU arm, produced by the oracle
    n_sample = args.batch
    cfg = _abi.make_config(batch=n_sample, n_grid=len(mp.s_ref))
    x0, x, u = perturbed_batch(x_nom, xp, up, n_sample, seed=0, cfg=cfg)
    times = []
    for i in range(args.warmup + args.steps):
        _, dt = cpu_oracle_rate(x0, x, u, mp, n_sample)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    val = n_sample / (ms * 1e-3)
    sample = f"all {n_sample} problems of one step, OpenMP over problems"
    emit({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "solves/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": n_sample, "integrator": "IRK-GL4 (controller.py:305-307)",
                   "note": "acados/HPIPM not installable here: CPU restatement (oracle/ihm2_oracle.cpp), not acados"},
        "cpu_baseline": {"value": val, "unit": "solves/s", "cores": CPU_THREADS, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def other_configs(torch, dev):
    """Driver-timed records of the other BASELINE.json configurations, on reduced budgets so that the default run still
    ends within a few minutes: configs[4] (DPC rollouts + gradients, 2^20), configs[2] (closed-loop laps with the
    soft-constrained controller, 3072 cars instead of 65536) -- device time by CUDA events, rank 0 only."""
    from ihm2_b200 import _abi, dpc
    from ihm2_b200.engine import Engine
    from ihm2_b200.track import load_packaged_track

    out = {}
    # ---- configs[4]: dpc/main.py rollout + gradient, B = 2^20, fp64
    B = 1 << 20
    g = np.random.default_rng(3)
    mp = dpc.MotionPlanner("fsds_competition_1", v_ref=5.0)
    s0 = g.uniform(0.0, mp.lap_length, B)
    x0 = np.stack([np.interp(s0, mp.s_ref, mp.X_ref) + g.normal(0, 0.3, B), np.interp(s0, mp.s_ref, mp.Y_ref) + g.normal(0, 0.3, B),
                   np.interp(s0, mp.s_ref, mp.phi_ref) + g.normal(0, 0.1, B), np.abs(5.0 + g.normal(0, 1.0, B))], 1)
    ref = mp.plan(s0, x0[:, 2])
    u = np.stack([g.uniform(-dpc.T_max, dpc.T_max, (B, dpc.Nf)), g.uniform(-dpc.delta_max, dpc.delta_max, (B, dpc.Nf))], -1)
    w = dpc.DPCWeights(q_lon=1.0, q_lat=10.0, q_phi=5.0, q_v=1.0, r_T=1e-4, r_delta=2.0, q_lon_f=2.0, q_lat_f=20.0, q_phi_f=10.0, q_v_f=2.0)
    dx0, du, dref = [torch.tensor(a, dtype=torch.float64, device=dev) for a in (x0, u, ref)]
    for _ in range(3):
        dpc.rollout_grad(dx0, du, dref, w)
    ms = []
    for _ in range(8):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); dpc.rollout_grad(dx0, du, dref, w); e1.record()     # 2.8 GB of inputs / outputs: larger than L2
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    bytes_per = 8 * (4 + 2 * dpc.Nf + 4 * (dpc.Nf + 1) + 1 + 2 * dpc.Nf)
    m = float(np.median(ms))
    out["config5_dpc"] = {"workload": "configs[4]: DPC kin4 rollout N=40 + IHM1.5 cost + gradient in the controls, B = 2^20, fp64, fsds_competition_1, seed 3",
                          "value": B / (m * 1e-3), "unit": "rollouts+grads/s", "ms_per_batch": m,
                          "algorithmic_bytes_per_rollout": bytes_per, "algorithmic_GBps": B * bytes_per / (m * 1e-3) / 1e9}
    del dx0, du, dref

    # ---- configs[2]: closed-loop laps, reduced to 3072 cars (the set-up that laps, scripts/run_config3.py)
    B, n_steps = 3072, 1100
    names = ["fsds_competition_1", "fsds_competition_2", "fsds_competition_3"]
    mps = [load_packaged_track(n) for n in names]
    s_ref = np.stack([t.s_ref for t in mps]); kappa_ref = np.stack([t.kappa_ref for t in mps])
    widths = (np.stack([t.right_widths for t in mps]), np.stack([t.left_widths for t in mps]))
    track = (np.arange(B) % 3).astype(np.int32)
    lap = np.array([t.lap_length for t in mps])
    gen = torch.Generator().manual_seed(1)
    r = torch.rand(B, 8, generator=gen, dtype=torch.float64).numpy()
    p = np.tile(_abi.default_params(), (B, 1))
    f_m = 0.9 + 0.2 * r[:, 0]
    p[:, _abi.P_M] *= f_m; p[:, _abi.P_IZ] *= f_m
    p[:, _abi.P_DA] *= 0.85 + 0.3 * r[:, 1]; p[:, _abi.P_BA] *= 0.85 + 0.3 * r[:, 2]
    for j, c in enumerate((_abi.P_CR0, _abi.P_CR1, _abi.P_CR2)):
        p[:, c] *= 0.9 + 0.2 * r[:, 3 + j]
    p[:, _abi.P_CM0] *= 0.95 + 0.1 * r[:, 6]
    p[:, _abi.P_CROSS] = 0.0
    cfg = _abi.make_config(batch=B, n_grid=s_ref.shape[1], n_tracks=3, s_target=24.0, nonlinear_constraints=True,
                           qp_tolerate_maxiter=True)
    eng = Engine(cfg, s_ref, kappa_ref, track_index=track, params=p, lap_length=lap, widths=widths)
    x = torch.zeros(B, 8, dtype=torch.float64, device=dev); x[:, 0] = -6.0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = eng.launch_count
    e0.record(); res = eng.closed_loop(x, n_steps); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    done, fail, lap_t = res["done"].cpu().numpy(), res["fail"].cpu().numpy(), res["lap_time"].cpu().numpy()
    steps_run = np.where(fail >= 0, fail + 1, np.where(done >= 0, done + 1, n_steps))
    fin = (done >= 0) & np.isfinite(lap_t)
    out["config3_closed_loop"] = {
        "workload": f"configs[2] reduced: {B} closed-loop cars (of 65536) x <= {n_steps} control steps, fsds_competition_1/2/3, randomised "
                    "tyre / mass parameters (seed 1), soft-constrained controller, qp_tolerate_maxiter, straight slip angles in the plant",
        "value": float(steps_run.sum() / (ms * 1e-3)), "unit": "closed-loop steps/s (shift + RTI solve + plant step)",
        "device_ms": ms, "gpu_launches": int(eng.launch_count - l0), "laps_finished": int(fin.sum()), "dnf": int(B - fin.sum()),
        "stopped_on_solver_failure": int((fail >= 0).sum()),
        "lap_time_s_median": {n: (float(np.median(lap_t[fin & (track == i)])) if (fin & (track == i)).any() else None) for i, n in enumerate(names)}}
    eng.close()
    return out


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL prints its version line) write to fd 1; the contract is ONE JSON line on stdout.
    Point fd 1 at stderr for the rest of the process and keep the real stdout for emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=B_PER_GPU, help="problems per GPU (default: the config's 4096)")
    ap.add_argument("--no-extras", action="store_true", help="skip the records of the other BASELINE configs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist

    import __graft_entry__ as g
    if rank == 0:
        g.build()
    from ihm2_b200 import _abi
    from ihm2_b200.engine import Engine, measure_fp64_tflops
    from ihm2_b200.track import load_packaged_track
    from ihm2_b200.workloads import nominal_iterate, perturbed_batch

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
        os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()
    B = args.batch
    mp = load_packaged_track(TRACK)
    # synthetic inputs: nominal from the product path itself, instance i keyed by its global index
    x_nom, xp, up = nominal_iterate(TRACK, 20)
    cfg = _abi.make_config(batch=B, n_grid=len(mp.s_ref))
    x0_h, x_h, u_h = perturbed_batch(x_nom, xp, up, B, seed=0, first=rank * B, cfg=cfg)
    eng = Engine(cfg, mp.s_ref, mp.kappa_ref)
    x_ws = torch.as_tensor(x_h, device=dev); u_ws = torch.as_tensor(u_h, device=dev)
    x0_d = torch.as_tensor(x0_h, device=dev)
    u0_d = torch.empty(B, 2, dtype=torch.float64, device=dev); st_d = torch.empty(B, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB of L2

    def step():
        eng.set_iterate(x_ws, u_ws)          # restore the warm start (the solve updates the iterate in place)
        eng.solve(x0_d, u0_d, st_d)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    sync_all()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = eng.launch_count
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sync_all()
    for e0, e1 in ev:
        flush.fill_(1)                       # L2 flush between timed iterations, outside the event pair
        e0.record(); step(); e1.record()
    sync_all()
    launches = eng.launch_count - l0
    times = [e0.elapsed_time(e1) for e0, e1 in ev]
    total = torch.tensor([sum(times)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.MAX)
    ms_per_step = float(total.item()) / args.steps
    value = world * B / (ms_per_step * 1e-3)
    stats = eng.stats()
    status = st_d.cpu().numpy()
    n_ipm = float(stats[:, 0].mean().item())

    # end to end through the host entry point (numpy in, numpy out)
    eng.set_iterate(x_ws, u_ws)
    for _ in range(2):
        eng.set_iterate(x_ws, u_ws); eng.compute_control_host(x0_h, do_shift=False)
    sync_all()
    t_e2e = []
    for _ in range(max(3, args.steps // 2)):
        flush.fill_(1); torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.set_iterate(x_ws, u_ws)
        eng.compute_control_host(x0_h, do_shift=False)
        t_e2e.append(time.perf_counter() - t0)
    e2e_t = torch.tensor([sum(t_e2e) / len(t_e2e)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = world * B / float(e2e_t.item())
    # second end-to-end figure, the reference's full call shape: it also reads every stage of x_pred / u_pred back to
    # the host after each solve (new_python/controller.py:402-403); page-locked destination buffers
    xh_pin = torch.empty(B, N_HORIZON + 1, 8, dtype=torch.float64).pin_memory()
    uh_pin = torch.empty(B, N_HORIZON, 2, dtype=torch.float64).pin_memory()
    t_e2e2 = []
    for _ in range(max(3, args.steps // 2)):
        flush.fill_(1); torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.set_iterate(x_ws, u_ws)
        eng.compute_control_host(x0_h, do_shift=False)
        eng.get_iterate_host(xh_pin.numpy(), uh_pin.numpy())
        t_e2e2.append(time.perf_counter() - t0)
    e2e2_t = torch.tensor([sum(t_e2e2) / len(t_e2e2)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e2_t, op=dist.ReduceOp.MAX)
    e2e2_val = world * B / float(e2e2_t.item())
    clocks = sampler.stop()

    # the same steps on a handle WITHOUT launch-order history (what a first solve of fresh problems costs):
    # the timed loop above runs with schedule_by_history, i.e. each step launches its problems longest first
    # by the iteration counts of the previous step, as consecutive control steps of a closed loop do
    cfg_nh = _abi.make_config(batch=B, n_grid=len(mp.s_ref), schedule_by_history=False)
    eng_nh = Engine(cfg_nh, mp.s_ref, mp.kappa_ref)
    t_nh = []
    u0_nh = torch.empty_like(u0_d); st_nh = torch.empty_like(st_d)
    for i in range(3 + max(3, args.steps // 4)):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng_nh.set_iterate(x_ws, u_ws); eng_nh.solve(x0_d, u0_nh, st_nh); e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            t_nh.append(e0.elapsed_time(e1))
    nh_t = torch.tensor([sum(t_nh) / len(t_nh)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(nh_t, op=dist.ReduceOp.MAX)
    ms_no_history = float(nh_t.item())
    eng_nh.close()

    # per-kernel device times (synchronising events inside the library; separate untimed pass)
    eng.profile(True)
    for _ in range(5):
        flush.fill_(1); step()
    torch.cuda.synchronize()
    lin_ms, qp_ms = eng.profile_read()
    eng.profile(False)

    # ---- the solver set-up the reference itself configures (new_python/controller.py:301-308): SQP, 2 iterations,
    # MERIT_BACKTRACKING, on the same inputs; one controller call = up to 2 QPs + line-search linearisations
    cfg_sqp = _abi.make_config(batch=B, n_grid=len(mp.s_ref), nlp_solver=_abi.SQP, nlp_max_iter=2,
                               globalization=_abi.MERIT_BACKTRACKING)
    eng_sqp = Engine(cfg_sqp, mp.s_ref, mp.kappa_ref)
    t_sqp, l_sqp = [], 0
    u0_s = torch.empty_like(u0_d); st_s = torch.empty_like(st_d)
    for i in range(3 + max(3, args.steps // 4)):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l_sqp = eng_sqp.launch_count
        e0.record(); eng_sqp.set_iterate(x_ws, u_ws); eng_sqp.solve(x0_d, u0_s, st_s); e1.record()
        torch.cuda.synchronize()
        l_sqp = eng_sqp.launch_count - l_sqp
        if i >= 3:
            t_sqp.append(e0.elapsed_time(e1))
    sqp_stats = eng_sqp.stats()
    sqp_ms = torch.tensor([sum(t_sqp) / len(t_sqp)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(sqp_ms, op=dist.ReduceOp.MAX)
    reference_setup = {"solver": "SQP, nlp_solver_max_iter 2, MERIT_BACKTRACKING (new_python/controller.py:301-308)",
                       "ms_per_step": float(sqp_ms.item()), "controller_calls_per_s": world * B / (float(sqp_ms.item()) * 1e-3),
                       "qp_solves_per_call": float(sqp_stats[:, 1].mean().item()), "gpu_launches_per_call": int(l_sqp),
                       "status_counts": {str(k): int(v) for k, v in enumerate(np.bincount(st_s.cpu().numpy(), minlength=5)) if v}}
    eng_sqp.close()

    # ---- mixed precision (linearize_fp32: IRK collocation + sensitivities in fp32, QP in fp64), same inputs
    cfg_mp = _abi.make_config(batch=B, n_grid=len(mp.s_ref), linearize_fp32=True)
    eng_mp = Engine(cfg_mp, mp.s_ref, mp.kappa_ref)
    u0_mp = torch.empty_like(u0_d); st_mp = torch.empty_like(st_d)
    t_mp = []
    for i in range(3 + max(3, args.steps // 4)):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng_mp.set_iterate(x_ws, u_ws); eng_mp.solve(x0_d, u0_mp, st_mp); e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            t_mp.append(e0.elapsed_time(e1))
    both = (st_mp == 0) & (st_d == 0)
    dev_u0 = float(((u0_mp - u0_d)[both].abs().max() / u0_d[both].abs().max().clamp(min=1.0)).item())
    mp_ms = torch.tensor([sum(t_mp) / len(t_mp)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(mp_ms, op=dist.ReduceOp.MAX)
    mixed_precision = {"what": "linearize_fp32: IRK collocation, Newton solves and sensitivities in fp32; iterate, [B A], c and the QP in fp64",
                       "ms_per_step": float(mp_ms.item()), "value": world * B / (float(mp_ms.item()) * 1e-3), "unit": "solves/s",
                       "u0_max_rel_deviation_from_fp64": dev_u0, "stated_tolerance": 1e-4}
    eng_mp.close()

    # per-rank record: the time is the slowest rank's; the shards differ in their interior-point work
    mine = torch.tensor([sum(times) / args.steps, float(stats[:, 0].sum().item())], dtype=torch.float64, device=dev)
    per_rank = [torch.zeros_like(mine) for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, mine)
    else:
        per_rank = [mine]
    per_rank = [[float(t[0].item()), float(t[1].item())] for t in per_rank]

    extra = {}
    if rank == 0 and not args.no_extras:
        extra = other_configs(torch, dev)

    # final gather of controls / status over NCCL (the path's only collective), after the timed loop
    from ihm2_b200.dist import gather_results
    gu, gs, gst = gather_results(u0_d, [B] * world), gather_results(st_d, [B] * world), gather_results(stats, [B] * world)
    torch.cuda.synchronize()
    status_all = gs.cpu().numpy()
    assert gu.shape[0] == world * B and gst.shape[0] == world * B

    if rank == 0:
        fl = flops_per_solve(n_ipm)
        peak = measure_fp64_tflops()
        ach = B * fl["qp"] / (qp_ms * 1e-3) * 1e-12
        n_cpu = 1024
        cpu_val, cpu_t = cpu_oracle_rate(x0_h, x_h, u_h, mp, n_cpu)
        traffic, traffic_src = None, None   # dram bytes of the QP launch from the committed ncu --set full summary
        prof = os.path.join(ROOT, "profiles", "r02_qp_rti_kernel_ncu_summary.csv")
        if os.path.exists(prof) and B == B_PER_GPU:
            unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tb = 0.0
            for line in open(prof):
                f = line.strip().split(",")
                if len(f) == 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tb += float(f[2]) * unit.get(f[1], 1.0)
            if tb > 0:
                traffic, traffic_src = tb, "profiles/r02_qp_rti_kernel_ncu_summary.csv (dram__bytes_read.sum + dram__bytes_write.sum, one launch)"
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        hbm = json.load(open(peaks_path))["hbm_gbs"] if os.path.exists(peaks_path) else 6650.0
        out = {
            "metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "p50_batch_latency_ms": statistics.median(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "batch_per_gpu": B, "integrator": "IRK-GL4 (controller.py:305-307)", "qp": "IPM, tol 1e-9, <=50 iterations",
                       "l2": "flushed between timed steps (256 MiB write, outside the event pairs)",
                       "launch_order": "longest first by the IPM iteration counts of the handle's previous solve (scheduling only, results bit-identical; "
                                       "the previous solve here is the previous step on the same inputs, in a closed loop it is the previous control step)",
                       "ms_per_step_without_history": ms_no_history, "value_without_history": world * B / (ms_no_history * 1e-3),
                       "parallelism": f"{world} rank(s), problems sharded by index, final NCCL gather after the timed loop",
                       "ipm_iterations_mean": n_ipm},
            "status_counts": {str(k): int(v) for k, v in enumerate(np.bincount(status_all, minlength=5)) if v},
            "solved": int((status_all == 0).sum()), "failed": int((status_all != 0).sum()),
            "failed_note": "status 4: QPs of the hard-bounded OCP of python/mpc.py that the perturbation makes infeasible (each one checked "
                           "with an LP in tests/test_oracle_soft.py); they stop at the interior point's infeasibility exit and count as solves in `value`",
            "value_solved_only": value * float((status_all == 0).mean()),
            "per_rank": {"ms_per_step": [t[0] for t in per_rank], "ipm_iterations_per_step": [t[1] for t in per_rank]},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "solves/s", "h2d_bytes_per_step": B * 8 * 8, "d2h_bytes_per_step": B * (2 * 8 + 4)},
            "e2e_with_trajectories": {"value": e2e2_val, "unit": "solves/s", "h2d_bytes_per_step": B * 8 * 8,
                                      "d2h_bytes_per_step": B * (2 * 8 + 4) + B * ((N_HORIZON + 1) * 8 + N_HORIZON * 2) * 8,
                                      "note": "as e2e plus the read-back of x_pred / u_pred of every stage (new_python/controller.py:402-403) into page-locked host buffers"},
            "reference_setup": reference_setup,
            "mixed_precision": mixed_precision,
            "gpu_launches": int(launches),
            "kernels_ms": {"linearize_irk": lin_ms, "qp_rti": qp_ms},
            "roofline": {"bound": "tensor", "bound_note": "FP64: the kernel's serial sweeps run on the FP64 tensor pipe (DMMA, mma.m8n8k4.f64), the row passes on the FP64 CUDA cores; both peak at the same rate on B200 and the denominator is the larger of the two, measured in this process -- not the bf16 tensor figure and not HBM (SURVEY.md 8d)", "kernel": "qp_rti_kernel", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach / peak if peak else None, "traffic": traffic, "traffic_unit": "bytes per launch", "traffic_source": traffic_src,
                         "peak_source": "larger of a DFMA-chain and a DMMA-chain (mma.m8n8k4.f64) microbenchmark run in this process (MEASURED_PEAKS.json has no FP64 entry)",
                         "flops_per_solve": fl["total"], "flops_per_launch": B * fl["qp"],
                         "hbm_secondary": {"algorithmic_bytes_per_solve": 6800, "achieved_gbs": value / world * 6800 * 1e-9, "peak_gbs": hbm}},
            "cpu_baseline": {"value": cpu_val, "unit": "solves/s", "cores": CPU_THREADS, "kind": "port",
                             "sample": f"first {n_cpu} of the {B} problems of one step, OpenMP over problems, {cpu_t:.2f} s"},
        }
        out.update(extra)
        emit(out)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
