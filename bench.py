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


def reference_arm(args, rank: int):
    """--impl reference: the reference's own CPU implementation of the path cannot be built here
    (acados / HPIPM / CasADi absent), so this times the CPU oracle port, all host threads."""
    if rank != 0:
        return
    import __graft_entry__ as g
    g.build()
    from ihm2_b200.track import load_packaged_track
    from ihm2_b200.workloads import perturbed_batch

    from ihm2_b200 import _abi
    from oracle.lib import Oracle, initial_iterate

    mp = load_packaged_track(TRACK)
    # same nominal as the GPU arm (20 perfect-model RTI steps from the controller's initial guess)
    cfg1 = _abi.make_config(batch=1, n_grid=len(mp.s_ref))
    o1 = Oracle(cfg1, mp.s_ref, mp.kappa_ref)
    xi, ui = initial_iterate(cfg1)
    xs = np.zeros((1, 8)); xs[0, 0] = -6.0
    for _ in range(20):
        _, _, _, xi, ui = o1.solve(xs, xi, ui, do_shift=True)
        xs = xi[:, 1, :].copy()
    n_sample = 1024
    x0, x, u = perturbed_batch(xs[0], xi[0], ui[0], n_sample, seed=0)
    times = []
    for i in range(args.warmup + args.steps):
        _, dt = cpu_oracle_rate(x0, x, u, mp, n_sample)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    val = n_sample / (ms * 1e-3)
    sample = f"first {n_sample} of the {B_PER_GPU} problems per step, OpenMP over problems"
    emit({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "solves/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: 4096 perturbed initial states, single SQP-RTI step, data/skidpad, N=40, IRK-GL4",
                   "note": "acados/HPIPM not installable here: CPU restatement (oracle/ihm2_oracle.cpp), not acados"},
        "cpu_baseline": {"value": val, "unit": "solves/s", "cores": CPU_THREADS, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


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
    clocks = sampler.stop()

    # the same steps on a handle WITHOUT launch-order history (what a first solve of fresh problems costs):
    # the timed loop above runs with schedule_by_history, i.e. each step launches its problems longest first
    # by the iteration counts of the previous step, as consecutive control steps of a closed loop do
    cfg_nh = _abi.make_config(batch=B, n_grid=len(mp.s_ref), schedule_by_history=False)
    eng_nh = Engine(cfg_nh, mp.s_ref, mp.kappa_ref)
    t_nh = []
    for i in range(3 + max(3, args.steps // 4)):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng_nh.set_iterate(x_ws, u_ws); eng_nh.solve(x0_d, u0_d, st_d); e1.record()
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
        prof = os.path.join(ROOT, "profiles", "r01_qp_rti_kernel_ncu_summary.csv")
        if os.path.exists(prof) and B == B_PER_GPU:
            unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tb = 0.0
            for line in open(prof):
                f = line.strip().split(",")
                if len(f) == 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tb += float(f[2]) * unit.get(f[1], 1.0)
            if tb > 0:
                traffic, traffic_src = tb, "profiles/r01_qp_rti_kernel_ncu_summary.csv (dram__bytes_read.sum + dram__bytes_write.sum, one launch)"
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        hbm = json.load(open(peaks_path))["hbm_gbs"] if os.path.exists(peaks_path) else 6650.0
        out = {
            "metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "p50_batch_latency_ms": statistics.median(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[1]: 4096 perturbed initial states, single SQP-RTI step, data/skidpad, N=40",
                       "batch_per_gpu": B, "integrator": "IRK-GL4 (controller.py:305-307)", "qp": "IPM, tol 1e-9, <=50 iterations",
                       "l2": "flushed between timed steps (256 MiB write, outside the event pairs)",
                       "launch_order": "longest first by the IPM iteration counts of the handle's previous solve (scheduling only, results bit-identical; "
                                       "the previous solve here is the previous step on the same inputs, in a closed loop it is the previous control step)",
                       "ms_per_step_without_history": ms_no_history, "value_without_history": world * B / (ms_no_history * 1e-3),
                       "parallelism": f"{world} rank(s), problems sharded by index, final NCCL gather after the timed loop",
                       "ipm_iterations_mean": n_ipm,
                       "status_counts": {str(k): int(v) for k, v in enumerate(np.bincount(status_all, minlength=5)) if v}},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "solves/s", "h2d_bytes_per_step": B * 8 * 8, "d2h_bytes_per_step": B * (2 * 8 + 4)},
            "gpu_launches": int(launches),
            "kernels_ms": {"linearize_irk": lin_ms, "qp_rti": qp_ms},
            "roofline": {"bound": "fp64", "bound_note": "FP64 CUDA-core / FP64 tensor-core (DMMA) throughput, SURVEY.md 8d: neither HBM nor the bf16 tensor peak bounds this path", "kernel": "qp_rti_kernel", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach / peak if peak else None, "traffic": traffic, "traffic_unit": "bytes per launch", "traffic_source": traffic_src,
                         "peak_source": "larger of a DFMA-chain and a DMMA-chain (mma.m8n8k4.f64) microbenchmark run in this process (MEASURED_PEAKS.json has no FP64 entry)",
                         "flops_per_solve": fl["total"], "flops_per_launch": B * fl["qp"],
                         "hbm_secondary": {"algorithmic_bytes_per_solve": 6800, "achieved_gbs": value / world * 6800 * 1e-9, "peak_gbs": hbm}},
            "cpu_baseline": {"value": cpu_val, "unit": "solves/s", "cores": CPU_THREADS, "kind": "port",
                             "sample": f"first {n_cpu} of the {B} problems of one step, OpenMP over problems, {cpu_t:.2f} s"},
        }
        emit(out)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
