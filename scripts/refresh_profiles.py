"""gpurun_out/{launches.csv, qp_full.ncu-rep, bench1.json, bench_ref.json} -> profiles/rNN_* (tracked summaries).
usage: python scripts/refresh_profiles.py r01 "<note>" """
import collections, csv, os, shutil, subprocess, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
note = sys.argv[2] if len(sys.argv) > 2 else ""

rows = [r for r in csv.reader(open(os.path.join(G, "launches.csv"))) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui, gi = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit"), H.index("Grid Size")
raw = []
for r in rows[hdr + 1:]:
    v = float(r[vi].replace(",", "")); u = r[ui]
    ms = v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v
    raw.append((r[0], r[ki].split("(")[0], r[gi].replace(",", " "), ms))
with open(os.path.join(P, f"{tag}_launches_raw.csv"), "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c 120 python bench.py --steps 3 --warmup 3 ({note})\nid,kernel,grid,ms\n")
    for r in raw:
        f.write(f"{r[0]},{r[1]},{r[2]},{r[3]:.6f}\n")
big = collections.defaultdict(lambda: [0, 0.0])
for _, name, g, ms in raw:
    if ("qp_rti" in name and ms > 1.0) or ("linearize" in name and ms > 0.3):   # the B = 4096 launches
        big[name][0] += 1; big[name][1] += ms
tot = sum(v[1] for v in big.values())
with open(os.path.join(P, f"{tag}_launches_summary.csv"), "w") as f:
    f.write("# shares over the B=4096 launches of the ncu launch list (the B=1 launches of the nominal-trajectory set-up are in the raw list only)\n")
    f.write(f"# {note}\nkernel,launches,total_ms,ms_per_launch,share\n")
    for k, v in sorted(big.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k},{v[0]},{v[1]:.4f},{v[1] / v[0]:.4f},{v[1] / tot:.4f}\n")

out = subprocess.run(["ncu", "-i", os.path.join(G, "qp_full.ncu-rep"), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(out.splitlines()))
Hh, U, V = rr[0], rr[1], rr[2]
keys = """dram__bytes_read.sum dram__bytes_write.sum gpu__time_duration.sum launch__grid_size launch__registers_per_thread
launch__shared_mem_per_block_dynamic launch__occupancy_limit_shared_mem launch__occupancy_limit_registers launch__waves_per_multiprocessor
sm__warps_active.avg.per_cycle_active sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
smsp__issue_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed sm__cycles_elapsed.avg
gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed lts__t_sector_hit_rate.pct lts__throughput.avg.pct_of_peak_sustained_elapsed
sm__throughput.avg.pct_of_peak_sustained_elapsed smsp__average_warp_latency_per_inst_issued.ratio sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_tensor_op_dmma.avg.pct_of_peak_sustained_active sm__pipe_tensor_op_dmma_cycles_active.avg.pct_of_peak_sustained_active
smsp__sass_thread_inst_executed_op_dfma_pred_on.sum smsp__sass_thread_inst_executed_op_dadd_pred_on.sum smsp__sass_thread_inst_executed_op_dmul_pred_on.sum""".split()
with open(os.path.join(P, f"{tag}_qp_rti_kernel_ncu_summary.csv"), "w") as f:
    f.write(f"# ncu --set full --clock-control none, qp_rti_kernel<40>, B = 4096 (config 2), one launch; {note}\nmetric,unit,value\n")
    for i, h in enumerate(Hh):
        if h in keys or ("issue_stalled" in h and "per_issue_active" in h) or ("dmma" in h.lower() and "pct" in h):
            f.write(f"{h},{U[i]},{V[i]}\n")
for a, b in (("bench1.json", f"{tag}_bench_n1.json"), ("bench_ref.json", f"{tag}_bench_reference.json")):
    if os.path.exists(os.path.join(G, a)):
        shutil.copy(os.path.join(G, a), os.path.join(P, b))
print(open(os.path.join(P, f"{tag}_launches_summary.csv")).read())
print(open(os.path.join(P, f"{tag}_qp_rti_kernel_ncu_summary.csv")).read())
