"""ncu report (--set full) -> small CSV of the metrics the roofline discussion uses.
usage: python scripts/summarize_ncu.py <file.ncu-rep> <out.csv> "<note>"   (first kernel of the report)"""
import csv, subprocess, sys
rep, out, note = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(txt.splitlines()))
H, U, V = rr[0], rr[1], rr[2]
keys = set("""dram__bytes_read.sum dram__bytes_write.sum gpu__time_duration.sum launch__grid_size launch__block_size launch__registers_per_thread
launch__shared_mem_per_block_dynamic launch__occupancy_limit_shared_mem launch__occupancy_limit_registers launch__waves_per_multiprocessor
sm__warps_active.avg.per_cycle_active sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active
smsp__issue_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum smsp__thread_inst_executed_per_inst_executed.ratio
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum l1tex__data_pipe_lsu_wavefronts_mem_shared.sum sm__cycles_elapsed.avg
gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed lts__t_sector_hit_rate.pct lts__throughput.avg.pct_of_peak_sustained_elapsed
sm__throughput.avg.pct_of_peak_sustained_elapsed sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active
sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active
smsp__sass_thread_inst_executed_op_dfma_pred_on.sum smsp__sass_thread_inst_executed_op_dadd_pred_on.sum smsp__sass_thread_inst_executed_op_dmul_pred_on.sum
l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum""".split())
name = V[H.index("Kernel Name")] if "Kernel Name" in H else ""
with open(out, "w") as f:
    f.write(f"# ncu --set full --clock-control none, first captured launch of {name}; {note}\nmetric,unit,value\n")
    for i, h in enumerate(H):
        if h in keys or ("issue_stalled" in h and "per_issue_active" in h):
            f.write(f"{h},{U[i]},{V[i]}\n")
print(open(out).read())
