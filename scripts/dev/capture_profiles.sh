#!/bin/bash
# Round-end capture on the GPU box (one GPU): bench lines, ncu launch list of the bench command, one ncu --set full capture
# each of the QP kernel and the lineariser.  Every command runs plain first; ncu only after it exited 0.
# Outputs under gpurun_out/ (scripts/refresh_profiles.py turns them into profiles/rNN_*).
set -u
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err || { echo "bench failed"; tail -5 gpurun_out/bench1.err; exit 1; }
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err || echo "reference arm failed"
python bench.py --no-extras --steps 3 --warmup 3 > gpurun_out/plain_bench.log 2>&1 || { echo "plain bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --no-extras --steps 3 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
python scripts/profile_qp.py 4096 2 > gpurun_out/plain_qp.log 2>&1 || { echo "profile_qp failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:qp_rti_kernel -s 21 -c 1 -f -o gpurun_out/qp_full \
  python scripts/profile_qp.py 4096 2 > gpurun_out/ncu_qp.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:linearize_irk_kernel -s 21 -c 1 -f -o gpurun_out/lin_full \
  python scripts/profile_qp.py 4096 2 > gpurun_out/ncu_lin.log 2>&1
ls -la gpurun_out/*.ncu-rep | tail -3
cut -c1-300 gpurun_out/bench1.json
