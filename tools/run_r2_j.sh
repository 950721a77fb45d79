#!/bin/bash
# one GPU: final rows-kernel sweep, c4 launch list (ncu, duration only), full ncu capture of the c4 SpMV kernels and of the rows
# kernel on the two particle matrices
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
SWEEP_QUICK=1 timeout 900 python tools/spmv_sweep.py 160 > gpurun_out/j_spmv_sweep.txt 2> gpurun_out/j_spmv_sweep.err; echo "sweep rc=$?"
grep -E "==|best|vector|stream|library" gpurun_out/j_spmv_sweep.txt
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --graph 0"
$CMD > gpurun_out/j_plain.log 2>&1; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/j_launches_c4.csv $CMD > gpurun_out/j_ncu_list.log 2>&1; echo "ncu list rc=$?"
python tools/launch_summary.py gpurun_out/j_launches_c4.csv 30 > gpurun_out/j_launches_c4.txt; head -n 25 gpurun_out/j_launches_c4.txt
ncu --set full --clock-control none --import-source on -k regex:k_spmv_ --launch-skip 30 --launch-count 36 -o gpurun_out/j_ncu_full_c4 $CMD > gpurun_out/j_ncu_full.log 2>&1; echo "ncu full rc=$?"
python tools/profile_spmv.py 2.1 128 3 > gpurun_out/j_prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_spmv_rows -s 2 -c 1 -o gpurun_out/j_ncu_rows_r21 python tools/profile_spmv.py 2.1 128 3 > gpurun_out/j_ncu_r21.log 2>&1; echo "ncu r21 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_spmv_rows -s 2 -c 1 -o gpurun_out/j_ncu_rows_r31 python tools/profile_spmv.py 3.1 112 3 > gpurun_out/j_ncu_r31.log 2>&1; echo "ncu r31 rc=$?"
ls -la gpurun_out/j_*.ncu-rep
