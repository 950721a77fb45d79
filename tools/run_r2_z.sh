#!/bin/bash
# one GPU, coded columns on by default: the whole GPU suite, one bench line per single-GPU BASELINE config, the reference arm,
# the c4 launch list (ncu, duration only) and a full ncu capture of the c4 SpMV kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | head -2; nproc
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider -rf > gpurun_out/z_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/z_pytest.log | cut -c 1-400
timeout 600 python bench.py > gpurun_out/z_bench_c4.json 2> gpurun_out/z_bench_c4.err; echo "c4 rc=$?"
tail -n 1 gpurun_out/z_bench_c4.json | cut -c 1-3000
timeout 400 python bench.py --config c2 > gpurun_out/z_bench_c2.json 2> gpurun_out/z_bench_c2.err; echo "c2 rc=$?"
tail -n 1 gpurun_out/z_bench_c2.json | cut -c 1-1200
timeout 600 python bench.py --config c3 --re 2.1 --cpu-seconds 15 > gpurun_out/z_bench_c3_r21.json 2> gpurun_out/z_bench_c3_r21.err; echo "c3 2.1 rc=$?"
tail -n 1 gpurun_out/z_bench_c3_r21.json | cut -c 1-600
timeout 600 python bench.py --config c3 --re 3.1 --cpu-seconds 15 > gpurun_out/z_bench_c3_r31.json 2> gpurun_out/z_bench_c3_r31.err; echo "c3 3.1 rc=$?"
tail -n 1 gpurun_out/z_bench_c3_r31.json | cut -c 1-600
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/z_bench_ref.json 2> gpurun_out/z_bench_ref.err; echo "ref rc=$?"
tail -n 1 gpurun_out/z_bench_ref.json | cut -c 1-800
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --graph 0"
$CMD > gpurun_out/z_plain.log 2>&1; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/z_launches_c4.csv $CMD > gpurun_out/z_ncu_list.log 2>&1; echo "ncu list rc=$?"
python tools/launch_summary.py gpurun_out/z_launches_c4.csv 40 > gpurun_out/z_launches_c4.txt; head -n 30 gpurun_out/z_launches_c4.txt
ncu --set full --clock-control none --import-source on -k regex:k_spmv_ --launch-skip 30 --launch-count 24 -o gpurun_out/z_ncu_full_c4 $CMD > gpurun_out/z_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
