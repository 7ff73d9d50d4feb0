#!/bin/bash
# HQC (binary BP) GPU pass: parity tests, bench line, launch list, full captures of the check / bit passes.
tag=${1:-hqc}
out=gpurun_out
mkdir -p $out
# (parity tests and bench lines: scripts/gpu_check.sh; this script only profiles, on a 4096-trial batch)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/${tag}_launches_hqc128.csv \
  python bench.py --workload hqc128 --trials 4096 --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bp_cn -s 1 -c 1 -f -o $out/${tag}_bp_cn \
  python bench.py --workload hqc128 --trials 4096 --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_cn.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bp_vn -s 14 -c 4 -f -o $out/${tag}_bp_vn \
  python bench.py --workload hqc128 --trials 4096 --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_vn.log 2>&1
ls -la $out/${tag}_*
