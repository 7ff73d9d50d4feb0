#!/bin/bash
# HQC (binary BP) GPU pass: parity tests, bench line, launch list, full captures of the check / bit passes.
tag=${1:-hqc}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_bp.py -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 600 python bench.py --workload hqc128 --steps 3 --warmup 3 > $out/${tag}_bench_hqc128.json 2> $out/${tag}_bench_hqc128.err; echo "exit $?" >> $out/${tag}_bench_hqc128.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/${tag}_launches_hqc128.csv \
  python bench.py --workload hqc128 --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bp_cn -s 1 -c 1 -f -o $out/${tag}_bp_cn \
  python bench.py --workload hqc128 --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_cn.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bp_vn -s 8 -c 3 -f -o $out/${tag}_bp_vn \
  python bench.py --workload hqc128 --steps 1 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_vn.log 2>&1
tail -5 $out/${tag}_pytest.log; cat $out/${tag}_bench_hqc128.json; tail -3 $out/${tag}_bench_hqc128.err
