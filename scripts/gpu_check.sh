#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list and one full capture of the dominant kernel.
# Usage (from the repo root, under gpurun):  bash scripts/gpu_check.sh [tag]
tag=${1:-r1}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?" >> $out/${tag}_smoke.log
for wl in kyber512 kyber768 kyber1024; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?" >> $out/${tag}_bench_$wl.err
done
for wl in hqc128 hqc192 hqc256 decoder450; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?" >> $out/${tag}_bench_$wl.err
done
# ncu: launch list (cold-cache serialised) of the default bench command
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_k512.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_list.log 2>&1
# ncu: one full capture of the dominant kernel (skip the FIRST-iteration variant and warm-up launches)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_cn_fast -s 23 -c 1 -f -o $out/${tag}_cn_fast_k512 \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_vn -s 23 -c 1 -f -o $out/${tag}_vn_k512 \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_full_vn.log 2>&1
tail -3 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log; cat $out/${tag}_bench_kyber512.json
