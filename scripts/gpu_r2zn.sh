#!/bin/bash
# Round-2 closing pass of one build: every GPU test, smoke, bench lines of every workload, the reference arm, the ncu launch
# list of the default bench command and full captures of the four dominant kernels (summaries only come back).
tag=${1:-r2zn}
out=gpurun_out
mkdir -p $out
bash scripts/gpu_pass.sh $tag kyber768 kyber1024 hqc128 hqc192 hqc256 decoder450
timeout 900 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench_kyber512.json 2> $out/${tag}_bench_kyber512.err; echo "bench default exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "reference arm exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_k512.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sub --brief > $out/${tag}_ncu_list.log 2>&1
bash scripts/gpu_r2t_ncu.sh $tag > $out/${tag}_ncu.log 2>&1
ls $out | grep $tag | head -60
