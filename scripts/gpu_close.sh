#!/bin/bash
# Closing pass without ncu: every GPU test, smoke, the bench line of every workload, the reference arm.
tag=${1:-r2close}
out=gpurun_out; mkdir -p $out
bash scripts/gpu_pass.sh $tag kyber768 kyber1024 hqc128 hqc192 hqc256 decoder450
timeout 900 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench_kyber512.json 2> $out/${tag}_bench_kyber512.err; echo "bench default exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "reference arm exit $?"
