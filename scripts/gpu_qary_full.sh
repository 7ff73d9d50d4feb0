#!/bin/bash
# q-ary pass with the full default bench line (single-decode legs included, no CPU baseline, no kyber768 sub-object)
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_qary.py -m gpu -x -q > $out/${tag}_pytest_qary.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_qary.log
tail -3 $out/${tag}_pytest_qary.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sub > $out/${tag}_bench_kyber512.json 2> $out/${tag}_bench_kyber512.err; echo "bench exit $?"
