#!/bin/bash
# q-ary quick pass: q-ary GPU parity tests, then bench lines of the Kyber workloads named on the command line.
# Usage (under gpurun):  bash scripts/gpu_qary_quick.sh <tag> [workload ...]
tag=${1:-r2}; shift
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_qary.py -m gpu -x -q > $out/${tag}_pytest_qary.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_qary.log
for wl in "$@"; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?" >> $out/${tag}_bench_$wl.err
done
tail -3 $out/${tag}_pytest_qary.log
