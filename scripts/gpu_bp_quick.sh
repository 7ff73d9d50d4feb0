#!/bin/bash
# binary BP quick pass: its GPU parity tests, then bench lines of the HQC workloads named on the command line.
tag=${1:-r2}; shift
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_bp.py tests/test_zz_no_alloc.py -m gpu -x -q > $out/${tag}_pytest_bp.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_bp.log
tail -3 $out/${tag}_pytest_bp.log
for wl in "$@"; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?"
done
