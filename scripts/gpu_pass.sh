#!/bin/bash
# Short GPU-box pass: parity tests, smoke, bench lines of the workloads named on the command line.
# Usage (under gpurun):  bash scripts/gpu_pass.sh <tag> [workload ...]
tag=${1:-r2}; shift
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?" >> $out/${tag}_smoke.log
for wl in "$@"; do
  steps=10; case $wl in hqc*|decoder450) steps=3;; esac
  timeout 900 python bench.py --workload $wl --steps $steps --warmup 3 > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?" >> $out/${tag}_bench_$wl.err
done
tail -3 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log
