#!/bin/bash
tag=${1:-r2e}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q --durations=12 > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?" >> $out/${tag}_smoke.log
for wl in hqc128 hqc256; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?" >> $out/${tag}_bench_$wl.err
done
tail -15 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log; cat $out/${tag}_bench_hqc128.json | head -c 1500
