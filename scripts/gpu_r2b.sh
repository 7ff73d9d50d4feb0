#!/bin/bash
# Round-2 GPU pass b: all GPU tests (early termination, hqc192/256), default bench, ncu launch list of one verified step.
tag=${1:-r2b}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=12 > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?" >> $out/${tag}_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench_kyber512.json 2> $out/${tag}_bench_kyber512.err; echo "bench exit $?" >> $out/${tag}_bench_kyber512.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 260 -c 420 --csv --log-file $out/${tag}_launches_k512.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --brief --no-sub > $out/${tag}_ncu_list.log 2>&1
tail -5 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log; cat $out/${tag}_bench_kyber512.json; tail -3 $out/${tag}_bench_kyber512.err
