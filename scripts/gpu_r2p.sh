#!/bin/bash
# full GPU test pass + default benches + launch list of the default bench
tag=${1:-r2p}
out=gpurun_out
mkdir -p $out
timeout 1800 python -m pytest tests -m gpu -q --durations=8 > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/${tag}_smoke.log 2>&1
timeout 900 python bench.py > $out/${tag}_bench_kyber512.json 2> $out/${tag}_bench_kyber512.err
timeout 600 python bench.py --workload hqc128 --no-sub > $out/${tag}_bench_hqc128.json 2> $out/${tag}_bench_hqc128.err
timeout 600 python bench.py --workload hqc256 --no-sub > $out/${tag}_bench_hqc256.json 2> $out/${tag}_bench_hqc256.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/${tag}_launches_k512.csv python bench.py --steps 2 --warmup 1 --no-sub --brief > $out/${tag}_ncu.log 2>&1
tail -6 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log; wc -l $out/${tag}_launches_k512.csv
