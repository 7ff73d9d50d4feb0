#!/bin/bash
# Round-2 first GPU pass: parity tests, smoke, default bench (with the kyber768 sub-object), ncu of the exact kernel.
tag=${1:-r2a}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=15 > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?" >> $out/${tag}_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench_kyber512.json 2> $out/${tag}_bench_kyber512.err; echo "bench exit $?" >> $out/${tag}_bench_kyber512.err
tail -5 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log; cat $out/${tag}_bench_kyber512.json; tail -3 $out/${tag}_bench_kyber512.err
