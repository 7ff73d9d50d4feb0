#!/bin/bash
# N = 4 pass (under gpurun --gpus 4): bench lines of configs[4] / configs[1] and the Kyber-1024 sweep of configs[2]
N=4; tag=${1:-r2w}
out=gpurun_out; mkdir -p $out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 600 bash -c "$(declare -f run); N=$N; run 29511 bench.py --gpus $N --workload hqc256 --steps 3 --warmup 3 --no-cpu-baseline" > $out/${tag}_bench_hqc256_n$N.json 2> $out/${tag}_bench_hqc256_n$N.err; echo "exit $?" >> $out/${tag}_bench_hqc256_n$N.err
timeout 600 bash -c "$(declare -f run); N=$N; run 29512 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-sub" > $out/${tag}_bench_kyber512_n$N.json 2> $out/${tag}_bench_kyber512_n$N.err; echo "exit $?" >> $out/${tag}_bench_kyber512_n$N.err
timeout 600 bash -c "$(declare -f run); N=$N; run 29513 scripts/kyber_sweep.py --trials 100000" > $out/${tag}_kyber1024_sweep_n$N.jsonl 2> $out/${tag}_kyber1024_sweep_n$N.err; echo "exit $?" >> $out/${tag}_kyber1024_sweep_n$N.err
for f in $out/${tag}_bench_*_n$N.json; do echo $f; head -c 500 $f; echo; done; cat $out/${tag}_kyber1024_sweep_n$N.jsonl | cut -c1-400; tail -2 $out/${tag}_*_n$N.err
