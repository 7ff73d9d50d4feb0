#!/bin/bash
# Multi-GPU pass: bash scripts/gpu_multi.sh N tag  (under gpurun --gpus N)
N=${1:-2}; tag=${2:-r2g}
out=gpurun_out; mkdir -p $out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
timeout 900 bash -c "$(declare -f run); N=$N; run 29511 --workload hqc256 --steps 3 --warmup 3" > $out/${tag}_bench_hqc256_n$N.json 2> $out/${tag}_bench_hqc256_n$N.err; echo "exit $?" >> $out/${tag}_bench_hqc256_n$N.err
timeout 900 bash -c "$(declare -f run); N=$N; run 29512 --steps 10 --warmup 3" > $out/${tag}_bench_kyber512_n$N.json 2> $out/${tag}_bench_kyber512_n$N.err; echo "exit $?" >> $out/${tag}_bench_kyber512_n$N.err
if [ "$N" = "2" ]; then
timeout 900 bash -c "$(declare -f run); N=$N; run 29513 --impl reference --steps 2 --warmup 1" > $out/${tag}_bench_reference_n$N.json 2> $out/${tag}_bench_reference_n$N.err; echo "exit $?" >> $out/${tag}_bench_reference_n$N.err
fi
for f in $out/${tag}_bench_*_n$N.json; do echo $f; head -c 700 $f; echo; done; tail -q -n 3 $out/${tag}_bench_*_n$N.err
