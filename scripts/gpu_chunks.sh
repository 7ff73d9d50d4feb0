#!/bin/bash
# e2e of the prior-vector API against the number of pipeline chunks (SCALDPC_QARY_CHUNKS), kyber512 bench workload.
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
for n in 8 12 16 24; do
  SCALDPC_QARY_CHUNKS=$n timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sub --brief > $out/${tag}_chunks_$n.json 2> $out/${tag}_chunks_$n.err
  python - <<PY
import json
d=json.loads(open("$out/${tag}_chunks_$n.json").read().strip().splitlines()[-1])
print("chunks $n value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
PY
done
