#!/bin/bash
# bench lines of hqc128 / hqc256 only (tag = first argument)
tag=${1:-r2}
out=gpurun_out; mkdir -p $out
for wl in hqc128 hqc256; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline --brief > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err
  python - <<PY
import json
d=json.loads(open("$out/${tag}_bench_$wl.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("$wl", round(d["value"]), r["kernel"][:8], round(r["frac"],3), round(r["ms_per_launch"],3), "second", round(r["second_kernel"]["frac"],3))
PY
done
