#!/bin/bash
# Quick GPU pass while iterating on a kernel: q-ary parity tests + short bench lines.
# Usage: bash scripts/gpu_quick.sh tag [workloads...]
tag=${1:-q}; shift
wls=${@:-kyber512 kyber768}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_qary.py -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
for wl in $wls; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "bench $wl exit $?"
  python - <<PY
import json
d=json.load(open("$out/${tag}_bench_$wl.json"))
r=d["roofline"]
print("$wl", "obs value %.0f e2e %.0f |" % ((d.get("observed") or {}).get("value",0), ((d.get("observed") or {}).get("e2e") or {}).get("value",0)), "value %.0f e2e %.0f ms/step %.3f | cn %.4f ms frac %.3f | vn %.4f ms frac %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], r["ms_per_launch"], r["frac"], r["second_kernel"]["ms_per_launch"], r["second_kernel"]["frac"]))
PY
done
