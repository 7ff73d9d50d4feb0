#!/bin/bash
# Quick GPU pass for the binary BP path: parity tests + short bench lines (optionally A/B against the plain bit pass).
tag=${1:-qb}; shift
wls=${@:-hqc128}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_bp.py -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
for wl in $wls; do
  for mode in tma plain; do
    if [ $mode = tma ]; then export SCALDPC_BP_VN_TMA=1; else unset SCALDPC_BP_VN_TMA; fi
    timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_${wl}_$mode.json 2> $out/${tag}_bench_${wl}_$mode.err; echo "bench $wl $mode exit $?"
    python - <<PY
import json
d=json.load(open("$out/${tag}_bench_${wl}_$mode.json"))
r=d["roofline"]
print("$wl $mode", "value %.0f e2e %.0f ms/step %.3f succ %.4f it/trial %.3f | cn %.4f ms frac %.3f | vn %.4f ms frac %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["success_rate"], d["iterations_per_trial"], r["ms_per_launch"], r["frac"], r["second_kernel"]["ms_per_launch"], r["second_kernel"]["frac"]))
PY
  done
done
