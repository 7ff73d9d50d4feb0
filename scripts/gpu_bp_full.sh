#!/bin/bash
# binary BP pass with the full bench line of hqc128 (single-decode leg included)
tag=${1:-r2}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_bp.py tests/test_gpu_cli.py tests/test_zz_no_alloc.py -m gpu -x -q > $out/${tag}_pytest_bp.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_bp.log
tail -3 $out/${tag}_pytest_bp.log
timeout 900 python bench.py --workload hqc128 --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_hqc128.json 2> $out/${tag}_bench_hqc128.err; echo "bench exit $?"
python - <<PY
import json
d=json.loads(open("$out/${tag}_bench_hqc128.json").read().strip().splitlines()[-1])
print(round(d["value"]), round(d["e2e"]["value"]), d["single_decode"])
PY
