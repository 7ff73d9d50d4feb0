#!/bin/bash
tag=${1:-r2f}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 600 python scripts/hqc_decode_every.py --variant hqc128 --runs 1024 --decode-every 500 --rows 3000 --compare > $out/${tag}_hqc128_decode_every.csv 2>&1
python - <<'PY' > $out/${tag}_cli_hqc128.txt 2>&1
import pickle, numpy as np, time, sys
sys.path.insert(0, '.')
z = np.load('tests/golden/hqc_keys.npz')
open('/tmp/test-hqc128.key', 'wb').write(pickle.dumps((z['hqc128_pk'].tobytes(), z['hqc128_sk'].tobytes())))
from sca_ldpc_b200 import cli
t0 = time.time()
cli.main(['hqc_decode_bench', '--key-file', '/tmp/test-hqc128.key', '--code-weight', '50', '--decode-every', '500', '--max-checks', '4000',
          '--runs', '64', '--seed', '1', '--label', 'b200', '--csv-output', 'gpurun_out/%s_cli_hqc128.csv' % sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/r2f_cli_hqc128.csv'])
print('# wall %.1f s' % (time.time() - t0))
PY
tail -12 $out/${tag}_pytest.log; cat $out/${tag}_hqc128_decode_every.csv; tail -5 $out/${tag}_cli_hqc128.txt
