#!/usr/bin/env python
"""Times the two passes of the binary BP (scaldpc_bp_time_kernel) on the first T trials of a resident batch, T = 32 .. full,
and prints microseconds per tile: flat = HBM-bound at every size; falling at small T = an L2-resident schedule would pay
(it does not: small launches are latency-bound, profiles/r2q_l2probe_*.csv).  Also the A/B harness of bit-pass variants
(SCALDPC_BP_VN2=0/1 ...).
    python scripts/bp_l2_probe.py [hqc128|hqc256] [full batch]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from sca_ldpc_b200.ldpc import bp_decoder  # noqa: E402
from sca_ldpc_b200.workloads import HqcWorkload  # noqa: E402

variant = sys.argv[1] if len(sys.argv) > 1 else "hqc128"
full = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
sizes = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [32, 64, 96, 128, 192, 256, 512, 1024, 2048, full]
wl = HqcWorkload(variant=variant, seed=0)
dec = bp_decoder(csr=(wl.row_ptr, wl.col), shape=(wl.m, wl.n), max_iter=4, channel_probs=wl.channel_probs)
synd = np.random.RandomState(1).randint(0, 2, (full, wl.m)).astype(np.uint8)
dec.decode_batch(synd, early_stop=False)
mb_tile = dec.E * 128 / 1e6
print(f"# {variant}: E = {dec.E}, {mb_tile:.1f} MB of messages per tile; env " +
      " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("SCALDPC_")))
print("trials,tiles,MB,cn_us,vn_us,cn_us_per_tile,vn_us_per_tile,TBps_cn,TBps_vn")
for T in sizes:
    if T > full:
        continue
    cn = dec.time_kernel(0, T, reps=30) * 1e3
    vn = dec.time_kernel(1, T, reps=30) * 1e3
    tiles = (T + 31) // 32
    b = 2 * tiles * mb_tile * 1e6
    print(f"{T},{tiles},{tiles * mb_tile:.0f},{cn:.1f},{vn:.1f},{cn / tiles:.2f},{vn / tiles:.2f},{b / cn / 1e6:.2f},{b / vn / 1e6:.2f}")
