#!/usr/bin/env python
"""The kernels of ONE reference-shaped call, min_sum(channel_output, channel_output_sum) on the kyber512 bench workload, for
an ncu launch list:  SCALDPC_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --csv --log-file x.csv python scripts/single_call_probe.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sca_ldpc_b200 import QaryDecoder  # noqa: E402
from sca_ldpc_b200.workloads import KyberWorkload  # noqa: E402

w = KyberWorkload("kyber512", sum_weight=6, check_blocks=1, accuracy=0.98, nbits=4, nbits_sum=5, seed=0)
b = w.generate(4, seed=1)
d = QaryDecoder(w.H, 10, **w.decoder_kwargs)
for i in range(3):
    t0 = time.perf_counter()
    out = d.min_sum(b.priors[i], b.priors_sum[i])
    print(i, (time.perf_counter() - t0) * 1e3, "ms", sum(out[: w.bvars] == b.secret[i]) if hasattr(out, "__len__") else "")
