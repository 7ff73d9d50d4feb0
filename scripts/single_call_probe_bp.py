#!/usr/bin/env python
"""The kernels of ONE reference-shaped binary call, bp_decoder.decode(syndrome) on the hqc128 bench workload, for an ncu launch
list:  ncu --metrics gpu__time_duration.sum --csv --log-file x.csv python scripts/single_call_probe_bp.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sca_ldpc_b200.ldpc import bp_decoder  # noqa: E402
from sca_ldpc_b200.workloads import HqcWorkload  # noqa: E402

w = HqcWorkload("hqc128", W=50, R=3000, rho=0.9942, seed=0)
b = w.generate(4, seed=1)
bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=100, channel_probs=w.channel_probs)
for i in range(3):
    t0 = time.perf_counter()
    out = bpd.decode(b.syndromes[i])
    print(i, (time.perf_counter() - t0) * 1e3, "ms", bpd.iter, bpd.converge)
