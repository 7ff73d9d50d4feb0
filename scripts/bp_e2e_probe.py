#!/usr/bin/env python
"""Host-buffer decode_batch of the hqc128 bench batch: ms per call with / without the early copy of finished rows."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sca_ldpc_b200 import _lib  # noqa: E402
from sca_ldpc_b200.ldpc import bp_decoder  # noqa: E402
from sca_ldpc_b200.workloads import HqcWorkload  # noqa: E402

variant = sys.argv[1] if len(sys.argv) > 1 else "hqc128"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
wl = HqcWorkload(variant, seed=0)
b = wl.generate(T, seed=1)
dec = bp_decoder(csr=(wl.row_ptr, wl.col), shape=(wl.m, wl.n), max_iter=100, channel_probs=wl.channel_probs)
h_in = torch.from_numpy(b.syndromes).pin_memory().numpy()
h_out = torch.empty((T, wl.n), dtype=torch.uint8).pin_memory().numpy()
d_in = torch.from_numpy(b.syndromes).cuda()
rows = C.c_int()
for name, fn in (("device", lambda: dec.decode_batch(d_in)), ("host pinned", lambda: dec.decode_batch(h_in, out=h_out)),
                 ("host pageable out", lambda: dec.decode_batch(h_in))):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    _lib.check(_lib.lib().scaldpc_bp_last_early_copy(dec._h, C.byref(rows)))
    print(f"{name:18s} {(time.perf_counter() - t0) / 5 * 1e3:8.2f} ms per call, early-copy rows {rows.value}, "
          f"(iterations, compactions) {dec.last_run()}")
