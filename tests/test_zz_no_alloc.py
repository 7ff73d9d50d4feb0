"""scaldpc_bp_decode_static: the allocation-free variant of scaldpc_bp_decode (no counterpart in the reference).
Collected LAST on purpose (file name).

Checked through the C ABI against the existing entry point on the same inputs: decisions, iteration counts and convergence
flags must be IDENTICAL (integer outputs: the bar the other binary-BP tests use), on two calls in a row that share the static
scratch (the second with other inputs and another batch size, so stale contents of the first would show), and a request
beyond the array's size must be refused with a clear error."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _code(rs, m, n_msg, w):
    """H = [Hin | I]: m rows of weight w over n_msg message bits plus the identity, CSR with ascending columns."""
    cols = np.stack([np.sort(rs.choice(n_msg, w, replace=False)) for _ in range(m)])
    col = np.concatenate((cols, n_msg + np.arange(m)[:, None]), axis=1).astype(np.int32)
    row_ptr = (np.arange(m + 1) * (w + 1)).astype(np.int32)
    return row_ptr, col.reshape(-1), cols


def _syndromes(rs, cols, n_msg, T, weight, noise):
    m = cols.shape[0]
    synd = np.zeros((T, m), np.uint8)
    for t in range(T):
        e = np.zeros(n_msg, np.uint8)
        e[rs.choice(n_msg, weight, replace=False)] = 1
        synd[t] = (e[cols].sum(1) & 1) ^ (rs.rand(m) < noise)
    synd[T // 2:] = rs.randint(0, 2, (T - T // 2, m))  # half of the batch never converges
    return synd


def test_static_scratch_decode_equals_the_allocating_decode():
    import torch

    from sca_ldpc_b200 import _lib

    L = _lib.lib()
    rs = np.random.RandomState(5)
    m, n_msg, w = 300, 600, 6
    n = n_msg + m
    row_ptr, col, cols = _code(rs, m, n_msg, w)
    h = C.c_void_p()
    _lib.check(L.scaldpc_bp_create(m, n, row_ptr.ctypes.data, col.ctypes.data, 25, 0, C.byref(h)))
    try:
        probs = torch.tensor(np.concatenate((np.full(n_msg, 8 / n_msg), np.full(m, 0.02))), dtype=torch.float64, device="cuda:0")
        st = torch.cuda.current_stream().cuda_stream
        for T, weight in ((70, 8), (33, 6)):  # two calls in a row on the shared scratch
            synd = torch.tensor(_syndromes(rs, cols, n_msg, T, weight, 0.02), device="cuda:0")
            out = {}
            for name in ("ref", "static"):
                dec = torch.full((T, n), 7, dtype=torch.uint8, device="cuda:0")
                it = torch.full((T,), -1, dtype=torch.int32, device="cuda:0")
                cv = torch.full((T,), -1, dtype=torch.int32, device="cuda:0")
                torch.cuda.synchronize()
                if name == "ref":
                    rc = L.scaldpc_bp_decode(h, T, synd.data_ptr(), probs.data_ptr(), dec.data_ptr(), None, it.data_ptr(),
                                             cv.data_ptr(), _lib.DEVICE_PTRS, C.c_void_p(st))
                else:
                    before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
                    rc = L.scaldpc_bp_decode_static(h, T, synd.data_ptr(), probs.data_ptr(), dec.data_ptr(), it.data_ptr(),
                                                    cv.data_ptr(), C.c_void_p(st))
                    torch.cuda.synchronize()
                    assert torch.cuda.memory_allocated() == before
                    assert torch.cuda.mem_get_info()[0] >= free_before  # nothing was taken from the device either
                _lib.check(rc)
                torch.cuda.synchronize()
                out[name] = (dec.cpu().numpy(), it.cpu().numpy(), cv.cpu().numpy())
            for a, b in zip(out["ref"], out["static"]):
                assert np.array_equal(a, b)
            conv = out["static"][2]
            assert 0 < conv.sum() < T and set(np.unique(conv)) <= {0, 1}          # both kinds of trial were exercised
            assert out["static"][0].max() <= 1 and out["static"][1].min() >= 1
        # beyond the array: refused, with the sizes in the message
        big = int(_lib.BP_STATIC_BYTES // (128 * (m * (w + 1))) + 2) * 32
        rc = L.scaldpc_bp_decode_static(h, big, synd.data_ptr(), probs.data_ptr(), dec.data_ptr(), None, None, C.c_void_p(st))
        assert rc != 0
        msg = L.scaldpc_last_error().decode()
        assert "static" in msg and str(_lib.BP_STATIC_BYTES) in msg
    finally:
        L.scaldpc_bp_destroy(h)
