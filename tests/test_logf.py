"""sc_logf (sca_ldpc_b200/csrc/sc_logf.cuh) must be glibc's logf bit for bit: the reference's into_llr is
`(max / p).ln()` in f32 (decoder.rs:687), i.e. libm's logf.  Host evaluation of the device source (test only)."""
import ctypes as C
import ctypes.util

import numpy as np

import util


def _libm_logf(x):
    libm = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.logf.restype = C.c_float
    libm.logf.argtypes = [C.c_float]
    return np.array([libm.logf(float(v)) for v in x], dtype=np.float32)


def _sc_logf(x):
    L = util.hostmodel()
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty_like(x)
    L.hostmodel_logf(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), C.c_long(x.size))
    return y


def test_logf_bit_identical_to_libm():
    rs = np.random.RandomState(5)
    bits = rs.randint(0, 0x7F800000, size=200_000, dtype=np.int64).astype(np.uint32)  # all positive finite floats
    ratios = (1.0 / rs.random_sample(100_000)).astype(np.float32)                      # max/p, what into_llr feeds it
    edge = np.array([0.0, -0.0, 1.0, np.inf, -1.0, np.nan, 1e-45, 1.1754942e-38, 1.17549435e-38, 3.4028235e38,
                     np.nextafter(np.float32(1), np.float32(2)), np.nextafter(np.float32(1), np.float32(0)), 2.0, 0.5],
                    dtype=np.float32)
    x = np.concatenate((bits.view(np.float32), ratios, edge))
    got, ref = _sc_logf(x), _libm_logf(x)
    assert np.array_equal(got.view(np.uint32)[~np.isnan(ref)], ref.view(np.uint32)[~np.isnan(ref)])
    assert np.array_equal(np.isnan(got), np.isnan(ref))
