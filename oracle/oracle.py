"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the CPU oracle (oracle/liboracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The product package (sca_ldpc_b200) never does.

Classes mirror the reference's Python-facing decoder objects so parity tests read like the reference's own:
  QaryOracle            <-> simulate_rs.DecoderN*R*V*C*B* / DecoderN*R*SW*   (pydecoder.rs:24-65, 96-145)
  BinaryBpOracle        <-> ldpc.bp_decoder(..., bp_method="product_sum")     (hqc.py:694-708)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

ERRORS = {
    -1: "bad argument",
    -2: "probabilities do not sum to 1 +- 0.001 (decoder.rs:683-684)",
    -3: "no maximum probability found (decoder.rs:680)",
    -4: "no valid configuration for a check (decoder.rs:618)",
    -5: "node degree exceeds DV/DC (decoder.rs:465-473)",
    -6: "parity-check matrix is not [H' | I]-shaped / BSUM % B != 0 (decoder_special.rs:388-395,513-522)",
    -7: "sum value out of [-BSUM, BSUM] (decoder_special.rs:546)",
}

ALGO_ENUMERATE, ALGO_EXACT_DP, ALGO_FWD_BWD, ALGO_FAST, ALGO_SPLIT = 0, 1, 2, 3, 4


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__(f"oracle error {code}: {ERRORS.get(code, '?')}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile oracle/liboracle.so with gcc (make).  Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in ("qary_oracle.c", "qary_impl.inc", "bp_oracle.c", "Makefile")]
    if force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs
    ):
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, ip, i = C.c_void_p, C.POINTER(C.c_int), C.c_int
        L.oracle_graph_build.argtypes = [vp, i, i, i, i, i, i, C.POINTER(vp)]
        L.oracle_graph_build.restype = i
        L.oracle_graph_free.argtypes = [vp]
        L.oracle_graph_free.restype = None
        L.oracle_graph_edges.argtypes = [vp]
        L.oracle_graph_edges.restype = i
        for suf in ("f32", "f64"):
            f = getattr(L, f"oracle_into_llr_{suf}")
            f.argtypes = [vp, i, i, vp]
            f.restype = i
            f = getattr(L, f"oracle_min_sum_{suf}")
            f.argtypes = [vp, i, i, vp, vp, vp, vp, vp]
            f.restype = i
        L.oracle_min_sum_batch_f32.argtypes = [vp, i, i, i, i, vp, vp, vp, vp, vp, vp, i]
        L.oracle_min_sum_batch_f32.restype = i
        L.oracle_max_threads.restype = i
        L.oracle_has_openmp.restype = i
        L.bp_graph_build.argtypes = [i, i, vp, vp, C.POINTER(vp)]
        L.bp_graph_build.restype = i
        L.bp_graph_free.argtypes = [vp]
        L.bp_graph_free.restype = None
        L.bp_product_sum_decode.argtypes = [vp, vp, vp, i, vp, vp, ip]
        L.bp_product_sum_decode.restype = i
        L.bp_product_sum_decode_batch.argtypes = [vp, i, vp, vp, i, i, vp, vp, vp, vp, i]
        L.bp_product_sum_decode_batch.restype = None
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def into_llr(probs: np.ndarray, dtype=np.float32) -> np.ndarray:
    """decoder.rs:668-692: llr[q] = ln(max_p / p[q]); raises where the reference asserts."""
    probs = np.ascontiguousarray(probs, dtype=np.float32)
    rows, q = probs.shape
    out = np.empty((rows, q), dtype=dtype)
    fn = lib().oracle_into_llr_f32 if dtype == np.float32 else lib().oracle_into_llr_f64
    rc = fn(_ptr(probs), rows, q, _ptr(out))
    if rc:
        raise OracleError(rc)
    return out


class QaryOracle:
    """CPU restatement of Decoder (bsum=0) / DecoderSpecial (bsum>0).

    QaryOracle(H, iterations, B=, DV=, DC=)            generic, single alphabet  (decoder.rs)
    QaryOracle(H, iterations, B=, BSUM=, DV=, DC=)     special, H = [H' | I]     (decoder_special.rs)
    """

    def __init__(self, H, iterations, B, BSUM=0, DV=None, DC=None, algo=ALGO_ENUMERATE):
        H = np.ascontiguousarray(H)
        if H.dtype != np.int8:
            raise TypeError("parity check matrix must be int8 (pydecoder.rs:24)")
        self.R, self.N = H.shape
        self.B, self.BSUM = int(B), int(BSUM)
        self.iterations = int(iterations)
        self.algo = algo
        nz = H != 0
        self.DV = int(DV) if DV is not None else int(nz.sum(axis=0).max())
        self.DC = int(DC) if DC is not None else int(nz.sum(axis=1).max())
        g = C.c_void_p()
        rc = lib().oracle_graph_build(_ptr(H), self.R, self.N, self.DV, self.DC, self.B, self.BSUM, C.byref(g))
        if rc:
            raise OracleError(rc)
        self._g = g
        self.E = lib().oracle_graph_edges(g)
        self.bvars = self.N - self.R if self.BSUM else self.N
        self.Q, self.QS = 2 * self.B + 1, 2 * self.BSUM + 1

    def __del__(self):
        if getattr(self, "_g", None) and _lib is not None:
            _lib.oracle_graph_free(self._g)
            self._g = None

    # -- reference-shaped entry point: probabilities in, list of ints out (pydecoder.rs:53-65,125-145)
    def min_sum(self, channel_output, channel_output_sum=None):
        hard, _, _ = self.decode_probs(channel_output, channel_output_sum)
        return [int(x) for x in hard]

    def decode_probs(self, channel_output, channel_output_sum=None, dtype=np.float32):
        llr = into_llr(np.asarray(channel_output)[: self.bvars, : self.Q], dtype)
        llr_sum = None
        if self.BSUM:
            llr_sum = into_llr(np.asarray(channel_output_sum)[: self.R, : self.QS], dtype)
        return self.decode_llr(llr, llr_sum, dtype)

    def decode_llr(self, llr, llr_sum=None, dtype=np.float32, algo=None):
        """Returns (hard int8[N], posterior[bvars,Q], posterior_sum[R,QS] or None)."""
        algo = self.algo if algo is None else algo
        llr = np.ascontiguousarray(llr, dtype=dtype)
        assert llr.shape == (self.bvars, self.Q), llr.shape
        if self.BSUM:
            llr_sum = np.ascontiguousarray(llr_sum, dtype=dtype)
            assert llr_sum.shape == (self.R, self.QS), llr_sum.shape
        hard = np.zeros(self.N, dtype=np.int8)
        post = np.empty((self.bvars, self.Q), dtype=dtype)
        post_sum = np.empty((self.R, self.QS), dtype=dtype) if self.BSUM else None
        fn = lib().oracle_min_sum_f32 if dtype == np.float32 else lib().oracle_min_sum_f64
        rc = fn(self._g, self.iterations, algo, _ptr(llr), _ptr(llr_sum), _ptr(hard), _ptr(post), _ptr(post_sum))
        if rc:
            raise OracleError(rc)
        return hard, post, post_sum

    def decode_batch(self, x, x_sum=None, from_probs=False, algo=None, threads=0, want_post=True):
        """T decodes, one per thread (the reference's ThreadPoolExecutor model, decode.py:247-262).
        Returns (hard[T,N], post[T,bvars,Q], post_sum[T,R,QS] or None, status[T])."""
        algo = self.algo if algo is None else algo
        x = np.ascontiguousarray(x, dtype=np.float32)
        T = x.shape[0]
        assert x.shape == (T, self.bvars, self.Q), x.shape
        if self.BSUM:
            x_sum = np.ascontiguousarray(x_sum, dtype=np.float32)
            assert x_sum.shape == (T, self.R, self.QS), x_sum.shape
        hard = np.zeros((T, self.N), dtype=np.int8)
        post = np.empty((T, self.bvars, self.Q), dtype=np.float32) if want_post else None
        post_sum = np.empty((T, self.R, self.QS), dtype=np.float32) if (self.BSUM and want_post) else None
        status = np.zeros(T, dtype=np.int32)
        lib().oracle_min_sum_batch_f32(
            self._g, self.iterations, algo, T, int(from_probs), _ptr(x), _ptr(x_sum), _ptr(hard), _ptr(post),
            _ptr(post_sum), _ptr(status), int(threads),
        )
        return hard, post, post_sum, status


def max_threads() -> int:
    return lib().oracle_max_threads()


def has_openmp() -> bool:
    return bool(lib().oracle_has_openmp())


def dense_to_csr(H):
    H = np.asarray(H)
    m, n = H.shape
    rows, cols = np.nonzero(H % 2)
    row_ptr = np.zeros(m + 1, dtype=np.int32)
    np.add.at(row_ptr, rows + 1, 1)
    row_ptr = np.cumsum(row_ptr).astype(np.int32)
    return row_ptr, cols.astype(np.int32)


class BinaryBpOracle:
    """fp64 restatement of ldpc.bp_decoder (product_sum).  Constructor keywords follow the call sites
    hqc.py:694-699 and decode.py:155-161.  PARITY UNPINNED (see bp_oracle.c header)."""

    def __init__(self, H=None, error_rate=None, max_iter=0, bp_method="product_sum", channel_probs=None,
                 csr=None, shape=None):
        if bp_method not in ("product_sum", "ps", 0):
            raise ValueError("only bp_method='product_sum' is restated")
        if csr is not None:
            row_ptr, col = csr
            self.m, self.n = shape
        else:
            H = np.asarray(H)
            self.m, self.n = H.shape
            row_ptr, col = dense_to_csr(H)
        self.row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int32)
        self.col = np.ascontiguousarray(col, dtype=np.int32)
        g = C.c_void_p()
        rc = lib().bp_graph_build(self.m, self.n, _ptr(self.row_ptr), _ptr(self.col), C.byref(g))
        if rc:
            raise OracleError(rc)
        self._g = g
        self.max_iter = int(max_iter) if max_iter else self.n
        if channel_probs is not None and len(channel_probs) and channel_probs[0] is not None:
            cp = np.asarray(channel_probs, dtype=np.float64)
            if cp.shape[0] != self.n:
                raise ValueError("channel_probs has the wrong length")
        elif error_rate is not None:
            cp = np.full(self.n, float(error_rate), dtype=np.float64)
        else:
            raise ValueError("either error_rate or channel_probs is required")
        self.channel_probs = np.ascontiguousarray(cp)
        self.log_prob_ratios = np.zeros(self.n)
        self.iter = 0
        self.converge = 0

    def __del__(self):
        if getattr(self, "_g", None) and _lib is not None:
            _lib.bp_graph_free(self._g)
            self._g = None

    def _syndrome(self, v):
        s = np.zeros(self.m, dtype=np.uint8)
        rows = np.repeat(np.arange(self.m), np.diff(self.row_ptr))
        np.add.at(s, rows, v[self.col].astype(np.uint8))
        return s & 1

    def decode(self, vec):
        vec = np.asarray(vec).astype(np.uint8)
        if vec.shape[0] == self.n:
            synd = self._syndrome(vec)
            received = vec
        elif vec.shape[0] == self.m:
            synd = vec & 1
            received = None
        else:
            raise ValueError("input vector has the wrong length")
        synd = np.ascontiguousarray(synd, dtype=np.uint8)
        dec = np.zeros(self.n, dtype=np.uint8)
        it = C.c_int(0)
        self.converge = lib().bp_product_sum_decode(
            self._g, _ptr(synd), _ptr(self.channel_probs), self.max_iter, _ptr(dec), _ptr(self.log_prob_ratios),
            C.byref(it),
        )
        self.iter = it.value
        if received is not None:
            dec = dec ^ received
        return dec.astype(int)

    def decode_batch(self, synd, channel_probs=None, threads=0):
        """synd (T,m) uint8; channel_probs None (shared), or (T,n).  Returns dict of arrays."""
        synd = np.ascontiguousarray(synd, dtype=np.uint8)
        T = synd.shape[0]
        if channel_probs is None:
            cp, stride = self.channel_probs, 0
        else:
            cp = np.ascontiguousarray(channel_probs, dtype=np.float64)
            stride = self.n if cp.ndim == 2 else 0
        dec = np.zeros((T, self.n), dtype=np.uint8)
        lpr = np.zeros((T, self.n), dtype=np.float64)
        iters = np.zeros(T, dtype=np.int32)
        conv = np.zeros(T, dtype=np.int32)
        lib().bp_product_sum_decode_batch(
            self._g, T, _ptr(synd), _ptr(cp), stride, self.max_iter, _ptr(dec), _ptr(lpr), _ptr(iters), _ptr(conv),
            int(threads),
        )
        return {"decoding": dec, "log_prob_ratios": lpr, "iters": iters, "converged": conv}
