"""Host-side mirror of the reference's q-ary decoder objects over the libscaldpc C ABI.

Reference interface mirrored (paths relative to simulate-with-python/):
  simulate_rs/src/pydecoder.rs:13-71    class DecoderN{n}R{r}V{v}C{c}B{b}:  __new__(H, iterations), min_sum(probs)
  simulate_rs/src/pydecoder.rs:85-150   class DecoderN{n}R{r}SW{sw}:        __new__(H, iterations),
                                                                           min_sum(probs, probs_sum)
  simulate_rs/src/lib.rs:32-75          the registered concrete shapes
The compute runs on the GPU through scaldpc_qary_decode; nothing here computes messages on the host.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib

try:  # torch is plumbing only (device buffers / streams); numpy-only use works without it
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_torch(x) -> bool:
    return torch is not None and isinstance(x, torch.Tensor)


_STATUS_TEXT = {
    _lib.ERR_PROB_SUM: "a prior row does not sum to 1 +- 0.001 (decoder.rs:683-684)",
    _lib.ERR_NO_MAX: "No maximum probability found (decoder.rs:680)",
    _lib.ERR_NO_CONFIG: "a check has no valid configuration (decoder.rs:618)",
}


class QaryDecoder:
    """q-ary min-sum decoder on one B200.

    QaryDecoder(H, iterations, B=b, DV=v, DC=c)                 <-> Decoder<N,R,DV,DC,2B+1,B,i8>   (decoder.rs)
    QaryDecoder(H, iterations, B=b, BSUM=bsum, DV=v, DC=sw+1)   <-> DecoderSpecial<...>            (decoder_special.rs)
    """

    def __init__(self, H, iterations, B, BSUM=0, DV=0, DC=0, shape=None, device=0, csr=None):
        self._h = C.c_void_p()
        self._lock = threading.Lock()
        if csr is not None:
            # sparse constructor (scaldpc_qary_create_csr): csr = (row_ptr, col, val), `shape` = (R, N); H is ignored
            if int(iterations) < 0:
                raise OverflowError("iterations must be a non-negative integer (u32, pydecoder.rs:24)")
            rp, col, val = (np.ascontiguousarray(a, dtype=dt) for a, dt in zip(csr, (np.int32, np.int32, np.int8)))
            R, N = (int(x) for x in shape)
            if rp.shape != (R + 1,) or col.shape != val.shape or col.shape[0] != rp[-1]:
                raise ValueError("csr = (row_ptr[R+1], col[nnz], val[nnz])")
            _lib.check(_lib.lib().scaldpc_qary_create_csr(R, N, rp.ctypes.data, col.ctypes.data, val.ctypes.data, int(B),
                                                          int(BSUM), int(DV), int(DC), int(iterations), int(device),
                                                          C.byref(self._h)))
            self._finish_init(B, BSUM, device)
            return
        H = np.asarray(H) if not _is_torch(H) else H.cpu().numpy()
        if H.dtype != np.int8:
            # pyo3: PyReadonlyArray2<i8> extraction failure (pydecoder.rs:24, 96)
            raise TypeError(f"parity check matrix must be a 2-D int8 array, got dtype {H.dtype}")
        if H.ndim != 2:
            raise TypeError(f"parity check matrix must be 2-D, got {H.ndim}-D")
        if int(iterations) < 0:
            raise OverflowError("iterations must be a non-negative integer (u32, pydecoder.rs:24)")
        if shape is not None:
            # the reference copies py_parity_check[(row, col)] for the class's const R x N (pydecoder.rs:28-33):
            # a larger array is cropped, a smaller one panics on the out-of-bounds index
            if H.shape[0] < shape[0] or H.shape[1] < shape[1]:
                raise IndexError(f"parity check matrix is {H.shape[0]}x{H.shape[1]}, smaller than the decoder "
                                 f"class's {shape[0]}x{shape[1]} (ndarray index out of bounds, pydecoder.rs:31)")
            H = H[: shape[0], : shape[1]]
        R, N = H.shape
        H = np.ascontiguousarray(H)
        _lib.check(_lib.lib().scaldpc_qary_create(H.ctypes.data, R, N, int(B), int(BSUM), int(DV), int(DC),
                                                  int(iterations), int(device), C.byref(self._h)))
        self._finish_init(B, BSUM, device)

    def _finish_init(self, B, BSUM, device):
        vals = [C.c_int() for _ in range(7)]
        _lib.check(_lib.lib().scaldpc_qary_info(self._h, *[C.byref(v) for v in vals]))
        self.R, self.N, self.bvars, self.Q, self.QS, self.E, self.iterations = [v.value for v in vals]
        self.B, self.BSUM, self.device = int(B), int(BSUM), int(device)
        self.special = self.BSUM > 0

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib._lib is not None:
            _lib._lib.scaldpc_qary_destroy(h)

    # ------------------------------------------------------------------------------------------------------
    # reference-shaped entry point (pydecoder.rs:53-65, 125-145): probabilities in, list of ints out
    # ------------------------------------------------------------------------------------------------------
    def min_sum(self, channel_output, channel_output_sum=None):
        co = self._one(channel_output, self.bvars, self.Q, "channel_output")
        cos = None
        if self.special:
            if channel_output_sum is None:
                raise TypeError("min_sum() missing 1 required positional argument: 'channel_output_sum'")
            cos = self._one(channel_output_sum, self.R, self.QS, "channel_output_sum")
        elif channel_output_sum is not None:
            raise TypeError("min_sum() takes 1 positional argument but 2 were given")
        # Concurrent callers are COMBINED inside the library (scaldpc_qary_decode_one): the reference calls one decoder object
        # from a ThreadPoolExecutor (simulate/decode.py:247-262, GIL released in pydecoder.rs:55); ctypes releases the GIL for
        # the call, the first thread in decodes every request queued so far as one batch (a tile holds 32 trials, so 32
        # concurrent decodes cost what one costs) and the others wait for their row of the result.
        hard = np.empty(self.N, np.int8)
        rc = _lib.lib().scaldpc_qary_decode_one(self._h, co.ctypes.data, cos.ctypes.data if cos is not None else None,
                                                hard.ctypes.data)
        if rc in _STATUS_TEXT:  # the reference panics in THAT call only (decoder.rs:618, 680, 683-684)
            raise ValueError(_STATUS_TEXT[rc])
        _lib.check(rc)
        return hard.tolist()  # Python ints, like the pyo3 return value

    def min_sum_batch_f64(self, priors, priors_sum=None, *, llr=False, want_status=False):
        """fp64 verification mode (scaldpc_qary_decode_f64): the decode of min_sum_batch with every LLR, message and posterior in
        double.  priors (T, bvars, q) float32 [+ priors_sum (T, R, qs)]; returns hard int8 (T, N), posteriors float64
        (T, bvars, q), posteriors_sum float64 (T, R, qs) or None [+ status].  A checker for the f32 path, not a fast path."""
        x = np.ascontiguousarray(priors, dtype=np.float32)
        T = x.shape[0]
        if x.shape != (T, self.bvars, self.Q):
            raise ValueError(f"priors must be (T, {self.bvars}, {self.Q}), got {x.shape}")
        xs = None
        if self.special:
            if priors_sum is None:
                raise TypeError("priors_sum is required for the special decoder")
            xs = np.ascontiguousarray(priors_sum, dtype=np.float32)
            if xs.shape != (T, self.R, self.QS):
                raise ValueError(f"priors_sum must be (T, {self.R}, {self.QS}), got {xs.shape}")
        hard = np.empty((T, self.N), np.int8)
        post = np.empty((T, self.bvars, self.Q), np.float64)
        posts = np.empty((T, self.R, self.QS), np.float64) if self.special else None
        status = np.zeros(T, np.int32)
        rc = _lib.lib().scaldpc_qary_decode_f64(self._h, T, x.ctypes.data, xs.ctypes.data if xs is not None else None,
                                                hard.ctypes.data, post.ctypes.data,
                                                posts.ctypes.data if posts is not None else None, status.ctypes.data,
                                                _lib.IN_LLR if llr else 0)
        return self._finish(rc, hard, post, posts, status, want_status)

    def combining_stats(self):
        """(single decodes served, batches they were decoded in) by min_sum() on this decoder."""
        a, b = C.c_longlong(), C.c_longlong()
        _lib.check(_lib.lib().scaldpc_qary_one_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    @staticmethod
    def _one(x, rows, q, name):
        x = np.asarray(x)
        if x.dtype != np.float32:
            raise TypeError(f"{name} must be a float32 array (pydecoder.rs:53), got {x.dtype}")
        if x.ndim != 2 or x.shape[0] < rows or x.shape[1] < q:
            raise ValueError(f"{name} must be at least {rows}x{q}, got {x.shape}")
        return np.ascontiguousarray(x[:rows, :q])

    # ------------------------------------------------------------------------------------------------------
    # batched fast path: T independent decodes resident on the device
    # ------------------------------------------------------------------------------------------------------
    def min_sum_batch(self, priors, priors_sum=None, *, llr=False, want_posterior=True, exact=False,
                      generic=False, want_status=False, out=None, verify=True, early_stop=False):
        """Decode T trials.

        priors      (T, bvars, Q)  float32 probabilities (or LLRs with llr=True); numpy array or torch CUDA tensor
        priors_sum  (T, R, QS)     float32, special decoder only
        Returns (hard int8 (T, N), posterior (T, bvars, Q) or None, posterior_sum (T, R, QS) or None)
        [+ status int32 (T,) with want_status=True instead of raising on per-trial input errors].
        numpy in -> numpy out (host buffers, copies inside the call); torch CUDA in -> torch CUDA out
        (asynchronous on the current stream).
        out: optional preallocated C-contiguous int8 (T, N) numpy array that receives the hard decisions (host
        path only) -- e.g. a view of pinned memory, so the device-to-host copy is a true asynchronous DMA.
        exact=True: every trial with the exact-order check node (the reference's rounding, bit for bit).
        verify (default True): the fast check node's results are verified -- trials whose decisions could depend on
        the rounding order are re-decoded with the exact-order check node (SCALDPC_NO_VERIFY in scaldpc.h), so hard
        decisions equal the reference's per trial; `last_run` tells how many were re-decoded.
        early_stop (default False; an extension, the reference always runs every iteration): freeze every trial in
        the iteration its decisions first satisfy all parity checks (SCALDPC_QARY_EARLY_STOP); `last_iterations(T)`
        returns the per-trial iteration counts.
        """
        flags = self._flags(llr, exact, generic, verify, early_stop)
        on_dev = _is_torch(priors) and priors.is_cuda
        if self.special and priors_sum is None:
            raise TypeError("the special decoder needs priors_sum")
        if on_dev:
            return self._batch_torch(priors, priors_sum, flags, want_posterior, want_status)
        x = np.ascontiguousarray(priors, dtype=np.float32)
        T = x.shape[0]
        if x.shape != (T, self.bvars, self.Q):
            raise ValueError(f"priors must be (T, {self.bvars}, {self.Q}), got {x.shape}")
        xs = None
        if self.special:
            xs = np.ascontiguousarray(priors_sum, dtype=np.float32)
            if xs.shape != (T, self.R, self.QS):
                raise ValueError(f"priors_sum must be (T, {self.R}, {self.QS}), got {xs.shape}")
        if out is None:
            hard = np.empty((T, self.N), np.int8)
        else:
            hard = out
            if not (isinstance(hard, np.ndarray) and hard.dtype == np.int8 and hard.shape == (T, self.N)
                    and hard.flags.c_contiguous and hard.flags.writeable):
                raise ValueError(f"out must be a writable C-contiguous int8 array of shape ({T}, {self.N})")
        post = np.empty((T, self.bvars, self.Q), np.float32) if want_posterior else None
        posts = np.empty((T, self.R, self.QS), np.float32) if (want_posterior and self.special) else None
        status = np.zeros(T, np.int32)
        p = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        with self._lock:
            rc = _lib.lib().scaldpc_qary_decode(self._h, T, p(x), p(xs), p(hard), p(post), p(posts), p(status), flags,
                                                None)
        return self._finish(rc, hard, post, posts, status, want_status)

    @staticmethod
    def _flags(llr=False, exact=False, generic=False, verify=True, early_stop=False):
        return ((_lib.IN_LLR if llr else 0) | (_lib.CN_EXACT if exact else 0) | (_lib.CN_GENERIC if generic else 0)
                | (0 if verify else _lib.NO_VERIFY) | (_lib.QARY_EARLY_STOP if early_stop else 0))

    def min_sum_observed(self, table, obs, table_sum=None, obs_sum=None, *, llr=False, want_posterior=False,
                         exact=False, generic=False, want_status=False, out=None, verify=True, early_stop=False):
        """Decode T trials whose priors are rows of small tables (scaldpc_qary_decode_observed): the posterior of a
        variable given the oracle's bit pattern y is `table[y]` (max_likelihood.py:87-105, kyber.py:362-376), so only
        the row indices travel.  Bit-identical to min_sum_batch(table[obs], table_sum[obs_sum]).

        table      (n_rows, Q) float32      obs      (T, bvars) uint16
        table_sum  (n_rows_sum, QS) float32 obs_sum  (T, R) uint16            (special decoder)
        numpy in -> numpy out; torch CUDA tensors in (obs int16/uint16 storage) -> torch CUDA out.
        """
        flags = self._flags(llr, exact, generic, verify, early_stop)
        if self.special and (table_sum is None or obs_sum is None):
            raise TypeError("the special decoder needs table_sum and obs_sum")
        on_dev = _is_torch(obs) and obs.is_cuda
        if on_dev:
            dev = torch.device("cuda", self.device)
            i16 = (torch.int16, getattr(torch, "uint16", torch.int16))

            def chk(t, shape, dts, name):
                if t.device != dev or t.dtype not in dts or (shape is not None and tuple(t.shape) != shape):
                    raise ValueError(f"{name} must be a {dts[0]} tensor of shape {shape} on {dev}")
                return t.contiguous()

            T = obs.shape[0]
            obs = chk(obs, (T, self.bvars), i16, "obs")
            table = chk(table, None, (torch.float32,), "table")
            n_rows = table.shape[0]
            n_rows_sum = 0
            if self.special:
                obs_sum = chk(obs_sum, (T, self.R), i16, "obs_sum")
                table_sum = chk(table_sum, None, (torch.float32,), "table_sum")
                n_rows_sum = table_sum.shape[0]
            hard = torch.empty((T, self.N), dtype=torch.int8, device=dev)
            post = torch.empty((T, self.bvars, self.Q), dtype=torch.float32, device=dev) if want_posterior else None
            posts = (torch.empty((T, self.R, self.QS), dtype=torch.float32, device=dev)
                     if (want_posterior and self.special) else None)
            status = torch.zeros(T, dtype=torch.int32, device=dev) if want_status else None
            p = lambda t: None if t is None else t.data_ptr()  # noqa: E731
            stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            flags |= _lib.DEVICE_PTRS
        else:
            table = np.ascontiguousarray(table, dtype=np.float32)
            obs = np.ascontiguousarray(obs, dtype=np.uint16)
            T = obs.shape[0]
            if table.ndim != 2 or table.shape[1] != self.Q or obs.shape != (T, self.bvars):
                raise ValueError(f"table must be (n_rows, {self.Q}) and obs (T, {self.bvars})")
            n_rows, n_rows_sum = table.shape[0], 0
            if self.special:
                table_sum = np.ascontiguousarray(table_sum, dtype=np.float32)
                obs_sum = np.ascontiguousarray(obs_sum, dtype=np.uint16)
                if table_sum.ndim != 2 or table_sum.shape[1] != self.QS or obs_sum.shape != (T, self.R):
                    raise ValueError(f"table_sum must be (n_rows_sum, {self.QS}) and obs_sum (T, {self.R})")
                n_rows_sum = table_sum.shape[0]
            if out is None:
                hard = np.empty((T, self.N), np.int8)
            else:
                hard = out
                if not (isinstance(hard, np.ndarray) and hard.dtype == np.int8 and hard.shape == (T, self.N)
                        and hard.flags.c_contiguous and hard.flags.writeable):
                    raise ValueError(f"out must be a writable C-contiguous int8 array of shape ({T}, {self.N})")
            post = np.empty((T, self.bvars, self.Q), np.float32) if want_posterior else None
            posts = np.empty((T, self.R, self.QS), np.float32) if (want_posterior and self.special) else None
            status = np.zeros(T, np.int32)
            p = lambda a: None if a is None else a.ctypes.data  # noqa: E731
            stream = None
        with self._lock:
            rc = _lib.lib().scaldpc_qary_decode_observed(
                self._h, T, p(table), n_rows, p(table_sum if self.special else None), n_rows_sum, p(obs),
                p(obs_sum if self.special else None), p(hard), p(post), p(posts), p(status), flags, stream)
        if rc == _lib.ERR_ARG and T > 0:  # shapes were validated above: a row index outside its table (per trial)
            if want_status:
                return hard, post, posts, status
            raise ValueError(_lib.lib().scaldpc_last_error().decode())
        return self._finish(rc, hard, post, posts, status, want_status)

    def simulate(self, T, seed, table, table_sum, code, code_sum, nbits, nbits_sum, accuracy, eta, *, first_trial=0,
                 want_trials=False, exact=False, generic=False, verify=True, early_stop=False):
        """T Kyber attack-simulation trials synthesised, decoded and scored on the device (scaldpc_qary_simulate,
        SURVEY.md 8f.1): secret ~ centred binomial(eta) (kyber.py:35-47), ssum = H' s (kyber.py:85-92), every bit of
        the code word of a value observed correctly with probability `accuracy` (max_likelihood.py:9-18), priors =
        table[observed pattern], min_sum, success = decoded secret == s and decoded sums == -ssum.

        code / code_sum: one code word per value (index value + B / value + BSUM), bit k = oracle bit k.
        Returns a dict of counters; with want_trials also the per-trial arrays secret, ssum, obs, obs_sum, hard.
        Trial g = first_trial + t depends only on (seed, g)."""
        table = np.ascontiguousarray(table, dtype=np.float32)
        table_sum = np.ascontiguousarray(table_sum, dtype=np.float32)
        code = np.ascontiguousarray(code, dtype=np.uint16)
        code_sum = np.ascontiguousarray(code_sum, dtype=np.uint16)
        if table.ndim != 2 or table.shape[1] != self.Q or table_sum.ndim != 2 or table_sum.shape[1] != self.QS:
            raise ValueError(f"table must be (n_rows, {self.Q}) and table_sum (n_rows_sum, {self.QS})")
        if code.shape != (self.Q,) or code_sum.shape != (self.QS,):
            raise ValueError(f"code must have {self.Q} entries and code_sum {self.QS}")
        T = int(T)
        flags = self._flags(False, exact, generic, verify, early_stop)
        counters = np.zeros(5, np.int64)
        tr = {}
        if want_trials:
            tr = dict(secret=np.empty((T, self.bvars), np.int8), ssum=np.empty((T, self.R), np.int16),
                      obs=np.empty((T, self.bvars), np.uint16), obs_sum=np.empty((T, self.R), np.uint16),
                      hard=np.empty((T, self.N), np.int8))
        p = lambda k: tr[k].ctypes.data if k in tr else None  # noqa: E731
        with self._lock:
            rc = _lib.lib().scaldpc_qary_simulate(
                self._h, T, int(seed), int(first_trial), table.ctypes.data, table.shape[0], table_sum.ctypes.data,
                table_sum.shape[0], code.ctypes.data, int(nbits), code_sum.ctypes.data, int(nbits_sum), float(accuracy),
                int(eta), counters.ctypes.data, p("secret"), p("ssum"), p("obs"), p("obs_sum"), p("hard"), flags, None)
        if rc:
            msg = _lib.lib().scaldpc_last_error().decode()
            if rc == _lib.ERR_ARG:
                raise ValueError(msg)
            raise _lib.ScaldpcError(rc, msg)
        res = dict(zip(("trials", "successes", "secret_successes", "symbol_errors", "sum_symbol_errors"),
                       (int(c) for c in counters)))
        res.update(tr)
        return res

    def _batch_torch(self, x, xs, flags, want_posterior, want_status):
        dev = torch.device("cuda", self.device)
        if x.device != dev or x.dtype != torch.float32:
            raise ValueError(f"priors must be a float32 tensor on {dev}")
        x = x.contiguous()
        T = x.shape[0]
        if tuple(x.shape) != (T, self.bvars, self.Q):
            raise ValueError(f"priors must be (T, {self.bvars}, {self.Q}), got {tuple(x.shape)}")
        if self.special:
            if xs.device != dev or xs.dtype != torch.float32 or tuple(xs.shape) != (T, self.R, self.QS):
                raise ValueError(f"priors_sum must be a float32 (T, {self.R}, {self.QS}) tensor on {dev}")
            xs = xs.contiguous()
        hard = torch.empty((T, self.N), dtype=torch.int8, device=dev)
        post = torch.empty((T, self.bvars, self.Q), dtype=torch.float32, device=dev) if want_posterior else None
        posts = (torch.empty((T, self.R, self.QS), dtype=torch.float32, device=dev)
                 if (want_posterior and self.special) else None)
        status = torch.zeros(T, dtype=torch.int32, device=dev) if want_status else None
        p = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        stream = torch.cuda.current_stream(dev).cuda_stream
        with self._lock:
            rc = _lib.lib().scaldpc_qary_decode(self._h, T, p(x), p(xs if self.special else None), p(hard), p(post),
                                                p(posts), p(status), flags | _lib.DEVICE_PTRS, C.c_void_p(stream))
        return self._finish(rc, hard, post, posts, status, want_status)

    @staticmethod
    def _finish(rc, hard, post, posts, status, want_status):
        if want_status and rc in (_lib.ERR_PROB_SUM, _lib.ERR_NO_MAX, _lib.ERR_NO_CONFIG):
            return hard, post, posts, status
        if rc in (_lib.ERR_PROB_SUM, _lib.ERR_NO_MAX, _lib.ERR_NO_CONFIG):
            # the reference panics here (decoder.rs:618, 680, 683-684) -> PanicException; we raise ValueError
            raise ValueError(_lib.lib().scaldpc_last_error().decode())
        _lib.check(rc)
        return (hard, post, posts, status) if want_status else (hard, post, posts)

    # ------------------------------------------------------------------------------------------------------
    def set_iterations(self, iterations: int):
        _lib.check(_lib.lib().scaldpc_qary_set_max_iter(self._h, int(iterations)))
        self.iterations = int(iterations)

    @property
    def launch_count(self) -> int:
        return int(_lib.lib().scaldpc_qary_launch_count(self._h))

    @property
    def last_run(self) -> dict:
        """Trials of the last batch call and how many of them the verification re-decoded (scaldpc_qary_last_run)."""
        t, r, w = C.c_longlong(), C.c_longlong(), C.c_longlong()
        _lib.check(_lib.lib().scaldpc_qary_last_run(self._h, C.byref(t), C.byref(r), C.byref(w)))
        return {"trials": t.value, "redecoded": r.value, "trial_iterations": w.value}

    def last_iterations(self, T: int) -> np.ndarray:
        """Per-trial iteration counts of the last batch call, which must have used early_stop=True."""
        out = np.empty(int(T), np.int32)
        _lib.check(_lib.lib().scaldpc_qary_last_iterations(self._h, out.ctypes.data, int(T), 0, None))
        return out

    def set_risk_margin(self, eps: float):
        _lib.check(_lib.lib().scaldpc_qary_set_risk_margin(self._h, float(eps)))

    def time_kernel(self, which: int, T: int, reps: int = 20, exact=False, generic=False, stream=None) -> float:
        """Average duration (ms) of one check-node (which=0) or variable-node (which=1) launch over T trials,
        CUDA events on `stream`, on the message state left by the previous decode (bench.py roofline leg)."""
        ms = C.c_float()
        flags = (_lib.CN_EXACT if exact else 0) | (_lib.CN_GENERIC if generic else 0)
        _lib.check(_lib.lib().scaldpc_qary_time_kernel(self._h, which, T, flags, reps,
                                                       C.c_void_p(stream) if stream else None, C.byref(ms)))
        return ms.value

    # algorithmic HBM bytes of SURVEY.md 8(d): per trial-iteration 16*sum_e Q_e + 4*sum_v Q_v
    @property
    def msg_slots(self) -> int:
        return (self.E - (self.R if self.special else 0)) * self.Q + (self.R * self.QS if self.special else 0)

    @property
    def prior_slots(self) -> int:
        return self.bvars * self.Q + (self.R * self.QS if self.special else 0)

    def algorithmic_bytes_per_trial_iteration(self) -> int:
        return 16 * self.msg_slots + 4 * self.prior_slots
