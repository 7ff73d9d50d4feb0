"""Drop-in for `ldpc.bp_decoder` (third-party PyPI ldpc==0.1.3, simulate-with-python/requirements.txt:8) as the
reference uses it:

    simulate/hqc.py:694-708     bpd = bp_decoder(H, max_iter=100, bp_method="product_sum", channel_probs=...)
                                decoded = bpd.decode(msg)            # len(msg) == n: received vector
    simulate/decode.py:155-171  bpd = bp_decoder(H, error_rate=..., max_iter=n, bp_method="product_sum",
                                                 channel_probs=[None]*n or [...])
                                decoding = bpd.decode(syndrome)      # len == m: syndrome

plus the batched entry point `decode_batch` (T syndromes resident on one B200, per-trial early stop).
The message passing runs on the GPU through scaldpc_bp_decode; nothing here computes messages on the host.
"""
from __future__ import annotations

import ctypes as C
import types

import numpy as np

from . import _lib

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


def _csr_from_any(H):
    """dense ndarray / scipy.sparse matrix -> (m, n, row_ptr int32, col int32) with entries taken mod 2."""
    if hasattr(H, "tocsr"):
        S = H.tocsr().copy()
        S.data = S.data % 2
        S.eliminate_zeros()
        S.sort_indices()
        return S.shape[0], S.shape[1], S.indptr.astype(np.int32), S.indices.astype(np.int32)
    H = np.asarray(H)
    if H.ndim != 2:
        raise ValueError("parity check matrix must be 2-D")
    m, n = H.shape
    rows, cols = np.nonzero(H % 2)
    row_ptr = np.zeros(m + 1, dtype=np.int64)
    np.add.at(row_ptr, rows + 1, 1)
    return m, n, np.cumsum(row_ptr).astype(np.int32), cols.astype(np.int32)


class bp_decoder:
    """bp_decoder(H, error_rate=None, max_iter=0, bp_method="product_sum", channel_probs=[None])

    H may be a dense 0/1 array, a scipy.sparse matrix, or `csr=(row_ptr, col), shape=(m, n)`.
    Attributes after decode(): log_prob_ratios, iter, converge (as upstream).
    """

    def __init__(self, H=None, error_rate=None, max_iter=0, bp_method="product_sum", channel_probs=None, *,
                 csr=None, shape=None, device=0, **unused):
        if bp_method not in ("product_sum", "ps", 0):
            raise ValueError("only bp_method='product_sum' is implemented (the method the reference uses)")
        if csr is not None:
            self.m, self.n = (int(x) for x in shape)
            row_ptr, col = (np.ascontiguousarray(a, dtype=np.int32) for a in csr)
        else:
            self.m, self.n, row_ptr, col = _csr_from_any(H)
        self.row_ptr, self.col = np.ascontiguousarray(row_ptr), np.ascontiguousarray(col)
        cp = None
        if channel_probs is not None and len(channel_probs) and channel_probs[0] is not None:
            cp = np.asarray(channel_probs, dtype=np.float64)
            if cp.shape != (self.n,):
                raise ValueError(f"channel_probs must have length n = {self.n}")
        elif error_rate is not None:
            cp = np.full(self.n, float(error_rate), dtype=np.float64)
        else:
            raise ValueError("either error_rate or channel_probs is required")
        self.channel_probs = np.ascontiguousarray(cp)
        self.device = int(device)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().scaldpc_bp_create(self.m, self.n, self.row_ptr.ctypes.data, self.col.ctypes.data,
                                                int(max_iter), self.device, C.byref(self._h)))
        vals = [C.c_int() for _ in range(4)]
        _lib.check(_lib.lib().scaldpc_bp_info(self._h, *[C.byref(v) for v in vals]))
        self.E, self.max_iter = vals[2].value, vals[3].value
        self.log_prob_ratios = np.zeros(self.n)
        self.iter = 0
        self.converge = 0

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib._lib is not None:
            _lib._lib.scaldpc_bp_destroy(h)

    # ---- upstream-shaped single decode ---------------------------------------------------------------------
    def syndrome_of(self, v):
        v = np.asarray(v).astype(np.uint8) & 1
        s = np.zeros(self.m, dtype=np.uint8)
        rows = np.repeat(np.arange(self.m), np.diff(self.row_ptr))
        np.bitwise_xor.at(s, rows, v[self.col])
        return s

    def decode(self, vec):
        vec = np.asarray(vec).astype(np.uint8) & 1
        # upstream ldpc 0.1.x tests the SYNDROME length first, so for a square H (m == n) the vector is a syndrome
        if vec.shape[0] == self.m:       # syndrome: return the error estimate
            synd, received = vec, None
        elif vec.shape[0] == self.n:     # received vector: decode its syndrome, return the corrected vector
            synd, received = self.syndrome_of(vec), vec
        else:
            raise ValueError("input vector has the wrong length")
        out = self.decode_batch(synd[None], want_llr=True)
        self.log_prob_ratios = out["log_prob_ratios"][0].astype(np.float64)
        self.iter = int(out["iters"][0])
        self.converge = int(out["converged"][0])
        dec = out["decoding"][0]
        if received is not None:
            dec = dec ^ received
        return dec.astype(int)

    # ---- batched fast path -------------------------------------------------------------------------------
    def decode_batch(self, syndromes, channel_probs=None, *, want_llr=False, early_stop=True, compact=True, out=None,
                     n_checks=None):
        """syndromes (T, m) uint8; channel_probs None (the constructor's), (n,) or (T, n) float64.
        numpy in -> numpy out; torch CUDA uint8 tensor in -> torch CUDA tensors out.
        Returns dict(decoding (T,n) uint8, log_prob_ratios (T,n) float32 or None, iters (T,), converged (T,)).
        out: optional preallocated C-contiguous uint8 (T, n) numpy array for `decoding` (host path only), e.g. a view
        of pinned memory so the device-to-host copy of the T*n decision bytes is a true asynchronous DMA.
        n_checks: optional (T,) int32 (host path): nested graphs (scaldpc_bp_decode_nested, hqc.py:953-984): trial t
        decodes on the first n_checks[t] rows of H = [Hin | I_m] as if the later rows did not exist."""
        flags = 0 if early_stop else _lib.BP_NO_EARLY_STOP
        if not compact:
            flags |= _lib.BP_NO_COMPACT
        if torch is not None and isinstance(syndromes, torch.Tensor) and syndromes.is_cuda:
            if n_checks is not None:
                raise NotImplementedError("n_checks is supported for numpy (host) inputs")
            return self._batch_torch(syndromes, channel_probs, flags, want_llr)
        synd = np.ascontiguousarray(syndromes, dtype=np.uint8)
        T = synd.shape[0]
        if synd.shape != (T, self.m):
            raise ValueError(f"syndromes must be (T, {self.m}), got {synd.shape}")
        cp = self.channel_probs if channel_probs is None else np.ascontiguousarray(channel_probs, dtype=np.float64)
        if cp.ndim == 2:
            if cp.shape != (T, self.n):
                raise ValueError(f"channel_probs must be ({T}, {self.n})")
            flags |= _lib.BP_PROBS_PER_TRIAL
        elif cp.shape != (self.n,):
            raise ValueError(f"channel_probs must have length {self.n}")
        if out is None:
            dec = np.empty((T, self.n), np.uint8)
        else:
            dec = out
            if not (isinstance(dec, np.ndarray) and dec.dtype == np.uint8 and dec.shape == (T, self.n)
                    and dec.flags.c_contiguous and dec.flags.writeable):
                raise ValueError(f"out must be a writable C-contiguous uint8 array of shape ({T}, {self.n})")
        llr = np.empty((T, self.n), np.float32) if want_llr else None
        iters = np.empty(T, np.int32)
        conv = np.empty(T, np.int32)
        p = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        nc = self._n_checks(n_checks, T)
        if nc is not None:
            _lib.check(_lib.lib().scaldpc_bp_decode_nested(self._h, T, p(nc), p(synd), p(cp), p(dec), p(llr), p(iters),
                                                           p(conv), flags, None))
        else:
            _lib.check(_lib.lib().scaldpc_bp_decode(self._h, T, p(synd), p(cp), p(dec), p(llr), p(iters), p(conv), flags,
                                                    None))
        return dict(decoding=dec, log_prob_ratios=llr, iters=iters, converged=conv)

    def _n_checks(self, n_checks, T):
        if n_checks is None:
            return None
        nc = np.ascontiguousarray(n_checks, dtype=np.int32)
        if nc.shape != (T,) or (nc < 0).any() or (nc > self.m).any():
            raise ValueError(f"n_checks must be ({T},) with values in 0 .. {self.m}")
        return nc

    STAT_NAMES = ("unsatisfied", "good_flips", "bad_flips", "found_bad_satisfied_checks",
                  "found_bad_unsatisfied_checks", "success")

    def decode_scored(self, syndromes, truth, channel_probs=None, *, want_decoding=False, early_stop=True, n_checks=None):
        """decode_batch followed, on the device, by the statistics hqc.decode computes from the decoded vector
        (hqc.py:709-758; scaldpc_bp_decode_scored): 24 bytes per trial come back instead of n.  H must be [Hin | I_m].
        syndromes (T, m) uint8, truth (T, weight) int32 = support of the true message of every trial (y_sparse).
        n_checks: optional (T,) int32, nested graphs as decode_batch(n_checks=...): trial t uses its first n_checks[t] rows.
        Returns dict(stats (T, 6) int32 in STAT_NAMES order, iters, converged, decoding or None).
        `decoding` (want_decoding=True) is the ERROR ESTIMATE decode_batch returns; hqc.decode itself works on the
        corrected word msg ^ error (bpd.decode of a length-n vector, hqc.py:703-708) and the statistics are scored on
        that word, as the reference does (found_bad_unsatisfied = check bit 1 AND a decoded measurement error)."""
        flags = 0 if early_stop else _lib.BP_NO_EARLY_STOP
        synd = np.ascontiguousarray(syndromes, dtype=np.uint8)
        T = synd.shape[0]
        if synd.shape != (T, self.m):
            raise ValueError(f"syndromes must be (T, {self.m}), got {synd.shape}")
        truth = np.ascontiguousarray(truth, dtype=np.int32)
        if truth.ndim != 2 or truth.shape[0] != T:
            raise ValueError(f"truth must be ({T}, weight)")
        cp = self.channel_probs if channel_probs is None else np.ascontiguousarray(channel_probs, dtype=np.float64)
        if cp.ndim == 2:
            if cp.shape != (T, self.n):
                raise ValueError(f"channel_probs must be ({T}, {self.n})")
            flags |= _lib.BP_PROBS_PER_TRIAL
        elif cp.shape != (self.n,):
            raise ValueError(f"channel_probs must have length {self.n}")
        nc = self._n_checks(n_checks, T)
        stats = np.empty((T, 6), np.int32)
        dec = np.empty((T, self.n), np.uint8) if want_decoding else None
        iters = np.empty(T, np.int32)
        conv = np.empty(T, np.int32)
        p = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        _lib.check(_lib.lib().scaldpc_bp_decode_scored(self._h, T, self.n - self.m, p(nc), p(synd), p(cp), p(truth), truth.shape[1],
                                                       p(stats), p(dec), p(iters), p(conv), flags, None))
        return dict(stats=stats, iters=iters, converged=conv, decoding=dec)

    def _batch_torch(self, synd, channel_probs, flags, want_llr):
        dev = torch.device("cuda", self.device)
        synd = synd.contiguous()
        T = synd.shape[0]
        if synd.dtype != torch.uint8 or tuple(synd.shape) != (T, self.m) or synd.device != dev:
            raise ValueError(f"syndromes must be a uint8 (T, {self.m}) tensor on {dev}")
        if channel_probs is None:
            cp = getattr(self, "_cp_dev", None)
            if cp is None:
                cp = self._cp_dev = torch.from_numpy(self.channel_probs).to(dev)
        else:
            cp = channel_probs.contiguous()
            if cp.dtype != torch.float64 or cp.device != dev:
                raise ValueError("channel_probs must be a float64 tensor on the decoder's device")
            if cp.dim() == 2:
                flags |= _lib.BP_PROBS_PER_TRIAL
        dec = torch.empty((T, self.n), dtype=torch.uint8, device=dev)
        llr = torch.empty((T, self.n), dtype=torch.float32, device=dev) if want_llr else None
        iters = torch.empty(T, dtype=torch.int32, device=dev)
        conv = torch.empty(T, dtype=torch.int32, device=dev)
        p = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib().scaldpc_bp_decode(self._h, T, p(synd), p(cp), p(dec), p(llr), p(iters), p(conv),
                                                flags | _lib.DEVICE_PTRS, C.c_void_p(stream)))
        return dict(decoding=dec, log_prob_ratios=llr, iters=iters, converged=conv)

    @property
    def launch_count(self) -> int:
        return int(_lib.lib().scaldpc_bp_launch_count(self._h))

    def last_run(self):
        """(iterations launched, active-trial compactions) of the last decode on this decoder."""
        a, b = C.c_int(), C.c_int()
        _lib.check(_lib.lib().scaldpc_bp_last_run(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def time_kernel(self, which: int, T: int, reps: int = 20, stream=None) -> float:
        ms = C.c_float()
        _lib.check(_lib.lib().scaldpc_bp_time_kernel(self._h, which, T, 0, reps,
                                                     C.c_void_p(stream) if stream else None, C.byref(ms)))
        return ms.value

    def algorithmic_bytes_per_trial_iteration(self) -> int:
        """SURVEY.md 8(d), binary case: 16*E + 4*n + m/8 bytes."""
        return 16 * self.E + 4 * self.n + (self.m + 7) // 8


def _rep_code(n):
    """ldpc.codes.rep_code(n): (n-1) x n parity-check matrix of the repetition code (main.py:41-42)."""
    H = np.zeros((n - 1, n), dtype=int)
    idx = np.arange(n - 1)
    H[idx, idx] = 1
    H[idx, idx + 1] = 1
    return H


codes = types.ModuleType("ldpc.codes")
codes.rep_code = _rep_code


class bp_decoder_multi:
    """One decoder over several UNRELATED parity-check graphs of the same shape (segmented CSR; scaldpc_bp_create_multi).

    The reference draws a new code for every simulated attack (hqc.py:1056) and calls `bp_decoder(H, ...)` once per
    decode (hqc.py:694-699); a campaign of attacks is a batch whose trials live on different graphs.  Here:

        dec = bp_decoder_multi([(row_ptr_0, col_0), (row_ptr_1, col_1), ...], shape=(m, n), max_iter=100,
                               channel_probs=[...])
        out = dec.decode_batch(graph_of, syndromes)       # graph_of[t] = which graph trial t is decoded on

    Every trial's result equals bp_decoder(H_g, ...).decode_batch on its own graph."""

    STAT_NAMES = bp_decoder.STAT_NAMES

    def __init__(self, graphs, shape, max_iter=0, channel_probs=None, error_rate=None, device=0, _circulant=None):
        self.m, self.n = (int(x) for x in shape)
        if _circulant is not None:  # from_circulants: the graphs are built on the device
            first_col, rows, n_msg = _circulant
            self.G = first_col.shape[0]
            self._set_probs(channel_probs, error_rate)
            self.device = int(device)
            self._h = C.c_void_p()
            _lib.check(_lib.lib().scaldpc_bp_create_circulant(self.G, int(n_msg), first_col.shape[1], self.m,
                                                              first_col.ctypes.data, rows.ctypes.data, int(max_iter),
                                                              self.device, C.byref(self._h)))
            vals = [C.c_int() for _ in range(4)]
            _lib.check(_lib.lib().scaldpc_bp_info(self._h, *[C.byref(v) for v in vals]))
            self.E, self.max_iter = vals[2].value, vals[3].value
            return
        self.G = len(graphs)
        if self.G < 1:
            raise ValueError("at least one graph is needed")
        rps, cols = [], []
        for g in graphs:
            if isinstance(g, tuple) and len(g) == 2 and np.ndim(g[0]) == 1:
                rp, col = (np.ascontiguousarray(a, dtype=np.int32) for a in g)
            else:
                m, n, rp, col = _csr_from_any(g)
                if (m, n) != (self.m, self.n):
                    raise ValueError(f"every graph must be {self.m} x {self.n}")
            if rp.shape != (self.m + 1,):
                raise ValueError(f"row_ptr of every graph must have m + 1 = {self.m + 1} entries")
            rps.append(rp)
            cols.append(col)
        row_ptr = np.ascontiguousarray(np.concatenate(rps), dtype=np.int32)
        col = np.ascontiguousarray(np.concatenate(cols), dtype=np.int32)
        self._set_probs(channel_probs, error_rate)
        self.device = int(device)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().scaldpc_bp_create_multi(self.G, self.m, self.n, row_ptr.ctypes.data, col.ctypes.data,
                                                      int(max_iter), self.device, C.byref(self._h)))
        vals = [C.c_int() for _ in range(4)]
        _lib.check(_lib.lib().scaldpc_bp_info(self._h, *[C.byref(v) for v in vals]))
        self.E, self.max_iter = vals[2].value, vals[3].value

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib._lib is not None:
            _lib._lib.scaldpc_bp_destroy(h)

    def _set_probs(self, channel_probs, error_rate):
        if channel_probs is not None:
            cp = np.asarray(channel_probs, dtype=np.float64)
        elif error_rate is not None:
            cp = np.full(self.n, float(error_rate), dtype=np.float64)
        else:
            raise ValueError("either error_rate or channel_probs is required")
        if cp.shape != (self.n,):
            raise ValueError(f"channel_probs must have length n = {self.n}")
        self.channel_probs = np.ascontiguousarray(cp)

    @classmethod
    def from_circulants(cls, first_cols, rows, n_msg, **kw):
        """G graphs H_g = [rows of circulant(first_cols[g]) | I] built ON THE DEVICE (scaldpc_bp_create_circulant): what
        make_random_ldpc_parity_check_matrix (make_code.py:220-244) + add_check (hqc.py:885-908) + hqc.py:680 build as dense
        matrices.  first_cols (G, W) int32 = support of every first column (r1_sparse, hqc.py:1059); rows (G, m) int32 = the
        measured positions in the order they were added."""
        first_cols = np.ascontiguousarray(first_cols, dtype=np.int32)
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        if first_cols.ndim != 2 or rows.ndim != 2 or rows.shape[0] != first_cols.shape[0]:
            raise ValueError("first_cols must be (G, W) and rows (G, m)")
        m = rows.shape[1]
        return cls(None, (m, int(n_msg) + m), _circulant=(first_cols, rows, int(n_msg)), **kw)

    @classmethod
    def from_prefixes(cls, row_ptr, col, shape, levels, **kw):
        """The decode-every ladder of ONE growing H (hqc.py:953-984) as a family of graphs: graph k = the first levels[k]
        rows of H = [Hin | I] (the later rows have no edges, the later identity columns are isolated).  A trial decoded on
        graph k then costs what a decode on H_k costs -- bp_decoder.decode_batch(n_checks=...) on the full graph gives
        the same results but makes every trial pay for all the rows.  Pass `graph_of` = level index and zero the
        syndrome bits of the rows a trial does not use (decode_batch(..., n_checks=levels[graph_of]) does that)."""
        m, n = (int(x) for x in shape)
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int32)
        col = np.ascontiguousarray(col, dtype=np.int32)
        graphs = []
        for R in levels:
            R = int(R)
            if not 0 <= R <= m:
                raise ValueError(f"level {R} is outside 0 .. m = {m}")
            rp = row_ptr.copy()
            rp[R:] = row_ptr[R]
            graphs.append((rp, col[: row_ptr[R]]))
        dec = cls(graphs, shape, **kw)
        dec.levels = np.asarray(levels, dtype=np.int32)
        return dec

    def decode_batch(self, graph_of, syndromes, channel_probs=None, *, truth=None, n_checks=None, want_llr=False,
                     want_decoding=True, early_stop=True):
        """graph_of (T,) int32; syndromes (T, m) uint8; channel_probs None / (n,) / (T, n) float64; n_checks optional (T,)
        int32 (nested prefixes, as bp_decoder.decode_batch); truth optional (T, weight) int32 -> `stats` (T, 6) as
        bp_decoder.decode_scored.  Returns dict(decoding, log_prob_ratios, iters, converged[, stats])."""
        flags = 0 if early_stop else _lib.BP_NO_EARLY_STOP
        synd = np.ascontiguousarray(syndromes, dtype=np.uint8)
        T = synd.shape[0]
        if synd.shape != (T, self.m):
            raise ValueError(f"syndromes must be (T, {self.m}), got {synd.shape}")
        if n_checks is not None and getattr(self, "levels", None) is not None:
            # prefix graphs: a row beyond the trial's level has no edges, its syndrome bit must not count as unsatisfied
            synd = np.where(np.arange(self.m)[None, :] < np.asarray(n_checks)[:, None], synd, 0).astype(np.uint8)
        gof = np.ascontiguousarray(graph_of, dtype=np.int32)
        if gof.shape != (T,) or (T and (gof.min() < 0 or gof.max() >= self.G)):
            raise ValueError(f"graph_of must be ({T},) with values in 0 .. {self.G - 1}")
        cp = self.channel_probs if channel_probs is None else np.ascontiguousarray(channel_probs, dtype=np.float64)
        if cp.ndim == 2:
            if cp.shape != (T, self.n):
                raise ValueError(f"channel_probs must be ({T}, {self.n})")
            flags |= _lib.BP_PROBS_PER_TRIAL
        elif cp.shape != (self.n,):
            raise ValueError(f"channel_probs must have length {self.n}")
        nc = None
        if n_checks is not None:
            nc = np.ascontiguousarray(n_checks, dtype=np.int32)
            if nc.shape != (T,) or (nc < 0).any() or (nc > self.m).any():
                raise ValueError(f"n_checks must be ({T},) with values in 0 .. {self.m}")
        stats = tr = None
        weight = 0
        if truth is not None:
            tr = np.ascontiguousarray(truth, dtype=np.int32)
            if tr.ndim != 2 or tr.shape[0] != T:
                raise ValueError(f"truth must be ({T}, weight)")
            weight = tr.shape[1]
            stats = np.empty((T, 6), np.int32)
        dec = np.empty((T, self.n), np.uint8) if (want_decoding or truth is None) else None
        llr = np.empty((T, self.n), np.float32) if want_llr else None
        iters = np.empty(T, np.int32)
        conv = np.empty(T, np.int32)
        p = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        _lib.check(_lib.lib().scaldpc_bp_decode_multi(self._h, T, p(gof), p(nc), p(synd), p(cp), p(tr), weight, p(stats),
                                                      p(dec), p(llr), p(iters), p(conv), flags, None))
        out = dict(decoding=dec, log_prob_ratios=llr, iters=iters, converged=conv)
        if stats is not None:
            out["stats"] = stats
        return out

    @property
    def launch_count(self) -> int:
        return int(_lib.lib().scaldpc_bp_launch_count(self._h))
