"""Shared helpers of the test suite (fixture loading, the CPU oracle, the host model of the kernels)."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name))


def dense(z, prefix="H", with_vals=True):
    R, N = (int(x) for x in z[f"{prefix}_shape"])
    H = np.zeros((R, N), np.int8)
    H[z[f"{prefix}_rows"], z[f"{prefix}_cols"]] = z[f"{prefix}_vals"] if with_vals else 1
    return H


def reference_doctests():
    with open(os.path.join(GOLDEN, "reference_doctests.json")) as f:
        return json.load(f)


def same_nonfinite(a, b):
    """NaN / inf masks agree (NaN and +inf both mean "value excluded", but the reference keeps them apart)."""
    return (np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isposinf(a), np.isposinf(b))
            and np.array_equal(np.isneginf(a), np.isneginf(b)))


def max_abs_diff(a, b):
    m = np.isfinite(a) & np.isfinite(b)
    return float(np.abs(a[m].astype(np.float64) - b[m].astype(np.float64)).max()) if m.any() else 0.0


# ---- host model of the CUDA kernels (tests/hostmodel/hostmodel.cpp), TEST ONLY ---------------------------------
_HM = None


def hostmodel():
    global _HM
    if _HM is None:
        src = os.path.join(ROOT, "tests", "hostmodel", "hostmodel.cpp")
        out = os.path.join(ROOT, "tests", "hostmodel", "libhostmodel.so")
        deps = [src] + [os.path.join(ROOT, "sca_ldpc_b200", "csrc", f)
                        for f in ("qary_kernels.cuh", "qary_cn_fast.cuh", "qary_cn_exact.cuh", "qary_graph.h", "bp_kernels.cuh", "bp_graph.h", "tma.cuh")]
        deps = [d for d in deps if os.path.exists(d)]
        if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
            subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
                            "-Wno-unused-function", "-o", out, src], check=True, capture_output=True)
        _HM = C.CDLL(out)
        _HM.hostmodel_qary_decode.restype = C.c_int
    return _HM


def hostmodel_qary(H, b, bsum, max_iter, llr, llr_sum=None, flags=0, dv=0, dc=0):
    """Run the q-ary kernels' logic on the host.  llr (T,bvars,Q), llr_sum (T,R,QS).  Returns
    (rc, hard, post, post_sum, status, used_fast)."""
    L = hostmodel()
    H = np.ascontiguousarray(H, np.int8)
    R, N = H.shape
    llr = np.ascontiguousarray(llr, np.float32)
    T = llr.shape[0]
    hard = np.zeros((T, N), np.int8)
    post = np.zeros_like(llr)
    posts = None
    if bsum:
        llr_sum = np.ascontiguousarray(llr_sum, np.float32)
        posts = np.zeros_like(llr_sum)
    status = np.zeros(T, np.int32)
    used = C.c_int(0)
    err = C.create_string_buffer(512)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = L.hostmodel_qary_decode(p(H), R, N, b, bsum, dv, dc, max_iter, T, p(llr), p(llr_sum), p(hard), p(post),
                                 p(posts), p(status), C.c_uint(flags), C.byref(used), err, 512)
    return rc, hard, post, posts, status, bool(used.value)


def hostmodel_last_iterations(T):
    out = np.zeros(T, np.int32)
    hostmodel().hostmodel_last_iterations(out.ctypes.data_as(C.c_void_p), int(T))
    return out


def hostmodel_last_work():
    L = hostmodel()
    L.hostmodel_last_work.restype = C.c_longlong
    return int(L.hostmodel_last_work())


def hostmodel_bp(row_ptr, col, m, n, max_iter, synd, probs, early_stop=True, want_llr=True, compact=True):
    """Run the binary-BP kernels' logic on the host.  Returns dict(decoding, llr, iters, converged)."""
    L = hostmodel()
    L.hostmodel_bp_decode.restype = C.c_int
    row_ptr = np.ascontiguousarray(row_ptr, np.int32)
    col = np.ascontiguousarray(col, np.int32)
    synd = np.ascontiguousarray(synd, np.uint8)
    T = synd.shape[0]
    probs = np.ascontiguousarray(probs, np.float64)
    per_trial = int(probs.ndim == 2)
    dec = np.zeros((T, n), np.uint8)
    llr = np.zeros((T, n), np.float32) if want_llr else None
    iters = np.zeros(T, np.int32)
    conv = np.zeros(T, np.int32)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    ncomp = C.c_int(0)
    rc = L.hostmodel_bp_decode(m, n, p(row_ptr), p(col), max_iter, T, p(synd), p(probs), per_trial, int(early_stop),
                               int(compact), p(dec), p(llr), p(iters), p(conv), C.byref(ncomp))
    assert rc == 0, rc
    return dict(decoding=dec, llr=llr, iters=iters, converged=conv, compactions=ncomp.value)


def hostmodel_bp_multi(graphs, m, n, max_iter, graph_of, synd, probs, early_stop=True, want_llr=True):
    """Segmented-CSR binary BP through the modelled kernels.  graphs: list of (row_ptr, col)."""
    L = hostmodel()
    L.hostmodel_bp_decode_multi.restype = C.c_int
    row_ptr = np.ascontiguousarray(np.concatenate([np.asarray(g[0], np.int32) for g in graphs]), np.int32)
    col = np.ascontiguousarray(np.concatenate([np.asarray(g[1], np.int32) for g in graphs]), np.int32)
    graph_of = np.ascontiguousarray(graph_of, np.int32)
    synd = np.ascontiguousarray(synd, np.uint8)
    T = synd.shape[0]
    probs = np.ascontiguousarray(probs, np.float64)
    dec = np.zeros((T, n), np.uint8)
    llr = np.zeros((T, n), np.float32) if want_llr else None
    iters = np.zeros(T, np.int32)
    conv = np.zeros(T, np.int32)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = L.hostmodel_bp_decode_multi(len(graphs), m, n, p(row_ptr), p(col), max_iter, T, p(graph_of), p(synd), p(probs),
                                     int(early_stop), p(dec), p(llr), p(iters), p(conv))
    assert rc == 0, rc
    return dict(decoding=dec, llr=llr, iters=iters, converged=conv)


def llr_close(a, b, atol=2e-3, rtol=2e-4):
    """fp32 LLR-domain posteriors against the fp64 ratio-domain oracle: same infinities, finite values close."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    fin = np.isfinite(a) & np.isfinite(b)
    if not np.array_equal(np.isposinf(a), np.isposinf(b)) or not np.array_equal(np.isneginf(a), np.isneginf(b)):
        return False
    return bool(np.all(np.abs(a[fin] - b[fin]) <= atol + rtol * np.abs(b[fin])))


def early_stop_model(w, llr, llrs, iters, verify, eps=1e-2):
    """Per-trial early termination restated with the oracle: plain forward/backward flooding in the fast kernel's order of
    additions (algo 4, the sum-variables' messages make the full round trip); the trial stops in
    the first iteration whose hard decisions satisfy every row and whose posteriors are clear of ties; a trial that never
    does gets the reference's fixed-iteration decode (exact-order) in verified mode, else its last decisions."""
    T = llr.shape[0]
    Hs = w.H.astype(np.int32)
    hard = np.zeros((T, w.N), np.int8)
    its = np.zeros(T, np.int32)
    open_ = np.ones(T, bool)

    def margin(p):
        q = np.sort(np.where(np.isnan(p), np.inf, p), axis=-1)
        with np.errstate(invalid="ignore"):
            m = q[..., 1] - q[..., 0]
        return np.where(np.isnan(m), -1.0, m).min(axis=1)

    from oracle import oracle as O

    for k in range(1, iters + 1):
        d = O.QaryOracle(w.H, k, B=w.B, BSUM=w.BSUM, DC=w.sum_weight + 1)
        h, p, ps, _ = d.decode_batch(llr, llrs, from_probs=False, algo=O.ALGO_SPLIT)
        ok = ((h.astype(np.int32) @ Hs.T) == 0).all(1) & (np.minimum(margin(p), margin(ps)) > eps)
        newly = open_ & ok
        hard[newly], its[newly] = h[newly], k
        open_ &= ~ok
        if k == iters:
            hard[open_], its[open_] = h[open_], k
    if verify and open_.any():
        d = O.QaryOracle(w.H, iters, B=w.B, BSUM=w.BSUM, DC=w.sum_weight + 1)
        h, _, _, _ = d.decode_batch(llr[open_], llrs[open_], from_probs=False, algo=O.ALGO_EXACT_DP)
        hard[open_] = h
    return hard, its, open_
