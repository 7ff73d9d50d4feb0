"""Kernel LOGIC against the oracle without a GPU: tests/hostmodel compiles sca_ldpc_b200/csrc/qary_kernels.cuh with
g++ and runs the threads one after another (test infrastructure only; the product has no CPU path).  The same
comparisons run on the real device in tests/test_gpu_qary.py."""
import json

import numpy as np
import pytest

import util
from oracle import oracle as O
from sca_ldpc_b200 import _lib

EXACT, GENERIC, NO_VERIFY = _lib.CN_EXACT, _lib.CN_GENERIC, _lib.NO_VERIFY


def _kyber(cb, idx):
    z = util.load(f"kyber768_cb{cb}.npz")
    H = util.dense(z)
    llr = np.stack([O.into_llr(z["channel_output"][i]) for i in idx])
    llrs = np.stack([O.into_llr(z["channel_output_sum"][i]) for i in idx])
    return z, H, llr, llrs


@pytest.mark.parametrize("cb", [1, 2])
def test_fast_kernels_match_fixture(cb):
    """k_cn_fast<6,2> + k_vn<5>: hard decisions identical, posteriors within 1e-4, same inf/NaN masks."""
    idx = [0, 3] if cb == 2 else [0, 1, 2, 3]
    z, H, llr, llrs = _kyber(cb, idx)
    for j, i in enumerate(idx):
        it = int(z["iterations"][i])
        rc, hard, post, posts, status, fast = util.hostmodel_qary(H, 2, 12, it, llr[j:j + 1], llrs[j:j + 1], dc=7)
        assert rc == 0 and fast
        assert np.array_equal(hard[0], z["hard"][i])
        assert util.same_nonfinite(post[0], z["post"][i]) and util.same_nonfinite(posts[0], z["post_sum"][i])
        assert util.max_abs_diff(post[0], z["post"][i]) < 1e-4
        assert util.max_abs_diff(posts[0], z["post_sum"][i]) < 1e-4


def test_fast_kernel_equals_fast_oracle_bitwise():
    """The fast kernel and the oracle's model of it (algo 3: forward/backward check node, sum-variable messages taken
    from the channel except for the last iteration) are the same arithmetic, bit for bit -- including the trial whose
    priors are deterministic (most entries +inf), where an excluded sum value changes the normalisation constant."""
    for cb, idx in ((1, [1, 3]), (2, [0, 3])):
        z, H, llr, llrs = _kyber(cb, idx)
        for j, i in enumerate(idx):
            it = int(z["iterations"][i])
            d = O.QaryOracle(H, it, B=2, BSUM=12, DC=7)
            h2, p2, ps2 = d.decode_llr(llr[j], llrs[j], algo=O.ALGO_FAST)
            rc, hard, post, posts, _, fast = util.hostmodel_qary(H, 2, 12, it, llr[j:j + 1], llrs[j:j + 1], dc=7,
                                                                 flags=NO_VERIFY)
            assert rc == 0 and fast and np.array_equal(hard[0], h2)
            assert np.array_equal(post[0], p2, equal_nan=True) and np.array_equal(posts[0], ps2, equal_nan=True)
    # one and two iterations: FIRST|LAST and FIRST|WSUM -> LAST|RSUM variants
    z, H, llr, llrs = _kyber(1, [0])
    for it in (1, 2, 3):
        d = O.QaryOracle(H, it, B=2, BSUM=12, DC=7)
        h2, p2, ps2 = d.decode_llr(llr[0], llrs[0], algo=O.ALGO_FAST)
        rc, hard, post, posts, _, fast = util.hostmodel_qary(H, 2, 12, it, llr, llrs, dc=7, flags=NO_VERIFY)
        assert rc == 0 and fast and np.array_equal(hard[0], h2)
        assert np.array_equal(post[0], p2, equal_nan=True) and np.array_equal(posts[0], ps2, equal_nan=True)
        if it == 1:  # a single iteration has no simplification at all: equals the plain split forward/backward model
            h3, p3, ps3 = d.decode_llr(llr[0], llrs[0], algo=O.ALGO_SPLIT)
            assert np.array_equal(p2, p3, equal_nan=True) and np.array_equal(ps2, ps3, equal_nan=True)


@pytest.mark.parametrize("flags", [EXACT, EXACT | GENERIC])
def test_exact_kernel_is_bit_identical_special(flags):
    """k_cn_exact<6,2> (registers) and k_cn_generic in exact-order mode (run-time shapes) both reproduce the reference's
    rounding bit for bit (decoder_special.rs:539-553)."""
    z, H, llr, llrs = _kyber(1, [0, 3])
    for j, i in enumerate([0, 3]):
        it = min(3, int(z["iterations"][i]))
        d = O.QaryOracle(H, it, B=2, BSUM=12, DC=7)
        h, p, ps = d.decode_llr(llr[j], llrs[j], algo=O.ALGO_EXACT_DP)
        rc, hard, post, posts, _, fast = util.hostmodel_qary(H, 2, 12, it, llr[j:j + 1], llrs[j:j + 1], flags=flags,
                                                             dc=7)
        assert rc == 0 and not fast
        assert np.array_equal(hard[0], h)
        assert np.array_equal(post[0], p, equal_nan=True) and np.array_equal(posts[0], ps, equal_nan=True)


@pytest.mark.parametrize("variant,sw,it", [("kyber512", 6, 1), ("kyber512", 6, 3), ("kyber1024", 8, 2), ("kyber1024", 4, 3)])
def test_exact_register_kernel_every_shape(variant, sw, it):
    """k_cn_exact<SW,B> for the other template shapes, with excluded values (+inf LLRs) on both kinds of variable and
    1 / 2 / 3 iterations (first = last, first + last, a middle one): bit-identical to the oracle's exact-order DP, which
    tests/golden/make_golden.py asserts bit-identical to the reference's enumeration."""
    from sca_ldpc_b200.workloads import KyberWorkload

    w = KyberWorkload(variant, sw, 1, accuracy=0.9, nbits=3, nbits_sum=4, seed=3)
    b = w.generate(3, seed=2)
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    llr[0, ::7, 0] = np.inf
    llrs[1, ::5, 3] = np.inf
    d = O.QaryOracle(w.H, it, B=w.B, BSUM=w.BSUM, DC=sw + 1)
    rh, rp, rps, st = d.decode_batch(llr, llrs, from_probs=False, algo=O.ALGO_EXACT_DP)
    rc, hard, post, posts, _, fast = util.hostmodel_qary(w.H, w.B, w.BSUM, it, llr, llrs, flags=EXACT, dc=sw + 1)
    assert rc == 0 and not fast and np.array_equal(hard, rh)
    assert np.array_equal(post, rp, equal_nan=True) and np.array_equal(posts, rps, equal_nan=True)


def test_verified_fast_path_equals_reference_order():
    """The DEFAULT path (fast check node + verification) against the reference's rounding order on a
    Kyber-512 batch with converging and non-converging trials: the last iteration flags the trials at risk, they are
    decoded again with the exact-order check node, and every trial's hard decisions equal the oracle's; posteriors within
    1e-4 (bit-identical for the re-decoded ones)."""
    from sca_ldpc_b200.workloads import KyberWorkload

    w = KyberWorkload("kyber512", 6, 1, accuracy=0.95, nbits=4, nbits_sum=5, seed=0)
    T, it = 12, 6
    b = w.generate(T, seed=1)
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    d = O.QaryOracle(w.H, it, B=w.B, BSUM=w.BSUM, DC=7)
    rh, rp, rps, st = d.decode_batch(llr, llrs, from_probs=False, algo=O.ALGO_EXACT_DP)
    rc, hard, post, posts, _, fast = util.hostmodel_qary(w.H, w.B, w.BSUM, it, llr, llrs, dc=7)
    n_re = util.hostmodel().hostmodel_last_redecoded()
    assert rc == 0 and fast and 0 < n_re < T
    assert np.array_equal(hard, rh)
    assert util.same_nonfinite(post, rp) and util.max_abs_diff(post, rp) < 1e-4
    assert util.same_nonfinite(posts, rps) and util.max_abs_diff(posts, rps) < 1e-4
    # the flagged trials are exactly those whose decisions violate a parity check (none has a near-tie here)
    synd = rh.astype(np.int32) @ w.H.T.astype(np.int32)
    assert n_re >= int((synd != 0).any(1).sum())


def test_generic_fwd_bwd_kernel_special():
    z, H, llr, llrs = _kyber(1, [2])
    it = 4
    d = O.QaryOracle(H, it, B=2, BSUM=12, DC=7)
    h, p, ps = d.decode_llr(llr[0], llrs[0], algo=O.ALGO_FWD_BWD)
    rc, hard, post, posts, _, fast = util.hostmodel_qary(H, 2, 12, it, llr, llrs, flags=GENERIC | NO_VERIFY, dc=7)
    assert rc == 0 and not fast and np.array_equal(hard[0], h)
    assert util.same_nonfinite(post[0], p) and util.max_abs_diff(post[0], p) < 2e-5
    assert util.same_nonfinite(posts[0], ps) and util.max_abs_diff(posts[0], ps) < 2e-5


def _corrupted(n):
    co = np.zeros((n, 15), np.float32)
    co[:, 7] = 1.0
    co[1, 7], co[1, 14] = 0.1, 0.9
    return co


@pytest.mark.parametrize("flags", [0, EXACT])
def test_generic_decoder_rust_unit_tests(flags):
    """decoder.rs:770-854 through the generic kernels."""
    z = util.load("rust_unit_tests.npz")
    cases = [(z["H6"], 6, "small")]
    if not flags & EXACT:  # the exact-order DP on the 450x150 instance takes ~90 s on the host model; the GPU tier runs it
        cases.append((util.dense(z, "Hm"), 450, "medium"))
    for H, n, key in cases:
        llr = O.into_llr(_corrupted(n))[None]
        rc, hard, post, _, _, fast = util.hostmodel_qary(H, 7, 0, 10, llr, flags=flags)
        assert rc == 0 and not fast and (hard == 0).all()
        if flags & EXACT:
            assert np.array_equal(post[0], z[f"{key}_post"], equal_nan=True)
        else:
            assert util.same_nonfinite(post[0], z[f"{key}_post"])
            assert util.max_abs_diff(post[0], z[f"{key}_post"]) < 1e-4


def test_generic_decoder_doctest_seed1():
    z = util.load("decode_doctest_seed1.npz")
    H = util.dense(z)
    llr = O.into_llr(z["channel_output"])[None]
    rc, hard, post, _, _, _ = util.hostmodel_qary(H, 1, 0, 5, llr, flags=EXACT, dv=3, dc=7)
    assert rc == 0 and np.array_equal(hard[0], z["hard"]) and np.array_equal(post[0], z["post"], equal_nan=True)


def test_ragged_batch_and_irregular_rows():
    """T not a multiple of 32 (ragged last tile) and a parity-check matrix with rows of different degree
    (decoder_special.rs:510-511 trailing Nones) -> runtime-shape kernel, compared with the oracle per trial."""
    rs = np.random.RandomState(5)
    R, bv, B, sw = 6, 9, 2, 3
    H = np.zeros((R, bv + R), np.int8)
    for r in range(R):
        deg = 2 + (r % 2)  # rows of degree 2 and 3
        cols = rs.choice(bv, deg, replace=False)
        H[r, cols] = rs.choice([-1, 1], deg)
        H[r, bv + r] = 1
    T = 37
    Q, QS = 2 * B + 1, 2 * B * sw + 1
    pri = rs.dirichlet(np.ones(Q), size=(T, bv)).astype(np.float32)
    pri[rs.rand(T, bv, Q) < 0.15] = 0.0  # zero-probability entries -> +inf LLRs
    pri[..., B] += 1e-3
    pri /= pri.sum(-1, keepdims=True)
    pris = rs.dirichlet(np.ones(QS), size=(T, R)).astype(np.float32)
    llr = np.stack([O.into_llr(x) for x in pri])
    llrs = np.stack([O.into_llr(x) for x in pris])
    d = O.QaryOracle(H, 4, B=B, BSUM=B * sw, DC=sw + 1)
    # exact-order mode: bit-identical; default (verified forward/backward): decisions identical to the ENUMERATION too;
    # unverified forward/backward: its own CPU model
    for flags, algo in ((EXACT, O.ALGO_ENUMERATE), (0, O.ALGO_ENUMERATE), (NO_VERIFY, O.ALGO_FWD_BWD)):
        rc, hard, post, posts, _, fast = util.hostmodel_qary(H, B, B * sw, 4, llr, llrs, flags=flags, dc=sw + 1)
        assert rc == 0 and not fast
        for t in range(T):
            h, p, ps = d.decode_llr(llr[t], llrs[t], algo=algo)
            assert np.array_equal(hard[t], h), (flags, t)
            if flags & EXACT:
                assert np.array_equal(post[t], p, equal_nan=True) and np.array_equal(posts[t], ps, equal_nan=True)
            else:
                assert util.same_nonfinite(post[t], p) and util.max_abs_diff(post[t], p) < 1e-4


def test_no_config_status():
    """decoder.rs:618: a check whose variables cannot sum to zero -> SCALDPC_ERR_NO_CONFIG for that trial only."""
    H = np.array([[1, 1, 0], [0, 1, 1]], np.int8)
    p_ok = np.array([[0.2, 0.6, 0.2]] * 3, np.float32)
    p_bad = np.array([[0, 0, 1], [0, 0, 1], [0.2, 0.6, 0.2]], np.float32)  # v0 = v1 = +1 cannot satisfy check 0
    llr = np.stack([O.into_llr(p_ok), O.into_llr(p_bad)])
    rc, hard, post, _, status, _ = util.hostmodel_qary(H, 1, 0, 2, llr)
    assert rc == _lib.ERR_NO_CONFIG and status.tolist() == [0, _lib.ERR_NO_CONFIG]
    with pytest.raises(O.OracleError) as e:
        O.QaryOracle(H, 2, B=1).decode_llr(llr[1])
    assert e.value.code == -4


@pytest.mark.parametrize("restart_min", ["1", "100000"])
def test_early_termination_rounds(monkeypatch, restart_min):
    """SCALDPC_QARY_EARLY_STOP through the modelled kernels (frozen lanes, skipped tiles, the compacted second round when
    at most 1/8 is still running, exact-order decode of what never converges) == the per-trial model above."""
    from sca_ldpc_b200.workloads import KyberWorkload

    monkeypatch.setenv("SCALDPC_ES_RESTART_MIN", restart_min)
    w = KyberWorkload("kyber768", 6, 1, accuracy=0.995, nbits=3, nbits_sum=4, seed=0)
    T, it = 70, 7
    b = w.generate(T, seed=1)
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    # two trials that cannot converge: the information of half of their variables is wiped out
    llr[5, ::2] = 0.0
    llr[40, 1::2] = 0.0
    for verify in (True, False):
        want_h, want_it, never = util.early_stop_model(w, llr, llrs, it, verify)
        assert never.sum() >= 2 and (want_it < it).sum() > T // 2
        flags = _lib.QARY_EARLY_STOP | (0 if verify else NO_VERIFY)
        rc, hard, post, posts, _, fast = util.hostmodel_qary(w.H, w.B, w.BSUM, it, llr, llrs, flags=flags, dc=7)
        got_it = util.hostmodel_last_iterations(T)
        assert rc == 0 and fast
        assert np.array_equal(got_it, want_it)
        assert np.array_equal(hard, want_h)
        assert util.hostmodel().hostmodel_last_redecoded() == (int(never.sum()) if verify else 0)
        # slot-iterations launched: the compacted second round makes it fewer than the fixed-iteration decode's T * it
        work = util.hostmodel_last_work()
        assert work < T * it if restart_min == "1" else work == T * it + (never.sum() * it if verify else 0)


@pytest.mark.parametrize("variant,sw,cb", [("kyber512", 6, 1), ("kyber768", 6, 1), ("kyber1024", 4, 1)])
def test_fast_kernel_with_excluded_values_equals_fast_oracle_bitwise(variant, sw, cb):
    """k_cn_fast does not apply "own entry non-finite => +inf out" to the B-variables' messages (qary_cn_fast.cuh, kCnMaskOwn):
    an excluded value stays excluded through the variable's total either way.  The oracle's algo 3 DOES apply it
    (decoder.rs:307-311); decisions and posteriors must agree bit for bit on priors in which a third of the entries of both
    kinds of variable are zero (LLR +inf), for every kernel-variant schedule (1, 2, 3 and 5 iterations)."""
    from sca_ldpc_b200 import workloads as W
    w = W.KyberWorkload(variant, sum_weight=sw, check_blocks=cb, accuracy=0.95, seed=11)
    rs = np.random.RandomState(5)
    T, Q, QS, bv, R = 3, 2 * w.B + 1, 2 * w.BSUM + 1, w.H.shape[1] - w.H.shape[0], w.H.shape[0]
    pri = rs.dirichlet(np.ones(Q), size=(T, bv)).astype(np.float32)
    pri[rs.rand(T, bv, Q) < 0.33] = 0.0
    pri[..., w.B] += 1e-3  # never a variable without any value
    pri /= pri.sum(-1, keepdims=True)
    pris = rs.dirichlet(np.ones(QS), size=(T, R)).astype(np.float32)
    pris[rs.rand(T, R, QS) < 0.33] = 0.0
    pris[..., w.BSUM] += 1e-3
    pris /= pris.sum(-1, keepdims=True)
    llr = np.stack([O.into_llr(x) for x in pri])
    llrs = np.stack([O.into_llr(x) for x in pris])
    assert np.isinf(llr).mean() > 0.2 and np.isinf(llrs).mean() > 0.2
    for it in (1, 2, 3, 5):
        d = O.QaryOracle(w.H, it, B=w.B, BSUM=w.BSUM, DC=sw + 1)
        rc, hard, post, posts, _, fast = util.hostmodel_qary(w.H, w.B, w.BSUM, it, llr, llrs, dc=sw + 1, flags=NO_VERIFY)
        assert rc == 0 and fast
        for t in range(T):
            h2, p2, ps2 = d.decode_llr(llr[t], llrs[t], algo=O.ALGO_FAST)
            assert np.array_equal(hard[t], h2), (it, t)
            assert np.array_equal(post[t], p2, equal_nan=True) and np.array_equal(posts[t], ps2, equal_nan=True), (it, t)
        assert np.isinf(post).any() and np.isfinite(post).any()


def test_fast_kernel_negative_sum_coefficient_equals_fast_oracle_bitwise():
    """H = [H' | -I]: the sum-variable enters its check with coefficient -1 (its vectors are mirrored in the check frame,
    decoder.rs:164-172).  Never the case for the reference's own matrices, but the fast kernel keeps the branch -- and the
    table it parks in the sum-variable's slots is stored in the CHECK's frame -- so it is compared with its CPU model bit for
    bit for every kernel-variant schedule, excluded values included."""
    from sca_ldpc_b200 import workloads as W
    w = W.KyberWorkload("kyber768", sum_weight=6, check_blocks=1, accuracy=0.95, seed=12)
    H = w.H.copy()
    R = H.shape[0]
    bv = H.shape[1] - R
    H[:, bv:] = -H[:, bv:]
    rs = np.random.RandomState(9)
    T, Q, QS = 2, 2 * w.B + 1, 2 * w.BSUM + 1
    pri = rs.dirichlet(np.ones(Q), size=(T, bv)).astype(np.float32)
    pris = rs.dirichlet(np.ones(QS), size=(T, R)).astype(np.float32)
    pris[rs.rand(T, R, QS) < 0.2] = 0.0
    pris[..., w.BSUM] += 1e-3
    pris /= pris.sum(-1, keepdims=True)
    llr = np.stack([O.into_llr(x) for x in pri])
    llrs = np.stack([O.into_llr(x) for x in pris])
    for it in (1, 2, 3, 5):
        d = O.QaryOracle(H, it, B=w.B, BSUM=w.BSUM, DC=7)
        rc, hard, post, posts, _, fast = util.hostmodel_qary(H, w.B, w.BSUM, it, llr, llrs, dc=7, flags=NO_VERIFY)
        assert rc == 0 and fast
        for t in range(T):
            h2, p2, ps2 = d.decode_llr(llr[t], llrs[t], algo=O.ALGO_FAST)
            assert np.array_equal(hard[t], h2), (it, t)
            assert np.array_equal(post[t], p2, equal_nan=True) and np.array_equal(posts[t], ps2, equal_nan=True), (it, t)
            h1, p1, ps1 = d.decode_llr(llr[t], llrs[t], algo=O.ALGO_EXACT_DP)
            assert util.same_nonfinite(posts[t], ps1) and util.max_abs_diff(posts[t], ps1) < 1e-4
