"""The CPU oracle against every golden vector the reference's own tests hold for the decoder path
 and against the committed fixtures (tests/golden/make_golden.py).  No GPU."""
import json

import numpy as np
import pytest

import util
from oracle import oracle as O


def test_into_llr_rust_unit_test():
    """decoder.rs:743-768: p = [0 x4, .14 x7, .02, 0 x3] -> [inf x4, 0 x7, 1.9459101, inf x3]."""
    z = util.load("rust_unit_tests.npz")
    llr = O.into_llr(z["llr_probs"])
    for row in llr:
        assert np.array_equal(row, z["llr_expected"])
    assert np.float32(1.9459101) == z["llr_expected"][11]


def test_into_llr_asserts():
    """decoder.rs:683-684 / :680: rows must sum to 1 +- 0.001 and contain a non-NaN entry."""
    with pytest.raises(O.OracleError) as e:
        O.into_llr(np.array([[0.5, 0.3, 0.1]], np.float32))
    assert e.value.code == -2
    with pytest.raises(O.OracleError) as e:
        O.into_llr(np.array([[np.nan, np.nan, np.nan]], np.float32))
    assert e.value.code == -3


def _corrupted(n):
    co = np.zeros((n, 15), np.float32)
    co[:, 7] = 1.0
    co[1, 7], co[1, 14] = 0.1, 0.9
    return co


@pytest.mark.parametrize("algo", [O.ALGO_ENUMERATE, O.ALGO_EXACT_DP, O.ALGO_FWD_BWD])
def test_small_decoder_instance(algo):
    """decoder.rs:770-799: Decoder<6,3,3,4,15,7>, 10 iterations, variable 1 = {0: .1, +7: .9} -> all zero."""
    z = util.load("rust_unit_tests.npz")
    d = O.QaryOracle(z["H6"], 10, B=7, DV=3, DC=4, algo=algo)
    assert d.min_sum(_corrupted(6)) == [0] * 6
    hard, post, _ = d.decode_probs(_corrupted(6))
    assert np.array_equal(hard, z["small_hard"])
    if algo != O.ALGO_FWD_BWD:
        assert np.array_equal(post, z["small_post"], equal_nan=True)


@pytest.mark.parametrize("algo", [O.ALGO_ENUMERATE, O.ALGO_EXACT_DP, O.ALGO_FWD_BWD])
def test_medium_decoder_instance(algo):
    """decoder.rs:818-854: Decoder<450,150,3,7,15,7>, H = benches/parity_check_150_450.txt -> all zero."""
    z = util.load("rust_unit_tests.npz")
    H = util.dense(z, "Hm")
    assert H.shape == (150, 450) and (H != 0).sum() == 1050
    d = O.QaryOracle(H, 10, B=7, DV=3, DC=7, algo=algo)
    hard, post, _ = d.decode_probs(_corrupted(450))
    assert (hard == 0).all() and np.array_equal(hard, z["medium_hard"])
    if algo != O.ALGO_FWD_BWD:
        assert np.array_equal(post, z["medium_post"], equal_nan=True)
    else:
        assert util.same_nonfinite(post, z["medium_post"]) and util.max_abs_diff(post, z["medium_post"]) < 1e-4


def test_decode_doctest_seed1():
    """simulate/decode.py:192-209: seed 1, (3,6)+I code 150x450, B=1, 5 iterations -> 1 success."""
    z = util.load("decode_doctest_seed1.npz")
    H = util.dense(z)
    d = O.QaryOracle(H, 5, B=1, DV=3, DC=7)
    out = d.min_sum(z["channel_output"])
    assert out == [0] * 450
    assert int(z["successes"]) == 1 and int(z["errs"]) >= 1
    assert util.reference_doctests()["decode.py:192-209 simulate_frame_error_rate_rust"] == 1


@pytest.mark.parametrize("cb", [1, 2])
def test_kyber768_fixture(cb):
    """DecoderSpecial (decoder_special.rs:471-617) on reference-produced Kyber-768 priors: the enumeration oracle
    reproduces the stored outputs; the exact-order DP is bit-identical; forward/backward -- plain (algo 2), split around
    the sum-variable (algo 4, the fast kernel's order of additions) and split with the sum-variables' messages staying home
    (algo 3, the fast kernel's schedule) -- gives identical hard decisions and posteriors within 1e-4 (the north-star tolerance)."""
    z = util.load(f"kyber768_cb{cb}.npz")
    H = util.dense(z)
    meta = json.loads(str(z["meta"]))
    for i, m in enumerate(meta):
        it = int(z["iterations"][i])
        d = O.QaryOracle(H, it, B=2, BSUM=12, DC=7)
        llr, llrs = O.into_llr(z["channel_output"][i]), O.into_llr(z["channel_output_sum"][i])
        algos = [O.ALGO_EXACT_DP, O.ALGO_FWD_BWD, O.ALGO_SPLIT, O.ALGO_FAST] + ([O.ALGO_ENUMERATE] if (cb == 1 and i == 0) else [])
        for algo in algos:
            hard, post, posts = d.decode_llr(llr, llrs, algo=algo)
            assert np.array_equal(hard, z["hard"][i]), (cb, i, algo)
            if algo in (O.ALGO_FWD_BWD, O.ALGO_SPLIT, O.ALGO_FAST):
                assert util.same_nonfinite(post, z["post"][i]) and util.same_nonfinite(posts, z["post_sum"][i])
                assert util.max_abs_diff(post, z["post"][i]) < 1e-4
                assert util.max_abs_diff(posts, z["post_sum"][i]) < 1e-4
            else:
                assert np.array_equal(post, z["post"][i], equal_nan=True)
                assert np.array_equal(posts, z["post_sum"][i], equal_nan=True)
        # end-to-end property of SURVEY.md 3.2
        ok = bool((z["hard"][i][: 768] == z["secret"][i]).all() and
                  (z["hard"][i][768:].astype(np.int16) == -z["ssum"][i]).all())
        assert ok == m["success"]


def test_fp64_verification_mode():
    z = util.load("kyber768_cb1.npz")
    H = util.dense(z)
    d = O.QaryOracle(H, 5, B=2, BSUM=12, DC=7)
    h64, p64, _ = d.decode_probs(z["channel_output"][0], z["channel_output_sum"][0], dtype=np.float64)
    assert np.array_equal(h64, z["hard"][0])
    assert util.max_abs_diff(p64.astype(np.float32), z["post"][0]) < 1e-4


def test_oracle_graph_errors():
    """decoder_special.rs:388-395, decoder.rs:465-473 construction panics -> error codes."""
    H = np.zeros((2, 5), np.int8)
    H[0, [0, 1, 3]] = 1
    H[1, [1, 2, 4]] = 1
    O.QaryOracle(H, 1, B=2, BSUM=4, DC=3)
    with pytest.raises(O.OracleError) as e:   # BSUM % B != 0
        O.QaryOracle(H, 1, B=2, BSUM=5, DC=3)
    assert e.value.code == -6
    with pytest.raises(O.OracleError) as e:   # degree > DC
        O.QaryOracle(H, 1, B=2, BSUM=4, DC=2)
    assert e.value.code == -5
    with pytest.raises(TypeError):
        O.QaryOracle(H.astype(np.int32), 1, B=2)


def test_hqc_toy_fixture():
    """hqc.py:1229-1274 toy: N=20, weight 3, y=[4,5,7,9], certainty 1.0 -> BP recovers y (pinned outcome)."""
    z = util.load("hqc_toy.npz")
    H = util.dense(z, with_vals=False)
    bpd = O.BinaryBpOracle(H, max_iter=100, bp_method="product_sum", channel_probs=z["channel_probs"])
    decoded = bpd.decode(z["msg"])
    assert sorted(np.nonzero(decoded[:20])[0].tolist()) == z["y_sparse"].tolist()
    assert np.array_equal(decoded, z["decoded"]) and bpd.iter == int(z["iters"])
    assert np.allclose(bpd.log_prob_ratios, z["log_prob_ratios"], rtol=0, atol=1e-12, equal_nan=True)
    assert util.reference_doctests()["hqc.py:1229-1236 test_hqc_decode_toy_example(0)"] is True
    assert util.reference_doctests()["hqc.py:1277-1284 test_hqc_decode_full_example(0)"] is True


def test_hqc_small_fixture():
    z = util.load("hqc_small.npz")
    H = util.dense(z, with_vals=False)
    bpd = O.BinaryBpOracle(H, max_iter=100, bp_method="product_sum", channel_probs=z["channel_probs"])
    decoded = bpd.decode(z["msg"])
    assert np.array_equal(decoded, z["decoded"]) and bpd.iter == int(z["iters"])
    assert bpd.converge == int(z["converged"])
    assert np.allclose(bpd.log_prob_ratios, z["log_prob_ratios"], rtol=1e-12, atol=1e-12, equal_nan=True)


def test_rep_code_doctest():
    """decode.py:139-149: product_sum on rep_code(13), p=0.05 -> every frame decodes (100/100 in the reference);
    replayed with a seeded stream of the same shape."""
    import refshim

    H = refshim.rep_code(13)
    rs = np.random.RandomState(0)
    bpd = O.BinaryBpOracle(H, error_rate=0.05, max_iter=13, bp_method="product_sum")
    ok = 0
    for _ in range(100):
        err = (rs.rand(13) < 0.05).astype(int)
        ok += int((bpd.decode(err) == 0).all())
    assert ok == 100
    assert util.reference_doctests()["decode.py:139-149 simulate_frame_error_rate(rep_code(13))"] == 100


@pytest.mark.parametrize("B,col_w", [(2, 2), (3, 3), (1, 4)])
def test_special_oracle_equals_pinned_generic_oracle_when_alphabets_coincide(B, col_w):
    """DecoderSpecial has no live test in the reference (decoder_special.rs:687-747 are commented
    out), the generic Decoder is pinned by the Rust goldens above.  With sum_weight 1 the sum-variables' alphabet
    [-BSUM, BSUM] = [-B, B] IS the B-variables' alphabet, so H = [H' | I] is an ordinary single-alphabet problem and both
    restatements must agree bit for bit on finite priors: same graph order (columns ascending inside a row: the
    sum-variable last, decoder_special.rs:507-522 / decoder.rs:507-539), same S = (0 + a_0) + a_sum, same total-minus-self
    variable node.  This ties the unpinned half of oracle/qary_impl.inc to the pinned half."""
    rs = np.random.RandomState(17 + B)
    bv, Q = 7, 2 * B + 1
    R = bv * col_w  # every check: ONE B-variable (+-1) and its own sum-variable; every B-variable sits in col_w checks
    H = np.zeros((R, bv + R), np.int8)
    for r in range(R):
        H[r, r % bv] = rs.choice([-1, 1])
        H[r, bv + r] = 1
    T, it = 9, 4
    pri = rs.dirichlet(np.ones(Q) * 0.7, size=(T, bv)).astype(np.float32) + np.float32(1e-4)
    pri /= pri.sum(-1, keepdims=True)
    pris = rs.dirichlet(np.ones(Q) * 0.7, size=(T, R)).astype(np.float32) + np.float32(1e-4)
    pris /= pris.sum(-1, keepdims=True)
    special = O.QaryOracle(H, it, B=B, BSUM=B, DC=2)
    generic = O.QaryOracle(H, it, B=B, DV=col_w, DC=2)
    assert special.bvars == bv and generic.bvars == bv + R
    for algo in (O.ALGO_ENUMERATE, O.ALGO_EXACT_DP):
        sh, sp, sps, st = special.decode_batch(pri, pris, from_probs=True, algo=algo)
        gh, gp, _, gst = generic.decode_batch(np.concatenate([pri, pris], axis=1), from_probs=True, algo=algo)
        assert (st == 0).all() and (gst == 0).all()
        assert np.array_equal(sh, gh)
        assert np.array_equal(sp, gp[:, :bv]) and np.array_equal(sps, gp[:, bv:])
    # and the two are not trivially equal to the priors' decisions: the checks did move something
    prior_dec = np.concatenate([pri, pris], axis=1).argmax(-1) - B
    assert (sh != prior_dec).any()


def test_split_forward_backward_equals_enumeration_in_exact_arithmetic():
    """cn_split (oracle algo 4: the order of additions of the GPU fast kernel since round 2) is the same minimum over the same
    configurations as the reference's enumeration (decoder_special.rs:531-554), associated differently.  On priors whose LLRs
    are small integers every f32 sum is exact, so the two must agree BIT FOR BIT -- hard decisions and posteriors of both kinds
    of variable -- for even and odd sum_weight, three alphabets, a negative coefficient, excluded values and 1..3 iterations."""
    rs = np.random.RandomState(21)
    for sw, B, R, bv in ((2, 1, 4, 6), (3, 2, 5, 9), (4, 1, 4, 8), (5, 1, 3, 9), (6, 1, 3, 9)):
        Q, BS = 2 * B + 1, sw * B
        QS = 2 * BS + 1
        Hp = np.zeros((R, bv), np.int8)
        for r in range(R):
            cols = rs.choice(bv, sw, replace=False)
            Hp[r, cols] = rs.choice([-1, 1], sw)
        H = np.concatenate([Hp, np.eye(R, dtype=np.int8)], axis=1)
        H[0, bv] = -1  # one sum-variable enters its check mirrored
        llr = rs.randint(0, 9, size=(bv, Q)).astype(np.float32)
        llr[rs.rand(bv, Q) < 0.15] = np.inf
        llr[:, B] = 0.0
        llrs = rs.randint(0, 9, size=(R, QS)).astype(np.float32)
        llrs[rs.rand(R, QS) < 0.15] = np.inf
        llrs[:, BS] = 0.0
        for it in (1, 2, 3):
            d = O.QaryOracle(H, it, B=B, BSUM=BS, DC=sw + 1)
            h0, p0, ps0 = d.decode_llr(llr, llrs, algo=O.ALGO_ENUMERATE)
            for algo in (O.ALGO_EXACT_DP, O.ALGO_FWD_BWD, O.ALGO_SPLIT):
                h, p, ps = d.decode_llr(llr, llrs, algo=algo)
                assert np.array_equal(h, h0), (sw, B, it, algo)
                assert np.array_equal(p, p0, equal_nan=True) and np.array_equal(ps, ps0, equal_nan=True), (sw, B, it, algo)
