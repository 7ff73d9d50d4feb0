#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/ (run in the build container only).

Inputs come from the REFERENCE's own Python helpers imported from /root/reference (simulate.make_code,
simulate.kyber, simulate.max_likelihood, simulate.decode -- via tests/refshim.py) and from the reference's
Rust test data; expected outputs come from the CPU oracle (algo 0 = the reference's enumeration), after the
oracle itself has been checked against the reference's pinned results (asserted below, and again by
tests/test_oracle_golden.py against the stored files).

    python tests/golden/make_golden.py

Outputs (all small, committed):
  rust_unit_tests.npz          decoder.rs:743-854 golden inputs/outputs (+ benches/parity_check_150_450.txt as COO)
  decode_doctest_seed1.npz     decode.py:192-209: H, the channel output drawn with RandomState(1), expected result
  kyber768_cb{1,2}.npz         H = kyber.gen_ldpc_matrix(6, 0, cb), seeded secrets, reference-computed priors for
                               several (accuracy, #patterns) settings, oracle hard decisions + posteriors
  hqc_toy.npz                  hqc.py:1229-1274 toy inputs + oracle outputs (outcome pinned by the reference: True)
  hqc_small.npz                HQC-shaped random instance (N=997) with noisy checks: oracle log_prob_ratios
  reference_doctests.json      results of the reference's own doctests run through the oracle
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import refshim  # noqa: E402
from oracle import oracle as O  # noqa: E402


class OracleAsSimulateRs:
    """simulate_rs stand-in backed by the oracle (class lookup by name, decode.py:227-229, lib.rs:32-75)."""

    def __getattr__(self, name):
        m = re.fullmatch(r"DecoderN(\d+)R(\d+)V(\d+)C(\d+)B(\d+)", name)
        if m:
            n, r, v, c, b = map(int, m.groups())
            return lambda H, it: O.QaryOracle(H, it, B=b, DV=v, DC=c)
        m = re.fullmatch(r"DecoderN(\d+)R(\d+)SW(\d+)", name)
        if m:
            n, r, sw = map(int, m.groups())
            return lambda H, it: O.QaryOracle(H, it, B=2, BSUM=2 * sw, DC=sw + 1)
        raise AttributeError(name)


def coo(H):
    r, c = np.nonzero(H)
    return dict(shape=np.array(H.shape, np.int32), rows=r.astype(np.int32), cols=c.astype(np.int32),
                vals=H[r, c].astype(np.int8))


def save(name, **arrays):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print(f"wrote {name}: {os.path.getsize(path)} bytes")


def main():
    sim = refshim.install(bp_decoder_cls=O.BinaryBpOracle, decoder_module=OracleAsSimulateRs())
    from simulate import decode, hqc, kyber, make_code, max_likelihood, utils

    results = {}

    # ---- 1. Rust unit tests (decoder.rs:743-854)
    probs = np.array([[0, 0, 0, 0, .14, .14, .14, .14, .14, .14, .14, .02, 0, 0, 0]] * 6, dtype=np.float32)
    llr = O.into_llr(probs)
    inf = np.inf
    expected_llr = np.array([inf] * 4 + [0.0] * 7 + [1.9459101] + [inf] * 3, dtype=np.float32)
    assert all((llr[i] == expected_llr).all() for i in range(6)), llr
    H6 = np.array([[1, 1, 1, 1, 0, 0], [0, 0, 1, 1, 0, 1], [1, 0, 0, 1, 1, 0]], dtype=np.int8)
    Hm = np.loadtxt(os.path.join(refshim.REF_PY, "simulate_rs/benches/parity_check_150_450.txt"), dtype=np.int8)

    def corrupted(n):
        co = np.zeros((n, 15), np.float32)
        co[:, 7] = 1.0
        co[1, 7], co[1, 14] = 0.1, 0.9
        return co

    small = O.QaryOracle(H6, 10, B=7, DV=3, DC=4)
    hs, ps, _ = small.decode_probs(corrupted(6))
    assert (hs == 0).all()
    med = O.QaryOracle(Hm, 10, B=7, DV=3, DC=7)
    hm, pm, _ = med.decode_probs(corrupted(450))
    assert (hm == 0).all()
    m = coo(Hm)
    save("rust_unit_tests.npz", llr_probs=probs, llr_expected=expected_llr, H6=H6, small_hard=hs, small_post=ps,
         Hm_shape=m["shape"], Hm_rows=m["rows"], Hm_cols=m["cols"], Hm_vals=m["vals"], medium_hard=hm,
         medium_post=pm)
    results["decoder.rs:743-768 into_llr"] = "ok"
    results["decoder.rs:770-799 small_decoder_instance"] = "all zero"
    results["decoder.rs:818-854 medium_decoder_instance"] = "all zero"

    # ---- 2. decode.py:192-209 doctest, inputs captured by replaying its RNG stream
    rng = utils.make_random_state(1)
    H = make_code.make_regular_ldpc_parity_check_matrix_identity(300, 150, 3, 6, rng)
    n, B, BB = H.shape[1], 1, 3
    p_linear = 1 / BB
    good = np.array([p_linear] * BB)
    bad = np.array([p_linear] * BB)
    good[[B, -1]] = np.array([1.75 * p_linear, 0.25 * p_linear])
    bad[[-1, B]] = np.array([1.75 * p_linear, 0.25 * p_linear])
    co = np.zeros((n, BB)).astype(np.float32)
    while True:  # decode.py:248-259
        errs = 0
        for i in range(n):
            if rng.rand() < 0.005:
                co[i, :] = bad
                errs += 1
            else:
                co[i, :] = good
        if errs:
            break
    dec = O.QaryOracle(H.astype(np.int8), 5, B=1, DV=3, DC=7)
    hard, post, _ = dec.decode_probs(co)
    assert (hard == 0).all()
    # and through the reference's own driver
    rng = utils.make_random_state(1)
    H2 = make_code.make_regular_ldpc_parity_check_matrix_identity(300, 150, 3, 6, rng)
    assert (H2 == H).all()
    res = decode.simulate_frame_error_rate_rust(H2, 1, 0.005, 1, rng, 1)
    assert res == 1, res
    c = coo(H.astype(np.int8))
    save("decode_doctest_seed1.npz", H_shape=c["shape"], H_rows=c["rows"], H_cols=c["cols"], H_vals=c["vals"],
         channel_output=co, errs=np.int32(errs), hard=hard, post=post, successes=np.int32(res))
    results["decode.py:192-209 simulate_frame_error_rate_rust"] = int(res)

    # ---- 3. Kyber-768 (SURVEY.md 3.2 composition)
    settings = [(0.995, 3, 3, 5), (0.95, 3, 4, 10), (0.9, 5, 4, 10), (1, 2, 2, 5)]
    for cb in (1, 2):
        Hk = kyber.gen_ldpc_matrix(6, 0, cb).astype(np.int8)
        R, N = Hk.shape
        dec_name = f"DecoderN{N}R{R}SW6"
        all_co, all_cos, all_hard, all_post, all_posts, all_s, all_ssum, meta = [], [], [], [], [], [], [], []
        for si, (acc, npat, npat_sum, iters) in enumerate(settings):
            rs = np.random.RandomState(1000 * cb + si)
            # seeded replacement for kyber.sample_secret_coefs (os.urandom, kyber.py:35-47)
            bits = rs.randint(0, 2, size=(kyber.num_blocks, kyber.block_len, 2 * kyber.ETA))
            s = (bits[..., : kyber.ETA].sum(-1) - bits[..., kyber.ETA:].sum(-1)).tolist()
            ssum = kyber.compute_ssum(s, Hk, cb)
            pats = kyber.get_patterns(kyber.ETA, acc, npat, npat_sum, 6)
            coding = kyber.coding_from_patterns(pats["pattern"])
            coding_sum = kyber.coding_from_patterns(pats["pattern_sum"], 6)
            orc = max_likelihood.SimpleOracle(acc)

            def observe(value, cod):
                x = cod[value]
                flips = rs.rand(len(x)) >= acc
                return tuple(int(b) ^ int(f) for b, f in zip(x, flips))

            cache, cache_sum = {}, {}

            def distr(y, cod, sw, cache):
                if y not in cache:
                    cache[y] = max_likelihood.s_distribution_from_hard_y(
                        y, orc, kyber.secret_range, cod, kyber.secret_distribution(sw), sw)
                return cache[y]

            s_distr = [[distr(observe(v, coding), coding, 1, cache) for v in blk] for blk in s]
            ssum_distr = [distr(observe(v, coding_sum), coding_sum, 6, cache_sum) for v in ssum]
            co, cos = kyber.get_channel_probabilities(s_distr, ssum_distr, 6, cb)
            d = O.QaryOracle(Hk, iters, B=2, BSUM=12, DC=7)
            hard, post, posts = d.decode_probs(co, cos)
            # exact-order DP must agree with the enumeration bit for bit
            h1, p1, ps1 = d.decode_llr(O.into_llr(co), O.into_llr(cos), algo=O.ALGO_EXACT_DP)
            assert (h1 == hard).all() and np.array_equal(p1, post, equal_nan=True) and np.array_equal(
                ps1, posts, equal_nan=True)
            flat_s = np.array(s).reshape(-1)
            ok = bool((hard[: N - R] == flat_s).all() and (hard[N - R:] == -np.array(ssum)).all())
            prior_err = int((np.argmax(co, 1) - 2 != flat_s).sum())
            meta.append(dict(accuracy=acc, num_patterns=npat, num_patterns_sum=npat_sum, iterations=iters,
                             prior_errors=prior_err, posterior_errors=int((hard[: N - R] != flat_s).sum()),
                             success=ok))
            print(dec_name, meta[-1])
            all_co.append(co); all_cos.append(cos); all_hard.append(hard); all_post.append(post)
            all_posts.append(posts); all_s.append(flat_s.astype(np.int8)); all_ssum.append(np.array(ssum, np.int16))
        c = coo(Hk)
        save(f"kyber768_cb{cb}.npz", H_shape=c["shape"], H_rows=c["rows"], H_cols=c["cols"], H_vals=c["vals"],
             channel_output=np.stack(all_co), channel_output_sum=np.stack(all_cos), hard=np.stack(all_hard),
             post=np.stack(all_post), post_sum=np.stack(all_posts), secret=np.stack(all_s),
             ssum=np.stack(all_ssum), iterations=np.array([m["iterations"] for m in meta], np.int32),
             meta=np.array(json.dumps(meta)))
        results[f"kyber768 cb={cb} ({dec_name})"] = meta

    # ---- 4. HQC toy (hqc.py:1229-1274) and the reference doctests through the oracle
    results["hqc.py:1229-1236 test_hqc_decode_toy_example(0)"] = bool(hqc.test_hqc_decode_toy_example(0))
    rng = utils.make_random_state(0)
    Hgen = make_code.make_random_ldpc_parity_check_matrix(20, 3, rng)
    y_sparse = [4, 5, 7, 9]
    r1 = [i for (i, x) in enumerate(Hgen[:, 0]) if x != 0]
    ytr1 = hqc.sparse_times_sparse(y_sparse, r1, 20)
    Hin = Hgen.copy()
    checks = np.array([int(i in ytr1) for i in range(20)], np.uint8)
    Hfull = np.concatenate((Hin, np.identity(20, dtype=int)), axis=1)
    cp = np.concatenate((np.full(20, 4 / 20), np.zeros(20)))
    bpd = O.BinaryBpOracle(Hfull, max_iter=100, bp_method="product_sum", channel_probs=cp)
    msg = np.concatenate((np.zeros(20, int), checks.astype(int)))
    decoded = bpd.decode(msg)
    assert sorted(np.nonzero(decoded[:20])[0].tolist()) == y_sparse
    c = coo(Hfull.astype(np.int8))
    save("hqc_toy.npz", H_shape=c["shape"], H_rows=c["rows"], H_cols=c["cols"], channel_probs=cp, msg=msg,
         decoded=decoded, log_prob_ratios=bpd.log_prob_ratios, iters=np.int32(bpd.iter), y_sparse=np.array(y_sparse))

    results["decode.py:139-149 simulate_frame_error_rate(rep_code(13))"] = int(
        decode.simulate_frame_error_rate(refshim.rep_code(13), decode.ErrorsProvider(0.05, None, utils.make_random_state(0)),
                                         100, utils.make_random_state(0)))
    # the doctest shares ONE rng between provider and simulation; replay exactly that
    rng = utils.make_random_state(0)
    results["decode.py:139-149 simulate_frame_error_rate(rep_code(13))"] = int(
        decode.simulate_frame_error_rate(refshim.rep_code(13), decode.ErrorsProvider(0.05, None, rng), 100, rng))
    assert results["decode.py:139-149 simulate_frame_error_rate(rep_code(13))"] == 100

    if os.environ.get("GOLDEN_FULL_HQC", "0") == "1":  # 2.5 GB dense circulant; ~1 min
        results["hqc.py:1277-1284 test_hqc_decode_full_example(0)"] = bool(hqc.test_hqc_decode_full_example(0))
    else:
        prev = os.path.join(HERE, "reference_doctests.json")
        if os.path.exists(prev):
            old = json.load(open(prev))
            k = "hqc.py:1277-1284 test_hqc_decode_full_example(0)"
            if k in old:
                results[k] = old[k]

    # ---- 5. HQC-shaped small instance with noisy checks (numerical regression fixture for the binary path)
    from simulate import distance_spectrum
    rs = np.random.RandomState(7)
    Nn, W, omega, Rr, rho = 997, 11, 12, 420, 0.97
    first = distance_spectrum.gen_array_ds_multiplicity(Nn, W, 1, rs)
    Hc_col0 = first  # scipy.linalg.circulant(first)[:, 0] == first
    supp_col0 = np.nonzero(Hc_col0)[0]
    y = np.sort(rs.choice(Nn, omega, replace=False))
    ytr1 = set(hqc.sparse_times_sparse(y.tolist(), supp_col0.tolist(), Nn))
    bits = rs.choice(Nn, Rr, replace=False)
    from scipy.linalg import circulant
    Hgen = circulant(first)
    Hin = Hgen[bits]
    true_checks = np.array([int(b in ytr1) for b in bits], np.uint8)
    noisy = true_checks ^ (rs.rand(Rr) >= rho).astype(np.uint8)
    Hfull = np.concatenate((Hin, np.identity(Rr, dtype=int)), axis=1)
    cp = np.concatenate((np.full(Nn, omega / Nn), np.full(Rr, 1 - rho)))
    bpd = O.BinaryBpOracle(Hfull, max_iter=100, bp_method="product_sum", channel_probs=cp)
    msg = np.concatenate((np.zeros(Nn, int), noisy.astype(int)))
    decoded = bpd.decode(msg)
    ok = sorted(np.nonzero(decoded[:Nn])[0].tolist()) == y.tolist()
    print("hqc_small: success", ok, "iters", bpd.iter, "converged", bpd.converge)
    c = coo(Hfull.astype(np.int8))
    save("hqc_small.npz", H_shape=c["shape"], H_rows=c["rows"], H_cols=c["cols"], channel_probs=cp, msg=msg,
         decoded=decoded, log_prob_ratios=bpd.log_prob_ratios, iters=np.int32(bpd.iter),
         converged=np.int32(bpd.converge), y_sparse=y)
    results["hqc_small (N=997, W=11, R=420, rho=0.97)"] = dict(success=ok, iters=int(bpd.iter))

    with open(os.path.join(HERE, "reference_doctests.json"), "w") as f:
        json.dump(results, f, indent=1, sort_keys=True)
    print(json.dumps(results, indent=1))


if __name__ == "__main__":
    main()
