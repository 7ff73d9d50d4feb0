"""Parity of the CUDA q-ary decoder (through the C ABI / the simulate_rs-shaped Python classes) with the CPU oracle
and the committed golden fixtures.  Bars: hard decisions bit-identical; posteriors bit-identical in exact-order
mode, within 1e-4 absolute with identical inf/NaN masks in the fast (forward/backward) mode -- the tolerance
BASELINE.json's north_star states."""
import json
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import util
from oracle import oracle as O
from sca_ldpc_b200 import QaryDecoder, simulate_rs
from sca_ldpc_b200.workloads import KyberWorkload

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _corrupted(n):
    co = np.zeros((n, 15), np.float32)
    co[:, 7] = 1.0
    co[1, 7], co[1, 14] = 0.1, 0.9
    return co


def test_rust_unit_tests_through_the_boundary():
    """decoder.rs:770-854 via simulate_rs.DecoderN6R3V3C4B7 and a synthesised DecoderN450R150V3C7B7."""
    z = util.load("rust_unit_tests.npz")
    d = simulate_rs.DecoderN6R3V3C4B7(z["H6"], 10)
    assert d.min_sum(_corrupted(6)) == [0] * 6
    dm = simulate_rs.DecoderN450R150V3C7B7(util.dense(z, "Hm"), 10)
    assert dm.min_sum(_corrupted(450)) == [0] * 450
    for exact in (True, False):
        hard, post, _ = dm.min_sum_batch(_corrupted(450)[None], exact=exact)
        assert np.array_equal(hard[0], z["medium_hard"])
        if exact:
            assert np.array_equal(post[0], z["medium_post"], equal_nan=True)
        else:
            assert util.same_nonfinite(post[0], z["medium_post"]) and util.max_abs_diff(post[0], z["medium_post"]) < TOL


def test_decode_doctest_seed1_threadpool():
    """simulate/decode.py:192-262: class looked up by name, min_sum called from a ThreadPoolExecutor."""
    z = util.load("decode_doctest_seed1.npz")
    H = util.dense(z)
    n, r, v, c, B = 450, 150, 3, 7, 1
    decoder = getattr(simulate_rs, f"DecoderN{n}R{r}V{v}C{c}B{B}")(H.astype(np.int8), 5)
    with ThreadPoolExecutor(max_workers=4) as ex:
        futs = [ex.submit(decoder.min_sum, z["channel_output"].copy()) for _ in range(8)]
        assert all(f.result() == [0] * n for f in futs)
    hard, post, _ = decoder.min_sum_batch(z["channel_output"][None], exact=True)
    assert np.array_equal(hard[0], z["hard"]) and np.array_equal(post[0], z["post"], equal_nan=True)


@pytest.mark.parametrize("cb", [1, 2])
def test_kyber768_fixture(cb):
    """DecoderN1024R256SW6 / DecoderN1280R512SW6 (lib.rs:54-75) on reference-produced priors."""
    z = util.load(f"kyber768_cb{cb}.npz")
    H = util.dense(z)
    R, N = H.shape
    cls = getattr(simulate_rs, f"DecoderN{N}R{R}SW6")
    for i in range(4):
        it = int(z["iterations"][i])
        d = cls(H, it)
        co, cos = z["channel_output"][i], z["channel_output_sum"][i]
        assert d.min_sum(co, cos) == z["hard"][i].tolist()
        # fast path (k_cn_fast<6,2>)
        hard, post, posts = d.min_sum_batch(co[None], cos[None])
        assert np.array_equal(hard[0], z["hard"][i])
        assert util.same_nonfinite(post[0], z["post"][i]) and util.same_nonfinite(posts[0], z["post_sum"][i])
        assert util.max_abs_diff(post[0], z["post"][i]) < TOL and util.max_abs_diff(posts[0], z["post_sum"][i]) < TOL
        # exact-order path from oracle LLRs: bit-identical posteriors
        llr, llrs = O.into_llr(co), O.into_llr(cos)
        hard, post, posts = d.min_sum_batch(llr[None], llrs[None], llr=True, exact=True)
        assert np.array_equal(hard[0], z["hard"][i])
        assert np.array_equal(post[0], z["post"][i], equal_nan=True)
        assert np.array_equal(posts[0], z["post_sum"][i], equal_nan=True)
        # runtime-shape forward/backward kernel
        hard, post, posts = d.min_sum_batch(llr[None], llrs[None], llr=True, generic=True)
        assert np.array_equal(hard[0], z["hard"][i]) and util.max_abs_diff(post[0], z["post"][i]) < TOL


def test_device_logf_vs_host():
    """into_llr on the device (sc_logf.cuh, glibc's logf restated) against the oracle's libm logf: bit-identical
    LLRs, hence bit-identical posteriors."""
    z = util.load("kyber768_cb1.npz")
    d = simulate_rs.DecoderN1024R256SW6(util.dense(z), 1)
    co, cos = z["channel_output"][1], z["channel_output_sum"][1]
    # one iteration with uninformative checks is not needed: compare posteriors of a 0-information decode instead
    llr = O.into_llr(co)
    h1, p1, _ = d.min_sum_batch(co[None], cos[None])
    h2, p2, _ = d.min_sum_batch(llr[None], O.into_llr(cos)[None], llr=True)
    assert np.array_equal(h1, h2) and np.array_equal(p1, p2, equal_nan=True)


# (variant, sum_weight, check_blocks, oracle accuracy, code bits, sum code bits, trials, iterations)
PARITY_CASES = [
    ("kyber512", 6, 1, 0.98, 4, 5, 256, 10),   # THE bench workload (BASELINE.json configs[1], bench.py WORKLOADS)
    ("kyber512", 6, 1, 0.90, 4, 5, 48, 10),    # the same shape where (almost) nothing converges
    ("kyber768", 6, 2, 0.95, 3, 4, 64, 10),    # bench.py's kyber768 workload (north_star's target shape)
    ("kyber768", 6, 2, 0.90, 3, 4, 40, 10),
    ("kyber1024", 8, 1, 0.98, 5, 6, 40, 6),    # bench.py's kyber1024 workload
    ("kyber1024", 8, 1, 0.95, 4, 5, 40, 6),    # undecodable at this accuracy
    ("kyber1024", 4, 1, 0.95, 4, 5, 45, 6),
    ("kyber1024", 4, 1, 0.90, 4, 5, 45, 3),
]


@pytest.mark.parametrize("variant,sw,cb,acc,nb,nbs,T,it", PARITY_CASES)
def test_batched_parity_with_oracle(variant, sw, cb, acc, nb, nbs, T, it):
    """Every BASELINE.json Kyber shape, decodable and undecodable workloads, against the REFERENCE's arithmetic: the
    oracle's exact-order DP (ALGO_EXACT_DP, asserted bit-identical to the enumeration of decoder_special.rs:531-554 in
    tests/golden/make_golden.py and tests/test_oracle_golden.py).  north_star's bar: hard decisions identical per trial,
    posteriors within 1e-4 absolute."""
    w = KyberWorkload(variant, sum_weight=sw, check_blocks=cb, accuracy=acc, nbits=nb, nbits_sum=nbs, seed=0 if T >= 256 else 1)
    b = w.generate(T, seed=1 if T >= 256 else 3)
    d = QaryDecoder(w.H, it, **w.decoder_kwargs)
    orc = O.QaryOracle(w.H, it, B=w.B, BSUM=w.BSUM, DC=sw + 1)
    # identical LLR inputs (the oracle's into_llr) -> strict comparison
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    rh, rp, rps, st = orc.decode_batch(llr, llrs, from_probs=False, algo=O.ALGO_EXACT_DP)
    assert (st == 0).all()
    # 1. the DEFAULT path: fast check node, verified (at-risk trials re-decoded in the reference's rounding order)
    hard, post, posts = d.min_sum_batch(llr, llrs, llr=True)
    run = d.last_run
    assert run["trials"] == T
    assert np.array_equal(hard, rh), f"{int((hard != rh).any(1).sum())} trials differ from the reference arithmetic"
    assert util.same_nonfinite(post, rp) and util.max_abs_diff(post, rp) < TOL
    assert util.same_nonfinite(posts, rps) and util.max_abs_diff(posts, rps) < TOL
    synd = rh.astype(np.int32) @ w.H.T.astype(np.int32)
    assert run["redecoded"] >= int((synd != 0).any(1).sum())  # every trial that did not converge was decoded again
    if acc >= 0.98:
        assert run["redecoded"] < T // 2
    # 2. the exact-order kernel on every trial: bit-identical posteriors
    eh, ep, eps_ = d.min_sum_batch(llr, llrs, llr=True, exact=True)
    assert np.array_equal(eh, rh)
    assert np.array_equal(ep, rp, equal_nan=True) and np.array_equal(eps_, rps, equal_nan=True)
    # 3. separately: the UNVERIFIED fast kernel is the arithmetic of its CPU model (oracle algo 3) bit for bit
    fh, fp, fps, st = orc.decode_batch(llr, llrs, from_probs=False, algo=O.ALGO_FAST)
    uh, up, ups = d.min_sum_batch(llr, llrs, llr=True, verify=False)
    assert d.last_run["redecoded"] == 0
    assert np.array_equal(uh, fh)
    assert np.array_equal(up, fp, equal_nan=True) and np.array_equal(ups, fps, equal_nan=True)
    assert np.array_equal(w.success(b, hard), w.success(b, rh))


@pytest.mark.parametrize("variant,sw,cb", [("kyber512", 6, 1), ("kyber768", 6, 2), ("kyber1024", 8, 1), ("kyber1024", 4, 1)])
def test_fast_kernel_with_excluded_values(variant, sw, cb):
    """Priors in which a third of the entries of both kinds of variable are zero (LLR +inf) on every template shape.
    k_cn_fast does not apply "own entry non-finite => +inf out" to the B-variables' messages (kCnMaskOwn): unverified it must
    still equal its CPU model, which applies it, bit for bit; the default path must give the reference's decisions
    (exact-order DP) and posteriors within 1e-4 with identical inf masks."""
    w = KyberWorkload(variant, sum_weight=sw, check_blocks=cb, accuracy=0.95, seed=11)
    rs = np.random.RandomState(5)
    T, Q, QS, bv, R = 40, 2 * w.B + 1, 2 * w.BSUM + 1, w.H.shape[1] - w.H.shape[0], w.H.shape[0]
    pri = rs.dirichlet(np.ones(Q), size=(T, bv)).astype(np.float32)
    pri[rs.rand(T, bv, Q) < 0.33] = 0.0
    pri[..., w.B] += 1e-3
    pri /= pri.sum(-1, keepdims=True)
    pris = rs.dirichlet(np.ones(QS), size=(T, R)).astype(np.float32)
    pris[rs.rand(T, R, QS) < 0.33] = 0.0
    pris[..., w.BSUM] += 1e-3
    pris /= pris.sum(-1, keepdims=True)
    llr = np.stack([O.into_llr(x) for x in pri])
    llrs = np.stack([O.into_llr(x) for x in pris])
    for it in (1, 2, 3, 5):
        d = QaryDecoder(w.H, it, **w.decoder_kwargs)
        orc = O.QaryOracle(w.H, it, **w.decoder_kwargs)
        fh, fp, fps, st = orc.decode_batch(llr, llrs, from_probs=False, algo=O.ALGO_FAST)
        assert (st == 0).all()
        uh, up, ups = d.min_sum_batch(llr, llrs, llr=True, verify=False)
        assert np.array_equal(uh, fh)
        assert np.array_equal(up, fp, equal_nan=True) and np.array_equal(ups, fps, equal_nan=True)
        n = 8  # the reference's arithmetic on a few trials (the exact-order DP is slow on 37-value alphabets)
        rh, rp, rps, st = orc.decode_batch(llr[:n], llrs[:n], from_probs=False, algo=O.ALGO_EXACT_DP)
        hard, post, posts = d.min_sum_batch(llr, llrs, llr=True)
        assert np.array_equal(hard[:n], rh)
        assert util.same_nonfinite(post[:n], rp) and util.max_abs_diff(post[:n], rp) < TOL
        assert util.same_nonfinite(posts[:n], rps) and util.max_abs_diff(posts[:n], rps) < TOL



@pytest.mark.parametrize("variant,sw,cb", [("kyber512", 6, 1), ("kyber768", 6, 2), ("kyber1024", 8, 1),
                                           ("kyber1024", 4, 1)])
def test_batched_api_variants(variant, sw, cb):
    """Seeded synthetic batches (T not a multiple of 32): the runtime-shape kernels, torch CUDA tensors in / out, device-side
    into_llr, caller-provided pinned result buffer -- all against the default path's results."""
    import torch

    w = KyberWorkload(variant, sum_weight=sw, check_blocks=cb, accuracy=0.95, nbits=4, nbits_sum=5, seed=1)
    T = 40
    b = w.generate(T, seed=3)
    d = QaryDecoder(w.H, 6, **w.decoder_kwargs)
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    hard, post, posts = d.min_sum_batch(llr, llrs, llr=True)
    # runtime-shape kernels (k_cn_generic forward/backward + k_vn, full sum-variable round trip; flagged trials go through
    # the exact-order kernel as well): same decisions, posteriors within the tolerance
    gh, gp, gps = d.min_sum_batch(llr, llrs, llr=True, generic=True)
    assert np.array_equal(gh, hard)
    assert util.same_nonfinite(gp, post) and util.max_abs_diff(gp, post) < TOL
    assert util.same_nonfinite(gps, posts) and util.max_abs_diff(gps, posts) < TOL
    # the two exact-order kernels (registers / run-time shapes) agree bit for bit
    d.set_iterations(2)
    e1 = d.min_sum_batch(llr[:8], llrs[:8], llr=True, exact=True)
    e2 = d.min_sum_batch(llr[:8], llrs[:8], llr=True, exact=True, generic=True)
    for a, c in zip(e1, e2):
        assert np.array_equal(a, c, equal_nan=True)
    d.set_iterations(6)
    # probabilities in (into_llr on the device), device-resident tensors; per-trial errors are returned, not dropped
    th, tp, tps = d.min_sum_batch(torch.from_numpy(b.priors).cuda(), torch.from_numpy(b.priors_sum).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(th.cpu().numpy(), hard), "device-side into_llr must equal libm's logf bit for bit"
    assert np.array_equal(tp.cpu().numpy(), post, equal_nan=True)
    assert np.array_equal(tps.cpu().numpy(), posts, equal_nan=True)
    bad = b.priors.copy()
    bad[3, 2] *= 0.5
    with pytest.raises(ValueError):  # decoder.rs:683-684 through the device-pointer path
        d.min_sum_batch(torch.from_numpy(bad).cuda(), torch.from_numpy(b.priors_sum).cuda())
    _, _, _, st = d.min_sum_batch(torch.from_numpy(bad).cuda(), torch.from_numpy(b.priors_sum).cuda(), want_status=True)
    assert st[3].item() == -2 and int((st != 0).sum().item()) == 1
    # host buffers with a caller-provided (pinned) result buffer
    pinned = torch.empty((T, w.N), dtype=torch.int8).pin_memory().numpy()
    oh, _, _ = d.min_sum_batch(b.priors, b.priors_sum, want_posterior=False, out=pinned)
    assert oh is pinned and np.array_equal(pinned, hard)
    with pytest.raises(ValueError):
        d.min_sum_batch(b.priors, b.priors_sum, out=np.empty((T, w.N + 1), np.int8))


def test_exact_mode_matches_enumeration_oracle_kyber768():
    """Bit-exact against the reference's brute-force enumeration (decoder_special.rs:531-554) on a small batch."""
    w = KyberWorkload("kyber768", 6, 1, accuracy=0.95, nbits=3, nbits_sum=4, seed=2)
    b = w.generate(3, seed=5)
    d = QaryDecoder(w.H, 3, **w.decoder_kwargs)
    orc = O.QaryOracle(w.H, 3, B=2, BSUM=12, DC=7, algo=O.ALGO_ENUMERATE)
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    rh, rp, rps, st = orc.decode_batch(llr, llrs, from_probs=False)
    hard, post, posts = d.min_sum_batch(llr, llrs, llr=True, exact=True)
    assert np.array_equal(hard, rh) and np.array_equal(post, rp, equal_nan=True)
    assert np.array_equal(posts, rps, equal_nan=True)


def test_irregular_rows_zero_probabilities_and_errors():
    rs = np.random.RandomState(5)
    R, bv, B, sw = 6, 9, 2, 3
    H = np.zeros((R, bv + R), np.int8)
    for r in range(R):
        deg = 2 + (r % 2)
        cols = rs.choice(bv, deg, replace=False)
        H[r, cols] = rs.choice([-1, 1], deg)
        H[r, bv + r] = 1
    T = 37
    Q, QS = 2 * B + 1, 2 * B * sw + 1
    pri = rs.dirichlet(np.ones(Q), size=(T, bv)).astype(np.float32)
    pri[rs.rand(T, bv, Q) < 0.15] = 0.0
    pri[..., B] += 1e-3
    pri /= pri.sum(-1, keepdims=True)
    pris = rs.dirichlet(np.ones(QS), size=(T, R)).astype(np.float32)
    d = QaryDecoder(H, 4, B=B, BSUM=B * sw, DC=sw + 1)
    orc = O.QaryOracle(H, 4, B=B, BSUM=B * sw, DC=sw + 1)
    rh, rp, rps, st = orc.decode_batch(pri, pris, from_probs=True, algo=O.ALGO_ENUMERATE)
    llr = np.stack([O.into_llr(x) for x in pri])
    llrs = np.stack([O.into_llr(x) for x in pris])
    hard, post, posts = d.min_sum_batch(llr, llrs, llr=True, exact=True)
    assert np.array_equal(hard, rh) and np.array_equal(post, rp, equal_nan=True)
    hard, post, posts = d.min_sum_batch(pri, pris)
    assert np.array_equal(hard, rh) and util.same_nonfinite(post, rp) and util.max_abs_diff(post, rp) < TOL
    # decoder.rs:683-684: a prior row that does not sum to 1 -> the reference panics; here ValueError / status
    bad = pri.copy()
    bad[3, 2] *= 0.5
    with pytest.raises(ValueError):
        d.min_sum_batch(bad, pris)
    _, _, _, status = d.min_sum_batch(bad, pris, want_status=True)
    assert status[3] == -2 and (np.delete(status, 3) == 0).all()
    with pytest.raises(TypeError):
        d.min_sum(pri[0].astype(np.float64), pris[0])
    # empty batch
    h0, _, _ = d.min_sum_batch(pri[:0], pris[:0])
    assert h0.shape == (0, bv + R)


def test_no_config_is_reported_per_trial():
    H = np.array([[1, 1, 0], [0, 1, 1]], np.int8)
    d = simulate_rs.DecoderN3R2V2C2B1(H, 2)
    p_ok = np.array([[0.2, 0.6, 0.2]] * 3, np.float32)
    p_bad = np.array([[0, 0, 1], [0, 0, 1], [0.2, 0.6, 0.2]], np.float32)
    _, _, _, status = d.min_sum_batch(np.stack([p_ok, p_bad]), want_status=True)
    assert status.tolist() == [0, -4]
    with pytest.raises(ValueError):
        d.min_sum(p_bad)


def test_full_size_properties_kyber512():
    """BASELINE.json configs[1] at full size (10^4 trials): size-independent properties instead of the oracle --
    decoded words satisfy every parity check (sum_j H_ij s_j + s_sum,i = 0) whenever decoding succeeds, results do
    not depend on the batch split, and a trial's result does not depend on its batch neighbours."""
    w = KyberWorkload("kyber512", 6, 1, accuracy=0.98, nbits=4, nbits_sum=5, seed=0)
    T = 10_000
    b = w.generate(T, seed=1)
    d = QaryDecoder(w.H, 10, **w.decoder_kwargs)
    hard, _, _ = d.min_sum_batch(b.priors, b.priors_sum, want_posterior=False)
    ok = w.success(b, hard)
    assert ok.mean() > 0.9
    synd = hard[ok].astype(np.int32) @ w.H.T.astype(np.int32)
    assert (synd == 0).all()
    h2, _, _ = d.min_sum_batch(b.priors[1000:1777], b.priors_sum[1000:1777], want_posterior=False)
    assert np.array_equal(h2, hard[1000:1777])
    perm = np.random.RandomState(0).permutation(T)[:512]
    h3, _, _ = d.min_sum_batch(b.priors[perm], b.priors_sum[perm], want_posterior=False)
    assert np.array_equal(h3, hard[perm])
    # the whole batch against the exact-order kernel (the reference's arithmetic, test_batched_parity_with_oracle):
    # hard decisions identical for EVERY trial, posteriors within 1e-4
    _, post, posts = d.min_sum_batch(b.priors[:4096], b.priors_sum[:4096])
    n_re = d.last_run["redecoded"]
    assert 0 < n_re < 4096 // 4
    eh, ep, eps_ = d.min_sum_batch(b.priors, b.priors_sum, exact=True)
    assert np.array_equal(hard, eh)
    assert util.same_nonfinite(post, ep[:4096]) and util.max_abs_diff(post, ep[:4096]) < TOL
    assert util.same_nonfinite(posts, eps_[:4096]) and util.max_abs_diff(posts, eps_[:4096]) < TOL


def test_observed_priors_equal_expanded_priors():
    """scaldpc_qary_decode_observed: row indices + tables in, bit-identical to the decode of the
    expanded prior vectors; host buffers (chunked pipeline), device tensors, ragged T, bad indices / bad table rows."""
    import torch

    w = KyberWorkload("kyber768", sum_weight=6, check_blocks=2, accuracy=0.95, nbits=3, nbits_sum=4, seed=4)
    T = 77
    b = w.generate_observed(T, seed=5)
    d = QaryDecoder(w.H, 5, **w.decoder_kwargs)
    rh, rp, rps = d.min_sum_batch(w.table[b.obs], w.table_sum[b.obs_sum])
    hard, post, posts = d.min_sum_observed(w.table, b.obs, w.table_sum, b.obs_sum, want_posterior=True)
    assert np.array_equal(hard, rh) and np.array_equal(post, rp, equal_nan=True) and np.array_equal(posts, rps, equal_nan=True)
    assert w.success(b, hard).mean() > 0.5
    # device tensors, asynchronous on the current stream
    cu = lambda a, dt=None: torch.from_numpy(a if dt is None else a.view(dt)).cuda()  # noqa: E731
    th, _, _ = d.min_sum_observed(cu(w.table), cu(b.obs, np.int16), cu(w.table_sum), cu(b.obs_sum, np.int16))
    torch.cuda.synchronize()
    assert np.array_equal(th.cpu().numpy(), rh)
    # exact-order check node through the same entry point
    d.set_iterations(2)
    eh, ep, _ = d.min_sum_observed(w.table, b.obs[:3], w.table_sum, b.obs_sum[:3], want_posterior=True, exact=True)
    xh, xp, _ = d.min_sum_batch(w.table[b.obs[:3]], w.table_sum[b.obs_sum[:3]], exact=True)
    assert np.array_equal(eh, xh) and np.array_equal(ep, xp, equal_nan=True)
    # an index outside the table -> that trial's status is ERR_ARG, the others decode normally
    bad = b.obs.copy()
    bad[5, 17] = w.table.shape[0]
    _, _, _, st = d.min_sum_observed(w.table, bad, w.table_sum, b.obs_sum, want_status=True)
    assert st[5] == -1 and (np.delete(st, 5) == 0).all()
    with pytest.raises(ValueError):
        d.min_sum_observed(w.table, bad, w.table_sum, b.obs_sum)
    # a table row that does not sum to 1 -> PROB_SUM for exactly the trials that use it (decoder.rs:683-684)
    tab = w.table.copy()
    tab[3] *= 0.5
    _, _, _, st = d.min_sum_observed(tab, b.obs, w.table_sum, b.obs_sum, want_status=True)
    assert np.array_equal(st == -2, (b.obs == 3).any(1)) and ((st == 0) | (st == -2)).all()
    # empty batch
    h0, _, _ = d.min_sum_observed(w.table, b.obs[:0], w.table_sum, b.obs_sum[:0])
    assert h0.shape == (0, w.N)


def test_device_path_bounds_its_workspace(monkeypatch):
    """Device-pointer batches larger than the resident-trial cap are decoded in passes; results do not change."""
    import torch

    w = KyberWorkload("kyber768", sum_weight=6, check_blocks=1, accuracy=0.95, nbits=3, nbits_sum=4, seed=2)
    b = w.generate(150, seed=8)
    pri, pris = torch.from_numpy(b.priors).cuda(), torch.from_numpy(b.priors_sum).cuda()
    ref = QaryDecoder(w.H, 4, **w.decoder_kwargs).min_sum_batch(pri, pris, want_status=True)
    monkeypatch.setenv("SCALDPC_MAX_RESIDENT_TRIALS", "64")
    got = QaryDecoder(w.H, 4, **w.decoder_kwargs).min_sum_batch(pri, pris, want_status=True)
    torch.cuda.synchronize()
    for a, c in zip(ref, got):
        assert np.array_equal(a.cpu().numpy(), c.cpu().numpy(), equal_nan=True)


def test_concurrent_min_sum_calls_are_combined():
    """simulate/decode.py:247-262 drives ONE decoder object from a thread pool; concurrent min_sum() calls are decoded
    as combined batches and every caller gets its own result (or its own error)."""
    w = KyberWorkload("kyber768", sum_weight=6, check_blocks=1, accuracy=0.95, nbits=3, nbits_sum=4, seed=6)
    b = w.generate(96, seed=2)
    d = simulate_rs.DecoderN1024R256SW6(w.H, 4)
    ref, _, _ = d.min_sum_batch(b.priors, b.priors_sum, want_posterior=False)
    bad = b.priors[5].copy()
    bad[3] *= 0.5  # decoder.rs:683-684: this call, and only this call, must fail

    def call(i):
        if i == 5:
            with pytest.raises(ValueError):
                d.min_sum(bad, b.priors_sum[5])
            return None
        return d.min_sum(b.priors[i], b.priors_sum[i])

    n0 = d.launch_count
    r0, b0 = d.combining_stats()
    with ThreadPoolExecutor(max_workers=16) as ex:
        out = list(ex.map(call, range(96)))
    r1, b1 = d.combining_stats()
    assert r1 - r0 == 96 and 0 < b1 - b0 < 96  # scaldpc_qary_decode_one: every call served, in fewer batches than calls
    for i, o in enumerate(out):
        if i != 5:
            assert o == ref[i].tolist(), i
    per_batch = 2 + 2 * 4 + 1  # k_load_priors x2, (check + variable pass) x4, k_untile
    assert (d.launch_count - n0) < 96 * per_batch  # fewer batches than calls: calls were combined


def test_early_termination_small_batch_equals_model(monkeypatch):
    """SCALDPC_QARY_EARLY_STOP (an extension; off = the reference's fixed iteration count) against the per-trial model
    built from the oracle (tests/util.py early_stop_model): decisions, per-trial iteration counts, the re-decode of what
    never converges; with and without the compacted second round."""
    w = KyberWorkload("kyber768", 6, 1, accuracy=0.995, nbits=3, nbits_sum=4, seed=0)
    T, it = 70, 7
    b = w.generate(T, seed=1)
    llr = np.stack([O.into_llr(x) for x in b.priors])
    llrs = np.stack([O.into_llr(x) for x in b.priors_sum])
    llr[5, ::2] = 0.0   # two trials that cannot converge
    llr[40, 1::2] = 0.0
    for restart_min in ("1", "100000"):
        monkeypatch.setenv("SCALDPC_ES_RESTART_MIN", restart_min)
        d = QaryDecoder(w.H, it, **w.decoder_kwargs)
        for verify in (True, False):
            want_h, want_it, never = util.early_stop_model(w, llr, llrs, it, verify)
            hard, _, _ = d.min_sum_batch(llr, llrs, llr=True, early_stop=True, verify=verify)
            assert np.array_equal(d.last_iterations(T), want_it)
            assert np.array_equal(hard, want_h)
            run = d.last_run
            assert run["redecoded"] == (int(never.sum()) if verify else 0)
            if restart_min == "1":
                assert run["trial_iterations"] < T * it
        # flag off: nothing changes
        h0, _, _ = d.min_sum_batch(llr, llrs, llr=True)
        assert d.last_run["trial_iterations"] == T * it + d.last_run["redecoded"] * it


@pytest.mark.parametrize("variant,cb,acc,nb,nbs", [("kyber512", 1, 0.98, 4, 5), ("kyber768", 2, 0.95, 3, 4)])
def test_early_termination_full_size(variant, cb, acc, nb, nbs):
    """10^4 trials of the bench workloads: early termination against the fixed-iteration (reference semantics) decode.
    Trials that never converge get the reference's result (verified mode); a frozen trial may differ only if its
    decisions would still have moved after every row test held -- counted, and rare."""
    w = KyberWorkload(variant, 6, cb, accuracy=acc, nbits=nb, nbits_sum=nbs, seed=0)
    T = 10_000
    b = w.generate(T, seed=1)
    d = QaryDecoder(w.H, 10, **w.decoder_kwargs)
    ref, _, _ = d.min_sum_batch(b.priors, b.priors_sum, want_posterior=False)
    hard, _, _ = d.min_sum_batch(b.priors, b.priors_sum, want_posterior=False, early_stop=True)
    its = d.last_iterations(T)
    run = d.last_run
    differ = (hard != ref).any(1)
    assert differ.sum() <= T // 100, int(differ.sum())
    assert not differ[its == 10].any()          # ran to the cap: re-decoded like the fixed-iteration path does
    synd = hard[its < 10].astype(np.int32) @ w.H.T.astype(np.int32)
    assert (synd == 0).all()                    # frozen trials satisfy every row
    assert 1 <= its.min() and its.max() == 10 and its.mean() < 6
    assert run["trial_iterations"] < 0.75 * T * 10
    assert np.array_equal(w.success(b, hard)[~differ], w.success(b, ref)[~differ])
    # observation-indexed priors and the on-device Monte-Carlo loop take the same flag
    ob = w.generate_observed(T, seed=1)   # the same trials as b (the generator's stream depends on T)
    oh, _, _ = d.min_sum_observed(w.table, ob.obs[:2048], w.table_sum, ob.obs_sum[:2048], early_stop=True)
    assert np.array_equal(oh, hard[:2048])
    sim_a = d.simulate(4096, 7, **w.simulate_kwargs)
    sim_b = d.simulate(4096, 7, early_stop=True, **w.simulate_kwargs)
    assert abs(sim_a["successes"] - sim_b["successes"]) <= 41


def test_host_pipeline_deferred_redecode_and_overflow():
    """Host buffers, chunked pipeline (T >= 4096): the chunks only list their at-risk trials and put their priors aside, the
    exact-order re-decode runs once at the end.  An undecodable workload flags (almost) every trial, far more than the
    retry buffers hold, so the overflow path (priors uploaded again) runs too.  Results must equal the exact-order decode
    of the whole batch, the device-pointer path, and the observation-indexed path."""
    import torch

    w = KyberWorkload("kyber1024", sum_weight=4, check_blocks=1, accuracy=0.9, nbits=4, nbits_sum=5, seed=1)
    T = 4200
    b = w.generate(T, seed=3)
    d = QaryDecoder(w.H, 3, **w.decoder_kwargs)
    eh, ep, eps_ = d.min_sum_batch(b.priors, b.priors_sum, exact=True)
    hard, post, posts = d.min_sum_batch(b.priors, b.priors_sum)
    run = d.last_run
    assert run["trials"] == T and run["redecoded"] > 2048   # more than two chunks' worth of flagged trials
    assert np.array_equal(hard, eh)
    assert util.same_nonfinite(post, ep) and util.max_abs_diff(post, ep) < TOL
    assert util.same_nonfinite(posts, eps_) and util.max_abs_diff(posts, eps_) < TOL
    th, tp, tps = d.min_sum_batch(torch.from_numpy(b.priors).cuda(), torch.from_numpy(b.priors_sum).cuda())
    torch.cuda.synchronize()
    assert d.last_run["redecoded"] == run["redecoded"]
    assert np.array_equal(th.cpu().numpy(), hard) and np.array_equal(tp.cpu().numpy(), post, equal_nan=True)
    assert np.array_equal(tps.cpu().numpy(), posts, equal_nan=True)
    ob = w.generate_observed(T, seed=3)
    oh, op, ops = d.min_sum_observed(w.table, ob.obs, w.table_sum, ob.obs_sum, want_posterior=True)
    assert d.last_run["redecoded"] == run["redecoded"]
    assert np.array_equal(oh, hard) and np.array_equal(op, post, equal_nan=True) and np.array_equal(ops, posts, equal_nan=True)


def test_single_decode_graph_path():
    """Host buffers with T <= 32 (what min_sum() and the thread-pool combiner produce) replay ONE CUDA graph per (T,
    iterations, flags); results must equal the plain-launch path (device tensors), on the first call (capture) and on
    replays, for converging and at-risk trials, with and without posteriors, after a change of the iteration count."""
    import torch

    w = KyberWorkload("kyber768", sum_weight=6, check_blocks=1, accuracy=0.95, nbits=3, nbits_sum=4, seed=2)
    b = w.generate(40, seed=8)
    good = KyberWorkload("kyber768", sum_weight=6, check_blocks=1, accuracy=0.999, nbits=3, nbits_sum=4, seed=2).generate(40, seed=9)
    d = QaryDecoder(w.H, 4, **w.decoder_kwargs)
    for it in (4, 6):
        d.set_iterations(it)
        for batch in (b, good):
            for T in (1, 7, 32):
                pri, pris = batch.priors[:T], batch.priors_sum[:T]
                ref = d.min_sum_batch(torch.from_numpy(pri).cuda(), torch.from_numpy(pris).cuda())
                torch.cuda.synchronize()
                n_re = d.last_run["redecoded"]
                for _ in range(3):  # capture, replay, replay
                    got = d.min_sum_batch(pri, pris)
                    assert d.last_run == {"trials": T, "redecoded": n_re, "trial_iterations": (T + n_re) * it}
                    for a, c in zip(ref, got):
                        assert np.array_equal(a.cpu().numpy(), c, equal_nan=True)
                h2, _, _ = d.min_sum_batch(pri, pris, want_posterior=False)
                assert np.array_equal(h2, got[0])
    # the reference-shaped call itself
    d.set_iterations(5)
    one = d.min_sum(good.priors[3], good.priors_sum[3])
    assert one == d.min_sum_batch(good.priors[3:4], good.priors_sum[3:4], want_posterior=False)[0][0].tolist()
    assert (np.array(one[: w.bvars]) == good.secret[3]).all()


def test_sparse_constructor_equals_dense():
    """scaldpc_qary_create_csr: a decoder built from the CSR of H gives bit-identical results to the one built from the dense
    int8 matrix (pydecoder.rs:24-45), special and generic decoder."""
    w = KyberWorkload("kyber768", sum_weight=6, check_blocks=2, accuracy=0.95, nbits=3, nbits_sum=4, seed=4)
    b = w.generate(50, seed=5)
    rows, cols = np.nonzero(w.H)
    rp = np.concatenate(([0], np.cumsum(np.bincount(rows, minlength=w.R)))).astype(np.int32)
    sparse = QaryDecoder(None, 5, shape=(w.R, w.N), csr=(rp, cols, w.H[rows, cols]), **w.decoder_kwargs)
    dense = QaryDecoder(w.H, 5, **w.decoder_kwargs)
    assert (sparse.R, sparse.N, sparse.E, sparse.Q, sparse.QS) == (dense.R, dense.N, dense.E, dense.Q, dense.QS)
    for a, c in zip(sparse.min_sum_batch(b.priors, b.priors_sum), dense.min_sum_batch(b.priors, b.priors_sum)):
        assert np.array_equal(a, c, equal_nan=True)
    z = util.load("rust_unit_tests.npz")
    Hm = util.dense(z, "Hm")
    rows, cols = np.nonzero(Hm)
    rp = np.concatenate(([0], np.cumsum(np.bincount(rows, minlength=150)))).astype(np.int32)
    g = QaryDecoder(None, 10, B=7, DV=3, DC=7, shape=(150, 450), csr=(rp, cols, Hm[rows, cols]))
    hard, post, _ = g.min_sum_batch(_corrupted(450)[None], exact=True)
    assert np.array_equal(hard[0], z["medium_hard"]) and np.array_equal(post[0], z["medium_post"], equal_nan=True)


@pytest.mark.parametrize("variant,sw,cb,acc,nb,nbs", [("kyber768", 6, 2, 0.95, 3, 4), ("kyber1024", 8, 1, 0.95, 5, 6),
                                                      ("kyber1024", 4, 1, 0.99, 5, 6)])
def test_default_path_equals_exact_kernel_at_scale(variant, sw, cb, acc, nb, nbs):
    """The other BASELINE.json shapes at batch size (2048 trials, a mix of converging and non-converging ones): the default
    path (fast check node + verification) against the exact-order kernel = the reference's arithmetic
    (decoder_special.rs:539-553; tied to the enumeration oracle by test_batched_parity_with_oracle): hard decisions identical
    for every trial, posteriors within 1e-4 absolute with the same inf/NaN masks."""
    w = KyberWorkload(variant, sw, cb, accuracy=acc, nbits=nb, nbits_sum=nbs, seed=3)
    T = 2048
    b = w.generate(T, seed=4)
    d = QaryDecoder(w.H, 10, **w.decoder_kwargs)
    hard, post, posts = d.min_sum_batch(b.priors, b.priors_sum)
    n_re = d.last_run["redecoded"]
    eh, ep, eps_ = d.min_sum_batch(b.priors, b.priors_sum, exact=True)
    assert np.array_equal(hard, eh)
    assert util.same_nonfinite(post, ep) and util.max_abs_diff(post, ep) < TOL
    assert util.same_nonfinite(posts, eps_) and util.max_abs_diff(posts, eps_) < TOL
    ok = w.success(b, hard)
    print(variant, sw, "re-decoded", n_re, "success", ok.mean())
    assert n_re > 0 and ok.mean() < 1.0          # the verification had something to do
    assert n_re < T or ok.mean() == 0.0          # ... and re-decoded everything only where nothing decodes


def test_fp64_verification_mode_on_the_gpu():
    """scaldpc_qary_decode_f64 (north_star: "fp64 option for verification") against the oracle's f64 instantiation
    (oracle/qary_oracle.c, REAL = double; exact-order DP, asserted bit-equal to the enumeration in f32): hard decisions
    identical, posteriors within 1e-9 (CUDA's log() is within 1 ulp of libm's, everything else is the same sequence of
    double operations); and the f32 product path lies within 1e-4 of it wherever no decision is at a tie."""
    # special decoder: reference-produced kyber768 priors (lib.rs:54-64) ...
    z = util.load("kyber768_cb1.npz")
    H = util.dense(z)
    it = int(z["iterations"][0])
    d = simulate_rs.DecoderN1024R256SW6(H, it)
    orc = O.QaryOracle(H, it, B=2, BSUM=12, DC=7, algo=O.ALGO_EXACT_DP)
    n = 4
    same_it = [i for i in range(len(z["iterations"])) if int(z["iterations"][i]) == it][:n]
    co, cos = z["channel_output"][same_it], z["channel_output_sum"][same_it]
    hard, post, posts = d.min_sum_batch_f64(co, cos)
    assert post.dtype == np.float64 and posts.dtype == np.float64
    for k, i in enumerate(same_it):
        h64, p64, ps64 = orc.decode_probs(co[k], cos[k], dtype=np.float64)
        assert np.array_equal(hard[k], h64)
        assert util.same_nonfinite(post[k], p64) and util.same_nonfinite(posts[k], ps64)
        assert util.max_abs_diff(post[k], p64) < 1e-9 and util.max_abs_diff(posts[k], ps64) < 1e-9
        assert np.array_equal(hard[k], z["hard"][i])                     # = the reference's own f32 decisions
        assert util.max_abs_diff(post[k].astype(np.float32), z["post"][i]) < TOL
    # ... and a kyber512 batch larger than one tile, against the f32 product path
    w = KyberWorkload("kyber512", 6, 1, accuracy=0.98, nbits=4, nbits_sum=5, seed=0)
    b = w.generate(70, seed=9)
    d5 = QaryDecoder(w.H, 6, **w.decoder_kwargs)
    h64, p64, ps64 = d5.min_sum_batch_f64(b.priors, b.priors_sum)
    h32, p32, ps32 = d5.min_sum_batch(b.priors, b.priors_sum)
    o5 = O.QaryOracle(w.H, 6, algo=O.ALGO_EXACT_DP, **w.decoder_kwargs)
    for t in (0, 33, 69):
        oh, op, ops = o5.decode_probs(b.priors[t], b.priors_sum[t], dtype=np.float64)
        assert np.array_equal(h64[t], oh) and util.max_abs_diff(p64[t], op) < 1e-9 and util.max_abs_diff(ps64[t], ops) < 1e-9
    agree = (h64 == h32).all(1)
    assert agree.mean() > 0.9                                             # ties between f32 and f64 are rare
    assert util.max_abs_diff(p64[agree].astype(np.float32), p32[agree]) < 2e-4
    # generic decoder (decoder.rs doctest input): per-trial errors are reported as in the f32 path
    z = util.load("decode_doctest_seed1.npz")
    dg = getattr(simulate_rs, "DecoderN450R150V3C7B1")(util.dense(z).astype(np.int8), 5)
    hg, pg, _ = dg.min_sum_batch_f64(z["channel_output"][None])
    assert np.array_equal(hg[0], z["hard"]) and util.max_abs_diff(pg[0].astype(np.float32), z["post"]) < TOL
    bad = np.stack([z["channel_output"], z["channel_output"]])
    bad[1, 3] *= 0.5
    _, _, _, st = dg.min_sum_batch_f64(bad, want_status=True)
    assert st.tolist() == [0, -2]
