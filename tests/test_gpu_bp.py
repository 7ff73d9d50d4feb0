"""Parity of the CUDA binary product-sum BP (through the C ABI / the ldpc.bp_decoder-shaped class) with the fp64 CPU
oracle and the committed fixtures.  Bars: hard decisions, iteration counts and convergence flags identical;
posterior LLRs within 2e-3 + 2e-4*|L| (fp32 LLR domain vs fp64 probability ratios; parity of this path is
"unpinned" upstream, see oracle/bp_oracle.c)."""
import numpy as np
import pytest

import util
from oracle import oracle as O
from sca_ldpc_b200.ldpc import bp_decoder
from sca_ldpc_b200.workloads import HqcWorkload

pytestmark = pytest.mark.gpu


def test_hqc_toy_through_the_boundary():
    """hqc.py:694-708 call shape on the toy of hqc.py:1229-1274 (outcome pinned by the reference: y recovered)."""
    z = util.load("hqc_toy.npz")
    H = util.dense(z, with_vals=False)
    bpd = bp_decoder(H, max_iter=100, bp_method="product_sum", channel_probs=z["channel_probs"])
    decoded = bpd.decode(z["msg"])
    assert sorted(np.nonzero(decoded[:20])[0].tolist()) == z["y_sparse"].tolist()
    assert np.array_equal(decoded, z["decoded"]) and bpd.iter == int(z["iters"]) and bpd.converge == 1
    assert util.llr_close(bpd.log_prob_ratios, z["log_prob_ratios"])


def test_hqc_small_fixture():
    z = util.load("hqc_small.npz")
    H = util.dense(z, with_vals=False)
    bpd = bp_decoder(H, max_iter=100, bp_method="product_sum", channel_probs=z["channel_probs"])
    decoded = bpd.decode(z["msg"])
    assert np.array_equal(decoded, z["decoded"]) and bpd.iter == int(z["iters"])
    assert bpd.converge == int(z["converged"])
    assert util.llr_close(bpd.log_prob_ratios, z["log_prob_ratios"])


def test_rep_code_fer_loop():
    """decode.py:139-171: error_rate + channel_probs=[None]*n, syndrome in -> error pattern out; 100/100."""
    from sca_ldpc_b200.ldpc import codes

    H = codes.rep_code(13)
    bpd = bp_decoder(H, error_rate=0.05, max_iter=13, bp_method="product_sum", channel_probs=[None] * 13)
    rs = np.random.RandomState(0)
    ok = 0
    for _ in range(100):
        err = (rs.rand(13) < 0.05).astype(int)
        ok += int((bpd.decode(H @ err % 2) == err).all())
    assert ok == 100


def test_batch_early_stop_freeze_and_torch():
    import torch

    w = HqcWorkload(variant=(499, 6), W=9, R=230, rho=0.97, seed=3)
    batch = w.generate(45, seed=11)
    synd = batch.syndromes.copy()
    synd[7] = np.random.RandomState(1).randint(0, 2, w.R)
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=20, channel_probs=w.channel_probs)
    ref = orc.decode_batch(synd)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=20, channel_probs=w.channel_probs)
    out = bpd.decode_batch(synd, want_llr=True)
    assert np.array_equal(out["iters"], ref["iters"]) and np.array_equal(out["converged"], ref["converged"])
    assert np.array_equal(out["decoding"], ref["decoding"])
    for t in range(45):
        assert util.llr_close(out["log_prob_ratios"][t], ref["log_prob_ratios"][t]), t
    tout = bpd.decode_batch(torch.from_numpy(synd).cuda(), want_llr=False)
    torch.cuda.synchronize()
    assert np.array_equal(tout["decoding"].cpu().numpy(), ref["decoding"])
    assert np.array_equal(tout["iters"].cpu().numpy(), ref["iters"])
    pinned = torch.empty((45, w.n), dtype=torch.uint8).pin_memory().numpy()
    pout = bpd.decode_batch(synd, out=pinned)
    assert pout["decoding"] is pinned and np.array_equal(pinned, ref["decoding"])
    with pytest.raises(ValueError):
        bpd.decode_batch(synd, out=np.empty((45, w.n), np.int8))


def test_per_trial_priors():
    w = HqcWorkload(variant=(307, 5), W=7, R=150, rho=0.98, seed=5)
    batch = w.generate(5, seed=2)
    rs = np.random.RandomState(9)
    probs = np.tile(w.channel_probs, (5, 1)) * rs.uniform(0.5, 1.5, size=(5, w.n))
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=15, channel_probs=w.channel_probs)
    ref = orc.decode_batch(batch.syndromes, channel_probs=probs)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=15, channel_probs=w.channel_probs)
    out = bpd.decode_batch(batch.syndromes, channel_probs=probs, want_llr=True)
    assert np.array_equal(out["iters"], ref["iters"]) and np.array_equal(out["decoding"], ref["decoding"])
    for t in range(5):
        assert util.llr_close(out["log_prob_ratios"][t], ref["log_prob_ratios"][t]), t
    out2 = bpd.decode_batch(batch.syndromes, channel_probs=probs, early_stop=False)
    assert (out2["iters"] == 15).all()


def test_hqc128_full_size():
    """BASELINE.json configs[3]: HQC-128 (n = 17669 + R), W=50, R=3000, rho=0.9942.  A few trials against the fp64
    oracle, then a larger batch through size-independent properties: converged trials reproduce their syndrome,
    successful trials recover y exactly."""
    w = HqcWorkload("hqc128", W=50, R=3000, rho=0.9942, seed=0)
    batch = w.generate(96, seed=1)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=100, channel_probs=w.channel_probs)
    out = bpd.decode_batch(batch.syndromes, want_llr=True)
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=100, channel_probs=w.channel_probs)
    ref = orc.decode_batch(batch.syndromes[:6])
    assert np.array_equal(out["iters"][:6], ref["iters"]) and np.array_equal(out["decoding"][:6], ref["decoding"])
    for t in range(6):
        assert util.llr_close(out["log_prob_ratios"][t], ref["log_prob_ratios"][t]), t
    conv = out["converged"].astype(bool)
    assert conv.mean() > 0.9
    dec = out["decoding"]
    rows = np.repeat(np.arange(w.m), np.diff(w.row_ptr))
    for t in np.nonzero(conv)[0][:16]:
        s = np.zeros(w.m, np.uint8)
        np.bitwise_xor.at(s, rows, dec[t][w.col])
        assert np.array_equal(s, batch.syndromes[t])
    assert w.success(batch, dec)[conv].mean() > 0.9


@pytest.mark.parametrize("variant,R", [("hqc192", 5000), ("hqc256", 8000)])
def test_hqc192_hqc256_full_size(variant, R):
    """BASELINE.json configs[4] (HQC-256, n = 57637 + R) and the HQC-192 shape at the sizes bench.py quotes: six trials
    against the fp64 oracle (decisions, iteration counts, convergence identical; posteriors within the stated tolerance),
    one of them made undecodable so that the 100-iteration tail and the compaction around it run, then the
    size-independent properties on the whole batch."""
    w = HqcWorkload(variant, W=50, R=R, rho=0.9942, seed=0)
    batch = w.generate(70, seed=1)
    synd = batch.syndromes.copy()
    synd[3] = np.random.RandomState(9).randint(0, 2, w.R)  # garbage check bits: never converges
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=100, channel_probs=w.channel_probs)
    out = bpd.decode_batch(synd, want_llr=True)
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=100, channel_probs=w.channel_probs)
    ref = orc.decode_batch(synd[:6])
    assert np.array_equal(out["iters"][:6], ref["iters"]) and np.array_equal(out["converged"][:6], ref["converged"])
    assert np.array_equal(out["decoding"][:6], ref["decoding"])
    assert ref["iters"][3] == 100 and not ref["converged"][3]
    for t in range(6):
        assert util.llr_close(out["log_prob_ratios"][t], ref["log_prob_ratios"][t]), t
    conv = out["converged"].astype(bool)
    assert conv.mean() > 0.9 and bpd.last_run()[1] >= 1   # the running trials were compacted at least once
    dec = out["decoding"]
    rows = np.repeat(np.arange(w.m), np.diff(w.row_ptr))
    for t in np.nonzero(conv)[0][:8]:
        s = np.zeros(w.m, np.uint8)
        np.bitwise_xor.at(s, rows, dec[t][w.col])
        assert np.array_equal(s, synd[t])
    # these shapes carry fewer checks per unknown than the hqc128 workload: about half of the converged words are the
    # true y (bench.py: success rate 0.52 / 0.43), the others are different sparse solutions of the same syndrome
    assert w.success(batch, dec)[conv].mean() > 0.25


def test_active_trial_compaction_is_transparent():
    """100 trials, five of them never converge, per-trial priors: the running trials are compacted into fewer tiles;
    outputs must be bit-identical to the run without compaction and agree with the fp64 oracle."""
    w = HqcWorkload(variant=(499, 6), W=9, R=230, rho=0.97, seed=3)
    batch = w.generate(100, seed=21)
    synd = batch.syndromes.copy()
    rs = np.random.RandomState(4)
    for t in (3, 40, 41, 77, 99):
        synd[t] = rs.randint(0, 2, w.R)
    probs = np.tile(w.channel_probs, (100, 1)) * rs.uniform(0.8, 1.25, size=(100, w.n))
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=25, channel_probs=w.channel_probs)
    a = bpd.decode_batch(synd, channel_probs=probs, want_llr=True)
    assert bpd.last_run()[1] >= 2
    b = bpd.decode_batch(synd, channel_probs=probs, want_llr=True, compact=False)
    assert bpd.last_run()[1] == 0
    for k in ("decoding", "iters", "converged"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["log_prob_ratios"], b["log_prob_ratios"], equal_nan=True)
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=25, channel_probs=w.channel_probs)
    ref = orc.decode_batch(synd, channel_probs=probs)
    assert np.array_equal(a["iters"], ref["iters"]) and np.array_equal(a["decoding"], ref["decoding"])
    for t in range(100):
        assert util.llr_close(a["log_prob_ratios"][t], ref["log_prob_ratios"][t]), t


def test_large_batch_compaction_hqc128():
    """1024 trials of the HQC-128 workload plus a few garbage syndromes that never converge: the batch is compacted
    several times; results of a sample of trials (converging and not) must match a run of just those trials without
    compaction."""
    w = HqcWorkload("hqc128", W=50, R=3000, rho=0.9942, seed=0)
    batch = w.generate(1024, seed=5)
    synd = batch.syndromes.copy()
    bad = [5, 100, 513, 1023]
    rs = np.random.RandomState(8)
    for t in bad:
        synd[t] = rs.randint(0, 2, w.R)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=40, channel_probs=w.channel_probs)
    out = bpd.decode_batch(synd)
    assert bpd.last_run() == (40, bpd.last_run()[1]) and bpd.last_run()[1] >= 2
    conv = out["converged"].astype(bool)
    assert conv.mean() > 0.9 and not conv[bad].any()
    pick = np.concatenate([np.array(bad), np.nonzero(conv)[0][:28]])
    sub = bpd.decode_batch(synd[pick], compact=False)
    assert np.array_equal(sub["iters"], out["iters"][pick])
    assert np.array_equal(sub["decoding"], out["decoding"][pick])
    assert (out["iters"][~conv] == 40).all()


def test_shared_prior_first_pass_is_bit_identical(monkeypatch):
    """Shared priors: the first check-to-bit messages differ between trials only in sign.  Three ways to run iteration 1
    must agree bit for bit (decisions, iteration counts, posterior LLRs): k_bp_first_table + the FIRST bit pass that forms
    +-tab[e] on the fly (default), k_bp_first_table + k_bp_first_fill + the ordinary bit pass (round 1's way), and
    k_bp_cn<FIRST>.  Large enough for the degree-class kernels (> 64 tiles) and small enough for the single-launch path; with
    nested row limits too (an inactive check sends +0)."""
    w = HqcWorkload(variant=(997, 12), W=11, R=420, rho=0.97, seed=7)
    d = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=30, channel_probs=w.channel_probs)
    for T in (70, 2100):
        b = w.generate(T, seed=3)
        n_checks = np.array([150, 300, 420], np.int32)[np.arange(T) % 3]
        for nc in (None, n_checks):
            monkeypatch.delenv("SCALDPC_BP_NO_FIRST_TABLE", raising=False)
            monkeypatch.delenv("SCALDPC_BP_FIRST_FILL", raising=False)
            a = d.decode_batch(b.syndromes, want_llr=True, n_checks=nc)
            monkeypatch.setenv("SCALDPC_BP_FIRST_FILL", "1")
            f = d.decode_batch(b.syndromes, want_llr=True, n_checks=nc)
            monkeypatch.setenv("SCALDPC_BP_NO_FIRST_TABLE", "1")
            c = d.decode_batch(b.syndromes, want_llr=True, n_checks=nc)
            for other in (f, c):
                assert np.array_equal(a["decoding"], other["decoding"]) and np.array_equal(a["iters"], other["iters"])
                assert np.array_equal(a["log_prob_ratios"], other["log_prob_ratios"], equal_nan=True)


def reference_stats(N, error, checks, y_sparse):
    """hqc.py:709-749, restated loop for loop: -> the 6 arguments of tracking.add_decoder_stats (:751-759).
    `error` is the decoder's error estimate; hqc.decode scores `decoded = bpd.decode(msg)` with msg = [0]*N ++ checks of
    length n, i.e. the CORRECTED vector msg ^ error (hqc.py:703-708)."""
    y_sparse = set(int(v) for v in y_sparse)
    decoded = np.asarray(error).copy()
    decoded[N:N + len(checks)] ^= np.asarray(checks, dtype=decoded.dtype)
    unsatisfied = good = bad = fbs = fbu = 0
    for i, x in enumerate(decoded[:N]):
        if x:
            if i in y_sparse:
                good += 1
            else:
                bad += 1
    for x, c in zip(decoded[N:], checks):
        if c:
            unsatisfied += 1
            if not x:
                fbu += 1
        elif x:
            fbs += 1
    unequal = 0
    for i, yip in enumerate(decoded[:N]):
        yi = i in y_sparse
        if yi or yip:
            unequal |= yi != bool(yip)
    return [unsatisfied, good, bad, fbs, fbu, int(not unequal)]


def test_decode_scored_equals_reference_stats():
    """scaldpc_bp_decode_scored: the device statistics equal hqc.decode's loops on the decoded vectors,
    for a workload where some trials fail (few checks) so every counter is exercised."""
    w = HqcWorkload("hqc128", W=50, R=1400, rho=0.98, seed=2)
    batch = w.generate(70, seed=4)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=40, channel_probs=w.channel_probs)
    plain = bpd.decode_batch(batch.syndromes)
    out = bpd.decode_scored(batch.syndromes, batch.y, want_decoding=True)
    assert np.array_equal(out["decoding"], plain["decoding"]) and np.array_equal(out["iters"], plain["iters"])
    want = np.array([reference_stats(w.N, plain["decoding"][t], batch.syndromes[t], batch.y[t]) for t in range(70)])
    assert np.array_equal(out["stats"], want)
    assert np.array_equal(out["stats"][:, 5].astype(bool), w.success(batch, plain["decoding"]))
    assert 0 < out["stats"][:, 5].sum() < 70 or out["stats"][:, 2].sum() > 0   # not a degenerate all-success case
    # without the decoded vectors (the production call): same statistics
    lean = bpd.decode_scored(batch.syndromes, batch.y)
    assert lean["decoding"] is None and np.array_equal(lean["stats"], want)
    with pytest.raises(ValueError):
        bpd.decode_scored(batch.syndromes, batch.y[:3])


def test_nested_graphs_equal_decodes_on_the_truncated_graphs():
    """scaldpc_bp_decode_nested (the decode_every loop of hqc.py:953-984): one batch whose trials use
    different prefixes of the rows == separate fp64-oracle decodes on H_k = [Hin[:R_k] | I_{R_k}]."""
    w = HqcWorkload(variant=(997, 12), W=11, R=420, rho=0.97, seed=7)
    T = 96
    batch = w.generate(T, seed=3)
    levels = np.array([100, 200, 300, 420])
    n_checks = levels[np.arange(T) % 4].astype(np.int32)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=30, channel_probs=w.channel_probs)
    out = bpd.decode_batch(batch.syndromes, want_llr=True, n_checks=n_checks)
    sc = bpd.decode_scored(batch.syndromes, batch.y, n_checks=n_checks, want_decoding=True)
    assert np.array_equal(sc["decoding"], out["decoding"]) and np.array_equal(sc["iters"], out["iters"])
    some_unconverged = False
    for R in levels:
        sel = np.nonzero(n_checks == R)[0]
        rp, col = w.row_ptr[: R + 1], w.col[: w.row_ptr[R]]
        probs = w.channel_probs[: w.N + R]
        ref = O.BinaryBpOracle(csr=(rp, col), shape=(R, w.N + R), max_iter=30, channel_probs=probs).decode_batch(
            np.ascontiguousarray(batch.syndromes[sel, :R]))
        assert np.array_equal(out["decoding"][sel, : w.N + R], ref["decoding"]), R
        assert np.array_equal(out["iters"][sel], ref["iters"]) and np.array_equal(out["converged"][sel], ref["converged"]), R
        assert (out["decoding"][sel, w.N + R:] == 0).all()
        for i, t in enumerate(sel):
            assert util.llr_close(out["log_prob_ratios"][t, : w.N + R], ref["log_prob_ratios"][i]), (R, t)
            want = reference_stats(w.N, ref["decoding"][i], batch.syndromes[t, :R], batch.y[t])
            assert sc["stats"][t].tolist() == want, (R, t)
        some_unconverged |= not ref["converged"].all()
    assert some_unconverged  # the short prefixes cannot all decode: the freeze / tail paths ran with masks too
    # full prefix for everybody == the plain decode; no checks at all == the priors' decisions after one iteration
    full = bpd.decode_batch(batch.syndromes, n_checks=np.full(T, w.m, np.int32))
    plain = bpd.decode_batch(batch.syndromes)
    assert np.array_equal(full["decoding"], plain["decoding"]) and np.array_equal(full["iters"], plain["iters"])
    none = bpd.decode_batch(batch.syndromes, n_checks=np.zeros(T, np.int32))
    assert (none["decoding"] == 0).all() and (none["iters"] == 1).all() and none["converged"].all()
    with pytest.raises(ValueError):
        bpd.decode_batch(batch.syndromes, n_checks=np.full(T, w.m + 1, np.int32))


def test_nested_graphs_hqc128_with_compaction():
    """Nested decode at HQC-128 size with enough trials for active-trial compaction: every trial equals its own
    single-level decode (same handle, all trials on that level)."""
    w = HqcWorkload("hqc128", W=50, R=3000, rho=0.9942, seed=0)
    T = 512
    batch = w.generate(T, seed=5)
    levels = np.array([1500, 2000, 2500, 3000])
    n_checks = levels[(np.arange(T) * 7 // 3) % 4].astype(np.int32)
    bpd = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=60, channel_probs=w.channel_probs)
    out = bpd.decode_batch(batch.syndromes, n_checks=n_checks)
    assert bpd.last_run()[1] >= 1  # compacted at least once
    for R in levels:
        sel = np.nonzero(n_checks == R)[0]
        one = bpd.decode_batch(np.ascontiguousarray(batch.syndromes[sel]), n_checks=np.full(len(sel), R, np.int32))
        assert np.array_equal(out["decoding"][sel], one["decoding"]), R
        assert np.array_equal(out["iters"][sel], one["iters"]), R
    # success improves with the number of checks (the point of the decode_every loop)
    ok = w.success(batch, out["decoding"])
    rates = [ok[n_checks == R].mean() for R in levels]
    assert rates[-1] > rates[0]


def test_segmented_csr_batch_of_unrelated_graphs():
    """SURVEY.md 8f.2 / hqc.py:1056: every simulated attack draws its OWN code.  One batch over G different random codes
    (scaldpc_bp_create_multi / scaldpc_bp_decode_multi, a graph per tile) == G single-graph fp64-oracle decodes: decisions,
    iteration counts, convergence, posteriors, and hqc.decode's statistics; with each attack's decode-every ladder as
    nested prefixes of ITS graph."""
    from sca_ldpc_b200.ldpc import bp_decoder_multi

    G = 9
    ws = [HqcWorkload(variant=(997, 12), W=11, R=420, rho=0.97, seed=30 + g) for g in range(G)]
    m, n = ws[0].m, ws[0].n
    levels = np.array([140, 280, 420], np.int32)
    per = [5, 1, 40, 3, 33, 2, 7, 32, 6]                       # trials per graph: ragged, one graph needs two tiles
    graph_of = np.concatenate([np.full(c, g, np.int32) for g, c in enumerate(per)])
    perm = np.random.RandomState(0).permutation(len(graph_of))  # the caller's order is arbitrary
    graph_of = graph_of[perm]
    T = len(graph_of)
    synd = np.zeros((T, m), np.uint8)
    truth = np.zeros((T, ws[0].omega), np.int32)
    for g, w in enumerate(ws):
        sel = np.nonzero(graph_of == g)[0]
        b = w.generate(len(sel), seed=50 + g)
        synd[sel], truth[sel] = b.syndromes, b.y
    n_checks = levels[np.arange(T) % 3]
    dec = bp_decoder_multi([(w.row_ptr, w.col) for w in ws], shape=(m, n), max_iter=30, channel_probs=ws[0].channel_probs)
    out = dec.decode_batch(graph_of, synd, truth=truth, n_checks=n_checks, want_llr=True)
    for g, w in enumerate(ws):
        for R in levels:
            sel = np.nonzero((graph_of == g) & (n_checks == R))[0]
            if not len(sel):
                continue
            rp, col = w.row_ptr[: R + 1], w.col[: w.row_ptr[R]]
            ref = O.BinaryBpOracle(csr=(rp, col), shape=(R, w.N + R), max_iter=30, channel_probs=w.channel_probs[: w.N + R]
                                   ).decode_batch(np.ascontiguousarray(synd[sel, :R]))
            assert np.array_equal(out["decoding"][sel, : w.N + R], ref["decoding"]), (g, R)
            assert np.array_equal(out["iters"][sel], ref["iters"]) and np.array_equal(out["converged"][sel], ref["converged"])
            for i, t in enumerate(sel):
                assert util.llr_close(out["log_prob_ratios"][t, : w.N + R], ref["log_prob_ratios"][i]), (g, R, t)
                assert out["stats"][t].tolist() == reference_stats(w.N, ref["decoding"][i], synd[t, :R], truth[t]), (g, R, t)
    assert 0 < out["stats"][:, 5].sum() < T
    # one graph behind the multi interface == the plain decoder
    one = bp_decoder_multi([(ws[2].row_ptr, ws[2].col)], shape=(m, n), max_iter=30, channel_probs=ws[2].channel_probs)
    sel = np.nonzero(graph_of == 2)[0]
    a = one.decode_batch(np.zeros(len(sel), np.int32), synd[sel])
    b = bp_decoder(csr=(ws[2].row_ptr, ws[2].col), shape=(m, n), max_iter=30, channel_probs=ws[2].channel_probs).decode_batch(synd[sel])
    assert np.array_equal(a["decoding"], b["decoding"]) and np.array_equal(a["iters"], b["iters"])
    with pytest.raises(ValueError):
        dec.decode_batch(np.full(T, G, np.int32), synd)


def test_segmented_csr_hqc128_full_size():
    """Four different weight-50 codes of HQC-128 size, 24 trials each in one batch; every graph's first trials against the
    fp64 oracle on that graph, the rest through the syndrome property."""
    from sca_ldpc_b200.ldpc import bp_decoder_multi

    ws = [HqcWorkload("hqc128", W=50, R=3000, rho=0.9942, seed=60 + g) for g in range(4)]
    m, n = ws[0].m, ws[0].n
    graph_of = np.tile(np.arange(4, dtype=np.int32), 24)
    synd = np.zeros((96, m), np.uint8)
    for g, w in enumerate(ws):
        synd[graph_of == g] = w.generate(24, seed=70 + g).syndromes
    dec = bp_decoder_multi([(w.row_ptr, w.col) for w in ws], shape=(m, n), max_iter=100, channel_probs=ws[0].channel_probs)
    out = dec.decode_batch(graph_of, synd)
    for g, w in enumerate(ws):
        sel = np.nonzero(graph_of == g)[0]
        ref = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(m, n), max_iter=100, channel_probs=w.channel_probs).decode_batch(synd[sel[:3]])
        assert np.array_equal(out["decoding"][sel[:3]], ref["decoding"]) and np.array_equal(out["iters"][sel[:3]], ref["iters"])
        rows = np.repeat(np.arange(m), np.diff(w.row_ptr))
        for t in sel[out["converged"][sel].astype(bool)][:6]:
            s = np.zeros(m, np.uint8)
            np.bitwise_xor.at(s, rows, out["decoding"][t][w.col])
            assert np.array_equal(s, synd[t])
    assert out["converged"].mean() > 0.8


def test_prefix_graphs_equal_nested_decode():
    """The decode-every ladder as a family of prefix graphs (bp_decoder_multi.from_prefixes: every level pays for its own
    rows only) == the same ladder as per-trial row limits on the full graph (scaldpc_bp_decode_nested)."""
    from sca_ldpc_b200.ldpc import bp_decoder_multi

    w = HqcWorkload(variant=(997, 12), W=11, R=420, rho=0.97, seed=7)
    T = 90
    batch = w.generate(T, seed=3)
    levels = np.array([100, 200, 300, 420], np.int32)
    lvl = (np.arange(T) % 4).astype(np.int32)
    n_checks = levels[lvl]
    full = bp_decoder(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=30, channel_probs=w.channel_probs)
    a = full.decode_scored(batch.syndromes, batch.y, n_checks=n_checks, want_decoding=True)
    lad = bp_decoder_multi.from_prefixes(w.row_ptr, w.col, (w.m, w.n), levels, max_iter=30, channel_probs=w.channel_probs)
    b = lad.decode_batch(lvl, batch.syndromes, truth=batch.y, n_checks=n_checks)
    assert np.array_equal(a["iters"], b["iters"]) and np.array_equal(a["converged"], b["converged"])
    assert np.array_equal(a["stats"], b["stats"])
    for t in range(T):
        R = n_checks[t]
        assert np.array_equal(a["decoding"][t, : w.N + R], b["decoding"][t, : w.N + R]), t


def test_graphs_built_on_the_device_equal_host_built_graphs():
    """scaldpc_bp_create_circulant: the graphs of a batch of simulated attacks built by kernels from
    (support of the circulant's first column, measured positions) == the same graphs built on the host from their CSR
    (which tests above tie to the fp64 oracle): identical decisions, iteration counts and posteriors, bit for bit; and the
    rows are the reference's: rows of scipy-style circulant(first column) (make_code.py:220-244) plus the identity."""
    from sca_ldpc_b200.ldpc import bp_decoder_multi

    rs = np.random.RandomState(11)
    G, N, W, m = 7, 997, 11, 420
    first_cols = np.stack([rs.choice(N, W, replace=False) for _ in range(G)]).astype(np.int32)      # unsorted on purpose
    rows = np.stack([rs.choice(N, m, replace=False) for _ in range(G)]).astype(np.int32)            # order of measurement
    rows[3, 5] = rows[3, 4]                                                                          # a position measured twice
    graphs = []
    for g in range(G):
        first = np.zeros(N, np.int8)
        first[first_cols[g]] = 1
        circ = np.stack([np.roll(first, k) for k in range(N)], axis=1)      # circulant(c)[i, j] = c[(i - j) mod N]
        H = np.concatenate((circ[rows[g]], np.identity(m, dtype=np.int8)), axis=1)
        r, c = np.nonzero(H)
        graphs.append((np.concatenate(([0], np.cumsum(np.bincount(r, minlength=m)))).astype(np.int32), c.astype(np.int32)))
    probs = np.concatenate((np.full(N, 12 / N), np.full(m, 0.03)))
    T = 150
    graph_of = rs.randint(0, G, T).astype(np.int32)
    synd = rs.randint(0, 2, (T, m)).astype(np.uint8)
    ys = np.stack([rs.choice(N, 12, replace=False) for _ in range(T)])
    for t in range(0, T, 2):   # half of the trials decodable: syndrome of a sparse y plus a little noise
        e = np.zeros(N + m, np.uint8)
        e[ys[t]] = 1
        rp, col = graphs[graph_of[t]]
        synd[t] = np.array([e[col[rp[i]:rp[i + 1]]].sum() & 1 for i in range(m)], np.uint8) ^ (rs.rand(m) < 0.02)
    host = bp_decoder_multi(graphs, shape=(m, N + m), max_iter=30, channel_probs=probs)
    dev = bp_decoder_multi.from_circulants(first_cols, rows, N, max_iter=30, channel_probs=probs)
    assert dev.E == host.E and dev.G == G
    n_checks = np.array([140, 280, 420], np.int32)[np.arange(T) % 3]
    for nc in (None, n_checks):
        a = host.decode_batch(graph_of, synd, n_checks=nc, want_llr=True)
        b = dev.decode_batch(graph_of, synd, n_checks=nc, want_llr=True)
        assert np.array_equal(a["decoding"], b["decoding"]) and np.array_equal(a["iters"], b["iters"])
        assert np.array_equal(a["converged"], b["converged"])
        assert np.array_equal(a["log_prob_ratios"], b["log_prob_ratios"], equal_nan=True)
    assert 0 < a["converged"].sum() < T
    # bad descriptions are refused
    bad = first_cols.copy()
    bad[2, 1] = bad[2, 0]
    with pytest.raises(Exception):
        bp_decoder_multi.from_circulants(bad, rows, N, max_iter=30, channel_probs=probs)
    bad_rows = rows.copy()
    bad_rows[1, 7] = N
    with pytest.raises(Exception):
        bp_decoder_multi.from_circulants(first_cols, bad_rows, N, max_iter=30, channel_probs=probs)


def test_early_copy_of_finished_rows_is_transparent():
    """Host-buffer decodes into page-locked outputs copy the finished rows home while the stragglers iterate
    (scaldpc_bp_last_early_copy): same decisions, LLRs, iteration counts as the plain path into pageable memory."""
    import ctypes as C

    import torch

    from sca_ldpc_b200 import _lib

    wl = HqcWorkload("hqc128", seed=0)
    T = 2048
    b = wl.generate(T, seed=5)
    dec = bp_decoder(csr=(wl.row_ptr, wl.col), shape=(wl.m, wl.n), max_iter=40, channel_probs=wl.channel_probs)
    plain = dec.decode_batch(b.syndromes, want_llr=True)
    rows = C.c_int()
    _lib.check(_lib.lib().scaldpc_bp_last_early_copy(dec._h, C.byref(rows)))
    assert rows.value == -1                      # pageable output: the ordinary copy
    out = torch.empty((T, wl.n), dtype=torch.uint8).pin_memory().numpy()
    out[:] = 7
    pinned = dec.decode_batch(b.syndromes, out=out)
    _lib.check(_lib.lib().scaldpc_bp_last_early_copy(dec._h, C.byref(rows)))
    assert 0 < rows.value <= 256                 # the early copy ran, with stragglers re-sent at the end
    assert 0 < (plain["converged"] == 0).sum()   # ... and there were stragglers to re-send
    assert np.array_equal(pinned["decoding"], plain["decoding"]) and pinned["decoding"] is out
    assert np.array_equal(pinned["iters"], plain["iters"]) and np.array_equal(pinned["converged"], plain["converged"])
