"""Binary product-sum BP kernel logic (k_bp_cn / k_bp_vn / k_bp_syndrome / k_bp_finish_iter run on the host by
tests/hostmodel) against the fp64 oracle of ldpc.bp_decoder.  The same comparisons run on the device in
tests/test_gpu_bp.py.  Tolerance: hard decisions, iteration counts and convergence flags identical; posterior LLRs
within 2e-3 + 2e-4*|L| (fp32 LLR domain vs fp64 probability ratios)."""
import numpy as np

import util
from oracle import oracle as O
from sca_ldpc_b200.workloads import HqcWorkload


def _fixture(name):
    z = util.load(name)
    H = util.dense(z, with_vals=False)
    row_ptr, col = O.dense_to_csr(H)
    return z, H, row_ptr, col


def test_hqc_toy():
    """hqc.py:1229-1274 toy instance (certainty-1.0 checks: prior probability 0 -> LLR +inf)."""
    z, H, row_ptr, col = _fixture("hqc_toy.npz")
    m, n = H.shape
    msg = z["msg"].astype(np.uint8)
    synd = (H @ msg % 2).astype(np.uint8)[None]
    out = util.hostmodel_bp(row_ptr, col, m, n, 100, synd, z["channel_probs"])
    decoded = out["decoding"][0] ^ msg
    assert np.array_equal(decoded, z["decoded"])
    assert sorted(np.nonzero(decoded[:20])[0].tolist()) == z["y_sparse"].tolist()
    assert int(out["iters"][0]) == int(z["iters"]) and int(out["converged"][0]) == 1
    assert util.llr_close(out["llr"][0], z["log_prob_ratios"])


def test_hqc_small_fixture():
    z, H, row_ptr, col = _fixture("hqc_small.npz")
    m, n = H.shape
    msg = z["msg"].astype(np.uint8)
    synd = (H @ msg % 2).astype(np.uint8)[None]
    out = util.hostmodel_bp(row_ptr, col, m, n, 100, synd, z["channel_probs"])
    assert np.array_equal(out["decoding"][0] ^ msg, z["decoded"])
    assert int(out["iters"][0]) == int(z["iters"]) and int(out["converged"][0]) == int(z["converged"])
    assert util.llr_close(out["llr"][0], z["log_prob_ratios"])


def test_batch_early_stop_and_freeze():
    """T = 45 trials (ragged second tile) with different iteration counts: converged trials freeze while their tile
    mates keep iterating; a non-converging trial runs to max_iter."""
    w = HqcWorkload(variant=(499, 6), W=9, R=230, rho=0.97, seed=3)
    batch = w.generate(45, seed=11)
    synd = batch.syndromes.copy()
    synd[7] = np.random.RandomState(1).randint(0, 2, w.R)  # garbage syndrome: will not converge
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=20, channel_probs=w.channel_probs)
    ref = orc.decode_batch(synd)
    out = util.hostmodel_bp(w.row_ptr, w.col, w.m, w.n, 20, synd, w.channel_probs)
    assert np.array_equal(out["iters"], ref["iters"])
    assert np.array_equal(out["converged"], ref["converged"])
    assert np.array_equal(out["decoding"], ref["decoding"])
    assert len(set(ref["iters"].tolist())) >= 2 and ref["converged"][7] == 0
    for t in range(45):
        assert util.llr_close(out["llr"][t], ref["log_prob_ratios"][t]), t


def test_per_trial_priors_and_no_early_stop():
    w = HqcWorkload(variant=(307, 5), W=7, R=150, rho=0.98, seed=5)
    batch = w.generate(5, seed=2)
    rs = np.random.RandomState(9)
    probs = np.tile(w.channel_probs, (5, 1)) * rs.uniform(0.5, 1.5, size=(5, w.n))
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=15, channel_probs=w.channel_probs)
    ref = orc.decode_batch(batch.syndromes, channel_probs=probs)
    out = util.hostmodel_bp(w.row_ptr, w.col, w.m, w.n, 15, batch.syndromes, probs)
    assert np.array_equal(out["iters"], ref["iters"]) and np.array_equal(out["decoding"], ref["decoding"])
    for t in range(5):
        assert util.llr_close(out["llr"][t], ref["log_prob_ratios"][t]), t
    # without early stop every trial runs max_iter iterations; converged = state of the final iteration
    out2 = util.hostmodel_bp(w.row_ptr, w.col, w.m, w.n, 15, batch.syndromes, probs, early_stop=False)
    assert (out2["iters"] == 15).all()


def test_rep_code_syndrome_decoding():
    """decode.py:139-171: rep_code(13), error_rate 0.05, syndrome in -> error pattern out.  With a uniform prior the
    posterior of an end bit is EXACTLY tied (ratio 1 / LLR 0) after one iteration, so fp32 and fp64 may take the
    decision in different iterations; the decoded pattern is pinned, iteration counts are compared on a
    non-degenerate prior."""
    import refshim

    H = refshim.rep_code(13)
    row_ptr, col = O.dense_to_csr(H)
    rs = np.random.RandomState(0)
    errs = (rs.rand(64, 13) < 0.05).astype(np.uint8)
    synd = (errs @ H.T % 2).astype(np.uint8)
    orc = O.BinaryBpOracle(H, error_rate=0.05, max_iter=13)
    ref = orc.decode_batch(synd)
    out = util.hostmodel_bp(row_ptr, col, 12, 13, 13, synd, np.full(13, 0.05))
    assert np.array_equal(out["decoding"], ref["decoding"]) and (out["decoding"] == errs).all()
    assert (out["converged"] == 1).all()
    probs = rs.uniform(0.03, 0.07, 13)
    orc = O.BinaryBpOracle(H, max_iter=13, channel_probs=probs)
    ref = orc.decode_batch(synd)
    out = util.hostmodel_bp(row_ptr, col, 12, 13, 13, synd, probs)
    assert np.array_equal(out["decoding"], ref["decoding"]) and np.array_equal(out["iters"], ref["iters"])
    for t in range(64):
        assert util.llr_close(out["llr"][t], ref["log_prob_ratios"][t]), t


def test_active_trial_compaction_is_transparent():
    """100 trials (4 tiles, the last ragged), a few with garbage syndromes that never converge, per-trial priors:
    the running trials are compacted into fewer tiles (twice or more); every output must be bit-identical to the
    run without compaction and agree with the oracle."""
    w = HqcWorkload(variant=(499, 6), W=9, R=230, rho=0.97, seed=3)
    batch = w.generate(100, seed=21)
    synd = batch.syndromes.copy()
    rs = np.random.RandomState(4)
    for t in (3, 40, 41, 77, 99):
        synd[t] = rs.randint(0, 2, w.R)
    probs = np.tile(w.channel_probs, (100, 1)) * rs.uniform(0.8, 1.25, size=(100, w.n))
    a = util.hostmodel_bp(w.row_ptr, w.col, w.m, w.n, 25, synd, probs, compact=True)
    b = util.hostmodel_bp(w.row_ptr, w.col, w.m, w.n, 25, synd, probs, compact=False)
    assert a["compactions"] >= 2 and b["compactions"] == 0
    for k in ("decoding", "iters", "converged"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["llr"], b["llr"], equal_nan=True)
    orc = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(w.m, w.n), max_iter=25, channel_probs=w.channel_probs)
    ref = orc.decode_batch(synd, channel_probs=probs)
    assert np.array_equal(a["iters"], ref["iters"]) and np.array_equal(a["decoding"], ref["decoding"])
    assert (ref["converged"][[3, 40, 41, 77, 99]] == 0).all()


def test_segmented_csr_batch_equals_single_graph_decodes():
    """a batch over several UNRELATED graphs of one shape (the reference draws a new code per simulated
    attack, hqc.py:1056) through the modelled kernels with per-tile graph tables == the fp64 oracle on every graph by
    itself.  Trials of the graphs are interleaved, group sizes are not multiples of 32, one trial never converges."""
    from sca_ldpc_b200.workloads import HqcWorkload

    ws = [HqcWorkload(variant=(499, 6), W=9, R=230, rho=0.97, seed=s) for s in (3, 4, 5)]
    m, n = ws[0].m, ws[0].n
    T = 75
    graph_of = (np.arange(T) * 7 % 3).astype(np.int32)          # interleaved; counts 25 / 25 / 25
    graph_of[60:] = 1                                            # -> uneven groups, graph 1 spills into a second tile
    synd = np.zeros((T, m), np.uint8)
    for g, w in enumerate(ws):
        sel = np.nonzero(graph_of == g)[0]
        synd[sel] = w.generate(len(sel), seed=20 + g).syndromes
    synd[7] = np.random.RandomState(1).randint(0, 2, m)          # garbage: runs to the iteration cap
    out = util.hostmodel_bp_multi([(w.row_ptr, w.col) for w in ws], m, n, 25, graph_of, synd, ws[0].channel_probs)
    seen_iters = set()
    for g, w in enumerate(ws):
        sel = np.nonzero(graph_of == g)[0]
        ref = O.BinaryBpOracle(csr=(w.row_ptr, w.col), shape=(m, n), max_iter=25, channel_probs=w.channel_probs).decode_batch(synd[sel])
        assert np.array_equal(out["decoding"][sel], ref["decoding"]), g
        assert np.array_equal(out["iters"][sel], ref["iters"]) and np.array_equal(out["converged"][sel], ref["converged"]), g
        for i, t in enumerate(sel):
            assert util.llr_close(out["llr"][t], ref["log_prob_ratios"][i]), (g, t)
        seen_iters |= set(ref["iters"].tolist())
    assert out["iters"][7] == 25 and not out["converged"][7] and len(seen_iters) >= 3
    # the graphs really differ: decoding graph 1's trials on graph 0 gives something else
    w0 = ws[0]
    sel = np.nonzero(graph_of == 1)[0][:8]
    other = O.BinaryBpOracle(csr=(w0.row_ptr, w0.col), shape=(m, n), max_iter=25, channel_probs=w0.channel_probs).decode_batch(synd[sel])
    assert not np.array_equal(other["decoding"], out["decoding"][sel])
