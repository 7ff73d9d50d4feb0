"""Host-side input producers (sca_ldpc_b200/workloads.py) against the reference's Python helpers: through committed
fixtures everywhere, and by importing the reference itself when /root/reference is present (build container)."""
import numpy as np
import pytest

import refshim
import util
from sca_ldpc_b200 import workloads as W


@pytest.mark.parametrize("cb", [1, 2])
def test_qc_parity_check_equals_reference_matrix(cb):
    """kyber.gen_ldpc_matrix(6, 0, cb) (kyber.py:67-74, make_code.py:50-94) stored in the Kyber-768 fixtures."""
    z = util.load(f"kyber768_cb{cb}.npz")
    H = W.qc_parity_check(256, 6, 3, 0, cb)
    assert np.array_equal(H, util.dense(z))
    assert ((H != 0).sum(1) == 7).all() and set(np.unique(H)) == {-1, 0, 1}


def test_kyber_workload_shapes_and_consistency():
    for variant, sw, cb, Q, QS in (("kyber512", 6, 1, 7, 37), ("kyber768", 6, 2, 5, 25), ("kyber1024", 8, 1, 5, 33)):
        w = W.KyberWorkload(variant, sw, cb, accuracy=0.95, nbits=3, nbits_sum=4, seed=0)
        b = w.generate(5, seed=1)
        assert b.priors.shape == (5, w.bvars, Q) and b.priors_sum.shape == (5, w.R, QS)
        assert b.priors.dtype == np.float32 and np.allclose(b.priors.sum(-1), 1, atol=1e-5)
        assert np.allclose(b.priors_sum.sum(-1), 1, atol=1e-5)
        # every row of H is satisfied by (s, -ssum): sum_j H_ij s_j + s_sum,i = 0   (kyber.py:85-92, 374-375)
        full = np.concatenate((b.secret.astype(np.int32), -b.ssum.astype(np.int32)), axis=1)
        assert (full @ w.H.T.astype(np.int32) == 0).all()
        assert np.abs(b.secret).max() <= w.B and np.abs(b.ssum).max() <= w.BSUM
        b2 = w.generate(5, seed=1)
        assert np.array_equal(b.priors, b2.priors) and np.array_equal(b.secret, b2.secret)


def test_hqc_workload_graph():
    w = W.HqcWorkload(variant=(499, 6), W=9, R=230, rho=0.97, seed=3)
    assert w.row_ptr[-1] == 230 * 10 and (np.diff(w.col.reshape(230, 10), axis=1) > 0).all()
    b = w.generate(3, seed=1)
    # noise-free check bits equal H[:, :N] . e_y mod 2
    e = np.zeros((3, w.N), np.uint8)
    np.put_along_axis(e, b.y, 1, axis=1)
    clean = np.stack([e[t][w._cols].sum(1) & 1 for t in range(3)])
    assert (clean != b.syndromes).mean() < 0.1


@pytest.mark.skipif(not refshim.available(), reason="/root/reference is not present on this machine")
def test_against_the_reference_python_helpers():
    refshim.install()
    from simulate import kyber, make_code, max_likelihood, utils

    H = make_code.make_qary_qc_parity_check_matrix(256, 6, 3, utils.make_random_state(7), 2)
    assert np.array_equal(W.qc_parity_check(256, 6, 3, 7, 2), H)
    H = make_code.make_qary_qc_parity_check_matrix(256, 8, 4, utils.make_random_state(3), 1)
    assert np.array_equal(W.qc_parity_check(256, 8, 4, 3, 1), H)
    # Bayes posterior table (max_likelihood.py:87-105) for the reference's own 3-pattern coding at accuracy 0.95
    pats = kyber.get_patterns(2, 0.95, 3, 4, 6)
    coding = kyber.coding_from_patterns(pats["pattern"])
    cod = np.array([coding[s] for s in range(-2, 3)], np.int8)
    table = W.posterior_table(cod, W.cbd_prior(2), 0.95)
    orc = max_likelihood.SimpleOracle(0.95)
    for y in range(8):
        bits = tuple((y >> k) & 1 for k in range(3))
        ref = max_likelihood.s_distribution_from_hard_y(bits, orc, kyber.secret_range, coding,
                                                        kyber.secret_distribution(1), 1)
        assert np.allclose(table[y], ref, rtol=1e-12, atol=1e-15)
    assert np.allclose(W.cbd_prior(2, 6), [kyber.secret_distribution(6)[s] for s in range(-12, 13)])


def test_observed_batches_expand_to_the_prior_vectors():
    """generate_observed is generate in index form: table[obs] reproduces the priors bit for bit."""
    w = W.KyberWorkload("kyber512", sum_weight=6, check_blocks=1, accuracy=0.97, nbits=4, nbits_sum=5, seed=2)
    a, b = w.generate(33, seed=9), w.generate_observed(33, seed=9)
    assert b.obs.dtype == np.uint16 and b.obs.shape == (33, w.bvars) and b.obs_sum.shape == (33, w.R)
    assert np.array_equal(w.table[b.obs], a.priors) and np.array_equal(w.table_sum[b.obs_sum], a.priors_sum)
    assert np.array_equal(a.secret, b.secret) and np.array_equal(a.ssum, b.ssum)
