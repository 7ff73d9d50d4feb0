"""The trial generator of scaldpc_qary_simulate: oracle/kyber_sim.py pinned by known-answer vectors
and by the distributions of the reference's sampling; the device generator and scoring against the oracle (gpu)."""
import numpy as np
import pytest

from oracle import kyber_sim as KS
from sca_ldpc_b200 import workloads as W


def test_philox_known_answers():
    """Random123 kat_vectors, philox4x32 10 rounds."""
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
        ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
        ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
         (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
    ]
    for ctr, key, want in kat:
        got = KS.philox4x32_10(np.array(ctr, dtype=np.uint64), key)
        assert tuple(int(x) for x in got) == want


def small_workload(acc=0.9):
    return W.KyberWorkload("kyber512", sum_weight=2, check_blocks=1, accuracy=acc, nbits=3, nbits_sum=4, seed=3)


def gen(wl, T, seed=5, first=0):
    k = wl.simulate_kwargs
    return KS.generate(wl.H, T, seed, first, k["code"], k["nbits"], k["code_sum"], k["nbits_sum"], k["accuracy"],
                       k["eta"], wl.B, wl.BSUM)


def test_generator_matches_reference_distributions():
    wl = small_workload(0.9)
    T = 400
    secret, ssum, obs, obs_sum = gen(wl, T)
    # kyber.py:35-47 / :60-64: centred binomial(eta)
    freq = np.bincount(secret.ravel().astype(np.int64) + wl.B, minlength=wl.Q) / secret.size
    assert np.abs(freq - W.cbd_prior(wl.eta)).max() < 0.005
    # kyber.py:85-92
    assert (ssum == secret.astype(np.int64) @ wl.Hs.astype(np.int64).T).all()
    # max_likelihood.py:9-18: every bit right with probability `accuracy`
    k = wl.simulate_kwargs
    sent = k["code"][secret.astype(np.int64) + wl.B]
    flips = np.unpackbits((obs ^ sent).astype(np.uint16).view(np.uint8)).sum() / (obs.size * k["nbits"])
    assert abs(flips - 0.1) < 0.005
    sent = k["code_sum"][ssum.astype(np.int64) + wl.BSUM]
    flips = np.unpackbits((obs_sum ^ sent).astype(np.uint16).view(np.uint8)).sum() / (obs_sum.size * k["nbits_sum"])
    assert abs(flips - 0.1) < 0.005
    assert obs.max() < 2 ** k["nbits"] and obs_sum.max() < 2 ** k["nbits_sum"]


def test_generator_is_partition_independent():
    wl = small_workload()
    whole = gen(wl, 24, seed=2 ** 40 + 9, first=2 ** 33 - 10)  # crosses a 32-bit trial-counter boundary
    a = gen(wl, 10, seed=2 ** 40 + 9, first=2 ** 33 - 10)
    b = gen(wl, 14, seed=2 ** 40 + 9, first=2 ** 33)
    for w, x, y in zip(whole, a, b):
        assert (w == np.concatenate((x, y))).all()
    other = gen(wl, 24, seed=2 ** 40 + 10, first=2 ** 33 - 10)
    assert (whole[0] != other[0]).any()


def test_accuracy_one_never_flips():
    wl = small_workload(1.0)
    secret, ssum, obs, obs_sum = gen(wl, 8)
    k = wl.simulate_kwargs
    assert (obs == k["code"][secret.astype(np.int64) + wl.B]).all()
    assert (obs_sum == k["code_sum"][ssum.astype(np.int64) + wl.BSUM]).all()


# ---- device ----------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("variant,sw,acc,nb,nbs", [("kyber512", 2, 0.9, 3, 4), ("kyber768", 6, 0.95, 3, 4),
                                                   ("kyber1024", 8, 0.995, 5, 6), ("kyber512", 4, 1.0, 3, 5)])
def test_device_trials_equal_oracle_and_decode(variant, sw, acc, nb, nbs):
    from sca_ldpc_b200 import QaryDecoder
    wl = W.KyberWorkload(variant, sum_weight=sw, check_blocks=1, accuracy=acc, nbits=nb, nbits_sum=nbs, seed=1)
    dec = QaryDecoder(wl.H, 6, **wl.decoder_kwargs)
    T, seed, first = 77, 2 ** 35 + 12345, 2 ** 32 - 40
    res = dec.simulate(T, seed, first_trial=first, want_trials=True, **wl.simulate_kwargs)
    secret, ssum, obs, obs_sum = gen(wl, T, seed, first)
    assert (res["secret"] == secret).all() and (res["ssum"] == ssum).all()
    assert (res["obs"] == obs).all() and (res["obs_sum"] == obs_sum).all()
    hard, _, _ = dec.min_sum_observed(wl.table, obs, wl.table_sum, obs_sum)
    assert (res["hard"] == hard).all()
    okb = (hard[:, : wl.bvars] == secret).all(1)
    oks = (hard[:, wl.bvars:].astype(np.int16) == -ssum).all(1)
    assert res["trials"] == T
    assert res["successes"] == int((okb & oks).sum())
    assert res["secret_successes"] == int(okb.sum())
    assert res["symbol_errors"] == int((hard[:, : wl.bvars] != secret).sum())
    assert res["sum_symbol_errors"] == int((hard[:, wl.bvars:].astype(np.int16) != -ssum).sum())


@pytest.mark.gpu
def test_device_campaign_is_partition_independent(monkeypatch):
    from sca_ldpc_b200 import QaryDecoder
    wl = W.KyberWorkload("kyber768", sum_weight=6, check_blocks=1, accuracy=0.99, nbits=3, nbits_sum=4, seed=0)
    dec = QaryDecoder(wl.H, 8, **wl.decoder_kwargs)
    kw = wl.simulate_kwargs
    whole = dec.simulate(1000, 7, **kw)
    parts = [dec.simulate(n, 7, first_trial=f, **kw) for f, n in ((0, 333), (333, 1), (334, 666))]
    for key in whole:
        assert whole[key] == sum(p[key] for p in parts)
    assert whole["trials"] == 1000 and 0 < whole["successes"] <= whole["secret_successes"] <= 1000
    # several passes of the resident-trial bound give the same counters
    monkeypatch.setenv("SCALDPC_MAX_RESIDENT_TRIALS", "256")
    dec2 = QaryDecoder(wl.H, 8, **wl.decoder_kwargs)
    assert dec2.simulate(1000, 7, **kw) == whole


@pytest.mark.gpu
def test_device_simulate_argument_errors():
    from sca_ldpc_b200 import QaryDecoder
    wl = small_workload()
    dec = QaryDecoder(wl.H, 3, **wl.decoder_kwargs)
    kw = wl.simulate_kwargs
    with pytest.raises(ValueError):
        dec.simulate(4, 0, **{**kw, "nbits": 2})          # code words wider than nbits
    with pytest.raises(ValueError):
        dec.simulate(4, 0, **{**kw, "accuracy": 1.5})
    with pytest.raises(ValueError):
        dec.simulate(4, 0, **{**kw, "table": kw["table"][:4]})   # fewer rows than 2^nbits
    assert dec.simulate(0, 0, **kw)["trials"] == 0
