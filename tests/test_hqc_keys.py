"""Key-file front end: the HQC round-3 secret-key expansion restated in sca_ldpc_b200/hqc_keys.py
against known answers and against the reference's own key files (fixture tests/golden/hqc_keys.npz), and the
main.py-shaped CLI's host-side logic.  No GPU."""
import csv
import io
import os
import pickle

import numpy as np
import pytest

import util
from sca_ldpc_b200 import cli
from sca_ldpc_b200 import hqc_keys as K


def test_aes256_known_answer():
    """FIPS-197 appendix C.3."""
    ct = K.AES256(bytes(range(32))).encrypt_block(bytes.fromhex("00112233445566778899aabbccddeeff"))
    assert ct.hex() == "8ea2b7ca516745bfeafc49904b496089"
    try:
        from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes
    except Exception:
        return
    rs = np.random.RandomState(0)
    for _ in range(8):
        k, b = rs.bytes(32), rs.bytes(16)
        assert Cipher(algorithms.AES(k), modes.ECB()).encryptor().update(b) == K.AES256(k).encrypt_block(b)


def test_seed_expander_layout():
    """rng.c seedexpander: block i = AES(key, diversifier || maxlen_be32 || i_be32); reads are a plain byte stream."""
    seed = bytes(range(40))
    e = K.SeedExpander(seed)
    a = e.read(5) + e.read(40) + e.read(3)
    aes = K.AES256(seed[:32])
    want = b"".join(aes.encrypt_block(seed[32:40] + (0xFFFFFFFF).to_bytes(4, "big") + i.to_bytes(4, "big")) for i in range(3))
    assert a == want and e.calls == 3


@pytest.mark.parametrize("name", ["hqc128", "hqc192", "hqc256"])
def test_secrets_from_the_reference_key_files(name):
    """test-hqc*.key (hqc.py:74-93): y = secrets_from_key(sk)[1] (hqc.py:1043-1046).  liboqs is not available to compare
    with, but the public-key equation s = x + h*y is checked on the key itself -- no wrong expansion satisfies it."""
    z = util.load("hqc_keys.npz")
    pk, sk = z[f"{name}_pk"].tobytes(), z[f"{name}_sk"].tobytes()
    n, omega = K.PARAMS[name]
    assert K.param_set_of(pk, sk) == name
    x, y = K.secrets_from_key(sk, pk, verify=True)
    assert np.array_equal(x, z[f"{name}_x"]) and np.array_equal(y, z[f"{name}_y"])
    assert len(set(y.tolist())) == omega and y.max() < n
    assert K.public_key_check(pk, x, y, n)
    # the check has teeth: swapping x and y, moving one position, or touching the public key breaks it
    assert not K.public_key_check(pk, y, x, n)
    y2 = y.copy()
    y2[0] = (y2[0] + 1) % n
    assert not K.public_key_check(pk, x, y2, n)
    bad = bytearray(sk)
    bad[K.SEED_BYTES + 100] ^= 1
    with pytest.raises(ValueError):
        K.secrets_from_key(bytes(bad))
    with pytest.raises(ValueError):
        K.secrets_from_key(sk, pk[:-1] + bytes([pk[-1] ^ 1]))


def test_keyfile_loader_refuses_code(tmp_path):
    z = util.load("hqc_keys.npz")
    pk, sk = z["hqc128_pk"].tobytes(), z["hqc128_sk"].tobytes()
    good = tmp_path / "k.key"
    good.write_bytes(pickle.dumps((pk, sk)))          # the layout hqc.py:74-93 writes
    assert K.load_keyfile(str(good)) == (pk, sk)
    evil = tmp_path / "e.key"
    evil.write_bytes(pickle.dumps((os.getcwd, sk)))   # a pickle that names a callable
    with pytest.raises(pickle.UnpicklingError):
        K.load_keyfile(str(evil))
    short = tmp_path / "s.key"
    short.write_bytes(pickle.dumps((pk[:100], sk)))
    with pytest.raises(ValueError):
        K.load_keyfile(str(short))


def test_cli_arguments_and_ladders(tmp_path):
    """main.py:59-136 option names / defaults; the decode-every ladder's inputs; csv header written once (main.py:150-156)."""
    a = cli.build_parser().parse_args(["hqc_decode_bench"])
    assert (a.runs, a.decode_every, a.code_weight, a.param_set, a.error_rate) == (100, 500, 20, "128", 0.0)
    a = cli.build_parser().parse_args(["hqc_simulate", "--seed", "3", "--key-file", "k", "--csv-output", "o.csv", "--label", "x",
                                       "--param-set", "192", "--error-rate", "nan"])
    assert a.seed == 3 and a.key_file == "k" and a.error_rate != a.error_rate
    rng = np.random.RandomState(1)
    n, y = 997, np.array([4, 50, 333, 700])
    graphs, levels, synd, cert = cli.hqc_ladders(y, n, 7, 100, 300, (1.0, 1.0), 5, rng, shared_code=True)
    assert len(graphs) == 1 and levels.tolist() == [100, 200, 300] and synd.shape == (5, 300) and (synd == synd[0]).all()
    rows = graphs[0][1].reshape(300, 8)      # no noise, one code: all runs agree
    assert (rows[:, 7] == n + np.arange(300)).all() and (np.diff(rows[:, :7], axis=1) > 0).all()
    e = np.zeros(n, np.uint8)
    e[y] = 1
    assert np.array_equal(synd[0], e[rows[:, :7]].sum(1) & 1)
    graphs, _, synd, _ = cli.hqc_ladders(y, n, 7, 100, 300, (1.0, 1.0), 4, np.random.RandomState(2))
    assert len(graphs) == 4 and not np.array_equal(graphs[0][1], graphs[1][1])   # a new code per run (hqc.py:1056)
    for g in range(4):
        assert np.array_equal(synd[g], e[graphs[g][1].reshape(300, 8)[:, :7]].sum(1) & 1)
    _, _, noisy, cert = cli.hqc_ladders(y, n, 7, 100, 300, (0.9, 1.0), 400, np.random.RandomState(1), shared_code=True)
    truth = noisy[:, :].mean(0)   # checks whose true value is 1 are never wrong (certainty 1), the others flip 10 % of the time
    assert set(np.unique(cert).tolist()) == {0.9, 1.0}
    assert np.all(truth[cert[0] == 1.0] == 1.0) and 0.05 < truth[cert[0] == 0.9].mean() < 0.15
    out = tmp_path / "o.csv"
    row = {c: i for i, c in enumerate(cli.HQC_COLUMNS)}
    cli._write_csv(str(out), cli.HQC_COLUMNS, [row])
    cli._write_csv(str(out), cli.HQC_COLUMNS, [row, row])
    got = list(csv.reader(io.StringIO(out.read_text())))
    assert got[0] == cli.HQC_COLUMNS and len(got) == 4
    # hqc.py:245-264 column order
    assert cli.HQC_COLUMNS[:5] == ["label", "alg", "weight", "epsilon0", "epsilon1"]
    assert cli.HQC_COLUMNS[5:] == ["checks", "oracle_calls", "unsatisfied", "good_flips", "bad_flips",
                                   "found_bad_satisfied_checks", "found_bad_unsatisfied_checks", "success"]
