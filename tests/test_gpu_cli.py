"""The main.py-shaped CLI end to end on the GPU: --key-file test-hqc128.key yields the real y, the
decode-every ladder runs as one nested-graph batch, the CSV has the reference's columns (hqc.py:222-264)."""
import csv
import pickle

import numpy as np
import pytest

import util
from sca_ldpc_b200 import cli
from sca_ldpc_b200 import hqc_keys as K

pytestmark = pytest.mark.gpu


def test_hqc_decode_bench_with_the_reference_key(tmp_path):
    z = util.load("hqc_keys.npz")
    key = tmp_path / "test-hqc128.key"
    key.write_bytes(pickle.dumps((z["hqc128_pk"].tobytes(), z["hqc128_sk"].tobytes())))   # hqc.py:74-93 layout
    out = tmp_path / "stats.csv"
    argv = ["hqc_decode_bench", "--key-file", str(key), "--code-weight", "50", "--decode-every", "1000", "--max-checks", "3000",
            "--runs", "6", "--seed", "5", "--label", "t", "--csv-output", str(out)]
    rows = cli.main(argv)
    got = list(csv.DictReader(out.open()))
    assert list(got[0].keys()) == cli.HQC_COLUMNS and len(got) == len(rows)
    assert {r["alg"] for r in got} == {"hqc128"} and {r["weight"] for r in got} == {"50"}
    assert {r["epsilon0"] for r in got} == {"0.9942"} and {r["epsilon1"] for r in got} == {"1.0"}
    # every run's ladder stops at its first success (hqc.py:972-982); with 3000 checks of weight 50 most runs get there
    succ = [r for r in rows if r["success"]]
    assert len(succ) >= 4
    for r in succ:   # a successful decode flipped exactly the omega = 66 bits of the key's y and nothing else
        assert r["good_flips"] == 66 and r["bad_flips"] == 0
    ladders = [r["checks"] for r in rows]
    assert set(ladders) <= {1000, 2000, 3000} and ladders[0] == 1000
    # a wrong --param-set for the key is refused
    with pytest.raises(SystemExit):
        cli.main(["hqc_decode_bench", "--key-file", str(key), "--param-set", "192", "--runs", "1"])
    # appending keeps one header
    cli.main(argv)
    assert sum(1 for line in out.open() if line.startswith("label,")) == 1


def test_kyber_simulate_command(tmp_path):
    out = tmp_path / "k.csv"
    rows = cli.main(["kyber_simulate", "--param-set", "768", "--accuracy", "0.995", "--runs", "2048", "--seed", "3",
                     "--check-blocks", "1", "--csv-output", str(out)])
    assert rows[0]["trials"] == 2048 and rows[0]["success_rate"] > 0.8 and rows[0]["symbol_error_rate"] < 1e-3
    assert list(csv.DictReader(out.open()))[0]["alg"] == "kyber768"
