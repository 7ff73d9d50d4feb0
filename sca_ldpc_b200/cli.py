"""main.py-shaped command line for the two decoder paths  (SURVEY.md 8f.4).

    python -m sca_ldpc_b200.cli hqc_decode_bench --key-file test-hqc128.key --code-weight 50 --decode-every 500 \
        --runs 64 --error-rate 0.0 --seed 1 --csv-output stats.csv --label b200
    python -m sca_ldpc_b200.cli kyber_simulate --param-set 768 --accuracy 0.95 --runs 100000 --csv-output kyber.csv

Reference interface mirrored (simulate-with-python/):
  main.py:59-136     the shared argparse options: --seed --runs --decode-every --key-file --csv-output --code-weight --label
                     --param-set --error-rate                       (same names, defaults and meaning)
  main.py:138-156    command_hqc_simulate: run the simulation, print the stats table, append to --csv-output (header only
                     when the file is new)
  simulate/hqc.py:222-264   the CSV columns, in this order: label, alg, weight, epsilon0, epsilon1, checks, oracle_calls,
                     unsatisfied, good_flips, bad_flips, found_bad_satisfied_checks, found_bad_unsatisfied_checks, success
  simulate/hqc.py:987-1190  simulate_hqc_idealized_oracle: key pair from --key-file, y = secrets_from_key(sk)[1], a random
                     weight-W circulant code, one measured parity check of y*r1 per oracle answer, a decode every
                     DECODE_EVERY checks until y is recovered
What is NOT here: the chosen-ciphertext search that turns decapsulation-oracle answers into those parity checks
(hqc.py:1082-1164) needs the liboqs fork (missing submodule).  `hqc_decode_bench` starts
where the decoder's input starts: check i is the true bit (y*r1)[b_i] of a measured position b_i, reported wrong with the
idealised oracle's error rates (hqc.py:1001-1008: epsilon = (0.9942, 1.0) * (1 - error_rate), indexed by the true value),
with certainty epsilon[true value] as hqc.py:684-690 feeds it to the decoder.  `oracle_calls` is therefore the number of
measured checks (one idealised answer each), not the reference's decapsulation count.
Every run of --runs is one simulated attack with ITS OWN random code (hqc.py:1056); the runs of a pass are ONE batch on
the GPU: a segmented-CSR handle holding the runs' graphs (scaldpc_bp_create_multi), every run's decode-every ladder as
nested prefixes of its graph (the lanes of that run's tile).
"""
from __future__ import annotations

import argparse
import csv
import os
import sys
from math import isnan

import numpy as np

HQC_COLUMNS = ["label", "alg", "weight", "epsilon0", "epsilon1", "checks", "oracle_calls", "unsatisfied", "good_flips",
               "bad_flips", "found_bad_satisfied_checks", "found_bad_unsatisfied_checks", "success"]
KYBER_COLUMNS = ["label", "alg", "sum_weight", "check_blocks", "accuracy", "iterations", "trials", "successes",
                 "success_rate", "symbol_error_rate"]


def make_random_state(seed):
    """simulate/utils.py make_random_state: a numpy RandomState, seeded when --seed is given."""
    return np.random.RandomState(seed) if seed is not None else np.random.RandomState()


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="sca_ldpc_b200.cli", description=__doc__.split("\n\n")[0])
    ap.add_argument("command", choices=["hqc_decode_bench", "hqc_simulate", "kyber_simulate"])
    # main.py:59-136, same names / defaults
    ap.add_argument("--seed", type=int, required=False, help="Use a fixed seed to make simulations reproducible.")
    ap.add_argument("--runs", type=int, default=100, help="Number of simulated attacks (hqc) / Monte-Carlo trials (kyber).")
    ap.add_argument("--decode-every", type=int, default=500,
                    help="Try to decode every 'decode-every' time a measured check is added.")
    ap.add_argument("--key-file", type=str, help="pickle((pk, sk)) as written by the reference (hqc.py:74-93).")
    ap.add_argument("--csv-output", type=str, help="Append decoding stats to this csv file.")
    ap.add_argument("--code-weight", type=int, default=20, help="Column weight of the constructed LDPC code (HQC).")
    ap.add_argument("--label", type=str, help="Label column of the csv output.")
    ap.add_argument("--param-set", type=str, default="128", help="HQC: 128 / 192 / 256.  Kyber: 512 / 768 / 1024.")
    ap.add_argument("--error-rate", type=float, default=0.0,
                    help="Error rate of the measurement channel; 'NaN' guarantees no errors (main.py:118-127).")
    # additions (the reference hard-codes these)
    ap.add_argument("--max-checks", type=int, default=0, help="HQC: stop the ladder after this many checks (default: n/2).")
    ap.add_argument("--max-iter", type=int, default=100, help="BP iteration cap (hqc.py:696: 100; kyber: 10).")
    ap.add_argument("--accuracy", type=float, default=0.95, help="Kyber: per-bit accuracy of the plaintext-checking oracle.")
    ap.add_argument("--sum-weight", type=int, default=6, help="Kyber: row weight of H' (kyber.py:67-74).")
    ap.add_argument("--check-blocks", type=int, default=2, help="Kyber: check blocks of H (kyber.py:67-74).")
    ap.add_argument("--shared-code", action="store_true",
                    help="HQC: one code for all runs (default: a new code per run, as every reference process draws its own, hqc.py:1056)")
    ap.add_argument("--device", type=int, default=0)
    return ap


def _write_csv(path, columns, rows):
    new = not os.path.exists(path)
    with open(path, "w" if new else "a", newline="") as f:  # main.py:150-156: header only for a new file
        wr = csv.writer(f)
        if new:
            wr.writerow(columns)
        wr.writerows([[r[c] for c in columns] for r in rows])


def _print_table(columns, rows, out=sys.stdout):
    widths = [max(len(str(c)), *(len(str(r[c])) for r in rows)) if rows else len(c) for c in columns]
    print(" ".join(str(c).rjust(w) for c, w in zip(columns, widths)), file=out)
    for r in rows:
        print(" ".join(str(r[c]).rjust(w) for c, w in zip(columns, widths)), file=out)


# ---------------------------------------------------------------------------------------------------------------------
def hqc_ladders(y, n, weight, decode_every, max_checks, eps, runs, rng, shared_code=False):
    """Inputs of `runs` simulated attacks on the secret support y (hqc.py:1043-1164 restated from the decoder's side).

    Per run (or once, with shared_code): r1 = support of a random weight-`weight` circulant's first column (hqc.py:1056-1059),
    measured positions b_0 < b_1 < ..., check i = (y * r1)[b_i] reported with the idealised oracle's error rates.
    Returns (graphs [(row_ptr, col)] one per run or one in all, levels, synd (runs, R), certainty (runs, R))."""
    from .workloads import sparse_first_row

    R = max_checks
    e = np.zeros(n, np.uint8)
    e[y] = 1
    row_ptr = (np.arange(R + 1) * (weight + 1)).astype(np.int32)
    graphs, synd, certainty = [], np.zeros((runs, R), np.uint8), np.zeros((runs, R))
    cols = None
    for r in range(runs):
        if cols is None or not shared_code:
            r1 = sparse_first_row(n, weight, rng)                           # make_code.py:220-244, distance_spectrum.py:47-87
            bits = np.sort(rng.choice(n, R, replace=False))                 # which rows of Hgen get measured
            cols = np.sort((bits[:, None] - r1[None, :]) % n, axis=1)       # row b of circulant(r1): columns (b - r) mod n
            col = np.concatenate((cols, n + np.arange(R)[:, None]), axis=1).astype(np.int32).reshape(-1)
            graphs.append((row_ptr, col))
            truth = (e[cols].sum(1) & 1).astype(np.uint8)                   # (y * r1)[b] mod 2, hqc.py:1062-1075, :911-940
            # the idealised oracle (hqc.py:809-871): the answer for a bit whose true value is v is wrong with probability
            # 1 - epsilon[v], and its certainty is epsilon[v]
            eps_of = np.where(truth == 1, eps[1], eps[0])
        synd[r] = truth ^ (rng.random_sample(R) > eps_of)
        certainty[r] = eps_of
    levels = np.arange(decode_every, R + 1, decode_every, dtype=np.int32)
    return graphs, levels, synd, certainty


def command_hqc_decode_bench(args, out=sys.stdout):
    from . import hqc_keys
    from .ldpc import bp_decoder_multi

    rng = make_random_state(args.seed)
    name = f"hqc{args.param_set}"
    if name not in hqc_keys.PARAMS:
        raise SystemExit(f"unknown --param-set {args.param_set} (128, 192, 256)")
    n, omega = hqc_keys.PARAMS[name]
    if args.key_file:
        pk, sk = hqc_keys.load_keyfile(args.key_file)
        if hqc_keys.param_set_of(pk, sk) != name:
            raise SystemExit(f"{args.key_file} is a {hqc_keys.param_set_of(pk, sk)} key, --param-set says {name}")
        _, y = hqc_keys.secrets_from_key(sk, pk, verify=True)               # hqc.py:1043-1046; public-key equation checked
    else:  # hqc.py:1289-1291 (the reference's own full-size decode test): a random weight-omega support
        y = np.sort(rng.choice(n, omega, replace=False))
    noise = 1.0 if isnan(args.error_rate) else 1.0 - args.error_rate
    eps = (1.0, 1.0) if isnan(args.error_rate) else (0.9942 * noise, 1.0 * noise)   # hqc.py:1001-1008
    R = args.max_checks or (n // 2 // args.decode_every) * args.decode_every
    R = max(args.decode_every, (R // args.decode_every) * args.decode_every)
    m, nn = R, n + R
    # hqc.py:684-690: omega/n for the message bits, 1 - certainty for every measured check (it depends on the check's true
    # value, so the priors are per trial)
    base = np.full(n, omega / n)
    rows, first_success = [], np.full(args.runs, -1)
    levels = np.arange(args.decode_every, R + 1, args.decode_every, dtype=np.int32)
    L = len(levels)
    per_pass = max(1, min(64, 2048 // L))                                    # runs (= graphs) per GPU batch
    for r0 in range(0, args.runs, per_pass):
        rr = min(per_pass, args.runs - r0)
        graphs, levels, synd, certainty = hqc_ladders(y, n, args.code_weight, args.decode_every, R, eps, rr, rng,
                                                      shared_code=args.shared_code)
        dec = bp_decoder_multi(graphs, shape=(m, nn), max_iter=args.max_iter,
                               channel_probs=np.concatenate((base, np.full(R, 0.5))), device=args.device)
        s = np.repeat(synd, L, axis=0)                                       # (run, level) pairs: a prefix of the run's checks each
        probs = np.concatenate((np.broadcast_to(base, (rr * L, n)), np.repeat(1.0 - certainty, L, axis=0)), axis=1)
        graph_of = np.repeat(np.zeros(rr, np.int32) if args.shared_code else np.arange(rr, dtype=np.int32), L)
        truth = np.ascontiguousarray(np.broadcast_to(y.astype(np.int32), (rr * L, omega)))
        sc = dec.decode_batch(graph_of, s, channel_probs=probs, truth=truth, n_checks=np.tile(levels, rr), want_decoding=False)
        del dec
        st = sc["stats"].reshape(rr, L, 6)
        for i in range(rr):
            ok = np.nonzero(st[i, :, 5])[0]
            stop = int(ok[0]) if len(ok) else L - 1                          # hqc.py:972-982: the run ends at its first success
            first_success[r0 + i] = int(levels[ok[0]]) if len(ok) else -1
            for k in range(stop + 1):
                u, g, b, fbs, fbu, succ = (int(v) for v in st[i, k])
                rows.append({"label": args.label or "", "alg": name, "weight": args.code_weight, "epsilon0": eps[0],
                             "epsilon1": eps[1], "checks": int(levels[k]), "oracle_calls": int(levels[k]), "unsatisfied": u,
                             "good_flips": g, "bad_flips": b, "found_bad_satisfied_checks": fbs,
                             "found_bad_unsatisfied_checks": fbu, "success": bool(succ)})
    _print_table(HQC_COLUMNS, rows if len(rows) <= 40 else rows[:20] + rows[-20:], out)
    done = first_success[first_success > 0]
    print(f"# {args.runs} simulated attacks on {name} (key file: {args.key_file or 'none, random y'}), code weight "
          f"{args.code_weight}, {'one code for all runs' if args.shared_code else 'a new code per run'}: {len(done)} recovered y; "
          f"checks needed: median {int(np.median(done)) if len(done) else 'n/a'}, max {int(done.max()) if len(done) else 'n/a'}",
          file=out)
    if args.csv_output:
        _write_csv(args.csv_output, HQC_COLUMNS, rows)
    return rows


def command_kyber_simulate(args, out=sys.stdout):
    """SURVEY.md 3.2 (the composition main.py never wired, kyber.py:379-380 raises NotImplementedError): secrets, check sums,
    noisy oracle patterns, Bayes priors, min-sum decode, success = decoded secret == s -- all on the device
    (scaldpc_qary_simulate); the decoder class is the one kyber.py:396-402 would import."""
    from .qary import QaryDecoder
    from .workloads import KyberWorkload

    variant = {"512": "kyber512", "768": "kyber768", "1024": "kyber1024"}.get(args.param_set)
    if variant is None:
        raise SystemExit(f"unknown --param-set {args.param_set} (512, 768, 1024)")
    iters = args.max_iter if args.max_iter != 100 else 10
    wl = KyberWorkload(variant, sum_weight=args.sum_weight, check_blocks=args.check_blocks, accuracy=args.accuracy,
                       nbits=3, nbits_sum=4, seed=args.seed or 0)
    dec = QaryDecoder(wl.H, iters, device=args.device, **wl.decoder_kwargs)
    r = dec.simulate(args.runs, args.seed or 0, **wl.simulate_kwargs)
    row = {"label": args.label or "", "alg": variant, "sum_weight": args.sum_weight, "check_blocks": args.check_blocks,
           "accuracy": args.accuracy, "iterations": iters, "trials": r["trials"], "successes": r["successes"],
           "success_rate": r["successes"] / max(1, r["trials"]),
           "symbol_error_rate": r["symbol_errors"] / max(1, r["trials"] * wl.bvars)}
    _print_table(KYBER_COLUMNS, [row], out)
    if args.csv_output:
        _write_csv(args.csv_output, KYBER_COLUMNS, [row])
    return [row]


def main(argv=None):
    args = build_parser().parse_args(argv)
    if args.command in ("hqc_decode_bench", "hqc_simulate"):
        return command_hqc_decode_bench(args)
    return command_kyber_simulate(args)


if __name__ == "__main__":
    main()
