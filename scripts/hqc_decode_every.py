#!/usr/bin/env python
"""The reference's decode_every loop (simulate/hqc.py:953-984) for many simulated attack runs at once.

In the reference one attack run appends measured checks to H and, every DECODE_EVERY rows, decodes on the rows it has
so far (hqc.py:972-982 -> decode, :680-758), stopping at the first success.  Here every (run, level) pair is one trial
of ONE nested-graph batch on a handle built for the largest H (scaldpc_bp_decode_nested + the device statistics of
scaldpc_bp_decode_scored); runs are sharded over the GPUs and the per-level sums all-reduced once.
Output: one CSV row per level with the columns of HqcSimulationTracking.decoder_stats (hqc.py:232-243), averaged over
the runs, plus `first_success` = fraction of runs whose FIRST successful decode is at this level (where the reference's
loop would have stopped).

    python scripts/hqc_decode_every.py --variant hqc128 --runs 2048 --decode-every 500 --rows 3000
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 \
        scripts/hqc_decode_every.py --runs 16384
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from sca_ldpc_b200.ldpc import bp_decoder, bp_decoder_multi  # noqa: E402
from sca_ldpc_b200.shard import block_range  # noqa: E402
from sca_ldpc_b200.workloads import HqcWorkload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="hqc128")
    ap.add_argument("--runs", type=int, default=2048)
    ap.add_argument("--rows", type=int, default=3000)
    ap.add_argument("--decode-every", type=int, default=500)
    ap.add_argument("--weight", type=int, default=50)
    ap.add_argument("--rho", type=float, default=0.9942, help="certainty of a measured check (1 - its error rate)")
    ap.add_argument("--max-iter", type=int, default=100)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--compare", action="store_true",
                    help="also time the same ladder (a) as per-trial row limits on the full graph and (c) as one decode per level")
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = HqcWorkload(args.variant, W=args.weight, R=args.rows, rho=args.rho, seed=args.seed)
    dec = bp_decoder(csr=(wl.row_ptr, wl.col), shape=(wl.m, wl.n), max_iter=args.max_iter, channel_probs=wl.channel_probs,
                     device=local)
    levels = np.arange(args.decode_every, args.rows + 1, args.decode_every, dtype=np.int32)
    lo, hi = block_range(args.runs, rank, world)
    runs = hi - lo
    batch = wl.generate(runs, seed=args.seed * 1000003 + 17 + lo)   # runs keyed by the first GLOBAL run index of the shard
    L = len(levels)
    synd = np.repeat(batch.syndromes, L, axis=0)                    # (run, level) pairs: the same measurements, a prefix each
    truth = np.repeat(np.ascontiguousarray(batch.y, dtype=np.int32), L, axis=0)
    n_checks = np.tile(levels, runs)
    lvl = np.tile(np.arange(L, dtype=np.int32), runs)
    # the ladder as a family of prefix graphs (segmented CSR): a level pays for its own rows only
    lad = bp_decoder_multi.from_prefixes(wl.row_ptr, wl.col, (wl.m, wl.n), levels, max_iter=args.max_iter,
                                         channel_probs=wl.channel_probs, device=local)
    lad.decode_batch(lvl[: 4 * L], synd[: 4 * L], truth=truth[: 4 * L], n_checks=n_checks[: 4 * L], want_decoding=False)  # warm-up
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    out = lad.decode_batch(lvl, synd, truth=truth, n_checks=n_checks, want_decoding=False)
    dt = time.perf_counter() - t0
    cmp_line = ""
    if args.compare:
        dec.decode_scored(synd[: 4 * L], truth[: 4 * L], n_checks=n_checks[: 4 * L])
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        full = dec.decode_scored(synd, truth, n_checks=n_checks)
        dt_full = time.perf_counter() - t0
        assert np.array_equal(full["stats"], out["stats"]) and np.array_equal(full["iters"], out["iters"])
        dt_sep = 0.0
        for k, R in enumerate(levels):  # one single-graph decode per level on H_k = [Hin[:R_k] | I]
            rp, col = wl.row_ptr[: R + 1], wl.col[: wl.row_ptr[R]]
            dk = bp_decoder(csr=(rp, col), shape=(int(R), wl.N + int(R)), max_iter=args.max_iter,
                            channel_probs=wl.channel_probs[: wl.N + int(R)], device=local)
            sk = np.ascontiguousarray(batch.syndromes[:, :R])
            dk.decode_scored(sk[:4], np.ascontiguousarray(batch.y[:4], dtype=np.int32))
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            one = dk.decode_scored(sk, np.ascontiguousarray(batch.y, dtype=np.int32))
            dt_sep += time.perf_counter() - t0
            assert np.array_equal(one["stats"], out["stats"].reshape(runs, L, 6)[:, k])
        cmp_line = (f"# same ladder, same statistics: prefix graphs {dt:.3f} s | row limits on the full graph {dt_full:.3f} s | "
                    f"sum of {L} separate per-level decodes {dt_sep:.3f} s")
    stats = out["stats"].reshape(runs, L, 6).astype(np.int64)
    ok = stats[:, :, 5].astype(bool)
    first = ok & (np.cumsum(ok, axis=1) == 1)
    sums = np.concatenate((stats.sum(0), first.sum(0)[:, None], out["iters"].reshape(runs, L).sum(0)[:, None]), axis=1)
    tot = torch.from_numpy(sums).to(dev)
    tmax = torch.tensor([dt], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    if rank == 0:
        tot = tot.cpu().numpy() / float(args.runs)
        print("label,alg,weight,epsilon0,epsilon1,checks,unsatisfied,good_flips,bad_flips,found_bad_satisfied_checks,"
              "found_bad_unsatisfied_checks,success,first_success,iterations")
        for k, R in enumerate(levels):
            print(f"b200,{args.variant},{args.weight},{args.rho},{args.rho},{R}," + ",".join(f"{v:.6g}" for v in tot[k]))
        print(f"# {args.runs} runs x {L} levels = {args.runs * L} nested decodes on {world} GPU(s) in {float(tmax):.3f} s "
              f"({args.runs * L / float(tmax):.0f} decodes/s, host buffers in, statistics out)")
        if cmp_line:
            print(cmp_line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
