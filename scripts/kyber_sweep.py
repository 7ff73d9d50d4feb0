#!/usr/bin/env python
"""BASELINE.json configs[2]: Kyber-1024 q-ary BP sweep over oracle-accuracy levels, 10^5 trials sharded across the GPUs.

    python scripts/kyber_sweep.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        scripts/kyber_sweep.py --trials 100000

Default: the whole Monte-Carlo loop runs on the device (scaldpc_qary_simulate): rank r synthesises,
decodes and scores the trials block_range(trials, r, world) of ONE global Philox stream, so the counters do not depend
on the number of GPUs.  --host-trials: every batch of 12 500 trials is generated with numpy from a seed keyed by its
GLOBAL index and fed to the decoder as oracle observations (scaldpc_qary_decode_observed).  Either way the only
exchange is one all-reduce of the counters per accuracy level (sca_ldpc_b200/shard.py).  Accuracy levels = the keys
of the reference's pattern database (simulate/kyber.py:115-164).  One JSON line per level on rank 0.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from sca_ldpc_b200 import QaryDecoder  # noqa: E402
from sca_ldpc_b200.shard import Counters, all_reduce_counters, block_range, run_campaign  # noqa: E402
from sca_ldpc_b200.workloads import KyberWorkload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=100_000)
    ap.add_argument("--batch", type=int, default=12_500)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--sum-weight", type=int, default=8)
    ap.add_argument("--accuracies", default="0.9,0.95,0.995,1.0")
    ap.add_argument("--host-trials", action="store_true", help="numpy trial generation + observations API")
    ap.add_argument("--seed", type=int, default=20260000)
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        dist.all_reduce(torch.zeros(1, device=dev))  # NCCL sets its communicator up on the first collective (~1.5 s):
        torch.cuda.synchronize(dev)                  # not part of any accuracy level's campaign time
    n_batches = (args.trials + args.batch - 1) // args.batch
    for acc in (float(a) for a in args.accuracies.split(",")):
        wl = KyberWorkload("kyber1024", sum_weight=args.sum_weight, check_blocks=1, accuracy=acc, nbits=5, nbits_sum=6, seed=0)
        dec = QaryDecoder(wl.H, args.iters, device=local, **wl.decoder_kwargs)
        pinned = torch.empty((args.batch, wl.N), dtype=torch.int8).pin_memory().numpy()
        gpu_s = [0.0]

        def decode(batch):
            t0 = time.perf_counter()
            hard, _, _ = dec.min_sum_observed(wl.table, batch.obs, wl.table_sum, batch.obs_sum, out=pinned[: batch.obs.shape[0]])
            gpu_s[0] += time.perf_counter() - t0
            ok = wl.success(batch, hard)
            T = hard.shape[0]
            return Counters(trials=T, successes=int(ok.sum()),
                            symbol_errors=int((hard[:, : wl.bvars] != batch.secret).sum()),
                            iterations=T * args.iters, edge_updates=T * args.iters * dec.E)

        t0 = time.perf_counter()
        if not args.host_trials:
            lo, hi = block_range(args.trials, rank, world)
            dec.simulate(min(hi - lo, 4096), args.seed, first_trial=lo, **wl.simulate_kwargs)  # untimed warm-up pass
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            r = dec.simulate(hi - lo, args.seed, first_trial=lo, **wl.simulate_kwargs)
            gpu_s[0] = time.perf_counter() - t0
            mine = Counters(trials=r["trials"], successes=r["successes"], symbol_errors=r["symbol_errors"],
                            iterations=r["trials"] * args.iters, edge_updates=r["trials"] * args.iters * dec.E)
            total = all_reduce_counters(mine, device=dev) if world > 1 else mine
        else:
            total = run_campaign(decode, lambda seed: wl.generate_observed(args.batch, seed=seed), n_batches, base_seed=int(acc * 1e4),
                                 rank=rank, world=world, device=dev)
        wall = time.perf_counter() - t0
        dt = torch.tensor([gpu_s[0], wall], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({
                "workload": f"kyber1024 sum_weight {args.sum_weight}, {args.iters} iterations, oracle accuracy {acc}",
                "trials_from": "host numpy + observations API" if args.host_trials else "device (scaldpc_qary_simulate)",
                "n_gpus": world, "trials": total.trials, "success_rate": total.successes / max(1, total.trials),
                "symbol_error_rate": total.symbol_errors / max(1, total.trials * wl.bvars),
                "decode_trials_per_s": total.trials / float(dt[0]),  # the decode / simulate calls alone (max over ranks)
                "campaign_trials_per_s": total.trials / float(dt[1]),  # the whole loop, trial generation included
                "edge_updates": total.edge_updates}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
