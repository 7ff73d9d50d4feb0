"""Multi-GPU sharding of Monte-Carlo decode campaigns.

Trials are independent, so the path shards with NO data-path collective: rank g of G takes a contiguous block of
the global batch indices, every batch is generated from a seed keyed by its GLOBAL index (results do not depend on
G), the graph is replicated, and the only exchange is one all-reduce(sum) of a handful of int64 counters
(torch.distributed: NCCL over NVLink on GPUs, gloo in the CPU tests).  The reference's counterpart is a bash loop of
independent processes (run-parallel-hqc-simulation.sh:10-45) and a ThreadPoolExecutor (simulate/decode.py:247-262).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

COUNTERS = ("trials", "successes", "symbol_errors", "iterations", "edge_updates")


def block_range(total: int, rank: int, world: int):
    """Contiguous, balanced partition of range(total): the first total % world ranks get one extra item."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    q, r = divmod(total, world)
    start = rank * q + min(rank, r)
    return start, start + q + (1 if rank < r else 0)


def batch_seed(base_seed: int, global_batch: int) -> int:
    """Seed of a batch, keyed by its global index only."""
    return (int(base_seed) * 1_000_003 + int(global_batch) * 7919 + 12345) % (2 ** 31 - 1)


@dataclass
class Counters:
    trials: int = 0
    successes: int = 0
    symbol_errors: int = 0
    iterations: int = 0
    edge_updates: int = 0
    extra: dict = field(default_factory=dict)

    def as_array(self):
        return np.array([getattr(self, k) for k in COUNTERS], dtype=np.int64)

    @classmethod
    def from_array(cls, a):
        return cls(**{k: int(v) for k, v in zip(COUNTERS, a)})

    def add(self, other: "Counters"):
        for k in COUNTERS:
            setattr(self, k, getattr(self, k) + getattr(other, k))
        return self


def all_reduce_counters(c: Counters, device=None) -> Counters:
    """Sum the counters over all ranks (no-op without an initialised process group)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return c
    t = torch.from_numpy(c.as_array())
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return Counters.from_array(t.cpu().numpy())


def run_campaign(decode_batch, generate, n_batches: int, base_seed: int = 0, rank: int = 0, world: int = 1,
                 device=None) -> Counters:
    """Decode `n_batches` batches sharded over `world` ranks and return the GLOBAL counters.

    generate(seed)      -> batch (any object)
    decode_batch(batch) -> Counters of that batch (runs on this rank's GPU)
    """
    lo, hi = block_range(n_batches, rank, world)
    local = Counters()
    for g in range(lo, hi):
        local.add(decode_batch(generate(batch_seed(base_seed, g))))
    return all_reduce_counters(local, device)


def kyber_decode_fn(decoder, workload):
    """decode_batch for a QaryDecoder + KyberWorkload pair (used by examples and the GPU tests)."""

    def fn(batch):
        hard, _, _ = decoder.min_sum_batch(batch.priors, batch.priors_sum, want_posterior=False)
        ok = workload.success(batch, hard)
        T = hard.shape[0]
        it = max(1, decoder.iterations)
        return Counters(trials=T, successes=int(ok.sum()),
                        symbol_errors=int((hard[:, : workload.bvars] != batch.secret).sum()),
                        iterations=T * it, edge_updates=T * it * decoder.E)

    return fn
