"""The N>1 path on CPU: world_size-2 gloo process group, sharded campaign, counter all-reduce.  The decode itself is
replaced by a deterministic function of the batch (no GPU here); what is under test is the sharding logic of
sca_ldpc_b200/shard.py: contiguous balanced ranges, seeds keyed by the GLOBAL batch index, results independent of
the world size, and the single collective of the path."""
import os
import socket

import numpy as np
import pytest

from sca_ldpc_b200 import shard


def test_block_range_partitions():
    for total in (0, 1, 7, 10, 100_000):
        for world in (1, 2, 3, 8):
            parts = [shard.block_range(total, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == total
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.block_range(10, 2, 2)


def _fake_generate(seed):
    return np.random.RandomState(seed).randint(0, 100, size=64)


def _fake_decode(batch):
    return shard.Counters(trials=len(batch), successes=int((batch < 90).sum()), symbol_errors=int(batch.sum()),
                          iterations=10 * len(batch), edge_updates=10 * len(batch) * 1792)


def _worker(rank, world, port, n_batches, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c = shard.run_campaign(_fake_decode, _fake_generate, n_batches, base_seed=5, rank=rank, world=world)
        q.put((rank, c.as_array().tolist()))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_matches_single_rank():
    import torch.multiprocessing as mp

    n_batches = 7  # odd: ranks get 4 and 3 batches
    single = shard.run_campaign(_fake_decode, _fake_generate, n_batches, base_seed=5)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_batches, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert results[0] == results[1] == single.as_array().tolist()
    assert single.trials == 7 * 64
