"""world_size-2 gloo tests (CPU) of the host-side multi-GPU logic: dataset sharding, ragged gather, and the 4-float
MAX all-reduce that makes adaptive step control global across shards."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bayesflow_b200.parallel import dataset_shard, gather_rows


def test_dataset_shard_partitions():
    for n in (0, 1, 7, 1000, 1001):
        for w in (1, 2, 3, 8):
            ranges = [dataset_shard(n, r, w) for r in range(w)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        dataset_shard(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = dataset_shard(n, rank, world)
        # "draws" of dataset i are filled with i, so the gathered tensor must be arange
        local = torch.arange(lo, hi, dtype=torch.float32).reshape(-1, 1, 1).expand(hi - lo, 3, 2).contiguous()
        full = gather_rows(local, n)
        assert full.shape == (n, 3, 2)
        assert torch.equal(full[:, 0, 0], torch.arange(n, dtype=torch.float32))
        # adaptive accumulators {err_max, max|state|, nan flag, non-nan flag}: MAX over ranks == global max over rows
        acc = torch.tensor([0.1 * (rank + 1), 3.0 - rank, float(rank == 1), 1.0])
        dist.all_reduce(acc, op=dist.ReduceOp.MAX)
        assert torch.allclose(acc, torch.tensor([0.1 * world, 3.0, 1.0, 1.0]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [5, 8])
def test_gather_and_step_control_allreduce_gloo_world2(n):
    mp.spawn(_worker, args=(2, _free_port(), n), nprocs=2, join=True)
