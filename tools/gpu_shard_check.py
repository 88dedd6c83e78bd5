"""2-GPU check (torchrun): sample_sharded draws == the same rows of a single-GPU Sampler call (shard-invariant Philox
noise), and data-parallel Standardize buffers == a single process seeing the whole batch.  Prints PASS/FAIL per check."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bayesflow_b200 as bf  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.manual_seed(0)
    D, Cd, N, S = 2, 32, 7, 300
    net = bf.FlowMatching(subnet_kwargs={"dropout": 0.0})
    net.build((None, D), (None, Cd))
    net.cuda()
    for p in net.parameters():
        dist.broadcast(p.data, src=0)
    cond = torch.randn(N, Cd, device="cuda")
    dist.broadcast(cond, src=0)
    full = bf.Sampler().sample(net, S, conditions=cond, seed=99, method="euler", steps=10)          # every rank: all rows
    mine = bf.sample_sharded(net, S, cond, seed=99, gather=True, method="euler", steps=10)           # sharded + gathered
    ok1 = torch.equal(full, mine)
    cm = bf.ConsistencyModel(100, subnet_kwargs={"dropout": 0.0, "widths": (128, 128)})
    cm.build((None, D), (None, Cd))
    cm.cuda()
    for p in cm.parameters():
        dist.broadcast(p.data, src=0)
    full_c = bf.Sampler().sample(cm, S, conditions=cond, seed=5, steps=3)
    mine_c = bf.sample_sharded(cm, S, cond, seed=5, gather=True, steps=3)
    ok2 = torch.equal(full_c, mine_c)
    # data-parallel Standardize: rank r sees rows [r*B, (r+1)*B) of the global batch
    B = 4096
    g = torch.Generator(device="cuda").manual_seed(3)
    xg = torch.randn(world * B, 5, device="cuda", generator=g) * 3 + 2
    dp = bf.Standardize(group=dist.group.WORLD)
    dp.build((None, 5))
    dp.cuda()
    dp(xg[rank * B:(rank + 1) * B], stage="training")
    one = bf.Standardize()
    one.build((None, 5))
    one.cuda()
    one(xg, stage="training")
    ok3 = (torch.allclose(dp._buf(dp.moving_mean, 0), one._buf(one.moving_mean, 0), rtol=1e-6, atol=1e-6)
           and torch.allclose(dp._buf(dp.moving_m2, 0), one._buf(one.moving_m2, 0), rtol=1e-5)
           and float(dp._buf(dp.count, 0)) == world * B)
    bufs = [torch.empty_like(dp._buf(dp.moving_m2, 0)) for _ in range(world)]
    dist.all_gather(bufs, dp._buf(dp.moving_m2, 0))
    ok4 = all(torch.equal(b, bufs[0]) for b in bufs)
    res = torch.tensor([ok1, ok2, ok3, ok4], device="cuda", dtype=torch.int32)
    dist.all_reduce(res, op=dist.ReduceOp.MIN)
    if rank == 0:
        names = ["sharded FlowMatching draws == single-GPU rows", "sharded ConsistencyModel draws == single-GPU rows",
                 "data-parallel Standardize == whole-batch Standardize", "Standardize buffers identical on all ranks"]
        for n, r in zip(names, res.tolist()):
            print(("PASS " if r else "FAIL ") + n)
    dist.destroy_process_group()
    sys.exit(0 if bool(res.min()) else 1)


if __name__ == "__main__":
    main()
