"""N-GPU checks (torchrun, N >= 2) of the §8e rows on hardware.  Prints PASS/FAIL per check; exit 1 on any failure.

  1. sharded ADAPTIVE sampling with the reference's GLOBAL step control (utils/integrate.py:117-118, 353-358): draws
     bit-identical to the single-GPU call, same accepted / rejected counts - fused engine (cfg2 net) and layer-wise
     engine (W = 512 net), tsit5 and rk45;
  2. log_prob_sharded == single-GPU log_prob (inference_network.py:127-129), fixed and adaptive steps;
  3. data-parallel training == the single-GPU whole-batch step: FlowMatching eager, FlowMatching CUDA-graph replay
     (with dropout), DiffusionModel (global low-discrepancy t, diffusion_model.py:159-161): post-step parameters equal to
     fp32 round-off (the gradient sums use fp32 atomics)."""
import copy
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bayesflow_b200 as bf  # noqa: E402
from bayesflow_b200.training import Adam, Trainer  # noqa: E402

RESULTS = []


def record(name, ok, detail=""):
    RESULTS.append((name, bool(ok), detail))


def sync_params(net):
    for p in net.parameters():
        dist.broadcast(p.data, src=0)


def make(cls, D, C, widths, **kw):
    net = cls(subnet_kwargs={"dropout": 0.0, "widths": widths} | kw.pop("subnet_kwargs", {}), **kw)
    net.build((None, D), (None, C))
    net.cuda()
    sync_params(net)
    return net


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.manual_seed(0)
    D, Cd = 2, 32

    # ---- 1. adaptive sampling, global step control ---------------------------------------------------------------
    for label, widths, N, S in (("fused", (256,) * 5, 2 * world + 1, 200), ("layerwise", (512,) * 2, 2 * world + 1, 64)):
        net = make(bf.FlowMatching, D, Cd, widths)
        cond = torch.randn(N, Cd, device="cuda")
        dist.broadcast(cond, src=0)
        for method, kw in (("tsit5", {}), ("rk45", {"min_steps": 3, "max_steps": 40})):
            full = bf.Sampler().sample(net, S, conditions=cond, seed=7, method=method, steps="adaptive", **kw)
            info_full = dict(net.last_integration_info)
            mine = bf.sample_sharded(net, S, cond, seed=7, gather=True, method=method, steps="adaptive", **kw)
            info_mine = dict(net.last_integration_info)
            same_counts = (info_full["steps"], info_full["rejected"]) == (info_mine["steps"], info_mine["rejected"])
            record(f"{label} {method} adaptive: sharded draws bit-identical, same step counts",
                   torch.equal(full, mine) and same_counts,
                   f"steps {info_full['steps']}/{info_mine['steps']} rejected {info_full['rejected']}/{info_mine['rejected']} "
                   f"maxdiff {float((full - mine).abs().max()):.2e}")

    # ---- 2. log_prob sharded -----------------------------------------------------------------------------------------
    net = make(bf.FlowMatching, 3, 5, (64,) * 2)
    N, S = 2 * world + 1, 16
    cond = torch.randn(N, 5, device="cuda")
    x = torch.randn(N, S, 3, device="cuda")
    dist.broadcast(cond, src=0)
    dist.broadcast(x, src=0)
    for kw in ({"method": "euler", "steps": 8}, {"method": "tsit5", "steps": "adaptive", "min_steps": 2, "max_steps": 12}):
        full = net.log_prob(x.reshape(N * S, 3), conditions=cond.repeat_interleave(S, dim=0), **kw).reshape(N, S)
        mine = bf.log_prob_sharded(net, x, cond, gather=True, **kw)
        record(f"log_prob_sharded == single-GPU log_prob ({kw['method']}, steps={kw['steps']})",
               torch.allclose(full, mine, rtol=1e-5, atol=1e-5), f"maxdiff {float((full - mine).abs().max()):.2e}")

    # ---- 3. data-parallel training == whole-batch training ---------------------------------------------------------------
    def grads_dp_vs_single(cls, label, dropout, **kw):
        """all-reduced shard gradients == whole-batch gradients (same global noise, same dropout masks)"""
        torch.manual_seed(11)
        net = make(cls, D, Cd, (128,) * 3, subnet_kwargs={"dropout": dropout}, **kw)
        B = 512
        g = torch.Generator(device="cuda").manual_seed(5)
        xg = torch.randn(world * B, D, device="cuda", generator=g)
        cg = torch.randn(world * B, Cd, device="cuda", generator=g)
        dist.broadcast(xg, src=0)
        dist.broadcast(cg, src=0)
        from bayesflow_b200.training import train_engine
        eng = train_engine(net)
        l_one = net._loss_and_grads(xg, cg, None, stage="training", seed=77, world=1, noise_seed=99, row_offset=0)
        g_one, l_one = eng.grad.clone(), float(l_one)
        l_dp = net._loss_and_grads(xg[rank * B:(rank + 1) * B].contiguous(), cg[rank * B:(rank + 1) * B].contiguous(), None,
                                   stage="training", seed=77, world=world, noise_seed=99, row_offset=rank * B,
                                   loss_into_slot=True)
        dist.all_reduce(eng.grad_ext, op=dist.ReduceOp.SUM)
        worst = 0.0
        for i, p in enumerate(eng.params):
            a, b = eng.grad_of(p), g_one[eng.offsets[i]:eng.offsets[i] + p.numel()].view(p.shape)
            worst = max(worst, float((a - b).norm() / b.norm().clamp_min(1e-20)))
        record(f"DP-{world} {label}: all-reduced gradients == whole-batch gradients",
               worst < 2e-5 and abs(float(eng.loss_slot) - l_one) <= 2e-6 * abs(l_one),
               f"worst rel-L2 {worst:.2e}; loss {float(eng.loss_slot):.7f} vs {l_one:.7f}")

    grads_dp_vs_single(bf.FlowMatching, "FlowMatching (dropout 0)", 0.0)
    grads_dp_vs_single(bf.FlowMatching, "FlowMatching (dropout 0.05, masks by global row)", 0.05)
    grads_dp_vs_single(bf.DiffusionModel, "DiffusionModel (global low-discrepancy t)", 0.0)

    def trainer_dp_vs_single(label, graph, dropout, steps=4):
        """Trainer end to end (noise, all-reduce, per-variable clipnorm Adam): loss trajectory of DP-N == whole batch"""
        torch.manual_seed(11)
        net = make(bf.FlowMatching, D, Cd, (128,) * 3, subnet_kwargs={"dropout": dropout})
        ref = copy.deepcopy(net)
        B = 512
        g = torch.Generator(device="cuda").manual_seed(5)
        xg = torch.randn(steps, world * B, D, device="cuda", generator=g)
        cg = torch.randn(steps, world * B, Cd, device="cuda", generator=g)
        dist.broadcast(xg, src=0)
        dist.broadcast(cg, src=0)
        tr = Trainer(net, Adam(net, 1e-3, clipnorm=1.5), world_size=world, cuda_graph=graph, seed=123)
        one = Trainer(ref, Adam(ref, 1e-3, clipnorm=1.5), world_size=1, cuda_graph=False, seed=123)
        losses = []
        for s in range(steps):
            l_dp = tr.train_step(xg[s, rank * B:(rank + 1) * B].contiguous(), cg[s, rank * B:(rank + 1) * B].contiguous())
            l_one = one.train_step(xg[s], cg[s])
            losses.append((float(l_dp), float(l_one)))
        # parameters: Adam normalises every gradient entry, so fp32 round-off in near-zero entries moves single weights by
        # O(lr); the trajectories agree when the losses do and the bulk of the weights does
        worst = 0.0
        for p, q in zip(net.parameters(), ref.parameters()):
            worst = max(worst, float((p - q).norm() / q.norm().clamp_min(1e-12)))
        loss_ok = all(abs(a - b) <= 1e-4 * abs(b) for a, b in losses)
        record(f"DP-{world} Trainer {label} == single-GPU whole batch over {steps} steps", worst < 2e-3 and loss_ok,
               f"worst rel-L2 param diff {worst:.2e}; losses " + ", ".join(f"{a:.5f}/{b:.5f}" for a, b in losses))

    trainer_dp_vs_single("FlowMatching eager, dropout 0.05", False, 0.05)
    trainer_dp_vs_single("FlowMatching CUDA-graph replay, dropout 0.05", True, 0.05)

    ok = torch.tensor([int(all(r[1] for r in RESULTS))], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if rank == 0:
        for name, r, detail in RESULTS:
            print(("PASS " if r else "FAIL ") + name + ("   [" + detail + "]" if detail else ""))
    dist.destroy_process_group()
    sys.exit(0 if int(ok) else 1)


if __name__ == "__main__":
    main()
