"""Multi-GPU layout of the sampling path: one process per GPU (torchrun), rows sharded BY DATASET, weights replicated,
no data-path collective.  The only exchanges are (i) an optional gather of the finished draws and (ii),
for adaptive solvers with the reference's GLOBAL step control (integrate.py:117-118, 353-358), a MAX all-reduce of 4
floats per trial step, which `networks._Run.run_adaptive` issues when `net.step_control_group` is set."""
from __future__ import annotations

import torch
import torch.distributed as dist


def dataset_shard(num_datasets: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous range [lo, hi) of datasets owned by `rank` (dataset-major order as samplers.py:59-61); sizes differ by
    at most one, earlier ranks take the remainder."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size: {rank}/{world_size}")
    base, rem = divmod(num_datasets, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def sample_sharded(net, num_samples: int, conditions: torch.Tensor, *, group=None, gather: bool = False,
                   global_step_control: bool = True, seed=None, **kwargs):
    """`Sampler.sample` (approximators/helpers/samplers.py:71-148) for one rank: draws for this rank's datasets only,
    shape (n_local, num_samples, D); with gather=True every rank returns all (N, num_samples, D) draws.

    conditions: (N, C) for ALL datasets (every rank passes the same tensor) — only the local slice is used."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    N = conditions.shape[0]
    lo, hi = dataset_shard(N, rank, world)
    local = conditions[lo:hi].contiguous()
    n_local = hi - lo
    prev = getattr(net, "step_control_group", None)
    if world > 1 and global_step_control:
        net.step_control_group = group if group is not None else dist.group.WORLD
    # the noise stream is a function of (seed, GLOBAL row): every rank passes the same seed and its first global row, so the
    # sharded draws are exactly the rows a single-GPU call with this seed produces (all ranks take part in the broadcast)
    if seed is None:
        seed_t = torch.randint(0, 2 ** 62, (1,), dtype=torch.int64)
        if world > 1:
            seed_t = seed_t.to(conditions.device) if dist.get_backend(group) == "nccl" else seed_t
            dist.broadcast(seed_t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        seed = int(seed_t.item())
    try:
        if n_local > 0:
            flat = net.sample((n_local * num_samples,), conditions=local, conditions_repeats=num_samples, seed=int(seed),
                              row_offset=lo * num_samples, **kwargs)
            out = flat.reshape(n_local, num_samples, -1)
        else:
            out = conditions.new_zeros((0, num_samples, net._D))
    finally:
        net.step_control_group = prev
    if not gather or world == 1:
        return out
    return gather_rows(out, N, group=group)


def log_prob_sharded(net, samples: torch.Tensor, conditions: torch.Tensor | None = None, *, group=None, gather: bool = False,
                     global_step_control: bool = True, **kwargs):
    """`InferenceNetwork.log_prob` (networks/inference/inference_network.py:127-129) for one rank.

    samples: (N, S, D) draws of N datasets (or (N, D)); conditions: (N, C) for ALL datasets (every rank passes the same
    tensors) - only this rank's datasets are evaluated.  Rows are independent (utils/integrate.py:31-44), except that an
    adaptive solver's step controller takes maxima over ALL rows (:117-118, 353-358): with global_step_control the two
    maxima are MAX-all-reduced per trial step, so every rank walks the step sequence of the single-GPU call.
    Returns (n_local, S) log-densities, or all (N, S) with gather=True."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    squeeze = samples.ndim == 2
    if squeeze:
        samples = samples.unsqueeze(1)
    N, S, D = samples.shape
    lo, hi = dataset_shard(N, rank, world)
    n_local = hi - lo
    prev = getattr(net, "step_control_group", None)
    if world > 1 and global_step_control:
        net.step_control_group = group if group is not None else dist.group.WORLD
    try:
        if n_local > 0:
            x = samples[lo:hi].reshape(n_local * S, D).contiguous()
            c = None if conditions is None else conditions[lo:hi].repeat_interleave(S, dim=0).contiguous()
            out = net.log_prob(x, conditions=c, **kwargs).reshape(n_local, S)
        else:
            out = samples.new_zeros((0, S))
            if world > 1 and global_step_control and kwargs.get("steps", net.integrate_kwargs.get("steps")) in ("adaptive", "dynamic"):
                raise ValueError("log_prob_sharded with global step control needs at least one dataset per rank")
    finally:
        net.step_control_group = prev
    if gather and world > 1:
        out = gather_rows(out, N, group=group)
    return out.squeeze(1) if squeeze else out


def gather_rows(local: torch.Tensor, num_datasets: int, *, group=None) -> torch.Tensor:
    """all_gather of ragged dataset shards (sizes from dataset_shard) into one (N, ...) tensor on every rank."""
    world = dist.get_world_size(group)
    sizes = [dataset_shard(num_datasets, r, world) for r in range(world)]
    biggest = max(hi - lo for lo, hi in sizes)
    pad = local.new_zeros((biggest,) + tuple(local.shape[1:]))
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)
