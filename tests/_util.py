"""Shared helpers for the parity tests: build a bayesflow_b200 network and the CPU oracle on IDENTICAL weights."""
import torch

import oracle

_PKEYS = ("merge", "film_use_gamma", "norm", "residual", "time_embedding_dim")


def make_pair(D, C, widths=(256,) * 5, dm=False, net_kwargs=None, model_kwargs=None, params_fn=None, **sk):
    """-> (bayesflow_b200 network on cuda:0, oracle on CPU), same perturbed weights.
    `params_fn(params) -> params` edits the shared weights first (range stress tests)."""
    import bayesflow_b200 as bf

    params = oracle.perturb_params(oracle.init_params(D, C, widths, **{k: v for k, v in sk.items() if k in _PKEYS}))
    if params_fn is not None:
        params = params_fn(params)
    kw = {"dropout": 0.0, "widths": widths} | sk
    mk = model_kwargs or {}
    if dm:
        net = bf.DiffusionModel(subnet_kwargs=kw, **mk, **(net_kwargs or {}))
        orc = oracle.OracleDiffusionModel(params, D, subnet_kwargs=kw, **mk)
    else:
        net = bf.FlowMatching(subnet_kwargs=kw, **mk, **(net_kwargs or {}))
        orc = oracle.OracleFlowMatching(params, D, subnet_kwargs=kw, **mk)
    net.build((None, D), (None, C) if C else None)
    res = net.load_state_dict(params, strict=False)
    assert not res.unexpected_keys and not [k for k in res.missing_keys if "base_distribution" not in k], res
    net.cuda()
    return net, orc


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def grouped_conditions(B, C, rows_per_group, seed=2345):
    g = torch.Generator().manual_seed(seed)
    G = (B + rows_per_group - 1) // rows_per_group
    return torch.randn(G, C, generator=g).repeat_interleave(rows_per_group, dim=0)[:B].contiguous()


def as_float64(orc):
    """The same oracle evaluated in float64 (weights up-cast): used where the reference's own fp32 algebra is
    ill-conditioned (e.g. prediction_type='score' at small t computes (z + sigma^2 pred - z)/sigma^2 with
    sigma^2 ~ 1e-7, diffusion_model.py:402-404, 489) so that the comparison measures the kernel, not fp32 cancellation."""
    import copy
    from collections import OrderedDict
    p64 = OrderedDict((k, v.double()) for k, v in orc.p.items())
    o = copy.copy(orc)
    o.p = p64
    o.subnet = copy.copy(orc.subnet)
    o.subnet.p = p64
    return o
