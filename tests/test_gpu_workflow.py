"""Approximator-shaped end-to-end check on the CUDA path: what a BasicWorkflow does with an inference network
(workflows/basic_workflow.py:514-526 optimizer; approximators/continuous_approximator.py: standardize -> compute_metrics
-> optimizer step; approximators/helpers/samplers.py:13-155: sample num_samples draws per dataset) on a conjugate
Gaussian model whose posterior is known in closed form:

    theta ~ N(0, I_2),  x | theta ~ N(theta, 0.5^2 I_2)   =>   theta | x ~ N(x / 1.25, 0.2 I_2)   (std 0.4472)

Online training on fresh simulations with the CUDA-graph trainer, then amortised sampling for unseen datasets through
`Sampler` (conditions (N, C), never repeated in memory), posterior mean / std against the closed form."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model", ["flow_matching", "diffusion"])
def test_train_then_amortized_posterior_matches_the_closed_form(model):
    import bayesflow_b200 as bf
    from bayesflow_b200.training import Adam, Trainer, cosine_decay
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    torch.manual_seed(7)
    kw = {"widths": (128,) * 3, "dropout": 0.0}
    if model == "flow_matching":
        net = bf.FlowMatching(subnet_kwargs=kw, integrate_kwargs={"method": "rk45", "steps": 20})
        steps, lr = 1500, 2e-3
    else:
        net = bf.DiffusionModel(subnet_kwargs=kw, integrate_kwargs={"method": "rk45", "steps": 40})
        steps, lr = 2500, 2e-3
    net.build((None, 2), (None, 2))
    net.cuda()
    std = bf.Standardize()
    std.build((None, 2))
    std.cuda()
    tr = Trainer(net, Adam(net, cosine_decay(lr, steps), clipnorm=1.5), world_size=1, cuda_graph=(model == "flow_matching"),
                 seed=3)
    gen = torch.Generator(device="cuda").manual_seed(11)
    B = 4096
    first = last = None
    for it in range(steps):
        theta = torch.randn(B, 2, device="cuda", generator=gen)
        x = theta + 0.5 * torch.randn(B, 2, device="cuda", generator=gen)
        xs = std(x, stage="training") if it < 50 else std(x, stage="inference")   # running moments, then frozen
        loss = tr.train_step(theta, xs)
        if it == 0:
            first = float(loss)
    last = float(loss)
    assert math.isfinite(last) and last < first
    # amortised inference for 64 unseen datasets, 4000 draws each
    N, S = 64, 4000
    x_obs = torch.randn(N, 2, device="cuda", generator=gen) * math.sqrt(1.25)
    draws = bf.Sampler().sample(net, S, conditions=std(x_obs, stage="inference"), seed=5)
    assert draws.shape == (N, S, 2) and torch.isfinite(draws).all()
    post_mean, post_std = x_obs / 1.25, math.sqrt(0.2)
    err_mean = (draws.mean(1) - post_mean).abs()
    err_std = (draws.std(1) / post_std - 1).abs()
    # a short training run: the flow-matching net lands within 0.27 posterior standard deviations on EVERY dataset; the
    # diffusion net (EDM schedule, 2500 steps) is held to that on average and to 0.6 standard deviations at worst
    if model == "flow_matching":
        assert float(err_mean.max()) < 0.12 and float(err_std.max()) < 0.15, (float(err_mean.max()), float(err_std.max()))
    else:
        assert float(err_mean.mean()) < 0.12 and float(err_mean.max()) < 0.27, (float(err_mean.mean()), float(err_mean.max()))
        assert float(err_std.mean()) < 0.15 and float(err_std.max()) < 0.35, (float(err_std.mean()), float(err_std.max()))
    # the density path on the same network: log q(theta | x) near the closed-form log-density at the posterior mean
    if model == "flow_matching":
        lp = net.log_prob(post_mean[:16].contiguous(), conditions=std(x_obs, stage="inference")[:16].contiguous(),
                          method="rk45", steps=20)
        exact = -math.log(2 * math.pi * 0.2)
        assert float((lp - exact).abs().max()) < 0.35, (lp, exact)
