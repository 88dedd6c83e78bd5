"""GPU parity tests for the optimal-transport pairing against oracle/optimal_transport.py, plus
the reference's own closed-form tests (tests/test_utils/test_optimal_transport.py) run on the kernels."""
import numpy as np
import pytest
import torch

import oracle.optimal_transport as oot

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,m,D,reg", [(10, 20, 2, 1.0), (30, 20, 3, 0.1), (100, 200, 3, 0.1), (257, 257, 64, 0.5),
                                       (1024, 1024, 2, 0.1), (1, 7, 2, 1.0)])
def test_log_plan_matches_oracle(n, m, D, reg):
    import bayesflow_b200 as bf
    g = torch.Generator().manual_seed(n + m)
    x1, x2 = torch.randn(n, D, generator=g), torch.randn(m, D, generator=g)
    info = {}
    ref = oot.log_sinkhorn_plan(x1.double(), x2.double(), regularization=reg, atol=1e-5, max_steps=300, info=info)
    lp = bf.log_sinkhorn_plan(x1.cuda(), x2.cuda(), regularization=reg, atol=1e-5, max_steps=300).cpu()
    steps = bf.log_sinkhorn_plan.last_info["steps"]
    assert abs(steps - info["steps"]) <= max(2, info["steps"] // 10), (steps, info["steps"])
    # plans agree where there is mass; log-plan absolute error is bounded by the 1e-5 stopping tolerance scale
    np.testing.assert_allclose(torch.exp(lp).numpy(), torch.exp(ref).numpy(), rtol=2e-3, atol=1e-7)
    if steps < 300:  # converged: marginals are uniform (test_optimal_transport.py:215-224)
        assert torch.allclose(torch.logsumexp(lp, 0), torch.full((m,), -np.log(float(m))), atol=1e-3)
        assert torch.allclose(torch.logsumexp(lp, 1), torch.full((n,), -np.log(float(n))), atol=1e-3)


def test_reference_closed_form_tests_on_the_kernels():
    import bayesflow_b200 as bf
    # test_optimal_transport.py:249-263
    plan = torch.exp(bf.log_sinkhorn_plan(torch.ones(16).cuda(), torch.ones(64).cuda(), regularization=0.1)).cpu()
    np.testing.assert_allclose(plan.numpy(), np.full((16, 64), 1 / (16 * 64)), rtol=1e-4)
    # :98-116 a shuffled copy is matched back
    y = torch.randn(16, 2)
    p = torch.randperm(16)
    _, y2, _, a = bf.optimal_transport(y[p].cuda(), y.cuda(), regularization=0.01, seed=0, max_steps=10_000,
                                       method="log_sinkhorn", return_assignments=True)
    assert int((a.cpu() == p).sum()) >= 14 and torch.equal(y2.cpu(), y[a.cpu()])
    # :119-131 finite on well-separated inputs
    x, z = torch.randn(64, 4) * 10, torch.randn(64, 4) * 10 + 100
    ox, oy, _, a = bf.optimal_transport(x.cuda(), z.cuda(), regularization=0.1, seed=0, max_steps=1000,
                                        method="log_sinkhorn", return_assignments=True)
    assert torch.isfinite(ox).all() and torch.isfinite(oy).all() and int(a.min()) >= 0 and int(a.max()) < 64
    with pytest.raises(ValueError):
        bf.log_sinkhorn_plan(x.cuda(), z.cuda(), regularization=0.0)
    with pytest.raises(ValueError):
        bf.log_sinkhorn_plan(x.cuda(), z.cuda(), partial_factor=1.5)
    with pytest.raises(RuntimeError):
        bf.log_sinkhorn_plan(x, z)  # no CPU fallback


@pytest.mark.parametrize("n,D,Cd", [(64, 2, 5), (512, 2, 32)])
def test_conditional_assignments_match_oracle(n, D, Cd):
    """FlowMatching defaults (regularization 0.1, max_steps 100, condition_ratio 0.01): same conditional weight, and the
    Gumbel-max draw over the same Philox stream picks the same partner for (nearly) every row."""
    import bayesflow_b200 as bf
    g = torch.Generator().manual_seed(n)
    x0, x1, c = torch.randn(n, D, generator=g), torch.randn(n, D, generator=g), torch.randn(n, Cd, generator=g)
    kw = dict(regularization=0.1, max_steps=100, atol=1e-5, condition_ratio=0.01)
    info = {}
    ref = oot.log_sinkhorn_plan(x0, x1, c, info=info, **kw)
    a_ref = oot.gumbel_argmax(ref, seed=11)
    _, x1p, cp, a = bf.optimal_transport(x0.cuda(), x1.cuda(), conditions=c.cuda(), seed=11, method="log_sinkhorn",
                                         partial_factor=1.0, return_assignments=True, **kw)
    w = bf.log_sinkhorn_plan.last_info["w"]
    assert abs(w - info["w"]) <= 1e-3 * max(1.0, abs(info["w"])), (w, info["w"])
    agree = float((a.cpu() == a_ref).float().mean())
    assert agree >= 0.97, agree
    assert torch.equal(x1p.cpu(), x1[a.cpu()]) and torch.equal(cp.cpu(), c[a.cpu()])
    # transport cost never gets worse than the given pairing (test_optimal_transport.py:63-95); with condition_ratio 0.01
    # on a small batch the conditional weight keeps every row with its own conditions, i.e. the identity pairing
    assert float(((x0 - x1p.cpu()) ** 2).sum(1).mean()) <= float(((x0 - x1) ** 2).sum(1).mean())


def test_flow_matching_trains_with_optimal_transport():
    import bayesflow_b200 as bf
    net = bf.FlowMatching(use_optimal_transport=True, subnet_kwargs={"widths": (128, 128), "dropout": 0.0})
    net.build((None, 2), (None, 6))
    net.cuda()
    x, c = torch.randn(256, 2).cuda(), torch.randn(256, 6).cuda()
    out = net.compute_metrics(x, conditions=c, stage="training", ot_seed=3)
    out["loss"].backward()
    assert torch.isfinite(out["loss"]) and all(p.grad is not None and torch.isfinite(p.grad).all() for p in net.parameters())
    plain = bf.FlowMatching(subnet_kwargs={"widths": (128, 128), "dropout": 0.0})
    assert plain.optimal_transport_kwargs["method"] == "log_sinkhorn" and not plain.use_optimal_transport


@pytest.mark.parametrize("s", [0.3, 0.8])
def test_partial_ot_matches_oracle_and_leaves_unmatched_mass(s):
    """log_sinkhorn.py:110-122, 169-170 + ot_utils.py:37-92; reference test test_optimal_transport.py:289-327."""
    import bayesflow_b200 as bf
    g = torch.Generator().manual_seed(42)
    x, y = torch.randn(20, 2, generator=g), torch.randn(20, 2, generator=g)
    ref = torch.exp(oot.log_sinkhorn_plan(x.double(), y.double(), regularization=0.1, max_steps=10_000, partial_factor=s))
    plan = torch.exp(bf.log_sinkhorn_plan(x.cuda(), y.cuda(), regularization=0.1, max_steps=10_000, partial_factor=s)).cpu()
    assert plan.shape == (20, 20)
    np.testing.assert_allclose(plan.numpy(), ref.numpy(), rtol=5e-3, atol=1e-7)
    assert bool(((plan.sum(1) - s / 20).abs() < 0.05).all()) and bool(((plan.sum(0) - s / 20).abs() < 0.05).all())
    assert abs(float(plan.sum()) - s) < 1e-3
    # assignments stay inside the real columns
    _, _, _, a = bf.optimal_transport(x.cuda(), y.cuda(), method="log_sinkhorn", regularization=0.1, partial_factor=s, seed=2,
                                      return_assignments=True)
    a_ref = oot.gumbel_argmax(oot.log_sinkhorn_plan(x, y, regularization=0.1, partial_factor=s), seed=2)
    assert int(a.min()) >= 0 and int(a.max()) < 20 and float((a.cpu() == a_ref).float().mean()) >= 0.9


@pytest.mark.parametrize("n,m,D,reg,partial", [(10, 20, 2, 1.0, 1.0), (100, 200, 3, 0.5, 1.0), (257, 257, 8, 1.0, 1.0),
                                               (64, 64, 2, 1.0, 0.8), (1, 7, 2, 1.0, 1.0)])
def test_plain_sinkhorn_plan_matches_oracle(n, m, D, reg, partial):
    """the multiplicative Sinkhorn-Knopp (utils/optimal_transport/sinkhorn.py:65-186, `method="sinkhorn"`) on the kernels"""
    import bayesflow_b200 as bf
    from bayesflow_b200.optimal_transport import sinkhorn_plan
    g = torch.Generator().manual_seed(3 * n + m)
    x1, x2 = torch.randn(n, D, generator=g), torch.randn(m, D, generator=g)
    info = {}
    ref = oot.sinkhorn_plan(x1.double(), x2.double(), regularization=reg, atol=1e-5, max_steps=300, partial_factor=partial,
                            info=info)
    plan = sinkhorn_plan(x1.cuda(), x2.cuda(), regularization=reg, atol=1e-5, max_steps=300, partial_factor=partial).cpu()
    steps = bf.log_sinkhorn_plan.last_info["steps"]
    assert abs(steps - info["steps"]) <= max(2, info["steps"] // 10), (steps, info["steps"])
    np.testing.assert_allclose(plan.numpy(), ref.numpy(), rtol=2e-3, atol=1e-7)
    if steps < 300 and partial == 1.0:
        np.testing.assert_allclose(plan.sum(0).numpy(), np.full(m, 1.0 / m), rtol=2e-3)
    # the dispatcher's default method is "sinkhorn" (optimal_transport.py:12-72)
    y = torch.randn(16, 2)
    p = torch.randperm(16)
    _, y2, _, a = bf.optimal_transport(y[p].cuda(), y.cuda(), regularization=0.05, seed=0, max_steps=5000,
                                       return_assignments=True)
    assert int((a.cpu() == p).sum()) >= 14 and torch.equal(y2.cpu(), y[a.cpu()])
