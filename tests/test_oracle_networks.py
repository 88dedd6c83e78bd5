"""Oracle self-checks restating the reference's own invariants for the networks:
  * FiLM with zero init == identity           /root/reference/tests/test_networks/test_embeddings.py:85-104
  * Fourier embedding shapes                  /root/reference/tests/test_networks/test_embeddings.py:8-41
  * FM cycle consistency forward∘inverse = id /root/reference/tests/test_networks/test_flow_matching/test_flow_matching.py:99-108
  * density vs explicit Jacobian              /root/reference/tests/test_networks/test_inference_networks.py:114-150
  * parameter / FLOP counts                   SURVEY.md §8d table
"""
import math

import numpy as np
import pytest
import torch

import oracle
from oracle.networks import OracleTimeMLP


def test_param_and_flop_counts():
    assert sum(v.numel() for v in oracle.init_params(2, 32).values()) == 516_114
    assert sum(v.numel() for v in oracle.init_params(2, 2).values()) == 508_434
    assert sum(v.numel() for v in oracle.init_params(4, 20).values()) == 514_068
    assert sum(v.numel() for v in oracle.init_params(64, 32, (1024,) * 8).values()) == 10_955_856
    assert oracle.count_flops_per_eval(2, 32, (256,) * 5) == 1_020_416
    assert oracle.count_flops_per_eval(2, 2, (256,) * 5) == 1_005_056
    assert oracle.count_flops_per_eval(4, 20, (256,) * 5) == 1_016_320
    assert oracle.count_flops_per_eval(64, 32, (1024,) * 8) == 21_839_872


def test_film_zero_init_is_identity():
    p = oracle.init_params(3, None, (8, 8))
    for k in list(p):
        if "film" in k:
            p[k] = torch.zeros_like(p[k])
    net = OracleTimeMLP(p, widths=(8, 8), norm=None, residual=False)
    x = torch.randn(5, 8)
    t_emb = net.time_emb(torch.rand(5, 1))
    assert t_emb.shape == (5, 33)
    out = net.block(0, x, t_emb)
    ref = net.act(x @ p["subnet.blocks.0.dense.kernel"] + p["subnet.blocks.0.dense.bias"])
    torch.testing.assert_close(out, ref)


def test_fourier_embedding_values():
    p = oracle.init_params(2, None, (8,))
    net = OracleTimeMLP(p, widths=(8,))
    t = torch.tensor([[0.25]])
    e = net.time_emb(t)
    w = p["subnet.time_emb.frequencies"]
    assert e.shape == (1, 33)
    np.testing.assert_allclose(e[0, 0], 0.25)
    np.testing.assert_allclose(e[0, 1:17], np.sin(0.25 * w.numpy() * 2 * math.pi * 30.0), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(e[0, 17:], np.cos(0.25 * w.numpy() * 2 * math.pi * 30.0), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("C", [None, 3])
def test_cycle_consistency(C):
    D = 2
    p = oracle.perturb_params(oracle.init_params(D, C, (8, 8)))
    fm = oracle.OracleFlowMatching(p, D, subnet_kwargs={"widths": (8, 8), "dropout": 0.0})
    x = torch.randn(2, D)
    c = torch.randn(2, C) if C else None
    z = fm.forward(x, c)
    x2 = fm.inverse(z, c)
    torch.testing.assert_close(x, x2, atol=1e-3, rtol=1e-3)
    z1, lp1 = fm.forward(x, c, density=True)
    x3, lp2 = fm.inverse(z1, c, density=True)
    torch.testing.assert_close(lp1, lp2, atol=1e-3, rtol=1e-3)


def test_density_numerically():
    D, C = 2, 3
    p = oracle.perturb_params(oracle.init_params(D, C, (8, 8)))
    fm = oracle.OracleFlowMatching(p, D, subnet_kwargs={"widths": (8, 8), "dropout": 0.0},
                                   integrate_kwargs={"method": "rk45", "steps": 50})
    x = torch.randn(3, D)
    c = torch.randn(3, C)
    z, lp = fm.forward(x, c, density=True)
    jac = torch.autograd.functional.jacobian(lambda a: fm.forward(a, c), x)  # (3,D,3,D)
    J = torch.stack([jac[i, :, i, :] for i in range(3)])
    expected = fm.base_distribution.log_prob(z) + torch.log(torch.abs(torch.linalg.det(J)))
    torch.testing.assert_close(lp, expected, atol=1e-3, rtol=1e-3)


def test_dm_schedule_scalars():
    ns = oracle.EDMNoiseSchedule()
    np.testing.assert_allclose(float(ns.log_snr_min), -2 * math.log(80.0), rtol=1e-6)
    np.testing.assert_allclose(float(ns.log_snr_max), -2 * math.log(1e-4), rtol=1e-6)
    t = torch.tensor([0.0, 0.3, 1.0])
    lam = ns.get_log_snr(t, training=False)
    np.testing.assert_allclose(lam[0], ns.log_snr_max, rtol=1e-5)
    np.testing.assert_allclose(lam[2], ns.log_snr_min, rtol=1e-5)
    np.testing.assert_allclose(ns.get_t_from_log_snr(lam, False), t, atol=1e-5)
    a, s = ns.get_alpha_sigma(lam)
    np.testing.assert_allclose(a ** 2 + s ** 2, np.ones(3), rtol=1e-6)
    cs = oracle.CosineNoiseSchedule()
    lam = cs.get_log_snr(t, training=False)
    np.testing.assert_allclose(lam[[0, 2]], [15.0, -15.0], rtol=1e-3)


def test_oracle_density_matches_the_numerical_jacobian():
    """Pin of the oracle's density path, as the reference pins its own (tests/test_networks/test_inference_networks.py:
    114-150): log_density from forward(density=True) == log N(z) + log|det dz/dx| with the Jacobian of the WHOLE
    integrated map taken numerically (autograd), FlowMatching and DiffusionModel."""
    import oracle
    torch.manual_seed(1)
    D, C = 2, 3
    p = oracle.perturb_params(oracle.init_params(D, C, (32, 32)))
    x, c = torch.randn(3, D), torch.randn(3, C)
    for orc in (oracle.OracleFlowMatching(p, D, subnet_kwargs={"widths": (32, 32), "dropout": 0.0}),
                oracle.OracleDiffusionModel(p, D, subnet_kwargs={"widths": (32, 32), "dropout": 0.0})):
        z, lp = orc.forward(x, c, density=True, method="rk45", steps=40)
        for i in range(x.shape[0]):
            J = torch.autograd.functional.jacobian(
                lambda xi: orc.forward(xi[None], c[i:i + 1], method="rk45", steps=40)[0], x[i])
            ref = orc.base_distribution.log_prob(z[i]) + torch.log(torch.abs(torch.det(J)))
            assert abs(float(lp[i] - ref)) < 2e-3 * max(1.0, abs(float(ref))), (i, float(lp[i]), float(ref))
