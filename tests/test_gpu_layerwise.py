"""GPU parity of the layer-wise sampler (bfk_sample_layerwise: networks with W > 256 or D > 8, BASELINE.json configs[3])
against the CPU oracle on identical weights and inputs: velocity and fixed-noise draws within 2e-3 relative L2."""
import pytest
import torch

from _util import make_pair, rel_l2, grouped_conditions

pytestmark = pytest.mark.gpu
TOL = 2e-3


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    import bayesflow_b200._lib as L
    L.lib()


@pytest.mark.parametrize("D,C,widths,B", [
    (64, 32, (1024,) * 8, 300),      # cfg4 network
    (64, 32, (1024,) * 2, 1),
    (16, 5, (512,) * 2, 777),
    (2, 6, (320,) * 3, 260),         # narrow state, width just above the fused limit
    (12, None, (128,) * 2, 129),     # D > 8 alone forces the layer-wise path; unconditional
])
def test_velocity(D, C, widths, B):
    net, orc = make_pair(D, C, widths)
    g = torch.Generator().manual_seed(3456)
    z = torch.randn(B, D, generator=g)
    c = grouped_conditions(B, C, 100) if C else None
    for t in (0.0, 0.37, 1.0):
        out = net.velocity(z.cuda(), t, conditions=c.cuda() if C else None)
        assert net._engine.layerwise
        assert rel_l2(out, orc.velocity(z, t, c)) < TOL, (t, rel_l2(out, orc.velocity(z, t, c)))


@pytest.mark.parametrize("method,steps", [("euler", 20), ("rk45", 4), ("tsit5", 4)])
def test_fixed_noise_draws_cfg4_shape(method, steps):
    net, orc = make_pair(64, 32, (1024,) * 3)
    z = torch.randn(500, 64)
    c = grouped_conditions(500, 32, 100)
    ref = orc.inverse(z, c, method=method, steps=steps)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps=steps)
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)
    assert net.last_integration_info["path"] == "layerwise"


def test_many_chunks_and_repeats_keyword(monkeypatch):
    """rows split into several L2 chunks that do not align with the datasets; conditions passed as (N, C) + repeats"""
    monkeypatch.setenv("BFK_LW_CHUNK_ROWS", "256")
    net, orc = make_pair(16, 5, (512,) * 2)
    N, S = 7, 130
    z = torch.randn(N * S, 16)
    c = torch.randn(N, 5)
    ref = orc.inverse(z, c.repeat_interleave(S, dim=0), method="euler", steps=10)
    out = net(z.cuda(), conditions=c.cuda(), conditions_repeats=S, inverse=True, method="euler", steps=10)
    assert rel_l2(out, ref) < TOL
    assert net.last_integration_info["chunk_rows"] == 256


@pytest.mark.parametrize("sk", [{"merge": "add"}, {"activation": "silu"}, {"norm": None}, {"residual": False}])
def test_options(sk):
    net, orc = make_pair(10, 6, (128,) * 2, **sk)
    z, c = torch.randn(300, 10), grouped_conditions(300, 6, 50)
    ref = orc.inverse(z, c, method="rk45", steps=3)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="rk45", steps=3)
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


def test_forward_direction_scheduled_and_target_mask():
    net, orc = make_pair(12, 4, (128,) * 2)
    x, c = torch.randn(200, 12), grouped_conditions(200, 4, 50)
    grid = [1.0, 0.8, 0.5, 0.1, 0.0]
    ref = orc.forward(x, c, method="euler", steps=grid)
    out = net(x.cuda(), conditions=c.cuda(), inverse=False, method="euler", steps=grid)
    assert rel_l2(out, ref) < TOL
    tm = torch.ones(12)
    tm[3:6] = 0
    out = net.velocity(x.cuda(), 0.4, conditions=c.cuda(), target_mask=tm.cuda())
    assert rel_l2(out, orc.velocity(x, 0.4, c) * tm) < TOL


def test_diffusion_model_probability_flow():
    net, orc = make_pair(12, 20, (512,) * 2, dm=True)
    z, c = torch.randn(300, 12), grouped_conditions(300, 20, 100)
    for t in (0.9, 0.5, 0.1):
        assert rel_l2(net.velocity(z.cuda(), t, conditions=c.cuda()), orc.velocity(z, t, conditions=c)) < TOL
    ref = orc.inverse(z, c, method="euler", steps=30)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=30)
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


@pytest.mark.parametrize("method", ["tsit5", "rk45"])
def test_adaptive_on_the_layerwise_path(method):
    """integrate_adaptive (integrate.py:291-406) with the reference's global controller; default FlowMatching sampling
    (tsit5, adaptive) must work for wide / high-dimensional networks too."""
    net, orc = make_pair(12, 5, (512,) * 2)
    z, c = torch.randn(260, 12), grouped_conditions(260, 5, 65)
    info = {}
    ref = orc.inverse(z, c, method=method, steps="adaptive", info=info)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps="adaptive")
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)
    got = net.last_integration_info
    assert abs(got["steps"] - info["steps"]) <= max(2, 0.1 * info["steps"]), (got, info)
    if method == "tsit5":  # the network default (defaults.py:52) takes the same route
        out2 = net(z.cuda(), conditions=c.cuda(), inverse=True)
        assert torch.equal(out2, out)  # bit-reproducible: the LayerNorm statistics are per-panel partials, no atomics
    with pytest.raises(ValueError):
        net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps="adaptive")
