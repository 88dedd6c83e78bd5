"""GPU parity tests of DiffusionModel's reverse-SDE samplers (diffusion_model.py:830-856; utils/integrate.py:495-743,
964-1164, 1254-1451) against the CPU oracle, with the SAME Gaussian increments fed to both (the role z_history plays in
the reference), so the comparison is exact up to the kernels' fp16-operand rounding: 2e-3 relative L2 on the draws."""
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


def _normals(n, shape, seed=9):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(shape, generator=g) for _ in range(n)]


@pytest.mark.parametrize("method,steps", [("euler_maruyama", 30), ("two_step_adaptive", 20), ("two_step_adaptive", "adaptive"),
                                          ("euler_maruyama", "adaptive")])
@pytest.mark.parametrize("sched", ["edm", "cosine"])
def test_fixed_noise_sde_draws(method, steps, sched):
    net, orc = make_pair(3, 6, (128,) * 2, dm=True, model_kwargs={"noise_schedule": sched})
    B = 400
    z, c = torch.randn(B, 3), grouped_conditions(B, 6, 100)
    Z = _normals(1100, (B, 3))
    info = {}
    ref = orc.inverse(z, c, method=method, steps=steps, noise_fn=lambda i, s: {"xz": Z[i]}, info=info)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps=steps, noise_fn=lambda i, shape: Z[i])
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)
    if steps == "adaptive":
        assert net.last_integration_info["steps"] == info["steps"]


def test_default_sampler_is_the_reverse_sde_and_is_seeded():
    """DiffusionModel() samples with two_step_adaptive by default (defaults.py:50); same seed -> same draws; the draws'
    first two moments agree with the oracle's own SDE run on different normals (statistical parity)."""
    net, orc = make_pair(2, 4, (128,) * 2, dm=True)
    B = 4000
    c = grouped_conditions(B, 4, 2000)
    a = net.sample((B,), conditions=c.cuda(), seed=5)
    b = net.sample((B,), conditions=c.cuda(), seed=5)
    assert net.last_integration_info["method"] == "two_step_adaptive"
    assert torch.equal(a, b)
    ref = orc.inverse(torch.randn(B, 2), c, generator=torch.Generator().manual_seed(1))
    for g in range(2):
        sl = slice(g * 2000, (g + 1) * 2000)
        assert (a[sl].mean(0).cpu() - ref[sl].mean(0)).abs().max() < 0.15 * ref[sl].std(0).max() + 0.05
        assert ((a[sl].std(0).cpu() / ref[sl].std(0)) - 1).abs().max() < 0.15


@pytest.mark.parametrize("method,steps,extra", [
    ("sea", 25, {}), ("shark", 25, {}), ("langevin", 40, {}), ("langevin", 15, {"step_size_factor": 0.02}),
    ("euler_maruyama", 20, {"corrector_steps": 2}), ("shark", 12, {"corrector_steps": 1, "step_size_factor": 0.05}),
    ("langevin", 12, {"corrector_steps": 1}), ("two_step_adaptive", "adaptive", {"corrector_steps": 1}),
])
@pytest.mark.parametrize("sched", ["edm", "cosine"])
def test_sea_shark_langevin_and_corrector(method, steps, extra, sched):
    """utils/integrate.py:746-961, 1167-1251: same normals in the reference's draw order (dW, aux, corrector draws)."""
    net, orc = make_pair(3, 6, (128,) * 2, dm=True, model_kwargs={"noise_schedule": sched})
    B = 300
    z, c = torch.randn(B, 3), grouped_conditions(B, 6, 100)
    Z = _normals(2200, (B, 3), seed=11)
    ref = orc.inverse(z, c, method=method, steps=steps, noise_fn=lambda i, s: {"xz": Z[i]}, **extra)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps=steps, noise_fn=lambda i, shape: Z[i], **extra)
    assert torch.isfinite(out).all()
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


def test_philox_increments_are_keyed_by_seed_draw_and_global_row():
    """no noise_fn: bfk_normal_fill keyed by (seed, draw, global row) - a shard of rows equals those rows of the big call
    for the fixed-step samplers (time is shared; nothing else couples the rows)."""
    net, _ = make_pair(2, 4, (128,) * 2, dm=True)
    B = 512
    c = grouped_conditions(B, 4, 128).cuda()
    for method, kw in (("euler_maruyama", {}), ("shark", {}), ("langevin", {"corrector_steps": 1})):
        full = net.sample((B,), conditions=c, seed=21, method=method, steps=8, **kw)
        again = net.sample((B,), conditions=c, seed=21, method=method, steps=8, **kw)
        other = net.sample((B,), conditions=c, seed=22, method=method, steps=8, **kw)
        part = net.sample((128,), conditions=c[256:384], seed=21, method=method, steps=8, row_offset=256, **kw)
        assert torch.equal(full, again) and not torch.equal(full, other)
        assert rel_l2(part, full[256:384]) < 1e-5, (method, rel_l2(part, full[256:384]))
        assert net.last_integration_info["draws"] == {"euler_maruyama": 8, "shark": 16, "langevin": 16}[method]


def test_method_errors_and_layerwise_path():
    net, _ = make_pair(2, None, (64,) * 2, dm=True)
    z = torch.randn(16, 2).cuda()
    for m in ("sea", "shark", "langevin"):
        with pytest.raises(ValueError):
            net(z, inverse=True, method=m, steps="adaptive")
    # corrector / langevin with a target mask: the reference's score_fn ignores the mask (diffusion_model.py:845-849)
    net3, orc3 = make_pair(3, 6, (128,) * 2, dm=True)
    B = 200
    zz, c = torch.randn(B, 3), grouped_conditions(B, 6, 100)
    Z = _normals(40, (B, 3), seed=13)
    mask, fixed = torch.tensor([1.0, 0.0, 1.0]), torch.tensor([0.0, 0.25, 0.0])
    for method, kw in (("langevin", {}), ("euler_maruyama", {"corrector_steps": 1})):
        ref = orc3.inverse(zz, c, method=method, steps=8, noise_fn=lambda i, s: {"xz": Z[i]}, target_mask=mask,
                           targets_fixed=fixed, **kw)
        out = net3(zz.cuda(), conditions=c.cuda(), inverse=True, method=method, steps=8, noise_fn=lambda i, shape: Z[i],
                   target_mask=mask.cuda(), targets_fixed=fixed.cuda(), **kw)
        assert rel_l2(out, ref) < TOL, (method, rel_l2(out, ref))
    # wide network: the same host loop drives the layer-wise velocity kernels
    netw, orcw = make_pair(12, 5, (512,) * 2, dm=True)
    zz, c = torch.randn(200, 12), grouped_conditions(200, 5, 50)
    Z = _normals(12, (200, 12))
    ref = orcw.inverse(zz, c, method="euler_maruyama", steps=10, noise_fn=lambda i, s: {"xz": Z[i]})
    out = netw(zz.cuda(), conditions=c.cuda(), inverse=True, method="euler_maruyama", steps=10, noise_fn=lambda i, shape: Z[i])
    assert rel_l2(out, ref) < TOL
