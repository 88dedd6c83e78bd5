"""GPU parity tests proper: the CUDA path (through the C ABI of libbfk.so) against the CPU oracle on identical weights
and inputs.  Tolerance from BASELINE.json north_star: 2e-3 relative L2 on velocities and on fixed-noise draws; posterior
mean / std within 1 %.  Sizes are chosen so that the oracle finishes in seconds; tile-unaligned on purpose."""
import pytest
import torch

from _util import make_pair, rel_l2, grouped_conditions, as_float64

pytestmark = pytest.mark.gpu
TOL = 2e-3


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    import bayesflow_b200._lib as L
    L.lib()  # fail loudly if libbfk.so is missing: there is no fallback


@pytest.mark.parametrize("D,C,widths,B", [
    (2, 32, (256,) * 5, 1), (2, 32, (256,) * 5, 127), (2, 32, (256,) * 5, 128), (2, 32, (256,) * 5, 4097),
    (2, 2, (256,) * 5, 1000), (4, 20, (256,) * 5, 513), (2, None, (256,) * 5, 300), (4, 20, (128,) * 3, 777),
    (2, 2, (64,) * 2, 257), (8, 5, (256,) * 2, 260), (3, 7, (96,) * 4, 99),
    (2, 6, (64, 96, 32), 150), (4, 20, (256, 128, 256), 300), (2, None, (32, 64), 77),   # unequal widths (projector)
])
def test_fm_velocity(D, C, widths, B):
    net, orc = make_pair(D, C, widths)
    g = torch.Generator().manual_seed(3456)
    z = torch.randn(B, D, generator=g)
    c = grouped_conditions(B, C, 100) if C else None
    for t in (0.0, 0.37, 1.0):
        ref = orc.velocity(z, t, c)
        out = net.velocity(z.cuda(), t, conditions=c.cuda() if C else None)
        assert rel_l2(out, ref) < TOL, (t, rel_l2(out, ref))


def test_velocity_ungrouped_conditions():
    """every row its own condition vector (no repeat structure to detect)"""
    net, orc = make_pair(2, 32)
    z, c = torch.randn(333, 2), torch.randn(333, 32)
    assert rel_l2(net.velocity(z.cuda(), 0.5, conditions=c.cuda()), orc.velocity(z, 0.5, c)) < TOL


@pytest.mark.parametrize("sk", [
    {"activation": "silu"}, {"activation": "relu"}, {"activation": "gelu"}, {"activation": "tanh"},
    {"merge": "add"}, {"film_use_gamma": True}, {"norm": None}, {"norm": "rms"}, {"residual": False},
    {"time_embedding_dim": 1}, {"time_embedding_dim": 8},
])
def test_fm_velocity_options(sk):
    net, orc = make_pair(2, 6, (128,) * 3, **sk)
    z = torch.randn(300, 2)
    c = grouped_conditions(300, 6, 50)
    ref = orc.velocity(z, 0.61, c)
    out = net.velocity(z.cuda(), 0.61, conditions=c.cuda())
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


def test_bf16_operands():
    net, orc = make_pair(2, 32, net_kwargs={"precision": "bf16"})
    z = torch.randn(500, 2)
    c = grouped_conditions(500, 32, 100)
    # bf16 operands (8-bit mantissa) are an opt-in; they are held to a looser bar and are NOT the benchmarked path
    assert rel_l2(net.velocity(z.cuda(), 0.3, conditions=c.cuda()), orc.velocity(z, 0.3, c)) < 1e-2


@pytest.mark.parametrize("method,steps", [("euler", 20), ("euler", 100), ("rk45", 7), ("tsit5", 7)])
@pytest.mark.parametrize("D,C", [(2, 32), (4, 20), (2, None)])
def test_fm_fixed_noise_draws(method, steps, D, C):
    net, orc = make_pair(D, C)
    B = 700
    z = torch.randn(B, D, generator=torch.Generator().manual_seed(3456))
    c = grouped_conditions(B, C, 100) if C else None
    ref = orc.inverse(z, c, method=method, steps=steps)
    out = net(z.cuda(), conditions=c.cuda() if C else None, inverse=True, method=method, steps=steps)
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


def test_fm_forward_direction_and_cycle():
    """flow_matching.py:254-294 integrates 1 -> 0; reference test_flow_matching.py:99-108 cycle consistency (1e-3)."""
    net, orc = make_pair(2, 3, (64, 64))
    x = torch.randn(64, 2)
    c = torch.randn(64, 3)
    ref = orc.forward(x, c, method="rk45", steps=20)
    z = net(x.cuda(), conditions=c.cuda(), method="rk45", steps=20)
    assert rel_l2(z, ref) < TOL
    back = net(z, conditions=c.cuda(), inverse=True, method="rk45", steps=20)
    torch.testing.assert_close(back.cpu(), x, atol=2e-3, rtol=2e-3)


def test_scheduled_steps():
    """integrate_scheduled, integrate.py:257-288: non-uniform grid given as a sequence / tensor."""
    net, orc = make_pair(2, 32)
    z = torch.randn(200, 2)
    c = grouped_conditions(200, 32, 100)
    grid = (torch.linspace(0, 1, 31) ** 2)
    for method in ("euler", "rk45", "tsit5"):
        ref = orc.inverse(z, c, method=method, steps=grid)
        out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps=grid.tolist())
        assert rel_l2(out, ref) < TOL, (method, rel_l2(out, ref))


def test_target_mask():
    net, orc = make_pair(4, 20)
    z = torch.randn(150, 4)
    c = grouped_conditions(150, 20, 50)
    tm = torch.tensor([1.0, 0.0, 1.0, 1.0])
    tf = torch.tensor([0.0, 0.7, 0.0, 0.0])
    ref = orc.inverse(z, c, method="euler", steps=10, target_mask=tm, targets_fixed=tf)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=10, target_mask=tm.cuda(),
              targets_fixed=tf.cuda())
    assert rel_l2(out, ref) < TOL
    torch.testing.assert_close(out[:, 1].cpu(), torch.full((150,), 0.7))


@pytest.mark.parametrize("pred", ["F", "velocity", "noise", "x", "score"])
def test_dm_velocity(pred):
    net, orc = make_pair(4, 20, dm=True, model_kwargs={"prediction_type": pred})
    z = torch.randn(500, 4)
    c = grouped_conditions(500, 20, 100)
    if pred == "score":  # the reference's fp32 algebra cancels catastrophically here: compare with its float64 evaluation
        orc, z, c = as_float64(orc), z.double(), c.double()
    for t in (0.95, 0.5, 0.05):
        ref = orc.velocity(z, t, conditions=c)
        out = net.velocity(z.float().cuda(), t, conditions=c.float().cuda())
        assert rel_l2(out, ref) < TOL, (t, rel_l2(out, ref))


@pytest.mark.parametrize("sched", ["edm", "cosine"])
@pytest.mark.parametrize("method,steps", [("euler", 50), ("rk45", 10)])
def test_dm_probability_flow_draws(sched, method, steps):
    net, orc = make_pair(4, 20, dm=True, model_kwargs={"noise_schedule": sched})
    z = torch.randn(600, 4)
    c = grouped_conditions(600, 20, 100)
    ref = orc.inverse(z, c, method=method, steps=steps)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps=steps)
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


def test_posterior_moments_within_one_percent():
    """per-parameter posterior mean and std within 1 % (of the posterior std / of the std) on one dataset"""
    net, orc = make_pair(2, 32)
    S = 4000
    z = torch.randn(S, 2, generator=torch.Generator().manual_seed(3456))
    c = grouped_conditions(S, 32, S)
    ref = orc.inverse(z, c, method="euler", steps=50)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=50).cpu()
    assert ((out.mean(0) - ref.mean(0)).abs() / ref.std(0) < 1e-2).all()
    assert ((out.std(0) / ref.std(0) - 1).abs() < 1e-2).all()


def test_full_size_properties():
    """cfg2 full size (10^6 rows): size-independent properties instead of the oracle — (i) the result is independent
    of how rows are tiled over CTAs: any 1000-row slice re-run on its own reproduces the same draws bit for bit;
    (ii) rows of one dataset with the same noise get the same draw; (iii) finite everywhere."""
    net, _ = make_pair(2, 32)
    N, S = 1000, 1000
    g = torch.Generator(device="cuda").manual_seed(3456)
    z = torch.randn(N * S, 2, device="cuda", generator=g)
    z[S + 1] = z[S]  # same dataset (row // S == 1), same noise
    cond = torch.randn(N, 32, device="cuda", generator=g)
    out = net(z, conditions=cond.repeat_interleave(S, dim=0), inverse=True, method="euler", steps=100)
    assert torch.isfinite(out).all()
    assert torch.equal(out[S], out[S + 1])
    for n in (0, 499, 999):
        sl = slice(n * S, (n + 1) * S)
        again = net(z[sl], conditions=cond[n:n + 1].expand(S, 32).contiguous(), inverse=True, method="euler", steps=100)
        assert torch.equal(again, out[sl])


def test_no_cpu_fallback():
    net, _ = make_pair(2, 32)
    with pytest.raises(RuntimeError):
        net.velocity(torch.randn(4, 2), 0.5, conditions=torch.randn(4, 32))


# ---------------------------------------------------------------------------------------------------- adaptive (I7)
@pytest.mark.parametrize("method", ["tsit5", "rk45"])
@pytest.mark.parametrize("D,C", [(2, 32), (4, 20)])
def test_fm_adaptive_draws(method, D, C):
    """integrate_adaptive (integrate.py:291-406) with the device-side controller: same draws within tolerance and the
    same number of accepted steps as the oracle's global controller on the same rows (+-2: the embedded error estimate
    sees the fp16-operand noise of the slopes)."""
    net, orc = make_pair(D, C)
    B = 600
    z = torch.randn(B, D, generator=torch.Generator().manual_seed(3456))
    c = grouped_conditions(B, C, 100)
    info = {}
    ref = orc.inverse(z, c, method=method, steps="adaptive", info=info)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method=method, steps="adaptive")
    got = net.last_integration_info
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)
    assert abs(got["steps"] - info["steps"]) <= 2, (got, info)
    assert abs(got["final_time"] - 1.0) < 1e-6


def test_fm_default_sample_is_adaptive_tsit5():
    """FlowMatching() with no integrate_kwargs samples with tsit5 / adaptive (defaults.py:52) through .sample()."""
    net, orc = make_pair(2, 2)
    c = grouped_conditions(300, 2, 100)
    out = net.sample((300,), conditions=c.cuda(), seed=7)
    assert out.shape == (300, 2) and torch.isfinite(out).all()
    assert net.last_integration_info["steps"] >= 50  # min_steps
    again = net.sample((300,), conditions=c.cuda(), seed=7)
    assert torch.equal(out, again)  # tests/test_networks/test_sample.py:6-54 seed determinism


def test_adaptive_tight_and_loose_tolerances():
    """atol/rtol flow through kwargs (integrate.py:316-317); a loose tolerance still takes min_steps, a tight one
    takes more trial steps; max_steps <= min_steps raises ValueError (:301-302)."""
    net, orc = make_pair(2, 32)
    z = torch.randn(300, 2)
    c = grouped_conditions(300, 32, 100)
    for kw in ({"atol": 1e-3, "rtol": 1e-2, "min_steps": 10}, {"atol": 1e-7, "rtol": 1e-5, "min_steps": 10}):
        info = {}
        ref = orc.inverse(z, c, method="rk45", steps="adaptive", info=info, **kw)
        out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="rk45", steps="adaptive", **kw)
        assert rel_l2(out, ref) < TOL, (kw, rel_l2(out, ref))
        assert net.last_integration_info["steps"] >= 10
    with pytest.raises(ValueError):
        net(z.cuda(), conditions=c.cuda(), inverse=True, method="rk45", steps="adaptive", min_steps=10, max_steps=10)
    with pytest.raises(ValueError):
        net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps="adaptive")


def test_dm_adaptive_probability_flow():
    net, orc = make_pair(4, 20, dm=True)
    z = torch.randn(400, 4)
    c = grouped_conditions(400, 20, 100)
    info = {}
    ref = orc.inverse(z, c, method="rk45", steps="adaptive", info=info)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="rk45", steps="adaptive")
    assert rel_l2(out, ref) < TOL, rel_l2(out, ref)


@pytest.mark.parametrize("scale", [5.0, 15.0])
def test_adaptive_with_rejections(scale):
    """A stiffer field (output projector scaled up) with min_steps=5 makes the controller reject trial steps and end
    with the exact-landing final step (integrate.py:365-378, 393-403)."""
    import oracle
    widths = (128,) * 3
    params = oracle.perturb_params(oracle.init_params(2, 32, widths))
    params["output_projector.kernel"] = params["output_projector.kernel"] * scale
    import bayesflow_b200 as bf
    kw = {"dropout": 0.0, "widths": widths}
    net = bf.FlowMatching(subnet_kwargs=kw)
    net.build((None, 2), (None, 32))
    net.load_state_dict(params, strict=False)
    net.cuda()
    orc = oracle.OracleFlowMatching(params, 2, subnet_kwargs=kw)
    z = torch.randn(300, 2, generator=torch.Generator().manual_seed(11))
    c = grouped_conditions(300, 32, 100)
    info = {}
    ref = orc.inverse(z, c, method="rk45", steps="adaptive", info=info, min_steps=5)
    out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="rk45", steps="adaptive", min_steps=5)
    got = net.last_integration_info
    assert info["rejected"] > 0 and got["rejected"] > 0, (info, got)
    assert abs(got["steps"] - info["steps"]) <= max(2, 0.15 * info["steps"]), (got, info)
    assert abs(got["final_time"] - 1.0) < 1e-6
    # the scaled field amplifies ANY perturbation (the solver's own rtol=1e-4 error included) by its Lipschitz constant:
    # this is a controller stress test, held to the parity bar times the amplification seen between two oracle runs
    assert rel_l2(out, ref) < (TOL if scale <= 5 else 3 * TOL), rel_l2(out, ref)
