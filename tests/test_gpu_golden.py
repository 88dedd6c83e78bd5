"""The CUDA path (through the C ABI) against the golden vectors of the reference's OWN classes (tests/golden/,
written by tools/make_golden.py from /root/reference on the keras stand-in) — the oracle is not in this loop.
Gate (BASELINE.json): velocity, fixed-noise draws and every gradient within 2e-3 relative L2 of the reference."""
import pytest
import torch

from _golden import FM_CASES, DM_CASES, load, subnet_kwargs
from _util import rel_l2

pytestmark = pytest.mark.gpu
TOL = 2e-3


def build(g, dm=False):
    import bayesflow_b200 as bf
    cls = bf.DiffusionModel if dm else bf.FlowMatching
    net = cls(subnet_kwargs=subnet_kwargs(g), **(g.get("model_kwargs") or {}))
    net.build((None, g["D"]), (None, g["C"]) if g["C"] else None)
    res = net.load_state_dict(g["params"], strict=False)
    assert not res.unexpected_keys and not [k for k in res.missing_keys if "base_distribution" not in k], res
    return net.cuda()


@pytest.mark.parametrize("name", FM_CASES)
def test_fm_cuda_equals_reference_golden(name):
    g = load(name)
    assert g["params_ok"], "torch's RNG stream changed: regenerate tests/golden with tools/make_golden.py"
    net = build(g)  # (fm_small_projector: unequal hidden widths -> residual projector, conditional_dense_block.py:152-156)
    z, c = g["z"].cuda(), g["conditions"].cuda()
    for t in (0.0, 0.37, 1.0):
        assert rel_l2(net.velocity(z, t, conditions=c), g[f"velocity_t{t}"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="euler", steps=20), g["euler20"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="rk45", steps=5), g["rk45_5"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="tsit5", steps=5), g["tsit5_5"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=False, method="euler", steps=10), g["forward_euler10"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="tsit5", steps="adaptive"), g["tsit5_adaptive"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="rk45", steps="adaptive", min_steps=3, max_steps=40),
                  g["rk45_adaptive_3_40"]) < TOL
    x0, x1, t = g["train_x0"].cuda(), g["train_x1"].cuda(), g["train_t"].cuda()
    tup = (x0, x1, t, t * x1 + (1 - t) * x0, x1 - x0)
    loss = net.compute_metrics(tup, conditions=c, stage="training")["loss"]
    loss.backward()
    assert abs(float(loss) - float(g["train_loss"])) <= TOL * abs(float(g["train_loss"]))
    named = dict(net.named_parameters())
    for k, n in g["train_grad_norms"].items():
        assert abs(float(named[k].grad.norm()) - n) <= 5e-3 * max(n, 1e-12), k
    for k, v in g["train_grads"].items():
        assert rel_l2(named[k].grad, v) < TOL, k
    if "log_prob_euler8" in g:
        lp = net.log_prob(z[:16], conditions=c[:16], method="euler", steps=8)
        assert rel_l2(lp, g["log_prob_euler8"]) < TOL


@pytest.mark.parametrize("name", DM_CASES)
def test_dm_cuda_equals_reference_golden(name):
    g = load(name)
    assert g["params_ok"]
    net = build(g, dm=True)
    z, c = g["z"].cuda(), g["conditions"].cuda()
    for t in (0.9, 0.5, 0.05):
        assert rel_l2(net.velocity(z, t, conditions=c), g[f"velocity_t{t}"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="euler", steps=50), g["euler50"]) < TOL
    assert rel_l2(net(z, conditions=c, inverse=True, method="rk45", steps=10), g["rk45_10"]) < TOL
    # training: the same (u0, eps) the reference drew through keras.random (diffusion_model.py:159-174)
    loss = net.compute_metrics(g["train_x"].cuda(), conditions=c, stage="training",
                               fixed=(g["train_u0"], g["train_eps"]))["loss"]
    loss.backward()
    assert abs(float(loss) - float(g["train_loss"])) <= TOL * abs(float(g["train_loss"]))
    named = dict(net.named_parameters())
    for k, v in g["train_grads"].items():
        assert rel_l2(named[k].grad, v) < TOL, k
