"""oracle/ against the committed golden vectors of the reference's own classes (tests/golden/, tools/make_golden.py).
Runs anywhere (no /root/reference needed): it is the portable half of the reference-backed pin."""
import pytest
import torch

import oracle
from _golden import FM_CASES, DM_CASES, load, subnet_kwargs


def rel(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def orc_grads(orc, fn):
    ps = {k: v.clone().requires_grad_(True) for k, v in orc.p.items()}
    old = orc.p, orc.subnet.p
    orc.p = orc.subnet.p = ps
    try:
        loss = fn()
        loss.backward()
    finally:
        orc.p, orc.subnet.p = old
    return loss.detach(), {k: v.grad for k, v in ps.items()}


@pytest.mark.parametrize("name", FM_CASES)
def test_fm_oracle_equals_reference_golden(name):
    g = load(name)
    assert g["params_ok"], "torch's RNG stream changed: regenerate tests/golden with tools/make_golden.py"
    orc = oracle.OracleFlowMatching(g["params"], g["D"], subnet_kwargs=subnet_kwargs(g))
    z, c = g["z"], g["conditions"]
    with torch.no_grad():
        for t in (0.0, 0.37, 1.0):
            assert rel(orc.velocity(z, t, c), g[f"velocity_t{t}"]) < 1e-6
        assert rel(orc.inverse(z, c, method="euler", steps=20), g["euler20"]) < 5e-6
        assert rel(orc.inverse(z, c, method="rk45", steps=5), g["rk45_5"]) < 5e-6
        assert rel(orc.inverse(z, c, method="tsit5", steps=5), g["tsit5_5"]) < 5e-6
        assert rel(orc.forward(z, c, method="euler", steps=10), g["forward_euler10"]) < 5e-6
        info = {}
        x = orc.inverse(z, c, method="tsit5", steps="adaptive", info=info)
        assert info == g["tsit5_adaptive_counts"] and rel(x, g["tsit5_adaptive"]) < 2e-5
        info = {}
        x = orc.inverse(z, c, method="rk45", steps="adaptive", min_steps=3, max_steps=40, info=info)
        assert info == g["rk45_adaptive_3_40_counts"] and rel(x, g["rk45_adaptive_3_40"]) < 2e-5
    x0, x1, t = g["train_x0"], g["train_x1"], g["train_t"]
    tup = (x0, x1, t, t * x1 + (1 - t) * x0, x1 - x0)
    loss, grads = orc_grads(orc, lambda: orc.compute_metrics(tup, conditions=c, stage="training")["loss"])
    assert abs(float(loss) - float(g["train_loss"])) <= 1e-6 * abs(float(g["train_loss"]))
    for k, n in g["train_grad_norms"].items():
        assert abs(float(grads[k].norm()) - n) <= 1e-5 * max(n, 1e-12), k
    for k, v in g["train_grads"].items():
        assert rel(grads[k], v) < 1e-5, k
    if "log_prob_euler8" in g:
        _, lp = orc.forward(z[:16], c[:16] if c is not None else None, density=True, method="euler", steps=8)
        assert rel(lp, g["log_prob_euler8"]) < 2e-5


@pytest.mark.parametrize("name", DM_CASES)
def test_dm_oracle_equals_reference_golden(name):
    g = load(name)
    assert g["params_ok"]
    orc = oracle.OracleDiffusionModel(g["params"], g["D"], subnet_kwargs=subnet_kwargs(g), **g["model_kwargs"])
    z, c = g["z"], g["conditions"]
    with torch.no_grad():
        for t in (0.9, 0.5, 0.05):
            assert rel(orc.velocity(z, t, conditions=c), g[f"velocity_t{t}"]) < 2e-5
        assert rel(orc.inverse(z, c, method="euler", steps=50), g["euler50"]) < 2e-5
        assert rel(orc.inverse(z, c, method="rk45", steps=10), g["rk45_10"]) < 2e-5
    loss, grads = orc_grads(orc, lambda: orc.compute_metrics(g["train_x"], conditions=c, stage="training",
                                                              fixed=(g["train_u0"], g["train_eps"]))["loss"])
    assert abs(float(loss) - float(g["train_loss"])) <= 2e-6 * abs(float(g["train_loss"]))
    for k, v in g["train_grads"].items():
        assert rel(grads[k], v) < 2e-5, k
