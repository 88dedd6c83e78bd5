"""Pins the oracle integrator with the reference's own closed-form tests
(/root/reference/tests/test_utils/test_integrate.py:15-25, 47-68, 71-93), restated on torch."""
import numpy as np
import pytest
import torch

from oracle.integrate import integrate, DETERMINISTIC_METHODS

TOLERANCE_ADAPTIVE = 1e-6
TOLERANCE_EULER = 1e-3


@pytest.mark.parametrize("method", DETERMINISTIC_METHODS)
def test_scheduled_integration(method):
    def fn(t, x):
        return {"x": t ** 2}

    steps = torch.arange(0.0, 1.0 + 1e-6, 0.01)
    result = integrate(fn, {"x": torch.tensor(0.0)}, steps=steps, method=method)["x"]
    np.testing.assert_allclose(result, (steps[-1] ** 3) / 3.0, atol=1e-1, rtol=1e-1)


@pytest.mark.parametrize("method, atol", [("euler", TOLERANCE_EULER), ("rk45", TOLERANCE_ADAPTIVE), ("tsit5", TOLERANCE_ADAPTIVE)])
def test_analytical_integration(method, atol):
    def fn(t, x):
        return {"x": torch.as_tensor([2.0 * t], dtype=torch.float32)}

    init = {"x": torch.tensor([1.0])}
    result = integrate(fn, init, start_time=0.0, stop_time=1.0, steps=100, method=method)["x"]
    result_adaptive = result if method == "euler" else integrate(
        fn, init, start_time=0.0, stop_time=1.0, steps="adaptive", method=method, max_steps=1_000)["x"]
    np.testing.assert_allclose(result, 2.0, atol=atol, rtol=0.1)
    np.testing.assert_allclose(result_adaptive, 2.0, atol=atol, rtol=0.1)
    # the reference tolerance (rtol=0.1) is loose; the tableaux are exact for a quadratic, so pin tighter
    if method != "euler":
        np.testing.assert_allclose(result, 2.0, atol=2e-6, rtol=0)
        np.testing.assert_allclose(result_adaptive, 2.0, atol=2e-6, rtol=0)
    else:
        np.testing.assert_allclose(result, 2.0 - 0.01, atol=1e-5, rtol=0)  # Euler-100 of dx/dt=2t is exactly 1+0.99


@pytest.mark.parametrize("method, atol", [("euler", TOLERANCE_EULER), ("rk45", TOLERANCE_ADAPTIVE), ("tsit5", TOLERANCE_ADAPTIVE)])
def test_analytical_backward_integration(method, atol):
    def fn(t, x):
        return {"x": torch.as_tensor([2.0 * t], dtype=torch.float32)}

    init = {"x": torch.tensor([2.0])}
    result = integrate(fn, init, start_time=1.0, stop_time=0.0, steps=100, method=method)["x"]
    result_adaptive = result if method == "euler" else integrate(
        fn, init, start_time=1.0, stop_time=0.0, steps="adaptive", method=method, max_steps=1_000)["x"]
    np.testing.assert_allclose(result, 1.0, atol=max(atol, 1.1e-2) if method == "euler" else atol, rtol=0.1)
    np.testing.assert_allclose(result_adaptive, 1.0, atol=max(atol, 1.1e-2) if method == "euler" else atol, rtol=0.1)


def test_scipy_integration():
    """tests/test_utils/test_integrate.py:28-44: d x / d t = exp(t) over [-1, 1] through method="scipy" (the dispatcher drops
    `scipy_kwargs`, integrate.py:478, so solve_ivp runs with its defaults)."""
    def fn(t, x):
        return {"x": torch.exp(t)}

    result = integrate(fn, {"x": torch.tensor(0.0)}, start_time=-1.0, stop_time=1.0, steps="adaptive", method="scipy",
                       scipy_kwargs={"atol": 1e-6, "rtol": 1e-6})["x"]
    np.testing.assert_allclose(float(result), np.exp(1.0) - np.exp(-1.0), atol=1e-6, rtol=1e-6)


def test_argument_errors():
    fn = lambda t, x: {"x": x}
    s = {"x": torch.ones(1)}
    with pytest.raises(ValueError):
        integrate(fn, s, start_time=0.0, stop_time=1.0, steps=0, method="euler")
    with pytest.raises(ValueError):
        integrate(fn, s, start_time=0.0, stop_time=1.0, steps="adaptive", method="euler")
    with pytest.raises(ValueError):
        integrate(fn, s, start_time=0.0, stop_time=1.0, steps=10, method="nope")
    with pytest.raises(ValueError):
        integrate(fn, s, steps=10, method="euler")
    with pytest.raises(ValueError):
        integrate(fn, s, start_time=0.0, stop_time=1.0, steps="adaptive", method="rk45", min_steps=10, max_steps=10)
    with pytest.raises(RuntimeError):
        integrate(fn, s, start_time=0.0, stop_time=1.0, steps=1.5, method="rk45")


def test_all_nan_raises():
    fn = lambda t, x: {"x": x * float("nan")}
    with pytest.raises(RuntimeError):
        integrate(fn, {"x": torch.ones(3, 2)}, start_time=0.0, stop_time=1.0, steps="adaptive", method="rk45")


# ---- stochastic half: the reference's Ornstein-Uhlenbeck weak-accuracy tests (tests/test_utils/test_integrate.py:96-222,
# TOL_MEAN / TOL_VAR as there) restated for the oracle's euler_maruyama / two_step_adaptive
@pytest.mark.parametrize("method,use_adapt", [("euler_maruyama", False), ("euler_maruyama", True),
                                              ("two_step_adaptive", False), ("two_step_adaptive", True)])
@pytest.mark.parametrize("backward", [False, True])
def test_oracle_sde_ou_weak_means_and_vars(method, use_adapt, backward):
    import numpy as np
    from oracle.integrate import integrate_stochastic
    a, sigma, x0, T, N = -1.0, 0.5, 1.2, 1.0, 10000
    gen = torch.Generator().manual_seed(42)
    out = integrate_stochastic(lambda t, x: {"x": a * x}, lambda t, x: {"x": torch.tensor([sigma])},
                               {"x": torch.ones(N) * x0}, start_time=T if backward else 0.0, stop_time=0.0 if backward else T,
                               steps="adaptive" if use_adapt else 200, generator=gen, method=method)["x"].numpy()
    s = -1.0 if backward else 1.0
    exp_mean = x0 * np.exp(s * a * T)
    exp_var = sigma ** 2 * (np.exp(2.0 * s * a * T) - 1.0) / (2.0 * s * a)
    np.testing.assert_allclose(out.mean(), exp_mean, atol=5e-2)
    np.testing.assert_allclose(out.var(), exp_var, atol=5e-2)
