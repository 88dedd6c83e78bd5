"""CPU-side tests (no GPU): the C-ABI library loads and exports every symbol include/bfk.h declares; the host mirror of
the reference surface keeps names, kwargs merge order and error types; the host-side ODE schedule equals the times the
oracle's integrator evaluates the network at."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import bayesflow_b200 as bf
from bayesflow_b200 import _lib
from bayesflow_b200.integrate import fixed_schedule, scheduled_schedule, integrate
import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "bfk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bfk_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/bfk.h but not exported by libbfk.so"
    assert L.bfk_version() >= 100


def test_config_queries_need_no_gpu():
    L = _lib.lib()
    cfg = _lib.NetConfig()
    cfg.D, cfg.C, cfg.W, cfg.L, cfg.T = 2, 32, 256, 5, 33
    cfg.activation, cfg.merge, cfg.norm, cfg.residual, cfg.precision = 0, 1, 1, 1, 0
    assert L.bfk_fused_supported(ctypes.byref(cfg)) == 1
    assert L.bfk_table_stride(ctypes.byref(cfg)) == 5 * 256 + 8
    assert L.bfk_packed_bytes(ctypes.byref(cfg)) > 6 * 256 * 256 * 2
    cfg.W = 250
    assert L.bfk_fused_supported(ctypes.byref(cfg)) == 0
    cfg.W, cfg.L = 256, 0
    assert L.bfk_packed_bytes(ctypes.byref(cfg)) == 0
    assert b"bad" in L.bfk_last_error_string()


def test_fixed_schedule_matches_oracle_evaluation_times():
    for method in ("euler", "rk45", "tsit5"):
        for (a, b, n) in [(0.0, 1.0, 7), (1.0, 0.0, 5)]:
            seen = []

            def fn(time, xz):
                seen.append(float(torch.as_tensor(time, dtype=torch.float32)))
                return {"xz": torch.zeros_like(xz)}
            oracle.integrate(fn, {"xz": torch.zeros(1, 1)}, start_time=a, stop_time=b, steps=n, method=method)
            times, hs = fixed_schedule(a, b, n, method)
            np.testing.assert_array_equal(times, np.asarray(seen, dtype=np.float32))
            assert hs.shape == (n,) and np.all(hs == np.float32((b - a) / n))


def test_scheduled_schedule_matches_oracle():
    grid = torch.linspace(0, 1, 9) ** 2
    for method in ("euler", "rk45"):
        seen = []

        def fn(time, xz):
            seen.append(float(torch.as_tensor(time, dtype=torch.float32)))
            return {"xz": torch.zeros_like(xz)}
        oracle.integrate(fn, {"xz": torch.zeros(1, 1)}, steps=grid, method=method)
        times, hs = scheduled_schedule(grid, method)
        np.testing.assert_allclose(times, np.asarray(seen, dtype=np.float32), rtol=0, atol=1e-7)
        np.testing.assert_allclose(hs, (grid[1:] - grid[:-1]).numpy())


def test_integrate_errors_match_reference():
    # integrate.py:226-239, 467-471, 489
    with pytest.raises(ValueError):
        integrate(None, torch.zeros(1, 1), start_time=None, stop_time=1.0, steps=10)
    with pytest.raises(ValueError):
        integrate(None, torch.zeros(1, 1), start_time=0.0, stop_time=1.0, steps=0)
    with pytest.raises(ValueError):
        integrate(None, torch.zeros(1, 1), start_time=0.0, stop_time=1.0, steps=3, method="heun")
    with pytest.raises(TypeError):
        integrate(None, torch.zeros(1, 1), start_time=0.0, stop_time=1.0, steps=3, method=3.0)
    with pytest.raises(RuntimeError):
        integrate(None, torch.zeros(1, 1), start_time=0.0, stop_time=1.0, steps=1.5)


def test_scipy_driver_host_logic():
    """`method="scipy"` (utils/integrate.py:409-452, the reference's own check tests/test_utils/test_integrate.py:28-44):
    the host driver with a stand-in engine - float64 flattened state, float32 time, last column back as float32, the two
    'has no effect' log lines of the dispatcher (:474-477)."""
    import logging

    class Engine:
        info = {}

        def _check_x(self, x):
            return x.to(torch.float32)

        def velocity(self, x, time):
            assert x.dtype == torch.float32 and isinstance(time, float) and time == float(np.float32(time))
            return torch.full_like(x, float(np.exp(time)))

    eng = Engine()
    records = []
    h = logging.Handler()
    h.emit = lambda r: records.append(r.getMessage())
    log = logging.getLogger("bayesflow_b200")
    log.addHandler(h)
    try:
        out = integrate(eng, torch.zeros(3, 2), start_time=-1.0, stop_time=1.0, steps="adaptive", method="scipy", max_steps=77)
    finally:
        log.removeHandler(h)
    assert out.dtype == torch.float32 and out.shape == (3, 2)
    np.testing.assert_allclose(out.numpy(), np.exp(1.0) - np.exp(-1.0), atol=1e-6, rtol=1e-6)
    assert eng.info["path"] == "scipy" and eng.info["evals"] > 6
    assert any("min_steps has no effect" in m for m in records) and any("max_steps has no effect" in m for m in records)
    # two-key state of the density path: keys flattened and concatenated in order
    from bayesflow_b200.integrate import solve_ivp_flat
    (a, b), res = solve_ivp_flat(lambda t, x, tr: (torch.ones_like(x) * 2.0 * t, torch.ones_like(tr)),
                                 [torch.ones(4, 3), torch.zeros(4, 1)], 0.0, 1.0)
    np.testing.assert_allclose(a.numpy(), 2.0, atol=1e-5)
    np.testing.assert_allclose(b.numpy(), 1.0, atol=1e-5)
    assert a.shape == (4, 3) and b.shape == (4, 1) and res.nfev >= 8


def test_surface_defaults_and_kwargs_merge():
    fm = bf.FlowMatching()
    assert fm.integrate_kwargs == {"method": "tsit5", "steps": "adaptive"}  # defaults.py:52
    assert fm.subnet.widths == (256,) * 5 and fm.subnet.activation == "mish" and fm.subnet.dropout == 0.05
    fm = bf.FlowMatching(subnet_kwargs={"widths": (64, 64)}, integrate_kwargs={"steps": 10})
    assert fm.subnet.widths == (64, 64) and fm.integrate_kwargs == {"method": "tsit5", "steps": 10}
    dm = bf.DiffusionModel()
    assert dm.integrate_kwargs == {"method": "two_step_adaptive", "steps": "adaptive"}  # defaults.py:50
    with pytest.raises(ValueError):
        bf.FlowMatching(time_power_law_alpha=-1.0)  # flow_matching.py:117-118
    with pytest.raises(ValueError):
        bf.DiffusionModel(prediction_type="nope")
    with pytest.raises(ValueError):
        bf.TimeMLP(widths=())
    for m in ("call", "sample", "log_prob", "compute_metrics", "build", "velocity"):
        assert callable(getattr(fm, m))


def test_state_dict_is_in_reference_variable_order():
    fm = bf.FlowMatching()
    fm.build((None, 2), (None, 32))
    ref = oracle.init_params(2, 32)
    keys = [k for k in fm.state_dict().keys() if "base_distribution" not in k]
    assert keys == list(ref.keys())
    for k, v in fm.state_dict().items():
        if k in ref:
            assert tuple(v.shape) == tuple(ref[k].shape), k
    assert sum(p.numel() for p in fm.parameters()) == 516_114


def test_no_cpu_fallback_on_cpu_tensors():
    fm = bf.FlowMatching()
    fm.build((None, 2), (None, 32))
    with pytest.raises(RuntimeError):
        fm(torch.randn(4, 2), conditions=torch.randn(4, 32), inverse=True, method="euler", steps=2)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "bayesflow_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle", src, flags=re.M), f


def test_training_workspace_query_and_config_gate():
    """bfk_train_workspace_bytes is pure host arithmetic: it must size the cfg2 / cfg4 nets and refuse what the
    layer-wise path does not cover (it never silently falls back)."""
    from bayesflow_b200.training import _bind
    L = _bind(_lib.lib())
    cfg = _lib.NetConfig()
    cfg.D, cfg.C, cfg.W, cfg.L, cfg.T = 2, 32, 256, 5, 33
    cfg.activation, cfg.merge, cfg.norm, cfg.residual, cfg.precision = 0, 1, 1, 1, 0
    fwd = L.bfk_train_workspace_bytes(ctypes.byref(cfg), 65536, 0)
    both = L.bfk_train_workspace_bytes(ctypes.byref(cfg), 65536, 1)
    assert 65536 * 256 * 4 * 10 < fwd < both < 4 << 30
    cfg.D, cfg.W, cfg.L = 64, 1024, 8
    assert L.bfk_train_workspace_bytes(ctypes.byref(cfg), 65536, 1) > both
    cfg.W = 250
    assert L.bfk_train_workspace_bytes(ctypes.byref(cfg), 1024, 1) == 0
    cfg.W, cfg.precision = 256, 1
    assert L.bfk_train_workspace_bytes(ctypes.byref(cfg), 1024, 1) == 0 and b"fp16" in L.bfk_last_error_string()


def test_cosine_decay_matches_keras_formula():
    from bayesflow_b200.training import cosine_decay
    lr = cosine_decay(5e-4, 1000)
    assert lr(0) == pytest.approx(5e-4) and lr(500) == pytest.approx(2.5e-4) and lr(1000) == pytest.approx(0.0, abs=1e-12)
    assert lr(2000) == lr(1000)
