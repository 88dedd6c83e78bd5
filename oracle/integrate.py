"""ODE integrators — CPU oracle (TEST INFRASTRUCTURE, never shipped).

Plain-PyTorch restatement of /root/reference/bayesflow/utils/integrate.py:24-489 (the deterministic half).
The reference is written on ``keras.ops``; on the torch backend ``fori_loop`` / ``while_loop`` / ``cond``
are Python control flow, python floats stay python doubles until they meet a tensor, and every reduction
(`max`, `norm`) is a torch op on fp32 tensors.  This file keeps exactly those semantics.

Pinned by the reference's own analytic tests (tests/test_utils/test_integrate.py:15-25, 47-93), restated in
/root/repo/tests/test_oracle_integrate.py.
"""
from __future__ import annotations

import math
from typing import Callable, Dict

import numpy as np
import torch

StateDict = Dict[str, torch.Tensor]

DETERMINISTIC_METHODS = ["euler", "rk45", "tsit5"]

# Dormand–Prince 5(4), integrate.py:74-96, 107-114
RK45_C = (0.0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0)
RK45_A = (
    (),
    (1 / 5,),
    (3 / 40, 9 / 40),
    (44 / 45, -56 / 15, 32 / 9),
    (19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729),
    (9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656),
)
RK45_B5 = (35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84)
RK45_B4 = (5179 / 57600, 0.0, 7571 / 16695, 393 / 640, -92097 / 339200, 187 / 2100, 1 / 40)

# Tsitouras 5(4), integrate.py:139-142, 146-179, 184-191, 201-209
TSIT5_C = (0.0, 0.161, 0.327, 0.9, 0.9800255409045097, 1.0, 1.0)
TSIT5_A = (
    (),
    (0.161,),
    (-0.0084806554923570, 0.3354806554923570),
    (2.897153057105494, -6.359448489975075, 4.362295432869581),
    (5.325864828439257, -11.74888356406283, 7.495539342889836, -0.09249506636175525),
    (5.86145544294270, -12.92096931784711, 8.159367898576159, -0.07158497328140100, -0.02826905039406838),
)
TSIT5_B = (0.09646076681806523, 0.01, 0.4798896504144996, 1.379008574103742, -3.290069515436081, 2.324710524099774)
# error = h * sum(e_i k_i), i = 1..7
TSIT5_E = (
    -0.00178001105222577714,
    -0.0008164344596567469,
    0.007880878010261995,
    -0.1447110071732629,
    0.5823571654525552,
    -0.45808210592918697,
    0.015151515151515152,
)


def _check_all_nans(state: StateDict) -> bool:
    # integrate.py:24-28
    return bool(all(bool(torch.all(torch.isnan(v))) for v in state.values()))


def euler_step(fn: Callable, state: StateDict, time, step_size, **_):
    # integrate.py:31-44
    k1 = fn(time, **state)
    new_state = dict(state)
    for key in k1.keys():
        new_state[key] = state[key] + step_size * k1[key]
    return new_state, time + step_size, None, 0.0


def add_scaled(state, ks, coeffs, h):
    # integrate.py:47-54
    out = {}
    for key, y in state.items():
        acc = torch.zeros_like(y)
        for c, k in zip(coeffs, ks):
            acc = acc + c * k[key]
        out[key] = y + h * acc
    return out


def _err_max(err_state: StateDict):
    # integrate.py:117-118 / 211-212: max over keys and rows of the row-wise L2 norm
    err_norm = torch.stack([torch.linalg.vector_norm(v, ord=2, dim=-1).reshape(-1).max() for v in err_state.values()])
    return err_norm.max()


def rk45_step(fn, state, time, step_size, k1=None, use_adaptive_step_size=True):
    # integrate.py:57-120
    h = step_size
    if k1 is None:
        k1 = fn(time, **state)
    k2 = fn(time + h * (1 / 5), **add_scaled(state, [k1], RK45_A[1], h))
    k3 = fn(time + h * (3 / 10), **add_scaled(state, [k1, k2], RK45_A[2], h))
    k4 = fn(time + h * (4 / 5), **add_scaled(state, [k1, k2, k3], RK45_A[3], h))
    k5 = fn(time + h * (8 / 9), **add_scaled(state, [k1, k2, k3, k4], RK45_A[4], h))
    k6 = fn(time + h, **add_scaled(state, [k1, k2, k3, k4, k5], RK45_A[5], h))

    new_state = {}
    for key in k1.keys():
        new_state[key] = state[key] + h * (
            35 / 384 * k1[key] + 500 / 1113 * k3[key] + 125 / 192 * k4[key] - 2187 / 6784 * k5[key] + 11 / 84 * k6[key]
        )
    new_time = time + h
    if not use_adaptive_step_size:
        return new_state, new_time, None, 0.0

    k7 = fn(time + h, **new_state)
    err_state = {}
    for key in k1.keys():
        y4 = state[key] + h * (
            5179 / 57600 * k1[key]
            + 7571 / 16695 * k3[key]
            + 393 / 640 * k4[key]
            - 92097 / 339200 * k5[key]
            + 187 / 2100 * k6[key]
            + 1 / 40 * k7[key]
        )
        err_state[key] = new_state[key] - y4
    return new_state, new_time, k7, _err_max(err_state)


def tsit5_step(fn, state, time, step_size, k1=None, use_adaptive_step_size=True):
    # integrate.py:123-214
    h = step_size
    c2, c3, c4, c5 = TSIT5_C[1:5]
    if k1 is None:
        k1 = fn(time, **state)
    k2 = fn(time + h * c2, **add_scaled(state, [k1], TSIT5_A[1], h))
    k3 = fn(time + h * c3, **add_scaled(state, [k1, k2], TSIT5_A[2], h))
    k4 = fn(time + h * c4, **add_scaled(state, [k1, k2, k3], TSIT5_A[3], h))
    k5 = fn(time + h * c5, **add_scaled(state, [k1, k2, k3, k4], TSIT5_A[4], h))
    k6 = fn(time + h, **add_scaled(state, [k1, k2, k3, k4, k5], TSIT5_A[5], h))

    new_state = {}
    for key in state.keys():
        new_state[key] = state[key] + h * (
            0.09646076681806523 * k1[key]
            + 0.01 * k2[key]
            + 0.4798896504144996 * k3[key]
            + 1.379008574103742 * k4[key]
            - 3.290069515436081 * k5[key]
            + 2.324710524099774 * k6[key]
        )
    new_time = time + h
    if not use_adaptive_step_size:
        return new_state, new_time, None, 0.0

    k7 = fn(time + h, **new_state)
    err_state = {}
    for key in state.keys():
        err_state[key] = h * (
            0.007880878010261995 * k3[key]
            + 0.5823571654525552 * k5[key]
            + 0.015151515151515152 * k7[key]
            - 0.00178001105222577714 * k1[key]
            - 0.0008164344596567469 * k2[key]
            - 0.1447110071732629 * k4[key]
            - 0.45808210592918697 * k6[key]
        )
    return new_state, new_time, k7, _err_max(err_state)


def _select_step(method, adaptive):
    # integrate.py:229-239 / 265-275 / 304-314
    if method == "euler":
        if adaptive:
            raise ValueError("Adaptive step sizing is not supported for the 'euler' method.")
        return euler_step
    if method == "rk45":
        return lambda fn, s, t, h, k1=None: rk45_step(fn, s, t, h, k1=k1, use_adaptive_step_size=adaptive)
    if method == "tsit5":
        return lambda fn, s, t, h, k1=None: tsit5_step(fn, s, t, h, k1=k1, use_adaptive_step_size=adaptive)
    if isinstance(method, str):
        raise ValueError(f"Unknown integration method name: {method!r}")
    raise TypeError(f"Invalid integration method: {method!r}")


def integrate_fixed(fn, state, start_time, stop_time, steps, method, **kwargs):
    # integrate.py:217-254
    if steps <= 0:
        raise ValueError("Number of steps must be positive.")
    step_fn = _select_step(method, adaptive=False)
    step_size = (stop_time - start_time) / steps
    time = start_time
    for _ in range(steps):
        state, time, _, _ = step_fn(fn, state, time, step_size)
    return state


def integrate_scheduled(fn, state, steps, method, **kwargs):
    # integrate.py:257-288
    step_fn = _select_step(method, adaptive=False)
    n = int(steps.shape[0]) if hasattr(steps, "shape") else len(steps)
    for i in range(n - 1):
        time = steps[i]
        step_size = steps[i + 1] - steps[i]
        state, _, _, _ = step_fn(fn, state, time, step_size)
    return state


def integrate_adaptive(fn, state, start_time, stop_time, min_steps, max_steps, method, info: dict | None = None,
                       **kwargs):
    # integrate.py:291-406
    if max_steps <= min_steps:
        raise ValueError("Maximum number of steps must be greater than minimum number of steps.")
    step_fn = _select_step(method, adaptive=True)

    f32 = torch.float32
    atol = torch.tensor(kwargs.get("atol", 1e-6), dtype=f32)
    rtol = torch.tensor(kwargs.get("rtol", 1e-4), dtype=f32)
    step_size = torch.tensor((stop_time - start_time) / float(min_steps), dtype=f32)
    step = torch.tensor(0.0, dtype=f32)
    count_not_accepted = 0.0
    time = start_time
    sign_dir = float(np.sign(stop_time - start_time))

    k1 = fn(start_time, **state)  # FSAL, :323

    def cond(_state, _time, _step_size, _step):
        time_remaining = sign_dir * (stop_time - (_time + _step_size))
        step_lt_min = bool(_step < float(min_steps))
        step_lt_max = bool(_step < float(max_steps))
        all_nans = _check_all_nans(_state)
        end_now = step_lt_min or (bool(torch.all(torch.as_tensor(time_remaining) > 0)) and step_lt_max)
        return (not all_nans) and end_now

    while cond(state, time, step_size, step):
        time_remaining = torch.abs(torch.as_tensor(stop_time - time, dtype=f32))
        min_step_size = time_remaining / (max_steps - step)
        max_step_size = time_remaining / torch.maximum(min_steps - step, torch.tensor(1.0))
        h = torch.sign(step_size) * torch.clip(torch.abs(step_size), min_step_size, max_step_size)

        new_state, new_time, new_k1, err = step_fn(fn, state, time, h, k1=k1)

        max_abs = None
        for v in state.values():
            m = torch.max(torch.abs(v))
            max_abs = m if max_abs is None else torch.maximum(max_abs, m)
        scale = atol + rtol * max_abs
        error_ratio = err / scale
        new_step_size = h * torch.clip(0.9 * (1.0 / (error_ratio + 1e-12)) ** 0.2, 0.2, 5.0)
        new_step_size = torch.sign(new_step_size) * torch.clip(torch.abs(new_step_size), min_step_size, max_step_size)

        too_big = bool(error_ratio > 1.0)
        at_min = bool(torch.abs(h) <= torch.abs(min_step_size))
        accepted = (not too_big) or at_min

        if accepted:
            state, time, k1 = new_state, new_time, new_k1
            step = step + 1.0
        else:
            count_not_accepted += 1.0
        step_size = new_step_size

    if _check_all_nans(state):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")

    time_diff = stop_time - time
    time_remaining = sign_dir * time_diff
    if bool(torch.all(torch.as_tensor(time_remaining) > 0)):
        state, time, _, _ = step_fn(fn, state, time, time_diff, k1=k1)
        step = step + 1.0

    if info is not None:
        info["steps"] = float(step)
        info["rejected"] = float(count_not_accepted)
    return state


def integrate_scipy(fn, state, start_time, stop_time, scipy_kwargs=None, info=None):
    # integrate.py:409-452.  The reference's Adapter().concatenate(keys, into="x", axis=-1).convert_dtype(float32, float64)
    # is: every key flattened, concatenated in key order, float64 on the way to scipy and float32 on the way back.
    import scipy.integrate
    scipy_kwargs = scipy_kwargs or {}
    keys = list(state.keys())
    shapes = {k: tuple(torch.as_tensor(state[k]).shape) for k in keys}
    sizes = {k: int(np.prod(shapes[k])) for k in keys}

    def state_to_vector(st):
        return np.concatenate([torch.as_tensor(st[k]).detach().cpu().numpy().reshape(-1) for k in keys],
                              axis=-1).astype(np.float32).astype(np.float64)

    def vector_to_state(x):
        out, o = {}, 0
        for k in keys:
            out[k] = torch.from_numpy(np.ascontiguousarray(x[o:o + sizes[k]].astype(np.float32))).reshape(shapes[k])
            o += sizes[k]
        return out

    def scipy_wrapper_fn(time, x):
        st = vector_to_state(x)
        time = torch.tensor(time, dtype=torch.float32)
        return state_to_vector(fn(time, **st))

    result = scipy.integrate.solve_ivp(scipy_wrapper_fn, (start_time, stop_time), state_to_vector(state), **scipy_kwargs)
    if info is not None:
        info["steps"], info["evals"] = int(result.t.shape[0]) - 1, int(result.nfev)
    return vector_to_state(result.y[:, -1])


def integrate(fn, state, start_time=None, stop_time=None, min_steps=50, max_steps=10_000, steps=100, method="rk45",
              **kwargs):
    # integrate.py:455-489
    if isinstance(steps, str) and steps in ["adaptive", "dynamic"]:
        if start_time is None or stop_time is None:
            raise ValueError(
                "Please provide start_time and stop_time for the integration, was "
                f"'start_time={start_time}', 'stop_time={stop_time}'."
            )
        if method == "scipy":  # :473-478 (the two "has no effect" warnings are log lines only)
            return integrate_scipy(fn, state, start_time, stop_time, info=kwargs.get("info"))
        return integrate_adaptive(fn, state, start_time, stop_time, min_steps, max_steps, method, **kwargs)
    elif isinstance(steps, int):
        if start_time is None or stop_time is None:
            raise ValueError(
                "Please provide start_time and stop_time for the integration, was "
                f"'start_time={start_time}', 'stop_time={stop_time}'."
            )
        return integrate_fixed(fn, state, start_time, stop_time, steps, method, **kwargs)
    elif isinstance(steps, (list, tuple, np.ndarray)) or torch.is_tensor(steps):
        return integrate_scheduled(fn, state, steps, method, **kwargs)
    else:
        raise RuntimeError(f"Type or value of `steps` not understood (steps={steps})")


# ---------------------------------------------------------------------------------------------------------------------
# Stochastic half (integrate.py:495-1451): euler_maruyama, two_step_adaptive, sea, shark, langevin and the corrector.
# `noise_fn(i, state) -> dict` replaces generate_noise (what the reference's z_history does on the JAX backend) so that a
# parity run can feed the SAME normals to the oracle and to the CUDA path; default: torch.randn from `generator`.
STOCHASTIC_METHODS = ["euler_maruyama", "sea", "shark", "two_step_adaptive", "langevin"]
TRACE = None  # tests may set this to a list: (time, step, E2, accepted) per two_step_adaptive trial


def generate_noise(state: StateDict, generator=None) -> StateDict:
    # integrate.py:495-499
    return {k: torch.randn(v.shape, dtype=v.dtype, generator=generator) for k, v in state.items()}


def stochastic_adaptive_step_size_controller(state, drift, adaptive_factor, min_step_size=-float("inf"),
                                             max_step_size=float("inf")):
    # integrate.py:502-537
    state_norm = torch.stack([torch.linalg.vector_norm(state[k], ord=2, dim=-1).max() for k in state])
    drift_norm = torch.stack([torch.linalg.vector_norm(drift[k], ord=2, dim=-1).max() for k in state])
    max_state_norm = torch.maximum(torch.tensor(1.0), state_norm.max() ** 2)
    max_drift_norm = torch.maximum(torch.tensor(1.0), drift_norm.max() ** 2)
    new_step_size = max_state_norm / max_drift_norm * adaptive_factor
    return torch.clip(new_step_size, min_step_size, max_step_size)


def euler_maruyama_step(drift_fn, diffusion_fn, state, time, step_size, noise, use_adaptive_step_size=False,
                        min_step_size=-float("inf"), max_step_size=float("inf"), **_):
    # integrate.py:540-596
    drift = drift_fn(time, **state)
    diffusion = diffusion_fn(time, **state)
    new_step_size = step_size
    if use_adaptive_step_size:
        sign_step = torch.sign(torch.as_tensor(step_size))
        new_step_size = stochastic_adaptive_step_size_controller(state, drift, max_step_size, min_step_size, max_step_size)
        new_step_size = sign_step * torch.abs(new_step_size)
    sqrt_step_size = torch.sqrt(torch.abs(torch.as_tensor(new_step_size, dtype=torch.float32)))
    new_state = {}
    for key, d in drift.items():
        base = state[key] + new_step_size * d
        if key in diffusion:
            base = base + diffusion[key] * sqrt_step_size * noise[key]
        new_state[key] = base
    if use_adaptive_step_size:
        return new_state, time + new_step_size, new_step_size, state
    return new_state, time + new_step_size, new_step_size


def two_step_adaptive_step(drift_fn, diffusion_fn, state, time, step_size, noise, last_state=None,
                           use_adaptive_step_size=True, min_step_size=-float("inf"), max_step_size=float("inf"),
                           e_rel=0.1, e_abs=None, r=0.9, adapt_safety=0.9, **_):
    # integrate.py:599-743
    state_euler, time_mid, _ = euler_maruyama_step(drift_fn, diffusion_fn, state, time, step_size, noise,
                                                   use_adaptive_step_size=False)
    drift_mid = drift_fn(time_mid, **state_euler)
    diffusion_mid = diffusion_fn(time_mid, **state_euler)
    sqrt_step_size = torch.sqrt(torch.abs(torch.as_tensor(step_size, dtype=torch.float32)))
    state_euler_mid = {}
    for key, d in drift_mid.items():
        base = state[key] + step_size * d
        if key in diffusion_mid:
            base = base + diffusion_mid[key] * sqrt_step_size * noise[key]
        state_euler_mid[key] = base
    state_heun = {key: 0.5 * (state_euler[key] + state_euler_mid[key]) for key in state}
    if not use_adaptive_step_size:
        return state_heun, time_mid, step_size
    if e_abs is None:
        e_abs = 0.02576
    at_min_step = bool(torch.abs(torch.as_tensor(step_size)) <= min_step_size)
    max_error = torch.tensor(0.0)
    for key in state:
        error_estimate = torch.abs(state_heun[key] - state_euler[key])
        delta = torch.maximum(torch.tensor(e_abs, dtype=torch.float32),
                              e_rel * torch.maximum(torch.abs(state_euler[key]), torch.abs(last_state[key])))
        max_error = torch.maximum(max_error, torch.max(error_estimate / (delta + 1e-10)))
    E2 = 1 * max_error
    accepted = bool(E2 <= 1.0) or at_min_step
    E2_safe = torch.maximum(E2, torch.tensor(1e-10))
    new_step_candidate = step_size * (adapt_safety * torch.pow(E2_safe, -r))
    new_step_size = torch.clip(torch.abs(new_step_candidate), min_step_size, max_step_size)
    new_step_size = torch.sign(torch.as_tensor(step_size)) * new_step_size
    if TRACE is not None:
        TRACE.append((float(time), float(step_size), float(E2), accepted))
    if accepted:
        return state_heun, time_mid, new_step_size, state_euler
    return state, time, new_step_size, state


def compute_levy_area(state, diffusion, noise, noise_aux, step_size):
    # integrate.py:746-762
    step_size_abs = torch.abs(torch.as_tensor(step_size, dtype=torch.float32))
    sqrt_step_size = torch.sqrt(step_size_abs)
    inv_sqrt3 = torch.tensor(1.0 / np.sqrt(3.0), dtype=torch.float32)
    H = {}
    for k in state:
        if k in diffusion:
            H[k] = 0.5 * step_size_abs * noise[k] + 0.5 * step_size_abs * sqrt_step_size * inv_sqrt3 * noise_aux[k]
        else:
            H[k] = torch.zeros_like(state[k])
    return H


def sea_step(drift_fn, diffusion_fn, state, time, step_size, noise, noise_aux, **_):
    # integrate.py:765-822 (shifted Euler for additive noise)
    diffusion = diffusion_fn(time, **state)
    sqrt_step_size = torch.sqrt(torch.abs(torch.as_tensor(step_size, dtype=torch.float32)))
    la = compute_levy_area(state, diffusion, noise, noise_aux, step_size)
    shifted = {k: (x + diffusion[k] * (0.5 * sqrt_step_size * noise[k] + la[k])) if k in diffusion else x
               for k, x in state.items()}
    drift_shifted = drift_fn(time, **shifted)
    new_state = {}
    for key, d in drift_shifted.items():
        base = state[key] + step_size * d
        if key in diffusion:
            base = base + diffusion[key] * sqrt_step_size * noise[key]
        new_state[key] = base
    return new_state, time + step_size, step_size


def shark_step(drift_fn, diffusion_fn, state, time, step_size, noise, noise_aux, **_):
    # integrate.py:825-921 (shifted additive-noise Runge-Kutta, two drift evaluations per step)
    h, t = step_size, time
    sqrt_h_mag = torch.sqrt(torch.abs(torch.as_tensor(h, dtype=torch.float32)))
    diffusion = diffusion_fn(t, **state)
    la = compute_levy_area(state, diffusion, noise, noise_aux, step_size)
    y_tilde_k = {k: (state[k] + diffusion[k] * la[k]) if k in diffusion else state[k] for k in state}
    f_tilde_k = drift_fn(t, **y_tilde_k)
    g_tilde_k = diffusion_fn(t, **y_tilde_k)
    y_tilde_mid = {}
    for k in state:
        drift_part = (5.0 / 6.0) * f_tilde_k[k] * h
        sto_part = (5.0 / 6.0) * g_tilde_k[k] * sqrt_h_mag * noise[k] if k in g_tilde_k else torch.zeros_like(state[k])
        y_tilde_mid[k] = y_tilde_k[k] + drift_part + sto_part
    f_tilde_mid = drift_fn(t + 5.0 / 6.0 * h, **y_tilde_mid)
    g_tilde_mid = diffusion_fn(t + 5.0 / 6.0 * h, **y_tilde_mid)
    new_state = {}
    for k in state:
        det = state[k] + (2.0 / 5.0) * f_tilde_k[k] * h + (3.0 / 5.0) * f_tilde_mid[k] * h
        sto1 = g_tilde_k[k] * ((2.0 / 5.0) * sqrt_h_mag * noise[k] + (6.0 / 5.0) * la[k]) if k in g_tilde_k else torch.zeros_like(det)
        sto2 = g_tilde_mid[k] * ((3.0 / 5.0) * sqrt_h_mag * noise[k] - (6.0 / 5.0) * la[k]) if k in g_tilde_mid else torch.zeros_like(det)
        new_state[k] = det + sto1 + sto2
    return new_state, t + h, h


def _apply_corrector(new_state, new_time, corrector_steps, score_fn, draw, step_size_factor=0.01, noise_schedule=None):
    # integrate.py:924-961 (annealed Langevin corrector, Song et al. 2020)
    for _ in range(corrector_steps):
        score = score_fn(new_time, **new_state)
        z_corr = draw(new_state)
        log_snr_t = noise_schedule.get_log_snr(torch.as_tensor(new_time, dtype=torch.float32), training=False)
        alpha_t, _ = noise_schedule.get_alpha_sigma(log_snr_t)
        for k in new_state:
            if k in score:
                z_norm = torch.linalg.vector_norm(z_corr[k], dim=-1, keepdim=True)
                score_norm = torch.clamp_min(torch.linalg.vector_norm(score[k], dim=-1, keepdim=True), 1e-8)
                e = 2.0 * alpha_t * (step_size_factor * z_norm / score_norm) ** 2
                new_state[k] = new_state[k] + e * score[k] + torch.sqrt(2.0 * e) * z_corr[k]
    return new_state


def integrate_langevin(state, start_time, stop_time, steps, score_fn, noise_schedule, draw, step_size_factor=0.01,
                       corrector_steps=0):
    # integrate.py:1167-1251 (annealed Langevin dynamics on a linear time grid)
    if steps <= 0:
        raise ValueError("Number of Langevin steps must be positive.")
    if score_fn is None or noise_schedule is None:
        raise ValueError("score_fn and noise_schedule must be provided.")
    dt = (stop_time - start_time) / float(steps)
    effective_factor = step_size_factor * 100 / np.sqrt(steps)
    i, time = 0, start_time
    while i < steps and not _check_all_nans(state):
        score = score_fn(time, **state)
        log_snr_t = noise_schedule.get_log_snr(torch.as_tensor(time, dtype=torch.float32), training=False)
        _, sigma_t = noise_schedule.get_alpha_sigma(log_snr_t)
        z = draw(state)
        new_state = {}
        for k in state:
            s_k = score.get(k)
            if s_k is None:
                new_state[k] = state[k]
                continue
            e = effective_factor * sigma_t ** 2
            new_state[k] = state[k] + e * s_k + torch.sqrt(2.0 * e) * z[k]
        time = time + dt
        if corrector_steps > 0:
            new_state = _apply_corrector(new_state, time, corrector_steps, score_fn, draw, step_size_factor, noise_schedule)
        state = new_state
        i += 1
    if _check_all_nans(state):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
    return state


def integrate_stochastic(drift_fn, diffusion_fn, state, start_time, stop_time, generator=None, steps=100,
                         method="euler_maruyama", min_steps=50, max_steps=1_000, noise_fn=None, info=None, score_fn=None,
                         corrector_steps=0, noise_schedule=None, step_size_factor=0.01, **kwargs):
    # integrate.py:1254-1451 (dispatch), :964-1046 (fixed), :1049-1164 (adaptive).  `noise_fn(k, state)` is the k-th draw
    # of standard normals of the call, in the reference's order (per step: dW, then the auxiliary normals of sea / shark,
    # then one draw per corrector step)
    is_adaptive = isinstance(steps, str) and steps in ["adaptive", "dynamic"]
    if is_adaptive:
        if start_time is None or stop_time is None:
            raise ValueError("Please provide start_time and stop_time for adaptive integration.")
        if min_steps <= 0 or max_steps <= 0 or max_steps < min_steps:
            raise ValueError("min_steps and max_steps must be positive, and max_steps >= min_steps.")
        initial_step = (stop_time - start_time) / float(min_steps)
        span_mag = abs(stop_time - start_time)
        min_step_size, max_step_size = span_mag / max_steps, span_mag / min_steps
    else:
        if steps <= 0:
            raise ValueError("Number of steps must be positive.")
        initial_step = (stop_time - start_time) / float(int(steps))
        min_step_size = max_step_size = initial_step
    if corrector_steps > 0 and (score_fn is None or noise_schedule is None):
        raise ValueError("Please provide both score_fn and noise_schedule when using corrector_steps > 0.")
    counter = [0]

    def draw(st):
        k = counter[0]
        counter[0] += 1
        return noise_fn(k, st) if noise_fn is not None else generate_noise(st, generator)

    if method == "euler_maruyama":
        step_fn = euler_maruyama_step
    elif method == "two_step_adaptive":
        step_fn = two_step_adaptive_step
    elif method in ("sea", "shark"):
        if is_adaptive:
            raise ValueError(f"{method.upper()} SDE solver does not support adaptive steps.")
        step_fn = sea_step if method == "sea" else shark_step
    elif method == "langevin":
        if is_adaptive:
            raise ValueError("Langevin sampling does not support adaptive steps.")
        return integrate_langevin(state, start_time, stop_time, int(steps), score_fn, noise_schedule, draw, step_size_factor,
                                  corrector_steps)
    else:
        raise TypeError(f"Invalid integration method: {method!r}")
    two_noise = method in ("sea", "shark")
    skw = {k: v for k, v in kwargs.items() if k in ("e_rel", "e_abs", "r", "adapt_safety")}
    f32 = torch.float32
    sign_dir = float(np.sign(stop_time - start_time))
    i, time, step, last = 0, start_time, initial_step, state
    if not is_adaptive:
        while i < int(steps) and not _check_all_nans(state):
            remaining = abs(stop_time - float(time))
            dt = float(np.sign(float(step))) * min(abs(float(step)), remaining)
            noise = draw(state)
            extra = {"noise_aux": draw(state)} if two_noise else {}
            state, time, _ = step_fn(drift_fn, diffusion_fn, state, time, dt, noise, use_adaptive_step_size=False,
                                     min_step_size=min_step_size, max_step_size=min(max_step_size, remaining), **extra, **skw)[:3]
            if corrector_steps > 0:
                state = _apply_corrector(state, time, corrector_steps, score_fn, draw, step_size_factor, noise_schedule)
            step = initial_step
            i += 1
        if _check_all_nans(state):
            raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
        if info is not None:
            info["steps"] = i
        return state
    while True:
        time_remaining = sign_dir * (stop_time - (float(time) + float(step)))
        if not (time_remaining > 0 and i < max_steps) or _check_all_nans(state):
            break
        remaining = abs(stop_time - float(time))
        dt = float(np.sign(float(step))) * min(abs(float(step)), remaining)
        state, time, step, last = step_fn(drift_fn, diffusion_fn, state, time, torch.tensor(dt, dtype=f32), draw(state),
                                          last_state=last, use_adaptive_step_size=True, min_step_size=min_step_size,
                                          max_step_size=min(max_step_size, remaining), **skw)
        if corrector_steps > 0:
            state = _apply_corrector(state, time, corrector_steps, score_fn, draw, step_size_factor, noise_schedule)
        i += 1
    if _check_all_nans(state):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
    time_diff = stop_time - float(time)
    if sign_dir * time_diff > 0:
        state = step_fn(drift_fn, diffusion_fn, state, time, time_diff, draw(state), last_state=last,
                        use_adaptive_step_size=False, min_step_size=min_step_size, max_step_size=sign_dir * time_diff, **skw)[0]
        i += 1
    if info is not None:
        info["steps"] = i
    return state
