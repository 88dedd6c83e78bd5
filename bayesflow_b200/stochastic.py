"""Reverse-SDE sampling for DiffusionModel: host-driven predictor steps over the CUDA velocity kernels.

  integrate_stochastic (dispatch, fixed, adaptive)  <- bayesflow/utils/integrate.py:1254-1451, 964-1046, 1049-1164
  euler_maruyama_step / step-size controller        <- utils/integrate.py:540-596, 502-537
  two_step_adaptive_step                            <- utils/integrate.py:599-743   (DiffusionModel's DEFAULT, defaults.py:50)
  DiffusionModel SDE branch, diffusion_term         <- networks/inference/diffusion/diffusion_model.py:830-856, 561-595

Every network evaluation is the same fused / layer-wise velocity kernel the ODE path uses (probability-flow velocity
v_ode); the reverse-SDE drift is 2*v_ode - fz*z and g = sqrt(beta) with (fz, beta) from `bfk_schedule_sde_terms`
(double).  State updates are `bfk_rk_combine`.  The host runs the reference's Python control flow (time is shared by all
rows), reads the error / norm maxima of the adaptive controllers once per step, and draws the Gaussian increments with
torch's device generator (parity with the reference is statistical, or exact when `noise_fn` feeds fixed normals).
Not built: sea, shark, langevin, corrector steps (remainder).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .density import _combine

BUILT = ("euler_maruyama", "two_step_adaptive")
TRACE = None  # debugging / tests: set to a list to record (time, step, E2, accepted) per two_step_adaptive trial


def _all_nan(x):
    return bool(torch.isnan(x).all())


class ReverseSDE:
    """drift(time, x), diffusion(time) for one sampling call."""

    def __init__(self, net, run, target_mask=None):
        self.run, self.L, self.sched = run, _lib.lib(), net._schedule_struct()
        self.L.bfk_schedule_sde_terms.argtypes = [C.POINTER(_lib.Schedule), C.c_double, C.POINTER(C.c_double),
                                                  C.POINTER(C.c_double)]
        self.mask = None
        if target_mask is not None:
            self.mask = torch.as_tensor(target_mask, dtype=torch.float32).reshape(-1)
        self.evals = 0

    def terms(self, time):
        fz, beta = C.c_double(), C.c_double()
        _lib.check(self.L.bfk_schedule_sde_terms(C.byref(self.sched), float(np.float32(time)), C.byref(fz), C.byref(beta)),
                   "bfk_schedule_sde_terms")
        return fz.value, beta.value

    def drift(self, time, x):
        """f - g^2 * score (diffusion_model.py:549-552) = 2 * v_ode - fz * x."""
        v = self.run.velocity(x, float(np.float32(time)))  # already zero where the target is fixed
        self.evals += 1
        fz, _ = self.terms(time)
        d = _combine(v, [v, x], ((0, 1.0), (1, -fz)), 1.0)
        return d if self.mask is None else d * self.mask.to(d.device)

    def diffusion(self, time, like):
        """g(t) as a python float, or a (D,) vector when a target mask zeroes some coordinates (:561-595)."""
        g = float(np.sqrt(np.float32(self.terms(time)[1])))
        return g if self.mask is None else g * self.mask.to(like.device)


def _em_update(x, drift, g, h, noise):
    """x + h*drift + g*sqrt(|h|)*noise (integrate.py:586-591)."""
    s = float(np.sqrt(np.float32(abs(h))))
    if isinstance(g, float):
        return _combine(x, [drift, noise], ((0, float(h)), (1, g * s)), 1.0)
    return _combine(x, [drift], ((0, 1.0),), float(h)) + (g * s) * noise


def _row_norm_max(x):
    return torch.linalg.vector_norm(x, ord=2, dim=-1).max()


def euler_maruyama_step(sde, x, time, h, noise, adaptive=False, min_h=-float("inf"), max_h=float("inf")):
    """integrate.py:540-596 -> (new_x, new_time, new_h[, prev_x])."""
    drift = sde.drift(time, x)
    g = sde.diffusion(time, x)
    new_h = h
    if adaptive:  # tamed-Euler controller, integrate.py:502-537 (one device read)
        sn, dn = (np.float32(v) for v in torch.stack([_row_norm_max(x), _row_norm_max(drift)]).tolist())
        cand = np.float32(max(np.float32(1.0), sn * sn) / max(np.float32(1.0), dn * dn) * np.float32(max_h))
        new_h = float(np.sign(h) * abs(np.clip(cand, min_h, max_h)))
    new_x = _em_update(x, drift, g, new_h, noise)
    if adaptive:
        return new_x, time + new_h, new_h, x
    return new_x, time + new_h, new_h


def two_step_adaptive_step(sde, x, time, h, noise, last=None, adaptive=True, min_h=-float("inf"), max_h=float("inf"),
                           e_rel=0.1, e_abs=None, r=0.9, adapt_safety=0.9):
    """integrate.py:599-743 (Jolicoeur-Martineau et al. 2021)."""
    x_euler, t_mid, _ = euler_maruyama_step(sde, x, time, h, noise)
    x_mid = _em_update(x, sde.drift(t_mid, x_euler), sde.diffusion(t_mid, x), h, noise)
    x_heun = _combine(x_euler, [x_euler, x_mid], ((0, -0.5), (1, 0.5)), 1.0)
    if not adaptive:
        return x_heun, t_mid, h
    if e_abs is None:
        e_abs = 0.02576
    at_min = bool(np.float32(abs(h)) <= np.float32(min_h))  # fp32 comparison as the reference's tensor op does
    delta = torch.clamp_min(e_rel * torch.maximum(x_euler.abs(), last.abs()), e_abs)
    E2 = np.float32(float(((x_heun - x_euler).abs() / (delta + 1e-10)).max()))  # the one device read per step
    accepted = bool(E2 <= 1.0) or at_min
    cand = np.float32(h) * np.float32(adapt_safety * np.power(max(E2, np.float32(1e-10)), np.float32(-r)))
    new_h = float(np.sign(h) * np.clip(abs(cand), min_h, max_h))
    if TRACE is not None:
        TRACE.append((float(time), float(h), float(E2), accepted))
    if accepted:
        return x_heun, t_mid, new_h, x_euler
    return x, time, new_h, x


def integrate_stochastic(sde, x, start_time, stop_time, steps=100, method="euler_maruyama", min_steps=50, max_steps=1_000,
                         seed=None, noise_fn=None, info=None, corrector_steps=0, **kwargs):
    """integrate.py:1254-1451 for the single-key state {"xz": x}."""
    is_adaptive = isinstance(steps, str) and steps in ["adaptive", "dynamic"]
    if is_adaptive:
        if start_time is None or stop_time is None:
            raise ValueError("Please provide start_time and stop_time for adaptive integration.")
        if min_steps <= 0 or max_steps <= 0 or max_steps < min_steps:
            raise ValueError("min_steps and max_steps must be positive, and max_steps >= min_steps.")
        initial_step = (stop_time - start_time) / float(min_steps)
        span = abs(stop_time - start_time)
        min_h, max_h = span / max_steps, span / min_steps
    else:
        if steps <= 0:
            raise ValueError("Number of steps must be positive.")
        initial_step = (stop_time - start_time) / float(int(steps))
        min_h = max_h = initial_step
    if corrector_steps:
        raise NotImplementedError("corrector steps are outside the accelerated path")
    if method == "euler_maruyama":
        step_fn, skw = euler_maruyama_step, {}
    elif method == "two_step_adaptive":
        step_fn = two_step_adaptive_step
        skw = {k: kwargs[k] for k in ("e_rel", "e_abs", "r", "adapt_safety") if k in kwargs}
    elif method in ("sea", "shark"):
        if is_adaptive:
            raise ValueError(f"{method.upper()} SDE solver does not support adaptive steps.")
        raise NotImplementedError(f"method={method!r} is not built on the accelerated path; built: {BUILT}")
    elif method == "langevin":
        if is_adaptive:
            raise ValueError("Langevin sampling does not support adaptive steps.")
        raise NotImplementedError(f"method='langevin' is not built on the accelerated path; built: {BUILT}")
    else:
        raise TypeError(f"Invalid integration method: {method!r}")
    gen = None
    if noise_fn is None:
        gen = seed if isinstance(seed, torch.Generator) else torch.Generator(device=x.device)
        if not isinstance(seed, torch.Generator):
            gen.manual_seed(int(seed) if seed is not None else torch.seed() & (2 ** 62 - 1))

    def noise_of(i):
        if noise_fn is not None:
            return noise_fn(i, x.shape).to(device=x.device, dtype=torch.float32).contiguous()
        return torch.randn(x.shape, device=x.device, dtype=torch.float32, generator=gen)

    sign_dir = float(np.sign(stop_time - start_time))
    i, time, step, last = 0, float(start_time), float(initial_step), x
    if not is_adaptive:
        while i < int(steps) and not _all_nan(x):
            remaining = abs(stop_time - time)
            dt = float(np.sign(step)) * min(abs(step), remaining)
            x, time = step_fn(sde, x, time, dt, noise_of(i), adaptive=False, min_h=min_h, max_h=min(max_h, remaining), **skw)[:2]
            i += 1
    else:
        while sign_dir * (stop_time - (time + step)) > 0 and i < max_steps and not _all_nan(x):
            remaining = abs(stop_time - time)
            dt = float(np.float32(np.sign(step) * min(abs(step), remaining)))
            args = dict(adaptive=True, min_h=min_h, max_h=min(max_h, remaining), **skw)
            if step_fn is two_step_adaptive_step:
                args["last"] = last
            x, time, step, last = step_fn(sde, x, time, dt, noise_of(i), **args)
            i += 1
    if _all_nan(x):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
    if is_adaptive:
        time_diff = stop_time - time
        if sign_dir * time_diff > 0:
            args = dict(adaptive=False, min_h=min_h, max_h=sign_dir * time_diff, **skw)
            if step_fn is two_step_adaptive_step:
                args["last"] = last
            x = step_fn(sde, x, time, time_diff, noise_of(i), **args)[0]
            i += 1
    if info is not None:
        info.update({"steps": i, "evals": sde.evals, "method": method})
    return x
