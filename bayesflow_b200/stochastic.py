"""Reverse-SDE sampling for DiffusionModel: host-driven predictor / corrector steps over the CUDA velocity kernels.

  integrate_stochastic (dispatch, fixed, adaptive)  <- bayesflow/utils/integrate.py:1254-1451, 964-1046, 1049-1164
  euler_maruyama_step / step-size controller        <- utils/integrate.py:540-596, 502-537
  two_step_adaptive_step                            <- utils/integrate.py:599-743   (DiffusionModel's DEFAULT, defaults.py:50)
  compute_levy_area, sea_step, shark_step           <- utils/integrate.py:746-762, 765-822, 825-921
  _apply_corrector, integrate_langevin              <- utils/integrate.py:924-961, 1167-1251
  DiffusionModel SDE branch, diffusion_term         <- networks/inference/diffusion/diffusion_model.py:830-856, 561-595

Every network evaluation is the same fused / layer-wise velocity kernel the ODE path uses (probability-flow velocity
v_ode); the reverse-SDE drift is 2*v_ode - fz*z, the score 2*(fz*z - v_ode)/beta and g = sqrt(beta), with (fz, beta) from
`bfk_schedule_sde_terms` and (alpha, sigma) from `bfk_schedule_alpha_sigma` (double).  State updates are `bfk_rk_combine`
and `bfk_sde_langevin_update`.  The host runs the reference's Python control flow (time is shared by all rows) and reads
the error / norm maxima of the adaptive controllers once per step.  The Gaussian increments are `bfk_normal_fill`
(Philox4x32-10) keyed by (seed, draw number, GLOBAL row): a sharded fixed-step run draws the rows one big call would;
`noise_fn(k, shape)` replaces the k-th draw for parity runs (draw order per step: dW, the auxiliary normals of sea / shark,
then one draw per corrector step - the order of the reference's keras.random calls).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .density import _combine

BUILT = ("euler_maruyama", "two_step_adaptive", "sea", "shark", "langevin")
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

    def alpha_sigma(self, time):
        a, s = C.c_double(), C.c_double()
        _lib.check(self.L.bfk_schedule_alpha_sigma(C.byref(self.sched), float(np.float32(time)), C.byref(a), C.byref(s)),
                   "bfk_schedule_alpha_sigma")
        return a.value, s.value

    def langevin_update(self, time, x, z, a, b, per_row):
        """x + e*score(time, x) + sqrt(2e)*z (integrate.py:949-959, 1228-1236); the score comes from one velocity call."""
        # with a target mask the reference's score_fn still sees the UNMASKED score (diffusion_model.py:845-849), so the
        # velocity comes from a run built without the mask (networks.py sets `score_run`)
        run = getattr(self, "score_run", None) or self.run
        if self.mask is not None and run is self.run:
            raise RuntimeError("ReverseSDE: a target mask needs `score_run` (the unmasked velocity) for Langevin updates")
        v = run.velocity(x, float(np.float32(time)))
        self.evals += 1
        fz, beta = self.terms(time)
        out = torch.empty_like(x)
        _lib.check(self.L.bfk_sde_langevin_update(_lib.ptr(x), _lib.ptr(v), _lib.ptr(z), x.shape[0], x.shape[1], fz, beta,
                                                  float(a), float(b), int(per_row), _lib.ptr(out), _lib.stream_ptr()),
                   "bfk_sde_langevin_update")
        return out

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


def _lin(x, terms):
    """x + sum_j c_j * t_j ; c_j python floats (one bfk_rk_combine launch) or (D,) vectors when a target mask is set."""
    if all(isinstance(c, float) for _, c in terms):
        return _combine(x, [t for t, _ in terms], tuple((j, c) for j, (_, c) in enumerate(terms)), 1.0)
    out = x
    for t, c in terms:
        out = out + c * t
    return out


def _levy(h):
    """H = hn*noise + ha*noise_aux, the space-time Levy area of integrate.py:746-762 -> (hn, ha, sqrt|h|)."""
    habs = np.float32(abs(h))
    sq = np.float32(np.sqrt(habs))
    return float(np.float32(0.5) * habs), float(np.float32(0.5) * habs * sq * np.float32(1.0 / np.sqrt(3.0))), float(sq)


def sea_step(sde, x, time, h, noise, noise_aux, **_):
    """integrate.py:765-822, shifted Euler for additive noise (one drift evaluation per step)."""
    g = sde.diffusion(time, x)
    hn, ha, sq = _levy(h)
    shifted = _lin(x, [(noise, g * (0.5 * sq + hn)), (noise_aux, g * ha)])
    drift = sde.drift(time, shifted)
    return _lin(x, [(drift, float(h)), (noise, g * sq)]), time + h, h


def shark_step(sde, x, time, h, noise, noise_aux, **_):
    """integrate.py:825-921, shifted additive-noise Runge-Kutta (two drift evaluations per step)."""
    g = sde.diffusion(time, x)
    hn, ha, sq = _levy(h)
    y_k = _lin(x, [(noise, g * hn), (noise_aux, g * ha)])
    f_k = sde.drift(time, y_k)
    t_mid = time + 5.0 / 6.0 * h
    y_mid = _lin(y_k, [(f_k, 5.0 / 6.0 * float(h)), (noise, g * (5.0 / 6.0 * sq))])
    f_mid = sde.drift(t_mid, y_mid)
    g_mid = sde.diffusion(t_mid, x)
    return _lin(x, [(f_k, 0.4 * float(h)), (f_mid, 0.6 * float(h)),
                    (noise, g * (0.4 * sq + 1.2 * hn) + g_mid * (0.6 * sq - 1.2 * hn)),
                    (noise_aux, (g - g_mid) * (1.2 * ha))]), time + h, h


def apply_corrector(sde, x, time, corrector_steps, draw, step_size_factor=0.01):
    """integrate.py:924-961: annealed Langevin corrector (Song et al. 2020), e = 2*alpha_t*(r*||z||/||score||)^2 per row."""
    for _ in range(int(corrector_steps)):
        z = draw()
        alpha, _ = sde.alpha_sigma(time)
        x = sde.langevin_update(time, x, z, alpha, step_size_factor, per_row=True)
    return x


def integrate_langevin(sde, x, start_time, stop_time, steps, draw, step_size_factor=0.01, corrector_steps=0):
    """integrate.py:1167-1251: annealed Langevin dynamics on a linear time grid, e = factor*100/sqrt(steps)*sigma_t^2."""
    if steps <= 0:
        raise ValueError("Number of Langevin steps must be positive.")
    dt = (stop_time - start_time) / float(steps)
    effective_factor = step_size_factor * 100 / np.sqrt(steps)
    i, time = 0, start_time
    while i < steps and not _all_nan(x):
        _, sigma = sde.alpha_sigma(time)
        z = draw()
        x = sde.langevin_update(time, x, z, effective_factor * float(np.float32(sigma)) ** 2, 0.0, per_row=False)
        time = time + dt
        if corrector_steps > 0:
            x = apply_corrector(sde, x, time, corrector_steps, draw, step_size_factor)
        i += 1
    if _all_nan(x):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
    return x, i


_MIX_A, _MIX_B, _M64 = 0x9E3779B97F4A7C15, 0xD1B54A32D192ED03, (1 << 64) - 1


def integrate_stochastic(sde, x, start_time, stop_time, steps=100, method="euler_maruyama", min_steps=50, max_steps=1_000,
                         seed=None, noise_fn=None, info=None, corrector_steps=0, step_size_factor=0.01, row_offset=0,
                         **kwargs):
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
    corrector_steps = int(corrector_steps or 0)

    if noise_fn is None:
        if isinstance(seed, torch.Generator):
            seed = int(seed.initial_seed())
        base_seed = int(seed) if seed is not None else torch.seed() & (2 ** 62 - 1)
    counter = [0]
    L = _lib.lib()

    def draw():
        """the k-th standard-normal draw of the call (Philox keyed by (seed, k), counter = global element index)"""
        k = counter[0]
        counter[0] += 1
        if noise_fn is not None:
            return noise_fn(k, x.shape).to(device=x.device, dtype=torch.float32).contiguous()
        out = torch.empty(x.shape, device=x.device, dtype=torch.float32)
        key = (base_seed * _MIX_A + (k + 1) * _MIX_B) & _M64
        _lib.check(L.bfk_normal_fill(_lib.ptr(out), x.shape[0], x.shape[1], int(row_offset), None, None, key, None,
                                     _lib.stream_ptr()), "bfk_normal_fill")
        return out

    if method == "euler_maruyama":
        step_fn, skw = euler_maruyama_step, {}
    elif method == "two_step_adaptive":
        step_fn = two_step_adaptive_step
        skw = {k: kwargs[k] for k in ("e_rel", "e_abs", "r", "adapt_safety") if k in kwargs}
    elif method in ("sea", "shark"):
        if is_adaptive:
            raise ValueError(f"{method.upper()} SDE solver does not support adaptive steps.")
        step_fn, skw = (sea_step if method == "sea" else shark_step), {}
    elif method == "langevin":
        if is_adaptive:
            raise ValueError("Langevin sampling does not support adaptive steps.")
        x, i = integrate_langevin(sde, x, start_time, stop_time, int(steps), draw, step_size_factor, corrector_steps)
        if info is not None:
            info.update({"steps": i, "evals": sde.evals, "method": method, "draws": counter[0]})
        return x
    else:
        raise TypeError(f"Invalid integration method: {method!r}")
    two_noise = method in ("sea", "shark")

    sign_dir = float(np.sign(stop_time - start_time))
    i, time, step, last = 0, float(start_time), float(initial_step), x
    if not is_adaptive:
        while i < int(steps) and not _all_nan(x):
            remaining = abs(stop_time - time)
            dt = float(np.sign(step)) * min(abs(step), remaining)
            noise = draw()
            extra = {"noise_aux": draw()} if two_noise else {}
            x, time = step_fn(sde, x, time, dt, noise, adaptive=False, min_h=min_h, max_h=min(max_h, remaining), **extra,
                              **skw)[:2]
            if corrector_steps > 0:
                x = apply_corrector(sde, x, time, corrector_steps, draw, step_size_factor)
            i += 1
    else:
        while sign_dir * (stop_time - (time + step)) > 0 and i < max_steps and not _all_nan(x):
            remaining = abs(stop_time - time)
            dt = float(np.float32(np.sign(step) * min(abs(step), remaining)))
            args = dict(adaptive=True, min_h=min_h, max_h=min(max_h, remaining), **skw)
            if step_fn is two_step_adaptive_step:
                args["last"] = last
            x, time, step, last = step_fn(sde, x, time, dt, draw(), **args)
            if corrector_steps > 0:
                # a rejected two_step_adaptive trial hands back ONE state as both the new and the previous state
                # (integrate.py:734-740) and the eager backends' corrector updates that dict in place (:959): the
                # controller's `last_state` is then the corrected state.  Kept, as oracle/integrate.py keeps it.
                same = last is x
                x = apply_corrector(sde, x, time, corrector_steps, draw, step_size_factor)
                if same:
                    last = x
            i += 1
    if _all_nan(x):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
    if is_adaptive:
        time_diff = stop_time - time
        if sign_dir * time_diff > 0:
            args = dict(adaptive=False, min_h=min_h, max_h=sign_dir * time_diff, **skw)
            if step_fn is two_step_adaptive_step:
                args["last"] = last
            x = step_fn(sde, x, time, time_diff, draw(), **args)[0]
            i += 1
    if info is not None:
        info.update({"steps": i, "evals": sde.evals, "method": method, "draws": counter[0]})
    return x
