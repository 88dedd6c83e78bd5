"""Density path: integrate {"xz", "trace"} with the exact (or Hutchinson) Jacobian trace of the velocity.

  FlowMatching._forward/_inverse(density=True)  <- flow_matching.py:272-284, 316-326
  FlowMatching._velocity_trace                  <- flow_matching.py:238-252
  jacobian_trace / _hutchinson / vjp            <- utils/jacobian/jacobian_trace.py:10-89, utils/jacobian/vjp.py:44-68
  integrate (dict state, all three ODE methods) <- utils/integrate.py:31-214, 217-406, 455-489

The reference runs one forward and D ``torch.autograd.grad`` passes per velocity evaluation.  Here the forward is
``bfk_train_step`` (forward only, activations kept in the workspace) and each reverse pass is ``bfk_train_vjp`` (the
dX chain of the training backward, tcgen05 GEMMs).  The Runge-Kutta combinations are ``bfk_rk_combine``.  The host
decides WHEN to evaluate (the reference's Python control flow, same float arithmetic for the step controller); torch
only supplies device memory, the max-reductions of the controller and the stream.  No CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .integrate import _check_method

# tableaux as the reference writes them (integrate.py:74-96, 107-114, 139-209)
_RK45_C = (0.0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0)
_RK45_A = ((), (1 / 5,), (3 / 40, 9 / 40), (44 / 45, -56 / 15, 32 / 9),
           (19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729),
           (9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656))
_RK45_B5 = ((0, 35 / 384), (2, 500 / 1113), (3, 125 / 192), (4, -2187 / 6784), (5, 11 / 84))
_RK45_B4 = ((0, 5179 / 57600), (2, 7571 / 16695), (3, 393 / 640), (4, -92097 / 339200), (5, 187 / 2100), (6, 1 / 40))
_TSIT5_C = (0.0, 0.161, 0.327, 0.9, 0.9800255409045097, 1.0)
_TSIT5_A = ((), (0.161,), (-0.0084806554923570, 0.3354806554923570),
            (2.897153057105494, -6.359448489975075, 4.362295432869581),
            (5.325864828439257, -11.74888356406283, 7.495539342889836, -0.09249506636175525),
            (5.86145544294270, -12.92096931784711, 8.159367898576159, -0.07158497328140100, -0.02826905039406838))
_TSIT5_B = ((0, 0.09646076681806523), (1, 0.01), (2, 0.4798896504144996), (3, 1.379008574103742),
            (4, -3.290069515436081), (5, 2.324710524099774))
# error = h * (e3 k3 + e5 k5 + e7 k7 + e1 k1 + e2 k2 + e4 k4 + e6 k6) in the reference's order (integrate.py:201-209)
_TSIT5_E = ((2, 0.007880878010261995), (4, 0.5823571654525552), (6, 0.015151515151515152),
            (0, -0.00178001105222577714), (1, -0.0008164344596567469), (3, -0.1447110071732629),
            (5, -0.45808210592918697))


def _combine(y: torch.Tensor, ks, pairs, h: float) -> torch.Tensor:
    """y + h * sum_j c_j k_j   (integrate.py:47-54) through bfk_rk_combine."""
    from .training import _bind
    L = _bind(_lib.lib())
    n = len(pairs)
    ptrs = (C.c_void_p * n)(*[ks[i].data_ptr() for i, _ in pairs])
    coef = (C.c_float * n)(*[float(np.float32(c)) for _, c in pairs])
    out = torch.empty_like(y)
    _lib.check(L.bfk_rk_combine(_lib.ptr(y), ptrs, coef, n, float(np.float32(h)), y.numel(), _lib.ptr(out),
                                _lib.stream_ptr()), "bfk_rk_combine")
    return out


class VelocityTrace:
    """(time, xz) -> (velocity, Jacobian trace) for one density call: forward + D reverse passes in libbfk."""

    def __init__(self, net, conditions, B: int, max_steps=None, target_mask=None, seed=None):
        from .training import train_engine
        if target_mask is not None:
            raise NotImplementedError("density evaluation with target_mask is outside the accelerated path")
        self.net, self.eng, self.B, self.D = net, train_engine(net), int(B), net._D
        self.cond = None
        if net._C:
            cond = conditions.detach().to(torch.float32)
            if cond.ndim != 2 or cond.shape != (B, net._C):
                raise ValueError(f"conditions must have shape ({B}, {net._C}), got {tuple(cond.shape)}")
            if net.unconditional_mode:
                cond = torch.zeros_like(cond)
            self.cond = cond.contiguous()
        self.max_steps = max_steps
        # Hutchinson probes (jacobian_trace.py:58-89): bfk_normal_fill (Philox) keyed by (seed, probe number, GLOBAL row),
        # so a dataset shard draws the probes of the same rows of the single-GPU call
        self.gen = None
        if max_steps is not None and self.D > max_steps:
            self.gen = (int(seed) if seed is not None else torch.initial_seed()) & (2 ** 62 - 1)
        self.probe, self.row_offset = 0, 0
        self.evals = 0

    def _probe(self, like):
        out = torch.empty(self.B, self.D, device=like.device, dtype=torch.float32)
        key = (self.gen * 0x9E3779B97F4A7C15 + (self.probe + 1) * 0xD1B54A32D192ED03) & (2 ** 64 - 1)
        self.probe += 1
        _lib.check(_lib.lib().bfk_normal_fill(_lib.ptr(out), self.B, self.D, int(self.row_offset), None, None, key, None,
                                              _lib.stream_ptr()), "bfk_normal_fill")
        return out

    def affine(self, time: float):
        """velocity = cZ * xz + cP * pred(xz, t_net): (cZ, cP, t_net).  FlowMatching: (0, 1, time)."""
        return 0.0, 1.0, time

    def __call__(self, time: float, xz: torch.Tensor):
        eng, D = self.eng, self.D
        cZ, cP, t_net = self.affine(float(time))
        tvec = torch.full((self.B,), float(np.float32(t_net)), device=xz.device, dtype=torch.float32)
        pred, args, ws = eng.forward_keep(xz, tvec, self.cond)
        trace = torch.zeros(self.B, device=xz.device, dtype=torch.float32)
        if self.gen is None:  # exact: one reverse pass per diagonal entry (jacobian_trace.py:44-46)
            for d in range(D):
                eng.vjp(args, ws, col=d, trace=trace)
        else:  # Hutchinson (jacobian_trace.py:58-89)
            for _ in range(self.max_steps):
                proj = self._probe(xz)
                dx = eng.vjp(args, ws, dout=proj, want_dx=True)
                trace += (dx * proj).sum(-1)
        self.evals += 1
        if cZ != 0.0 or cP != 1.0:
            pred = float(np.float32(cZ)) * xz + float(np.float32(cP)) * pred
            trace = float(np.float32(cZ)) * D + float(np.float32(cP)) * trace
        return pred, trace.unsqueeze(-1)


def _step(fn, state, time, h, method, k1=None, adaptive=False):
    """One euler / rk45 / tsit5 step on the (xz, trace) state (integrate.py:31-44, 57-120, 123-214).
    -> (new_state, new_time, k7 or None, err (0-d device tensor) or None)."""
    nkeys = len(state)
    if method == "euler":
        k = fn(time, state[0])
        return tuple(_combine(state[i], [k[i]], ((0, 1.0),), h) for i in range(nkeys)), time + h, None, None
    cs, A = (_RK45_C, _RK45_A) if method == "rk45" else (_TSIT5_C, _TSIT5_A)
    ks = [k1 if k1 is not None else fn(time, state[0])]
    for s in range(1, 6):
        pairs = tuple(enumerate(A[s]))
        stage = tuple(_combine(state[i], [k[i] for k in ks], pairs, h) for i in range(nkeys))
        ks.append(fn(time + h * cs[s], stage[0]))
    B5 = _RK45_B5 if method == "rk45" else _TSIT5_B
    new_state = tuple(_combine(state[i], [k[i] for k in ks], B5, h) for i in range(nkeys))
    new_time = time + h
    if not adaptive:
        return new_state, new_time, None, None
    k7 = fn(new_time, new_state[0])
    ks.append(k7)
    errs = []
    for i in range(nkeys):
        kk = [k[i] for k in ks]
        if method == "rk45":   # y5 - y4 (integrate.py:107-116)
            e = new_state[i] - _combine(state[i], kk, _RK45_B4, h)
        else:                  # h * sum e_i k_i (integrate.py:201-210)
            e = _combine(torch.zeros_like(state[i]), kk, _TSIT5_E, h)
        errs.append(torch.linalg.vector_norm(e, ord=2, dim=-1).max())
    return new_state, new_time, k7, torch.stack(errs).max()


def _all_nans(state):
    return all(bool(torch.isnan(v).all()) for v in state)


def integrate_density(fn, state, start_time=None, stop_time=None, min_steps: int = 50, max_steps: int = 10_000,
                      steps=100, method: str = "rk45", info: dict | None = None, group=None, **kwargs):
    """bayesflow.utils.integrate (integrate.py:455-489) for the two-key state (xz, trace)."""
    if isinstance(steps, str) and steps in ["adaptive", "dynamic"]:
        if start_time is None or stop_time is None:
            raise ValueError("Please provide start_time and stop_time for the integration, was "
                             f"'start_time={start_time}', 'stop_time={stop_time}'.")
        if method == "scipy":  # integrate.py:409-452: keys flattened and concatenated in order (xz, trace)
            from .integrate import _warn_scipy_step_args, solve_ivp_flat
            _warn_scipy_step_args(min_steps, max_steps)
            out, res = solve_ivp_flat(lambda t, xz, tr: fn(t, xz), list(state), start_time, stop_time)
            if info is not None:
                info.update({"nfev": int(res.nfev), "steps": int(res.t.shape[0]) - 1})
            return tuple(out)
        return _integrate_adaptive(fn, state, float(start_time), float(stop_time), int(min_steps), int(max_steps), method,
                                   info, atol=kwargs.get("atol", 1e-6), rtol=kwargs.get("rtol", 1e-4), group=group)
    if isinstance(steps, (int, np.integer)) and not isinstance(steps, bool):
        if start_time is None or stop_time is None:
            raise ValueError("Please provide start_time and stop_time for the integration, was "
                             f"'start_time={start_time}', 'stop_time={stop_time}'.")
        if steps <= 0:
            raise ValueError("Number of steps must be positive.")
        _check_method(method)
        h = (stop_time - start_time) / steps
        time = start_time
        for _ in range(int(steps)):
            state, time, _, _ = _step(fn, state, time, h, method)
        return state
    if isinstance(steps, (list, tuple, np.ndarray)) or torch.is_tensor(steps):
        _check_method(method)
        arr = np.asarray(steps.detach().cpu() if torch.is_tensor(steps) else steps, dtype=np.float32)
        for i in range(len(arr) - 1):
            state, _, _, _ = _step(fn, state, float(arr[i]), float(arr[i + 1] - arr[i]), method)
        return state
    raise RuntimeError(f"Type or value of `steps` not understood (steps={steps})")


def _integrate_adaptive(fn, state, start_time, stop_time, min_steps, max_steps, method, info, atol, rtol, group=None):
    """integrate.py:291-406 with its global (all rows, both keys) step controller, fp32 controller arithmetic."""
    if max_steps <= min_steps:
        raise ValueError("Maximum number of steps must be greater than minimum number of steps.")
    _check_method(method)
    if method == "euler":
        raise ValueError("Adaptive step sizing is not supported for the 'euler' method.")
    f32 = np.float32
    step_size = f32((stop_time - start_time) / float(min_steps))
    step, rejected, time = 0, 0, start_time
    sign_dir = float(np.sign(stop_time - start_time))
    k1 = fn(start_time, state[0])

    def cond():
        remaining = sign_dir * (stop_time - (time + float(step_size)))
        go = step < min_steps or (remaining > 0 and step < max_steps)
        return go and not _all_nans(state)

    while cond():
        time_remaining = f32(abs(f32(stop_time - time)))
        min_h = f32(time_remaining / f32(max_steps - step))
        max_h = f32(time_remaining / f32(max(min_steps - step, 1.0)))
        h = f32(np.sign(step_size) * np.clip(abs(step_size), min_h, max_h))
        new_state, new_time, new_k1, err = _step(fn, state, time, float(h), method, k1=k1, adaptive=True)
        max_abs = torch.stack([v.abs().max() for v in state]).max()
        ctl = torch.stack([err, max_abs])
        if group is not None:  # rows sharded over ranks: the reference's GLOBAL controller needs the global maxima
            torch.distributed.all_reduce(ctl, op=torch.distributed.ReduceOp.MAX, group=group)
        err_f, max_abs_f = (f32(v) for v in ctl.tolist())  # the one sync per trial step
        scale = f32(f32(atol) + f32(rtol) * max_abs_f)
        ratio = f32(err_f / scale)
        with np.errstate(all="ignore"):
            factor = f32(np.clip(f32(0.9) * f32(f32(1.0) / f32(ratio + f32(1e-12))) ** f32(0.2), 0.2, 5.0))
        new_h = f32(h * factor)
        new_h = f32(np.sign(new_h) * np.clip(abs(new_h), min_h, max_h))
        accepted = (not ratio > 1.0) or (abs(h) <= abs(min_h))
        if accepted:
            state, time, k1 = new_state, new_time, new_k1
            step += 1
        else:
            rejected += 1
        step_size = new_h
    if _all_nans(state):
        raise RuntimeError(f"All values are NaNs in state during integration at {time}.")
    time_diff = stop_time - time
    if sign_dir * time_diff > 0:
        state, time, _, _ = _step(fn, state, time, time_diff, method, k1=k1)  # k7/err of the reference's call are unused
        step += 1
    if info is not None:
        info["steps"], info["rejected"] = float(step), float(rejected)
    return state
