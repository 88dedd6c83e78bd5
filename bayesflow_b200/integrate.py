"""Host-side driver of the ODE loop: the reference's ``bayesflow.utils.integrate`` dispatcher
(bayesflow/utils/integrate.py:455-489) and the fixed / scheduled / adaptive drivers (:217-288, :291-406), with the
same argument names, defaults and error types.  It only computes WHEN the velocity is evaluated (python doubles, cast
to fp32 where the reference casts) and hands the whole loop to the CUDA engine — it never evaluates a network.
``method="scipy"`` (integrate.py:409-452) is the one host-driven loop: scipy's step controller on the host, every
right-hand side through the engine's velocity kernel."""
from __future__ import annotations

from collections.abc import Sequence

import numpy as np
import torch

DETERMINISTIC_METHODS = ["euler", "rk45", "tsit5"]

# stage abscissae c_i (integrate.py:74-89 and :139-142); the 7th entry is the FSAL evaluation at t+h
STAGE_C = {
    "euler": (0.0,),
    "rk45": (0.0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0),
    "tsit5": (0.0, 0.161, 0.327, 0.9, 0.9800255409045097, 1.0, 1.0),
}


def _check_method(method):
    # integrate.py:229-239
    if method in DETERMINISTIC_METHODS:
        return
    if isinstance(method, str):
        raise ValueError(f"Unknown integration method name: {method!r}")
    raise TypeError(f"Invalid integration method: {method!r}")


def fixed_schedule(start_time, stop_time, steps: int, method: str):
    """Evaluation times (execution order) and step sizes of integrate_fixed (integrate.py:217-254): the step size is a
    python double, time accumulates in double and is cast to fp32 only when it meets the network
    (flow_matching.py:226)."""
    if steps <= 0:
        raise ValueError("Number of steps must be positive.")
    _check_method(method)
    c = STAGE_C[method][:1 if method == "euler" else 6]
    h = (stop_time - start_time) / steps
    times = np.empty((steps, len(c)), dtype=np.float64)
    t = start_time
    for i in range(steps):
        for s, cs in enumerate(c):
            times[i, s] = t + h * cs if s else t
        t = t + h
    return times.astype(np.float32).reshape(-1), np.full((steps,), h, dtype=np.float64).astype(np.float32)


def scheduled_schedule(steps, method: str):
    """integrate_scheduled (integrate.py:257-288): t_i = steps[i], h_i = steps[i+1]-steps[i] in the dtype of `steps`."""
    _check_method(method)
    if torch.is_tensor(steps):
        arr = steps.detach().cpu().numpy()
    else:
        arr = np.asarray(steps, dtype=np.float64 if isinstance(steps, (list, tuple)) else None)
    if arr.ndim != 1 or arr.shape[0] < 2:
        raise ValueError("`steps` must be a 1-D sequence of at least two time points.")
    c = np.asarray(STAGE_C[method][:1 if method == "euler" else 6], dtype=arr.dtype)
    h = arr[1:] - arr[:-1]
    times = arr[:-1, None] + h[:, None] * c[None, :]
    times[:, 0] = arr[:-1]
    return times.astype(np.float32).reshape(-1), h.astype(np.float32)


def _warn_scipy_step_args(min_steps, max_steps):
    # integrate.py:474-477 (the dispatcher's own default is min_steps=50, so the first warning fires unless the caller
    # passes 10 - kept as is)
    import logging
    log = logging.getLogger("bayesflow_b200")
    if min_steps != 10:
        log.warning("Setting min_steps has no effect for method 'scipy'")
    if max_steps != 10_000:
        log.warning("Setting max_steps has no effect for method 'scipy'")


def solve_ivp_flat(fn, parts, start_time, stop_time, scipy_kwargs=None):
    """integrate_scipy (integrate.py:409-452): the state tensors are flattened, concatenated into ONE float64 vector in
    key order and handed to ``scipy.integrate.solve_ivp`` with its defaults (RK45, rtol 1e-3, atol 1e-6, one error norm
    over the whole batch); every right-hand-side evaluation converts the vector back to float32 tensors, the time to
    float32, and runs `fn(time, *parts) -> tuple of derivatives` ON THE DEVICE (the kernels evaluate the network; scipy only
    drives the step control on the host).  The final column of the solution comes back as float32."""
    import scipy.integrate
    shapes = [tuple(p.shape) for p in parts]
    sizes = [int(np.prod(sh)) for sh in shapes]
    dev = parts[0].device

    def to_vec(ts):
        return np.concatenate([t.detach().to(torch.float32).reshape(-1).cpu().numpy().astype(np.float64) for t in ts])

    def to_parts(y):
        out, o = [], 0
        for sh, n in zip(shapes, sizes):
            out.append(torch.from_numpy(np.ascontiguousarray(y[o:o + n], dtype=np.float32)).reshape(sh).to(dev))
            o += n
        return out

    def rhs(time, y):
        return to_vec(fn(float(np.float32(time)), *to_parts(y)))

    res = scipy.integrate.solve_ivp(rhs, (start_time, stop_time), to_vec(parts), **(scipy_kwargs or {}))
    return to_parts(res.y[:, -1]), res


def integrate(engine, x: torch.Tensor, start_time=None, stop_time=None, min_steps: int = 50, max_steps: int = 10_000,
              steps=100, method: str = "rk45", **kwargs) -> torch.Tensor:
    """Same dispatch as bayesflow.utils.integrate (integrate.py:455-489) for the single-key state {"xz": x}."""
    if isinstance(steps, str) and steps in ["adaptive", "dynamic"]:
        if start_time is None or stop_time is None:
            raise ValueError(
                "Please provide start_time and stop_time for the integration, was "
                f"'start_time={start_time}', 'stop_time={stop_time}'."
            )
        if method == "scipy":
            _warn_scipy_step_args(min_steps, max_steps)
            (out,), res = solve_ivp_flat(lambda t, xz: (engine.velocity(xz, t),), [engine._check_x(x)], start_time, stop_time)
            engine.info = {"steps": int(res.t.shape[0]) - 1, "evals": int(res.nfev), "rejected": None, "path": "scipy"}
            return out
        return engine.run_adaptive(x, float(start_time), float(stop_time), int(min_steps), int(max_steps), method,
                                   atol=kwargs.get("atol", 1e-6), rtol=kwargs.get("rtol", 1e-4))
    elif isinstance(steps, (int, np.integer)) and not isinstance(steps, bool):
        if start_time is None or stop_time is None:
            raise ValueError(
                "Please provide start_time and stop_time for the integration, was "
                f"'start_time={start_time}', 'stop_time={stop_time}'."
            )
        times, hs = fixed_schedule(float(start_time), float(stop_time), int(steps), method)
        return engine.run_fixed(x, times, hs, method)
    elif isinstance(steps, (Sequence, np.ndarray)) or torch.is_tensor(steps):
        times, hs = scheduled_schedule(steps, method)
        return engine.run_fixed(x, times, hs, method)
    else:
        raise RuntimeError(f"Type or value of `steps` not understood (steps={steps})")
