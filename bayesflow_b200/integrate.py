"""Host-side driver of the ODE loop: the reference's ``bayesflow.utils.integrate`` dispatcher
(bayesflow/utils/integrate.py:455-489) and the fixed / scheduled / adaptive drivers (:217-288, :291-406), with the
same argument names, defaults and error types.  It only computes WHEN the velocity is evaluated (python doubles, cast
to fp32 where the reference casts) and hands the whole loop to the CUDA engine — it never evaluates a network."""
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
            raise NotImplementedError("method='scipy' runs the network on the host; outside the accelerated path")
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
