"""Mini-batch optimal-transport pairing — bayesflow/utils/optimal_transport/optimal_transport.py:12-72
and log_sinkhorn.py:14-172 behind the same names and arguments, on the `bfk_ot_*` kernels (include/bfk.h).  No CPU path.

Supported: method "log_sinkhorn" / "log_sinkhorn_knopp" (the FlowMatching default, networks/defaults.py:54-63), balanced
and partial (`partial_factor < 1`, ot_utils.py:37-92) marginals, squared-Euclidean cost, conditional cost weighting
(`condition_ratio < 0.5`, ot_utils.py:95-164), and the multiplicative "sinkhorn" / "sinkhorn_knopp" method
(sinkhorn.py:14-186) on the same device sweeps."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

OPTIMAL_TRANSPORT_DEFAULTS = {"method": "log_sinkhorn", "regularization": 0.1, "max_steps": 100, "atol": 1e-5,
                              "partial_factor": 1.0, "condition_ratio": 0.01}  # networks/defaults.py:54-63
POLL_EVERY = 16  # Sinkhorn steps enqueued between two reads of the device-side loop state


def _flat2(x, what):
    if not torch.is_tensor(x) or not x.is_cuda:
        raise RuntimeError(f"bayesflow_b200 runs on CUDA tensors only (no CPU fallback): {what}")
    x = x.detach().to(torch.float32)
    return x.reshape(x.shape[0], -1).contiguous()


def _cost(a, b, cosine=False):
    n, m = a.shape[0], b.shape[0]
    M = torch.empty(n, m, device=a.device, dtype=torch.float32)
    _lib.check(_lib.lib().bfk_ot_cost(_lib.ptr(a), _lib.ptr(b), n, m, a.shape[1], 1 if cosine else 0, _lib.ptr(M),
                                      _lib.stream_ptr()), "bfk_ot_cost")
    return M


def search_for_conditional_weight(M, Cm, condition_ratio, initial_w=1.0, max_iter=10, abs_tol=1e-3, max_w=1e8):
    """ot_utils.py:95-164 — the ratio is an exact integer count on the device; the ten-step search runs on the host."""
    n = M.shape[0]
    count = torch.zeros(1, device=M.device, dtype=torch.int64)

    def get_ratio(w):
        _lib.check(_lib.lib().bfk_ot_ratio(_lib.ptr(M), _lib.ptr(Cm), n, float(w), C.c_void_p(count.data_ptr()),
                                           _lib.stream_ptr()), "bfk_ot_ratio")
        return float(torch.tensor(int(count.item()) / float(n * n), dtype=torch.float32))  # float32 mean

    r_t = float(torch.tensor(condition_ratio, dtype=torch.float32))
    if get_ratio(0.0) < r_t:
        return 0.0
    it, low, high = 0, 0.0, float(initial_w)
    while it < max_iter and get_ratio(high) > r_t and high < max_w:
        it, low, high = it + 1, high, high * 2.0
    high = min(high, max_w)
    it, best_w, best_r = 0, high, get_ratio(high)
    while it < max_iter and abs(best_r - r_t) > abs_tol:
        mid = (low + high) / 2.0
        r_mid = get_ratio(mid)
        if r_mid < r_t:
            high = mid
        else:
            low = mid
        if abs(r_mid - r_t) < abs(best_r - r_t):
            best_w, best_r = mid, r_mid
        it += 1
    return best_w


def log_sinkhorn_plan(x1, x2, conditions=None, regularization: float = 1.0, atol: float = 1e-5, max_steps: int = 1000,
                      condition_ratio: float = 0.5, partial_factor: float = 1.0, _parts=False, _plain=False, **kwargs):
    """log_sinkhorn.py:38-172 -> (n, m) log transport plan.  `_plain`: the multiplicative Sinkhorn-Knopp of
    sinkhorn.py:65-186 (same sweeps, its `tiny` terms and error metric; see csrc/bfk_ot.cuh)."""
    if not (0.0 < partial_factor <= 1.0):
        raise ValueError(f"s must be in (0, 1] for partial OT, got {partial_factor}")
    partial = partial_factor < 1.0
    if regularization <= 0.0:
        raise ValueError(f"regularization must be positive, got {regularization}")
    a, b = _flat2(x1, "x1"), _flat2(x2, "x2")
    n, m = a.shape[0], b.shape[0]
    L = _lib.lib()
    S = _cost(a, b)
    Cm, w = None, 0.0
    if conditions is not None and condition_ratio < 0.5:
        if n != m or conditions.shape[0] != n:
            raise ValueError("conditional optimal transport needs as many conditions as samples in x1 and x2")
        c = _flat2(conditions, "conditions")
        Cm = _cost(c, c, cosine=True)
        sk = {k: kwargs[k] for k in ("initial_w", "max_iter", "abs_tol", "max_w") if k in kwargs}  # filter_kwargs
        w = search_for_conditional_weight(S, Cm, condition_ratio, **sk)
    _lib.check(L.bfk_ot_scale(_lib.ptr(S), _lib.ptr(Cm) if w != 0.0 else None, float(w), float(regularization), S.numel(),
                              _lib.stream_ptr()), "bfk_ot_scale")
    dev = S.device
    n0, m0, log_a, log_b = n, m, None, None
    if partial:
        # augment_for_partial_ot (ot_utils.py:37-92): a dummy row / column of zeros, corner -A / reg with
        # A = max(cost) + 1 (or `dummy_cost`), marginals (1/n0, ..., 1 - s) and (1/m0, ..., 1 - s)
        dummy = kwargs.get("dummy_cost")
        corner = (S.min() - 1.0 / regularization) if dummy is None else torch.tensor(-float(dummy) / regularization, device=dev)
        aug = torch.zeros(n0 + 1, m0 + 1, device=dev, dtype=torch.float32)
        aug[:n0, :m0] = S
        aug[n0, m0] = corner
        S, n, m = aug, n0 + 1, m0 + 1
        one_minus_s = float(torch.tensor(1.0, dtype=torch.float32) - torch.tensor(partial_factor, dtype=torch.float32))
        log_a = torch.log(torch.cat([torch.full((n0,), 1.0 / n0), torch.tensor([one_minus_s])]).to(torch.float32)).to(dev)
        log_b = torch.log(torch.cat([torch.full((m0,), 1.0 / m0), torch.tensor([one_minus_s])]).to(torch.float32)).to(dev)
    ctrl = torch.empty(8, device=dev, dtype=torch.float32)
    u, v = torch.empty(n, device=dev, dtype=torch.float32), torch.empty(m, device=dev, dtype=torch.float32)
    work = torch.empty(max(1, L.bfk_ot_work_floats(n, m)), device=dev, dtype=torch.float32)
    _lib.check(L.bfk_ot_init(_lib.ptr(ctrl), _lib.ptr(u), _lib.ptr(v), n, m, float(atol), int(max_steps), _lib.stream_ptr()),
               "bfk_ot_init")
    if _plain:
        _lib.check(L.bfk_ot_plain_mode(_lib.ptr(S), S.numel(), _lib.ptr(ctrl), _lib.stream_ptr()), "bfk_ot_plain_mode")
    done = max_steps <= 0
    enq = 0
    while not done:
        k = min(POLL_EVERY, max_steps - enq)
        _lib.check(L.bfk_ot_log_sinkhorn_step(_lib.ptr(S), n, m, _lib.ptr(log_a), _lib.ptr(log_b), _lib.ptr(u), _lib.ptr(v),
                                              _lib.ptr(ctrl), _lib.ptr(work), k, _lib.stream_ptr()), "bfk_ot_log_sinkhorn_step")
        enq += k
        state = ctrl.tolist()
        done = state[2] != 0.0 or enq >= max_steps
    log_sinkhorn_plan.last_info = {"steps": int(ctrl[1].item()), "err": float(ctrl[0].item()), "w": w}
    if _parts:
        return (S, u, v, n0, m0, ctrl) if _plain else (S, u, v, n0, m0)
    if _plain:
        return (S - ctrl[7] + u.unsqueeze(1) + v.unsqueeze(0))[:n0, :m0]
    return (S + u.unsqueeze(1) + v.unsqueeze(0))[:n0, :m0]


def log_sinkhorn(x1, x2, conditions=None, seed=None, **kwargs):
    """log_sinkhorn.py:14-35 -> (n,) int64 indices into x2."""
    S, u, v, n, m = log_sinkhorn_plan(x1, x2, conditions=conditions, _parts=True, **kwargs)
    if isinstance(seed, torch.Generator):
        seed = int(torch.randint(0, 2 ** 62, (1,), generator=seed, device=seed.device).item())
    elif seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    out = torch.empty(n, device=S.device, dtype=torch.int64)
    # partial OT: rows / columns of the real samples only (log_sinkhorn.py:169-170)
    _lib.check(_lib.lib().bfk_ot_sample(_lib.ptr(S), S.shape[1], _lib.ptr(v), n, m, C.c_uint64(int(seed) & (2 ** 64 - 1)),
                                        C.c_void_p(out.data_ptr()), _lib.stream_ptr()), "bfk_ot_sample")
    return out


def sinkhorn_plan(x1, x2, conditions=None, **kwargs):
    """sinkhorn.py:65-186 -> (n, m) transport plan (probabilities)."""
    return torch.exp(log_sinkhorn_plan(x1, x2, conditions=conditions, _plain=True, **kwargs))


def sinkhorn(x1, x2, conditions=None, seed=None, **kwargs):
    """sinkhorn.py:14-62 -> (n,) int64 indices into x2, drawn from Categorical(log(plan + 1e-10))."""
    S, u, v, n, m, ctrl = log_sinkhorn_plan(x1, x2, conditions=conditions, _parts=True, _plain=True, **kwargs)
    if isinstance(seed, torch.Generator):
        seed = int(torch.randint(0, 2 ** 62, (1,), generator=seed, device=seed.device).item())
    elif seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    out = torch.empty(n, device=S.device, dtype=torch.int64)
    _lib.check(_lib.lib().bfk_ot_sample_plan(_lib.ptr(S), S.shape[1], _lib.ptr(u), _lib.ptr(v), _lib.ptr(ctrl), n, m,
                                             C.c_uint64(int(seed) & (2 ** 64 - 1)), C.c_void_p(out.data_ptr()),
                                             _lib.stream_ptr()), "bfk_ot_sample_plan")
    return out


methods = {"sinkhorn": sinkhorn, "sinkhorn_knopp": sinkhorn, "log_sinkhorn": log_sinkhorn,
           "log_sinkhorn_knopp": log_sinkhorn}


def optimal_transport(x1, x2, conditions=None, method="sinkhorn", return_assignments=False, **kwargs):
    """optimal_transport.py:12-72: x2 (and the conditions) re-indexed so that x2[i] is matched to x1[i]."""
    assignments = methods[method.lower()](x1, x2, conditions, **kwargs)
    x2 = x2.index_select(0, assignments)
    if conditions is not None:
        conditions = conditions.index_select(0, assignments)
    if return_assignments:
        return x1, x2, conditions, assignments
    return x1, x2, conditions
