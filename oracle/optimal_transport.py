"""CPU oracle for the optimal-transport pairing.  TEST INFRASTRUCTURE ONLY — see
oracle/__init__.py.

Torch restatement of bayesflow/utils/optimal_transport: ot_utils.py:6-34 (costs), :95-164 (conditional weight search),
log_sinkhorn.py:38-172 (log-stabilised Sinkhorn-Knopp plan).  Pinned by the reference's own closed-form tests
(tests/test_utils/test_optimal_transport.py:98-116 assignment of a shuffled copy, :215-224 marginals, :249-263 analytical
plan for identical points), restated in tests/test_oracle_edge.py.  The categorical draw restates the product's
Gumbel-max over the Philox stream (oracle/edge.py) because keras.random.categorical has no portable bit stream."""
from __future__ import annotations

import numpy as np
import torch

from .edge import philox_words, _u01


def squared_euclidean(x1, x2):
    # ot_utils.py:6-16
    x1, x2 = x1.reshape(x1.shape[0], -1), x2.reshape(x2.shape[0], -1)
    d2 = (x1 * x1).sum(1, keepdim=True) + (x2 * x2).sum(1, keepdim=True).T - 2.0 * x1 @ x2.T
    return d2.clamp_min(0.0)


def cosine_distance(x1, x2, eps=1e-8):
    # ot_utils.py:19-34
    x1, x2 = x1.reshape(x1.shape[0], -1), x2.reshape(x2.shape[0], -1)
    x1 = x1 / (x1.norm(dim=1, keepdim=True) + eps)
    x2 = x2 / (x2.norm(dim=1, keepdim=True) + eps)
    return 1.0 - (x1 @ x2.T).clamp(-1.0, 1.0)


def search_for_conditional_weight(M, Cm, condition_ratio, initial_w=1.0, max_iter=10, abs_tol=1e-3, max_w=1e8):
    # ot_utils.py:95-164
    delta = M.diagonal().unsqueeze(1) - M

    def get_ratio(w):
        return float((w * Cm <= delta).to(M.dtype).mean())

    r_t = float(torch.tensor(condition_ratio, dtype=M.dtype))
    if get_ratio(0.0) < r_t:
        return M, 0.0
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
    return M + best_w * Cm, best_w


def augment_for_partial_ot(cost_scaled, regularization, s, dummy_cost=None):
    # ot_utils.py:37-92
    if dummy_cost is None:
        dummy_cost = float((-cost_scaled * regularization).max()) + 1
    n0, m0 = cost_scaled.shape
    aug = torch.zeros(n0 + 1, m0 + 1, dtype=cost_scaled.dtype)
    aug[:n0, :m0] = cost_scaled
    aug[n0, m0] = -dummy_cost / regularization
    one_minus_s = 1.0 - torch.tensor(s, dtype=cost_scaled.dtype)
    a = torch.cat([torch.full((n0,), 1.0 / n0, dtype=cost_scaled.dtype), one_minus_s.reshape(1)])
    b = torch.cat([torch.full((m0,), 1.0 / m0, dtype=cost_scaled.dtype), one_minus_s.reshape(1)])
    return aug, a, b


def log_sinkhorn_plan(x1, x2, conditions=None, regularization=1.0, atol=1e-5, max_steps=1000, condition_ratio=0.5,
                      partial_factor=1.0, info=None, **kwargs):
    # log_sinkhorn.py:38-172
    if not (0.0 < partial_factor <= 1.0):
        raise ValueError(f"s must be in (0, 1] for partial OT, got {partial_factor}")
    partial = partial_factor < 1.0
    cost = squared_euclidean(x1, x2)
    if regularization <= 0.0:
        raise ValueError(f"regularization must be positive, got {regularization}")
    w = 0.0
    if conditions is not None and condition_ratio < 0.5:
        cost, w = search_for_conditional_weight(cost, cosine_distance(conditions, conditions), condition_ratio)
    S = -cost / regularization
    if partial:
        S, a, b = augment_for_partial_ot(S, regularization, partial_factor, kwargs.get("dummy_cost"))
        log_a, log_b = torch.log(a), torch.log(b)
        n, m = S.shape
    else:
        n, m = S.shape
        log_a = torch.full((n,), -float(np.log(n)), dtype=S.dtype)
        log_b = torch.full((m,), -float(np.log(m)), dtype=S.dtype)
    u, v = torch.zeros(n, dtype=S.dtype), torch.zeros(m, dtype=S.dtype)
    err, steps = 1e30, 0
    while err > atol and steps < max_steps:
        u_next = log_a - torch.logsumexp(S + v.unsqueeze(0), dim=1)
        v = log_b - torch.logsumexp(S + u_next.unsqueeze(1), dim=0)
        err = float((u_next - u).abs().max())
        u, steps = u_next, steps + 1
    if info is not None:
        info.update(steps=steps, err=err, w=w)
    log_plan = S + u.unsqueeze(1) + v.unsqueeze(0)
    return log_plan[:-1, :-1] if partial else log_plan


def sinkhorn_plan(x1, x2, conditions=None, regularization=1.0, max_steps=1000, atol=1e-5, condition_ratio=0.5,
                  partial_factor=1.0, info=None, **kwargs):
    # sinkhorn.py:65-186 (the multiplicative Sinkhorn-Knopp: Gaussian kernel exp(S - max S), scalings u, v, `tiny` = 1e-12)
    if not (0.0 < partial_factor <= 1.0):
        raise ValueError(f"s must be in (0, 1] for partial OT, got {partial_factor}")
    partial = partial_factor < 1.0
    cost = squared_euclidean(x1, x2)
    if regularization <= 0.0:
        raise ValueError(f"regularization must be positive, got {regularization}")
    w = 0.0
    if conditions is not None and condition_ratio < 0.5:
        cost, w = search_for_conditional_weight(cost, cosine_distance(conditions, conditions), condition_ratio)
    S = -cost / regularization
    if partial:
        S, a, b = augment_for_partial_ot(S, regularization, partial_factor, kwargs.get("dummy_cost"))
    else:
        n, m = S.shape
        a = torch.ones(n, dtype=S.dtype) / n
        b = torch.ones(m, dtype=S.dtype) / m
    plan = torch.exp(S - S.max())
    u, v = torch.ones_like(a), torch.ones_like(b)
    tiny = 1e-12
    err, steps = 1e30, 0
    while err > atol and steps < max_steps:
        u_new = a / (plan @ v + tiny)
        v_new = b / (plan.T @ u_new + tiny)
        du = float(torch.log((u_new + tiny) / (u + tiny)).abs().max())
        dv = float(torch.log((v_new + tiny) / (v + tiny)).abs().max())
        err, u, v, steps = max(du, dv), u_new, v_new, steps + 1
    if info is not None:
        info.update(steps=steps, err=err, w=w)
    plan = (u.unsqueeze(1) * plan) * v.unsqueeze(0)
    return plan[:-1, :-1] if partial else plan


def gumbel_argmax(log_plan, seed):
    """Categorical draw per row: argmax_j log_plan[i,j] - log(-log U_ij), U from Philox word i*m + j."""
    n, m = log_plan.shape
    U = _u01(philox_words(n * m, seed)).reshape(n, m)
    g = -np.log(-np.log(U))
    return torch.from_numpy(np.argmax(log_plan.double().numpy() + g, axis=1))


def optimal_transport(x1, x2, conditions=None, seed=0, return_assignments=False, **kwargs):
    # optimal_transport.py:12-72 with method = "log_sinkhorn"
    kwargs.pop("method", None)
    a = gumbel_argmax(log_sinkhorn_plan(x1, x2, conditions, **kwargs), seed)
    out = (x1, x2[a], conditions[a] if conditions is not None else None)
    return out + (a,) if return_assignments else out
