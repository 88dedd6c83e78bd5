"""CPU oracle for ConsistencyModel.  TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.

Torch restatement of bayesflow/networks/inference/consistency/consistency_model.py:153-416 on the restated TimeMLP
(oracle/networks.py).  The reference holds no golden vectors for this network (its tests check shapes, determinism under
a seed and serialization: tests/test_networks/test_consistency_model/test_consistency_model.py), so the anchors are the
closed-form properties of the consistency parameterisation — c_skip(eps) = 1, c_out(eps) = 0 (boundary condition
f(x, eps) = x, Song et al. 2023 eq. 5) — and the discretisation end points t_1 = eps, t_N = max_time; both restated in
tests/test_oracle_edge.py.  "Parity unpinned" at the Keras boundary, as for the other networks."""
from __future__ import annotations

import math

import torch

from .networks import OracleTimeMLP, TIME_MLP_DEFAULTS, _mask

GOLDEN = 0x9E3779B97F4A7C15


class OracleConsistencyModel:
    def __init__(self, params, D, total_steps, subnet_kwargs=None, max_time=80, sigma2=1.0, eps=0.001, s0=10, s1=150,
                 rho=7.0, p_mean=-1.1, p_std=2.0):
        kw = dict(TIME_MLP_DEFAULTS) | (subnet_kwargs or {})  # consistency_model.py:93-96
        self.p, self.D = params, D
        self.subnet = OracleTimeMLP(params, **kw)
        self.total_steps, self.max_time, self.sigma2, self.eps = float(total_steps), float(max_time), float(sigma2), float(eps)
        self.sigma = math.sqrt(self.sigma2)
        self.s0, self.s1, self.rho, self.p_mean, self.p_std = float(s0), float(s1), float(rho), float(p_mean), float(p_std)
        if self.total_steps < self.s0:
            raise ValueError(f"total_steps={self.total_steps} must be greater than or equal to s0={self.s0}.")
        self.c_huber = 0.00054 * math.sqrt(D)  # :198-199
        self.current_step = 0

    def schedule_discretization(self, step):
        # :153-162
        k_ = math.floor(self.total_steps / (math.log(math.floor(self.s1 / self.s0)) / math.log(2.0) + 1.0))
        return min(self.s0 * 2.0 ** math.floor(step / k_), self.s1) + 1.0

    def discretize_time(self, n_k, dtype=torch.float32):
        # :164-176
        idx = torch.arange(1, n_k + 1, dtype=dtype)
        r = 1.0 / self.rho
        return (self.eps ** r + (idx - 1.0) / (float(n_k) - 1.0) * (self.max_time ** r - self.eps ** r)) ** self.rho

    def consistency_function(self, x, t, conditions=None, training=False, dropout_masks=None):
        # :308-343
        t = torch.as_tensor(t, dtype=x.dtype)
        if t.ndim == 0:
            t = t.expand(x.shape[0], 1)
        subnet_out = self.subnet((x, t / self.max_time, conditions), training=training, dropout_masks=dropout_masks)
        f = subnet_out @ self.p["output_projector.kernel"].to(x.dtype) + self.p["output_projector.bias"].to(x.dtype)
        skip = self.sigma2 / ((t - self.eps) ** 2 + self.sigma2)
        out = self.sigma * (t - self.eps) / torch.sqrt(self.sigma2 + t ** 2)
        return skip * x + out * f

    def inverse(self, z, conditions=None, steps=None, noise_fn=None, target_mask=None, targets_fixed=None):
        """:222-306.  noise_fn(n, shape) supplies the step-n noise (the product's Philox stream in parity runs)."""
        steps = int(steps if steps is not None else self.s0 + 1)
        x = z.clone() * self.max_time
        dt = torch.flip(self.discretize_time(steps), dims=(-1,))
        if target_mask is not None:
            target_mask = torch.broadcast_to(target_mask, x.shape)
            targets_fixed = torch.broadcast_to(targets_fixed, x.shape)
            x = _mask(x, target_mask, targets_fixed)
        x = self.consistency_function(x, float(dt[0]), conditions)
        x = _mask(x, target_mask, targets_fixed) if target_mask is not None else x
        for n in range(1, steps):
            noise = noise_fn(n, x.shape) if noise_fn is not None else torch.randn(x.shape)
            x_n = x + math.sqrt(float(dt[n]) ** 2 - self.eps ** 2) * noise.to(x.dtype)
            x_n = _mask(x_n, target_mask, targets_fixed) if target_mask is not None else x_n
            x = self.consistency_function(x_n, float(dt[n]), conditions)
            x = _mask(x, target_mask, targets_fixed) if target_mask is not None else x
        return x

    def compute_metrics(self, x, conditions=None, sample_weight=None, stage="training", fixed=None, generator=None):
        # :345-416 (drop_cond_prob = drop_target_prob = 0: mask_x = 1)
        training = stage == "training"
        if training:
            self.current_step = min(self.current_step + 1, int(self.total_steps) - 1)
        dt = self.discretize_time(int(self.schedule_discretization(self.current_step)))
        if fixed is None:
            cdf = torch.erf((torch.log(dt.double()) - self.p_mean) / (math.sqrt(2.0) * self.p_std))
            p = (cdf[1:] - cdf[:-1]).float()
            idx = torch.multinomial(p, x.shape[0], replacement=True, generator=generator)
            noise = torch.randn(x.shape, generator=generator)
        else:
            idx, noise = fixed
        dt = dt.to(x.dtype)
        t1, t2 = dt[idx].unsqueeze(-1), dt[idx + 1].unsqueeze(-1)
        noise = noise.to(x.dtype)
        with torch.no_grad():
            teacher = self.consistency_function(x + t1 * noise, t1, conditions, training=False)
        student = self.consistency_function(x + t2 * noise, t2, conditions, training=False)
        lam = 1 / (t2 - t1)
        loss = lam * (torch.sqrt((teacher - student) ** 2 + self.c_huber ** 2) - self.c_huber)
        if sample_weight is not None:
            loss = loss * sample_weight.reshape(-1, 1).to(x.dtype)
        return {"loss": loss.mean()}


class OracleStableConsistencyModel:
    """stable_consistency_model.py:27-314 — the sampling side (consistency function :238-260, multistep sampling :163-236,
    time discretisation :111-122).  Pinned on the reference's own class (tests/test_reference_pin.py).  The
    continuous-time consistency TRAINING (:262-314) differentiates the subnet in forward mode (a JVP along (x, t)); it is
    not restated and not built on the CUDA path."""

    def __init__(self, params, D, subnet_kwargs=None, sigma=1.0):
        kw = dict(TIME_MLP_DEFAULTS) | (subnet_kwargs or {})  # :81-84
        self.p, self.D, self.sigma = params, D, float(sigma)
        self.subnet = OracleTimeMLP(params, **kw)

    @staticmethod
    def discretize_time(num_steps, rho=3.5):
        # :111-122 (fp32 like keras.ops)
        t = torch.linspace(0.0, math.pi / 2, num_steps, dtype=torch.float32)
        times = torch.exp((t - math.pi / 2) * rho) * math.pi / 2
        return torch.cat([torch.zeros(1), times[1:]])

    def consistency_function(self, x, t, conditions=None):
        # :238-260
        t = torch.as_tensor(t, dtype=x.dtype)
        if t.ndim == 0:
            t = t.expand(x.shape[0], 1)
        subnet_out = self.subnet((x / self.sigma, t, conditions), training=False)
        f = subnet_out @ self.p["output_projector.kernel"].to(x.dtype) + self.p["output_projector.bias"].to(x.dtype)
        return torch.cos(t) * x - torch.sin(t) * self.sigma * f

    def inverse(self, z, conditions=None, steps=15, rho=3.5, noise_fn=None, target_mask=None, targets_fixed=None):
        """:163-236.  noise_fn(n, shape) supplies the step-n noise."""
        x = z.clone() * self.sigma
        dt = torch.flip(self.discretize_time(steps, rho), dims=(-1,))
        t = torch.full(x.shape[:-1] + (1,), float(dt[0]), dtype=x.dtype)
        if target_mask is not None:
            target_mask = torch.broadcast_to(target_mask, x.shape)
            targets_fixed = torch.broadcast_to(targets_fixed, x.shape)
            x = _mask(x, target_mask, targets_fixed)
        x = self.consistency_function(x, t, conditions)
        x = _mask(x, target_mask, targets_fixed) if target_mask is not None else x
        for n in range(1, steps):
            noise = noise_fn(n, x.shape) if noise_fn is not None else torch.randn(x.shape)
            x_n = torch.cos(t) * x + torch.sin(t) * noise.to(x.dtype)
            t = torch.full_like(t, float(dt[n]))
            x_n = _mask(x_n, target_mask, targets_fixed) if target_mask is not None else x_n
            x = self.consistency_function(x_n, t, conditions)
            x = _mask(x, target_mask, targets_fixed) if target_mask is not None else x
        return x
