"""`ConsistencyModel` on the same kernels as FlowMatching: the TimeMLP subnet + output projector
evaluated by `bfk_velocity` / `bfk_sample_layerwise` (few-step sampling) and `bfk_train_step` with the pseudo-Huber loss
head (training).  Mirrors bayesflow/networks/inference/consistency/consistency_model.py:21-416 — same constructor
arguments, `consistency_function`, `_inverse` (multistep sampling, Song et al. 2023 Algorithm 1), `compute_metrics`
(improved consistency training, discretisation schedule N(k)).  No CPU path."""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib
from .networks import InferenceNetwork, TIME_MLP_DEFAULTS, find_network

_GOLDEN = 0x9E3779B97F4A7C15


def _lincomb(out, a, x, b=0.0, y=None, c=0.0, z=None):
    _lib.check(_lib.lib().bfk_lincomb3(_lib.ptr(out), float(a), _lib.ptr(x), float(b), _lib.ptr(y), float(c), _lib.ptr(z),
                                       x.numel(), _lib.stream_ptr()), "bfk_lincomb3")
    return out


class ConsistencyModel(InferenceNetwork):
    """consistency_model.py:21-416."""

    _wants_noise_offset = True

    def __init__(self, total_steps, subnet="time_mlp", max_time=80, sigma2: float = 1.0, eps: float = 0.001, s0=10, s1=150,
                 subnet_kwargs=None, drop_cond_prob: float = 0.0, **kwargs):
        self.rho = float(kwargs.pop("rho", 7.0))
        self.p_mean = float(kwargs.pop("p_mean", -1.1))
        self.p_std = float(kwargs.pop("p_std", 2.0))
        self.drop_target_prob = float(kwargs.pop("drop_target_prob", 0.0))
        super().__init__("normal", **kwargs)
        self.total_steps = float(total_steps)
        subnet_kwargs = subnet_kwargs or {}
        if subnet == "time_mlp":
            subnet_kwargs = TIME_MLP_DEFAULTS | subnet_kwargs
        self.subnet = find_network(subnet, **subnet_kwargs)
        self.output_projector = None
        self.sigma2 = float(sigma2)
        self.sigma = math.sqrt(self.sigma2)
        self.eps = float(eps)
        self.max_time = float(max_time)
        self.s0, self.s1 = float(s0), float(s1)
        if self.total_steps < self.s0:
            raise ValueError(f"total_steps={self.total_steps} must be greater than or equal to s0={self.s0}.")
        self.current_step = 0
        self.drop_cond_prob = drop_cond_prob
        self.unconditional_mode = False
        self.unique_n = None
        self.c_huber = None

    @property
    def student(self):
        return self.subnet

    def _schedule_struct(self):
        s = _lib.Schedule()
        s.model = 0  # the kernels return the raw projected subnet output f; c_skip / c_out are applied here
        return s

    def get_config(self):
        return {"total_steps": self.total_steps, "subnet": self.subnet.get_config(), "max_time": self.max_time,
                "sigma2": self.sigma2, "eps": self.eps, "s0": self.s0, "s1": self.s1, "rho": self.rho,
                "p_mean": self.p_mean, "p_std": self.p_std, "drop_cond_prob": self.drop_cond_prob,
                "drop_target_prob": self.drop_target_prob}

    # -- discretisation (consistency_model.py:153-176) ------------------------------------------------------------------
    def _schedule_discretization(self, step) -> float:
        k_ = math.floor(self.total_steps / (math.log(math.floor(self.s1 / self.s0)) / math.log(2.0) + 1.0))
        return min(self.s0 * 2.0 ** math.floor(step / k_), self.s1) + 1.0

    def _discretize_time(self, n_k: int) -> torch.Tensor:
        idx = torch.arange(1, n_k + 1, dtype=torch.float32)
        r = 1.0 / self.rho
        return (self.eps ** r + (idx - 1.0) / (float(n_k) - 1.0) * (self.max_time ** r - self.eps ** r)) ** self.rho

    def build(self, xz_shape, conditions_shape=None):
        if self.built:
            return
        self._build_common(xz_shape, conditions_shape)
        self.c_huber = 0.00054 * math.sqrt(self._D)  # :198-199
        max_n = int(self._schedule_discretization(self.total_steps))
        if max_n != self.s1 + 1:
            raise ValueError("The maximum number of discretization steps must be equal to s1 + 1.")
        self.unique_n = sorted({int(self._schedule_discretization(step)) for step in range(int(self.total_steps))})
        self.built = True

    # -- consistency function (:308-343) --------------------------------------------------------------------------------
    def _coefs(self, t):
        skip = self.sigma2 / ((t - self.eps) ** 2 + self.sigma2)
        out = self.sigma * (t - self.eps) / math.sqrt(self.sigma2 + t ** 2)
        return skip, out

    def consistency_function(self, x, t, conditions=None, training=False, **kwargs):
        """skip(t) * x + out(t) * projector(subnet(x, t / max_time, conditions)); t is shared by all rows (a float or a
        constant tensor) — per-row times only occur in training and go through `compute_metrics`."""
        if training:
            raise NotImplementedError("training-mode evaluation goes through compute_metrics")
        if torch.is_tensor(t):
            if t.numel() != 1 and not bool((t == t.reshape(-1)[0]).all()):
                raise NotImplementedError("per-row time is only used by the training path (compute_metrics)")
            t = float(t.reshape(-1)[0])
        run = self._run(conditions, x.shape[0], {k: v for k, v in kwargs.items() if k == "conditions_repeats"})
        return self._cf(run, run._check_x(x), float(np.float32(t)))

    def _cf(self, run, x, t: float):
        f = run.velocity(x, float(np.float32(t / self.max_time)))
        skip, out = self._coefs(t)
        return _lincomb(f, skip, x, out, f)

    def _forward(self, x, conditions=None, **kwargs):
        raise NotImplementedError("Consistency Models are not invertible")

    # -- multistep sampling (:222-306) ----------------------------------------------------------------------------------
    def _inverse(self, z, conditions=None, density=False, training=False, **kwargs):
        if density:
            raise NotImplementedError("Consistency Models are not invertible")
        seed = kwargs.pop("seed", None)
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        row_offset = int(kwargs.pop("noise_row_offset", 0))
        steps = int(kwargs.get("steps", self.s0 + 1))
        run = self._run(conditions, z.shape[0], {k: v for k, v in kwargs.items() if k == "conditions_repeats"})
        z = run._check_x(z)
        times = torch.flip(self._discretize_time(steps), dims=(-1,)).tolist()
        x = _lincomb(torch.empty_like(z), self.max_time, z)
        x = self._apply_targets(x, kwargs)
        x = self._apply_targets(self._cf(run, x, times[0]), kwargs)
        noise = torch.empty_like(x) if steps > 1 else None
        B, D = x.shape
        for n in range(1, steps):
            # fresh N(0, I) per step from the counter-based stream: key = seed + n * golden ratio, rows by global index
            _lib.check(_lib.lib().bfk_normal_fill(_lib.ptr(noise), B, D, row_offset, None, None,
                                                  _lib.C.c_uint64((int(seed) + n * _GOLDEN) & (2 ** 64 - 1)), None,
                                                  _lib.stream_ptr()), "bfk_normal_fill")
            scale = math.sqrt(times[n] ** 2 - self.eps ** 2)
            x_n = self._apply_targets(_lincomb(noise, 1.0, x, scale, noise), kwargs)
            x = self._apply_targets(self._cf(run, x_n, times[n]), kwargs)
            noise = x_n  # reuse the buffer
        return x

    # -- training (:345-416) --------------------------------------------------------------------------------------------
    def _loss_and_grads(self, x, conditions=None, sample_weight=None, stage="training", seed=None, world=1, backward=True,
                        fixed=None, **kwargs):
        """Teacher pass (forward only, at t1) and student pass (forward + backward, at t2) through `bfk_train_step`; the
        pseudo-Huber head gets a = c_out(t2), target = teacher - c_skip(t2) * input and row weight 1 / (t2 - t1).
        `fixed=(time_index (B,), noise (B,D))` bypasses the RNG for parity runs."""
        from .training import train_engine
        if self.drop_cond_prob > 0 or self.drop_target_prob > 0:
            raise NotImplementedError("drop_cond_prob / drop_target_prob > 0 are outside the accelerated training path")
        if (conditions is None) != (self._C is None):
            raise ValueError("network was built " + ("without" if self._C is None else "with") + " conditions")
        if not torch.is_tensor(x) or not x.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        eng = train_engine(self)
        training = stage == "training"
        if training:
            self.current_step = min(self.current_step + 1, int(self.total_steps) - 1)  # :354-356
        dt = self._discretize_time(int(self._schedule_discretization(self.current_step)))  # float32, host
        B, dev = x.shape[0], x.device
        if fixed is None:
            lt = torch.log(dt.double())
            cdf = torch.erf((lt - self.p_mean) / (math.sqrt(2.0) * self.p_std))
            p = (cdf[1:] - cdf[:-1]).clamp_min(0).to(torch.float32)  # :368-374
            idx = torch.multinomial(p, B, replacement=True).to(dev)
            noise = self.base_distribution.sample((B,))
        else:
            idx, noise = fixed[0].to(dev).long(), fixed[1].to(dev, torch.float32)
        dtd = dt.to(dev)
        t1, t2 = dtd[idx].unsqueeze(-1), dtd[idx + 1].unsqueeze(-1)
        x = x.detach().to(torch.float32)
        rate = float(self.subnet.dropout or 0.0) if training else 0.0
        dseed = self._dropout_seed(seed)
        s2, sg, e = self.sigma2, self.sigma, self.eps
        inp1, inp2 = x + t1 * noise, x + t2 * noise
        _, f1 = eng.step(inp1, (t1 / self.max_time).reshape(-1), conditions, None, dropout=rate, seed=dseed ^ _GOLDEN, backward=False,
                         want_pred=True, want_loss=False)
        teacher = s2 / ((t1 - e) ** 2 + s2) * inp1 + sg * (t1 - e) / torch.sqrt(s2 + t1 ** 2) * f1
        skip2, out2 = s2 / ((t2 - e) ** 2 + s2), sg * (t2 - e) / torch.sqrt(s2 + t2 ** 2)
        lam = (1.0 / (t2 - t1)).reshape(-1)
        if sample_weight is not None:
            # weighted_mean multiplies the (B, D) loss by the weights (tensor_utils.py:181-198)
            lam = lam * sample_weight.reshape(-1).to(torch.float32)
        loss, _ = eng.step(inp2, (t2 / self.max_time).reshape(-1), conditions, (teacher - skip2 * inp2).contiguous(),
                           pred_scale=out2.reshape(-1).contiguous(), row_weight=lam.contiguous(), loss="pseudo_huber",
                           huber_c=self.c_huber, dropout=rate, seed=dseed, global_batch=B * world, backward=backward)
        return loss


class StableConsistencyModel(InferenceNetwork):
    """stable_consistency_model.py:27-314 — the SAMPLING side on the CUDA velocity kernels: consistency function
    `cos(t) x - sin(t) sigma F(x / sigma, t, c)` (:238-260), time discretisation (:111-122), multistep sampling (:163-236;
    per-step noise from `bfk_normal_fill`, Philox keyed by (seed, step, GLOBAL row)).  The continuous-time consistency
    TRAINING (:262-314) needs a forward-mode derivative of the subnet along (x, t) and is not built: `compute_metrics`
    raises NotImplementedError (weights trained with the reference load through `load_state_dict`)."""

    _wants_noise_offset = True
    EPS_WARN = 0.1

    def __init__(self, subnet="time_mlp", sigma: float = 1.0, subnet_kwargs=None, weight_mlp_kwargs=None,
                 drop_cond_prob: float = 0.0, **kwargs):
        self.p_mean = float(kwargs.pop("p_mean", -1.0))
        self.p_std = float(kwargs.pop("p_std", 1.6))
        self.c = float(kwargs.pop("c", 0.1))
        self.drop_target_prob = float(kwargs.pop("drop_target_prob", 0.0))
        super().__init__("normal", **kwargs)
        subnet_kwargs = subnet_kwargs or {}
        if subnet == "time_mlp":
            subnet_kwargs = TIME_MLP_DEFAULTS | subnet_kwargs
        self.subnet = find_network(subnet, **subnet_kwargs)
        self.output_projector = None  # the reference's `subnet_projector` (its variable is named "output_projector", :134-138)
        self.weight_mlp_kwargs = dict(weight_mlp_kwargs or {})  # the adaptive-weight MLP only exists in training
        self.sigma = float(sigma)
        self.drop_cond_prob = drop_cond_prob
        self.unconditional_mode = False

    def _schedule_struct(self):
        s = _lib.Schedule()
        s.model = 0  # the kernels return the raw projected subnet output F
        return s

    def get_config(self):
        return {"subnet": self.subnet.get_config(), "sigma": self.sigma, "p_mean": self.p_mean, "p_std": self.p_std,
                "c": self.c, "drop_cond_prob": self.drop_cond_prob, "drop_target_prob": self.drop_target_prob}

    @staticmethod
    def _discretize_time(num_steps: int, rho: float = 3.5) -> torch.Tensor:
        # :111-122, fp32 like keras.ops; times[1] is read before the loop (steps = 1 fails in the reference too)
        if int(num_steps) < 2:
            raise IndexError("StableConsistencyModel needs steps >= 2 (stable_consistency_model.py:126 reads times[1])")
        t = torch.linspace(0.0, math.pi / 2, int(num_steps), dtype=torch.float32)
        times = torch.exp((t - math.pi / 2) * rho) * math.pi / 2
        return torch.cat([torch.zeros(1), times[1:]])

    def build(self, xz_shape, conditions_shape=None):
        if self.built:
            return
        self._build_common(xz_shape, conditions_shape)
        self.built = True

    def consistency_function(self, x, t, conditions=None, training=False, **kwargs):
        """cos(t) x - sin(t) sigma F(x / sigma, t, c); t shared by all rows (a float or a constant tensor)."""
        if training:
            raise NotImplementedError("StableConsistencyModel training (a JVP through the subnet) is not built")
        if torch.is_tensor(t):
            if t.numel() != 1 and not bool((t == t.reshape(-1)[0]).all()):
                raise NotImplementedError("per-row time only occurs in training, which is not built")
            t = float(t.reshape(-1)[0])
        run = self._run(conditions, x.shape[0], {k: v for k, v in kwargs.items() if k == "conditions_repeats"})
        return self._cf(run, run._check_x(x), float(np.float32(t)))

    def _cf(self, run, x, t: float):
        xs = x if self.sigma == 1.0 else _lincomb(torch.empty_like(x), 1.0 / self.sigma, x)
        f = run.velocity(xs, float(np.float32(t)))
        ct, st = float(np.cos(np.float32(t))), float(np.sin(np.float32(t)))
        return _lincomb(f, ct, x, -st * self.sigma, f)

    def _forward(self, x, conditions=None, **kwargs):
        raise NotImplementedError("Consistency Models are not invertible")

    def _inverse(self, z, conditions=None, density=False, training=False, **kwargs):
        if density:
            raise NotImplementedError("Consistency Models are not invertible")
        seed = kwargs.pop("seed", None)
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        row_offset = int(kwargs.pop("noise_row_offset", 0))
        noise_fn = kwargs.pop("noise_fn", None)
        steps, rho = int(kwargs.get("steps", 15)), float(kwargs.get("rho", 3.5))
        run = self._run(conditions, z.shape[0], {k: v for k, v in kwargs.items() if k == "conditions_repeats"})
        z = run._check_x(z)
        times = torch.flip(self._discretize_time(steps, rho), dims=(-1,)).tolist()
        x = _lincomb(torch.empty_like(z), self.sigma, z)
        x = self._apply_targets(x, kwargs)
        x = self._apply_targets(self._cf(run, x, times[0]), kwargs)
        B, D = x.shape
        t = times[0]
        for n in range(1, steps):
            if noise_fn is not None:
                noise = noise_fn(n, (B, D)).to(device=x.device, dtype=torch.float32).contiguous()
            else:
                noise = torch.empty_like(x)
                _lib.check(_lib.lib().bfk_normal_fill(_lib.ptr(noise), B, D, row_offset, None, None,
                                                      _lib.C.c_uint64((int(seed) + n * _GOLDEN) & (2 ** 64 - 1)), None,
                                                      _lib.stream_ptr()), "bfk_normal_fill")
            ct, st = float(np.cos(np.float32(t))), float(np.sin(np.float32(t)))
            x_n = self._apply_targets(_lincomb(torch.empty_like(x), ct, x, st, noise), kwargs)
            t = times[n]
            x = self._apply_targets(self._cf(run, x_n, t), kwargs)
        return x

    def _loss_and_grads(self, *args, **kwargs):
        raise NotImplementedError("StableConsistencyModel training (continuous-time consistency training, "
                                  "stable_consistency_model.py:262-314: a JVP through the subnet) is not built on the CUDA path")
