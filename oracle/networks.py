"""TimeMLP / FlowMatching / DiffusionModel — CPU oracle (TEST INFRASTRUCTURE, never shipped).

Plain-PyTorch FP32 restatement of the reference's Keras/torch path, statement by statement:

* TimeMLP.call                 /root/reference/bayesflow/networks/subnets/mlp/time_mlp.py:125-147
* ConditionalDenseBlock.call   /root/reference/bayesflow/networks/helpers/blocks/conditional_dense_block.py:114-131
* FiLM.call                    /root/reference/bayesflow/networks/helpers/embeddings/featurewise_linear_modulation.py:25-33
* FourierEmbedding.call        /root/reference/bayesflow/networks/helpers/embeddings/fourier_embedding.py:57-76
* FlowMatching                 /root/reference/bayesflow/networks/inference/flow_matching/flow_matching.py:132-178, 220-337
* DiffusionModel               /root/reference/bayesflow/networks/inference/diffusion/diffusion_model.py:142-232, 351-559, 711-717, 782-869
* noise schedules              /root/reference/bayesflow/networks/inference/diffusion/schedules/{noise_schedule,edm_noise_schedule,cosine_noise_schedule}.py
* DiagonalNormal               /root/reference/bayesflow/distributions/diagonal_normal.py:91-105

Keras (>=3.12, unpinned, not installed here) semantics ASSUMED — "parity unpinned" at this boundary:
  Dense: y = x @ kernel(in,out) + bias;  he_normal = truncated normal(±2σ), σ = sqrt(2/fan_in)/0.87962566103423978;
  LayerNormalization: last axis, epsilon=1e-3, biased variance, gamma*(x-mean)*rsqrt(var+eps)+beta;
  mish(x) = x*tanh(softplus(x));  silu/gelu(approximate=True, the keras 3 default)/relu/tanh/elu as in keras.activations;
  Dropout = inverted dropout, identity at inference;  "mse" = mean((y_pred-y_true)^2, axis=-1);
  random_normal initializer = N(0, 0.05^2).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Sequence

import torch
import torch.nn.functional as F

from .integrate import integrate

TIME_MLP_DEFAULTS = dict(  # networks/defaults.py:13-26
    widths=(256, 256, 256, 256, 256), time_embedding_dim=32, fourier_scale=30.0, activation="mish",
    kernel_initializer="he_normal", residual=True, dropout=0.05, norm="layer", merge="concat", film_use_gamma=False,
)


def activation_fn(name: str):
    if name == "mish":
        return lambda x: x * torch.tanh(F.softplus(x))
    if name in ("silu", "swish"):
        return F.silu
    if name == "gelu":
        return lambda x: F.gelu(x, approximate="tanh")  # keras.activations.gelu(x, approximate=True) is the keras 3 default
    if name == "relu":
        return F.relu
    if name == "tanh":
        return torch.tanh
    if name == "elu":
        return F.elu
    raise ValueError(f"unsupported activation {name!r}")


# --------------------------------------------------------------------------------------------------------------
# parameters, in the reference's variable creation order, kernels stored (in, out)
# --------------------------------------------------------------------------------------------------------------
def _he_normal(gen, fan_in, fan_out):
    std = math.sqrt(2.0 / fan_in) / 0.87962566103423978
    w = torch.empty(fan_in, fan_out, dtype=torch.float32)
    torch.nn.init.trunc_normal_(w, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
    return w


def _glorot_uniform(gen, fan_in, fan_out):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(fan_in, fan_out, generator=gen, dtype=torch.float32) * 2 - 1) * lim


def init_params(D: int, C: int | None, widths: Sequence[int] = TIME_MLP_DEFAULTS["widths"], *, time_embedding_dim=32,
                film_use_gamma=False, merge="concat", residual=True, norm="layer", seed=1234) -> "OrderedDict[str, torch.Tensor]":
    """Weights as Keras would create them (time_mlp.py:149-183, conditional_dense_block.py:133-156,
    fourier_embedding.py:78-87, flow_matching.py:188-199)."""
    g = torch.Generator().manual_seed(seed)
    p = OrderedDict()
    W0 = widths[0]
    T = time_embedding_dim + 1 if time_embedding_dim > 1 else 1
    if time_embedding_dim > 1:
        p["subnet.time_emb.frequencies"] = torch.randn(time_embedding_dim // 2, generator=g) * 0.05
    p["subnet.x_proj.kernel"] = _he_normal(g, D, W0)
    p["subnet.x_proj.bias"] = torch.zeros(W0)
    if C:
        p["subnet.c_proj.kernel"] = _he_normal(g, C, W0)
        p["subnet.c_proj.bias"] = torch.zeros(W0)
        K = 2 * W0 if merge == "concat" else W0
        p["subnet.merge_proj.kernel"] = _he_normal(g, K, W0)
        p["subnet.merge_proj.bias"] = torch.zeros(W0)
    prev = W0
    for i, w in enumerate(widths):
        pre = f"subnet.blocks.{i}."
        p[pre + "dense.kernel"] = _he_normal(g, prev, w)
        p[pre + "dense.bias"] = torch.zeros(w)
        if film_use_gamma:
            p[pre + "film.to_gamma.kernel"] = _he_normal(g, T, w)
            p[pre + "film.to_gamma.bias"] = torch.zeros(w)
        p[pre + "film.to_beta.kernel"] = _he_normal(g, T, w)
        p[pre + "film.to_beta.bias"] = torch.zeros(w)
        if norm in ("layer",):
            p[pre + "norm.gamma"] = torch.ones(w)
            p[pre + "norm.beta"] = torch.zeros(w)
        elif norm == "rms":
            p[pre + "norm.scale"] = torch.ones(w)
        if residual and prev != w:
            p[pre + "projector.kernel"] = _he_normal(g, prev, w)
            p[pre + "projector.bias"] = torch.zeros(w)
        prev = w
    p["output_projector.kernel"] = _glorot_uniform(g, prev, D)  # keras Dense default kernel init (flow_matching.py:188)
    p["output_projector.bias"] = torch.zeros(D)
    return p


def perturb_params(p, seed=4321, scale=0.1):
    """Perturb biases / gamma / beta so that a dropped term cannot hide."""
    g = torch.Generator().manual_seed(seed)
    q = OrderedDict()
    for k, v in p.items():
        if k.endswith("bias") or k.endswith("norm.gamma") or k.endswith("norm.beta") or k.endswith("norm.scale"):
            q[k] = v + scale * torch.randn(v.shape, generator=g)
        else:
            q[k] = v.clone()
    return q


def count_flops_per_eval(D, C, widths, T=33):
    """2*in*out per dense layer as the reference executes them (SURVEY.md §8d formula)."""
    W0 = widths[0]
    f = 2 * D * W0
    if C:
        f += 2 * C * W0 + 2 * (2 * W0) * W0
    prev = W0
    for w in widths:
        f += 2 * prev * w + 2 * T * w
        if prev != w:
            f += 2 * prev * w
        prev = w
    f += 2 * prev * D
    return f


# --------------------------------------------------------------------------------------------------------------
# integrate.py STOCHASTIC_METHODS (diffusion_model.py:733, 805, 830 test membership in THIS list, so method="scipy"
# takes the ODE branch)
_STOCHASTIC = ("euler_maruyama", "sea", "shark", "two_step_adaptive", "langevin")

class OracleTimeMLP:
    def __init__(self, params, *, widths, time_embedding_dim=32, fourier_scale=30.0, activation="mish", residual=True,
                 dropout=0.0, norm="layer", merge="concat", film_use_gamma=False, **_):
        self.p = params
        self.widths = tuple(widths)
        self.time_embedding_dim = time_embedding_dim
        self.fourier_scale = fourier_scale
        self.act = activation_fn(activation)
        self.residual = residual
        self.dropout = dropout
        self.norm = norm
        self.merge = merge
        self.film_use_gamma = film_use_gamma

    def time_emb(self, t):
        # fourier_embedding.py:57-76 (include_identity=True); Identity when time_embedding_dim == 1 (time_mlp.py:91-92)
        if self.time_embedding_dim == 1:
            return t
        w = self.p["subnet.time_emb.frequencies"]
        proj = t * w[None, :] * 2 * math.pi * self.fourier_scale
        return torch.cat([t, torch.sin(proj), torch.cos(proj)], dim=-1)

    def block(self, i, x, t_emb, training=False, dropout_masks=None):
        # conditional_dense_block.py:114-131
        pre = f"subnet.blocks.{i}."
        p = self.p
        h = x @ p[pre + "dense.kernel"] + p[pre + "dense.bias"]
        h = self.act(h)
        if training and self.dropout and self.dropout > 0:
            if dropout_masks is not None:  # externally supplied keep-mask (already scaled by 1/(1-rate))
                h = h * dropout_masks[i]
            else:
                h = F.dropout(h, self.dropout, training=True)
        beta = t_emb @ p[pre + "film.to_beta.kernel"] + p[pre + "film.to_beta.bias"]  # FiLM :25-33
        if self.film_use_gamma:
            gamma = t_emb @ p[pre + "film.to_gamma.kernel"] + p[pre + "film.to_gamma.bias"]
            h = (1.0 + gamma) * h + beta
        else:
            h = h + beta
        if self.residual:
            skip = x if (pre + "projector.kernel") not in p else x @ p[pre + "projector.kernel"] + p[pre + "projector.bias"]
            h = skip + h
        if self.norm == "layer":
            mean = h.mean(dim=-1, keepdim=True)
            var = h.var(dim=-1, unbiased=False, keepdim=True)
            h = (h - mean) * torch.rsqrt(var + 1e-3) * p[pre + "norm.gamma"] + p[pre + "norm.beta"]
        elif self.norm == "rms":
            # keras.layers.RMSNormalization: x * rsqrt(mean(x^2) + eps) * scale, eps = 1e-6
            h = h * torch.rsqrt(h.pow(2).mean(dim=-1, keepdim=True) + 1e-6) * p[pre + "norm.scale"]
        elif self.norm is not None:
            raise ValueError(f"unsupported norm {self.norm!r} in the oracle")
        return h

    def __call__(self, inputs, training=False, dropout_masks=None):
        # time_mlp.py:125-147
        x, t, conditions = inputs
        p = self.p
        h = x @ p["subnet.x_proj.kernel"] + p["subnet.x_proj.bias"]
        if conditions is not None and "subnet.c_proj.kernel" in p:
            hc = conditions @ p["subnet.c_proj.kernel"] + p["subnet.c_proj.bias"]
            if self.merge == "concat":
                h = torch.cat([h, hc], dim=-1)
            else:
                h = h + hc
            h = self.act(h) @ p["subnet.merge_proj.kernel"] + p["subnet.merge_proj.bias"]
        h = self.act(h)
        t_emb = self.time_emb(t)
        for i in range(len(self.widths)):
            h = self.block(i, h, t_emb, training=training, dropout_masks=dropout_masks)
        return h


class DiagonalNormal:
    # distributions/diagonal_normal.py:65-105 (mean 0, std 1 defaults)
    def __init__(self, dim, mean=0.0, std=1.0):
        self.dim = dim
        self._mean = torch.full((dim,), float(mean))
        self._std = torch.full((dim,), float(std))

    def log_prob(self, samples):
        result = -0.5 * torch.sum((samples - self._mean) ** 2 / self._std ** 2, dim=-1)
        return result + (-0.5 * self.dim * math.log(2.0 * math.pi) - torch.sum(torch.log(self._std)))

    def sample(self, batch_shape, generator=None):
        z = torch.randn(tuple(batch_shape) + (self.dim,), generator=generator)
        return self._mean + self._std * z


def _expand_time(time, xz):
    # flow_matching.py:226-228: convert_to_tensor(dtype of xz), expand_right_as, broadcast to (B,1)
    time = torch.as_tensor(time, dtype=xz.dtype)
    while time.ndim < xz.ndim:
        time = time.unsqueeze(-1)
    return torch.broadcast_to(time, xz.shape[:-1] + (1,))


def _mask(data, mask=None, replacement=None):
    # utils/tensor_utils.py:12-37
    if mask is None:
        return data
    if replacement is None:
        return mask * data
    return mask * data + (1 - mask) * replacement


def random_mask(shape, drop_prob, generator=None):
    # utils/tensor_utils.py:445-470: 1 (keep) where uniform > drop_prob; the python float 1.0 when drop_prob <= 0
    if drop_prob <= 0:
        return 1.0
    return (torch.rand(tuple(shape), generator=generator) > drop_prob).to(torch.float32)


def randomly_mask_along_axis(x, drop_prob, generator=None):
    # utils/tensor_utils.py:473-504 (axis 0): whole rows zeroed, the classifier-free-guidance dropout
    if drop_prob <= 0:
        return x
    return random_mask((x.shape[0],) + (1,) * (x.ndim - 1), drop_prob, generator) * x


class OracleFlowMatching:
    def __init__(self, params, D, subnet_kwargs=None, integrate_kwargs=None, time_power_law_alpha=0.0,
                 drop_cond_prob=0.0, drop_target_prob=0.0):
        kw = dict(TIME_MLP_DEFAULTS) | (subnet_kwargs or {})  # flow_matching.py:122-125
        self.p = params
        self.D = D
        self.subnet = OracleTimeMLP(params, **kw)
        self.integrate_kwargs = {"method": "tsit5", "steps": "adaptive"} | (integrate_kwargs or {})  # defaults.py:52
        self.base_distribution = DiagonalNormal(D)
        self.time_power_law_alpha = float(time_power_law_alpha)
        self.drop_cond_prob, self.drop_target_prob = float(drop_cond_prob), float(drop_target_prob)
        self.unconditional_mode = False

    def velocity(self, xz, time, conditions=None, training=False, dropout_masks=None, **kwargs):
        # flow_matching.py:220-236
        time = _expand_time(time, xz)
        subnet_out = self.subnet((xz, time, conditions), training=training, dropout_masks=dropout_masks)
        out = subnet_out @ self.p["output_projector.kernel"] + self.p["output_projector.bias"]
        if not training:
            out = _mask(out, kwargs.get("target_mask", None))
        return out

    def _prep(self, x, conditions, kwargs):
        target_mask = kwargs.get("target_mask", None)
        targets_fixed = kwargs.get("targets_fixed", None)
        if target_mask is not None:
            target_mask = torch.broadcast_to(target_mask, x.shape)
            targets_fixed = torch.broadcast_to(targets_fixed, x.shape)
            x = _mask(x, target_mask, targets_fixed)
        if self.unconditional_mode and conditions is not None:
            conditions = torch.zeros_like(conditions)
        return x, conditions

    def _velocity_trace(self, xz, time, conditions=None, **kwargs):
        # flow_matching.py:238-252 with utils/jacobian/jacobian_trace.py:10-55 (exact trace, D reverse passes)
        with torch.enable_grad():
            x = xz.detach().clone().requires_grad_(True)
            v = self.velocity(x, time=time, conditions=conditions, **kwargs)
            trace = torch.zeros(x.shape[:-1], dtype=x.dtype)
            for d in range(x.shape[-1]):
                (g,) = torch.autograd.grad(v[..., d].sum(), x, retain_graph=True)
                trace = trace + g[..., d]
        return v.detach(), trace.detach().unsqueeze(-1)

    def _integrate(self, x, conditions, density, start, stop, **kwargs):
        kwargs.pop("seed", None)
        integrate_kwargs = self.integrate_kwargs | kwargs
        x, conditions = self._prep(x, conditions, kwargs)
        mk = {k: v for k, v in kwargs.items() if k in ("target_mask",)}
        ik = {k: v for k, v in integrate_kwargs.items() if k in ("method", "steps", "min_steps", "max_steps", "atol", "rtol", "info")}
        if density:
            def deltas(time, xz, trace):
                v, tr = self._velocity_trace(xz, time=time, conditions=conditions, **mk)
                return {"xz": v, "trace": tr}
            state = {"xz": x, "trace": torch.zeros(x.shape[:-1] + (1,), dtype=x.dtype)}
            state = integrate(deltas, state, start_time=start, stop_time=stop, **ik)
            return state["xz"], state["trace"].squeeze(-1)

        def deltas(time, xz):
            return {"xz": self.velocity(xz, time=time, conditions=conditions, **mk)}
        state = integrate(deltas, {"xz": x}, start_time=start, stop_time=stop, **ik)
        return state["xz"]

    def forward(self, x, conditions=None, density=False, **kwargs):
        # flow_matching.py:254-294
        if density:
            z, trace = self._integrate(x, conditions, True, 1.0, 0.0, **kwargs)
            return z, self.base_distribution.log_prob(z) + trace
        return self._integrate(x, conditions, False, 1.0, 0.0, **kwargs)

    def inverse(self, z, conditions=None, density=False, **kwargs):
        # flow_matching.py:296-337
        if density:
            x, trace = self._integrate(z, conditions, True, 0.0, 1.0, **kwargs)
            return x, self.base_distribution.log_prob(z) - trace
        return self._integrate(z, conditions, False, 0.0, 1.0, **kwargs)

    def compute_metrics(self, x, conditions=None, sample_weight=None, stage="training", generator=None,
                        dropout_masks=None, cond_keep=None, target_keep=None):
        # flow_matching.py:132-178 (optimal transport at its default off).  `cond_keep` (B,1) / `target_keep` (B,D): fixed
        # 0 / 1 masks in place of the RNG draws of :167-170 (parity runs)
        if isinstance(x, (tuple, list)):
            x0, x1, t, x, target_velocity = x
        else:
            x1 = x
            x0 = self.base_distribution.sample(x1.shape[:-1], generator=generator)
            u = torch.rand((x0.shape[0],), generator=generator)
            t = u ** (1 + self.time_power_law_alpha)
            while t.ndim < x0.ndim:
                t = t.unsqueeze(-1)
            x = t * x1 + (1 - t) * x0
            target_velocity = x1 - x0
        if self.drop_cond_prob > 0 and conditions is not None:  # :167-168
            conditions = (cond_keep * conditions) if cond_keep is not None else \
                randomly_mask_along_axis(conditions, self.drop_cond_prob, generator)
        mask_x = target_keep if target_keep is not None else random_mask(x.shape, self.drop_target_prob, generator)  # :170
        if not isinstance(mask_x, float):
            x = mask_x * x + (1 - mask_x) * x1  # :171 maybe_mask_tensor(x, mask, replacement=x1)
        pred = self.velocity(x, time=t, conditions=conditions, training=stage == "training", dropout_masks=dropout_masks)
        loss = torch.mean((mask_x * pred - mask_x * target_velocity) ** 2, dim=-1)  # keras "mse" of the masked pair, :175
        loss = torch.mean(loss * sample_weight if sample_weight is not None else loss)  # weighted_mean
        return {"loss": loss}


# --------------------------------------------------------------------------------------------------------------
class EDMNoiseSchedule:
    # schedules/edm_noise_schedule.py:21-113
    def __init__(self, sigma_data=1.0, sigma_min=1e-4, sigma_max=80.0, variance_type="preserving"):
        self.sigma_data, self.sigma_min, self.sigma_max, self.rho = sigma_data, sigma_min, sigma_max, 7
        self.p_mean, self.p_std = -1.2, 1.2
        self._variance_type = variance_type
        self.log_snr_min = -2 * torch.log(torch.tensor(sigma_max, dtype=torch.float32))
        self.log_snr_max = -2 * torch.log(torch.tensor(sigma_min, dtype=torch.float32))
        self._log_snr_min_training = self.log_snr_min - 1
        self._log_snr_max_training = self.log_snr_max + 1

    def get_log_snr(self, t, training):
        if training:
            loc, scale = -2 * self.p_mean, 2 * self.p_std
            snr = loc + scale * torch.erfinv(2 * (1 - t) - 1) * math.sqrt(2)
            return torch.clip(snr, self._log_snr_min_training, self._log_snr_max_training)
        smin, smax = self.sigma_min ** (1 / self.rho), self.sigma_max ** (1 / self.rho)
        return -2 * self.rho * torch.log(smax + (1 - t) * (smin - smax))

    def get_t_from_log_snr(self, log_snr_t, training):
        if training:
            loc, scale = -2 * self.p_mean, 2 * self.p_std
            return 1 - 0.5 * (1 + torch.erf((log_snr_t - loc) / (scale * math.sqrt(2.0))))
        smin, smax = self.sigma_min ** (1 / self.rho), self.sigma_max ** (1 / self.rho)
        return 1 - ((torch.exp(-log_snr_t / (2 * self.rho)) - smax) / (smin - smax))

    def derivative_log_snr(self, log_snr_t, training=False):
        if training:
            raise NotImplementedError
        t = self.get_t_from_log_snr(log_snr_t, training)
        s_max, s_min = self.sigma_max ** (1 / self.rho), self.sigma_min ** (1 / self.rho)
        u = s_max + (1 - t) * (s_min - s_max)
        dsnr_dx = 2 * self.rho * (s_min - s_max) / u
        factor = torch.exp(-log_snr_t) / (1 + torch.exp(-log_snr_t))
        return -factor * dsnr_dx

    def get_weights_for_snr(self, log_snr_t):
        return 1 + torch.exp(-log_snr_t) / self.sigma_data ** 2

    # noise_schedule.py:70-119
    def get_drift(self, log_snr_t, x=None, training=False):
        beta = self.derivative_log_snr(log_snr_t, training)
        if x is None:
            return beta
        if self._variance_type == "preserving":
            return -0.5 * beta * x, beta
        return torch.zeros_like(beta), beta

    def get_alpha_sigma(self, log_snr_t):
        if self._variance_type == "preserving":
            return torch.sqrt(torch.sigmoid(log_snr_t)), torch.sqrt(torch.sigmoid(-log_snr_t))
        return torch.ones_like(log_snr_t), torch.sqrt(torch.exp(-log_snr_t))


class CosineNoiseSchedule(EDMNoiseSchedule):
    # schedules/cosine_noise_schedule.py:26-82
    def __init__(self, min_log_snr=-15.0, max_log_snr=15.0, shift=0.0, weighting="sigmoid"):
        self._variance_type = "preserving"
        self._shift, self._weighting = shift, weighting
        self.log_snr_min, self.log_snr_max = min_log_snr, max_log_snr
        self._t_min = self.get_t_from_log_snr(torch.tensor(float(max_log_snr)), True)
        self._t_max = self.get_t_from_log_snr(torch.tensor(float(min_log_snr)), True)

    def _truncated_t(self, t):
        return self._t_min + (self._t_max - self._t_min) * t

    def get_log_snr(self, t, training):
        return -2 * torch.log(torch.tan(math.pi * self._truncated_t(t) * 0.5)) + 2 * self._shift

    def get_t_from_log_snr(self, log_snr_t, training):
        return 2 / math.pi * torch.arctan(torch.exp((2 * self._shift - log_snr_t) * 0.5))

    def derivative_log_snr(self, log_snr_t, training=False):
        t = self.get_t_from_log_snr(log_snr_t, training)
        t_trunc = self._truncated_t(t)
        dsnr_dx = -(2 * math.pi) / torch.sin(math.pi * t_trunc)
        dsnr_dt = dsnr_dx * (self._t_max - self._t_min)
        factor = torch.exp(-log_snr_t) / (1 + torch.exp(-log_snr_t))
        return -factor * dsnr_dt

    def get_weights_for_snr(self, log_snr_t):
        if self._weighting == "sigmoid":
            return torch.sigmoid(-log_snr_t + 2)
        if self._weighting == "likelihood_weighting":
            return self.get_drift(log_snr_t) / torch.square(self.get_alpha_sigma(log_snr_t)[1])
        return torch.ones_like(log_snr_t)


class OracleDiffusionModel:
    def __init__(self, params, D, subnet_kwargs=None, integrate_kwargs=None, noise_schedule="edm",
                 prediction_type="F", loss_type="noise", schedule_kwargs=None, drop_cond_prob=0.0, drop_target_prob=0.0):
        self.drop_cond_prob, self.drop_target_prob = float(drop_cond_prob), float(drop_target_prob)
        kw = dict(TIME_MLP_DEFAULTS) | (subnet_kwargs or {})
        self.p, self.D = params, D
        self.subnet = OracleTimeMLP(params, **kw)
        self.noise_schedule = (EDMNoiseSchedule if noise_schedule == "edm" else CosineNoiseSchedule)(**(schedule_kwargs or {}))
        self.integrate_kwargs = {"method": "two_step_adaptive", "steps": "adaptive"} | (integrate_kwargs or {})
        self._prediction_type, self._loss_type = prediction_type, loss_type
        self.base_distribution = DiagonalNormal(D)

    def _transform_log_snr(self, log_snr):
        # diffusion_model.py:711-717
        ns = self.noise_schedule
        return 2 * ((log_snr - ns.log_snr_min) / (ns.log_snr_max - ns.log_snr_min)) - 1

    def convert_prediction_to_x(self, pred, z, alpha_t, sigma_t, log_snr_t, prediction_type):
        # diffusion_model.py:351-405
        if prediction_type == "velocity":
            return alpha_t * z - sigma_t * pred
        if prediction_type == "noise":
            return (z - sigma_t * pred) / alpha_t
        if prediction_type == "F":
            sd = getattr(self.noise_schedule, "sigma_data", 1.0)
            x1 = (sd ** 2 * alpha_t) / (torch.exp(-log_snr_t) + sd ** 2)
            x2 = torch.exp(-log_snr_t / 2) * sd / torch.sqrt(torch.exp(-log_snr_t) + sd ** 2)
            return x1 * z + x2 * pred
        if prediction_type == "x":
            return pred
        if prediction_type == "score":
            return (z + sigma_t ** 2 * pred) / alpha_t
        raise ValueError(prediction_type)

    def _subnet_pred(self, xz, norm_log_snr, conditions, training=False, dropout_masks=None):
        out = self.subnet((xz, norm_log_snr, conditions), training=training, dropout_masks=dropout_masks)
        return out @ self.p["output_projector.kernel"] + self.p["output_projector.bias"]

    def score(self, xz, log_snr_t, conditions=None, training=False):
        # diffusion_model.py:407-499 (no guidance)
        alpha_t, sigma_t = self.noise_schedule.get_alpha_sigma(log_snr_t)
        pred = self._subnet_pred(xz, self._transform_log_snr(log_snr_t), conditions, training)
        x_pred = self.convert_prediction_to_x(pred, xz, alpha_t, sigma_t, log_snr_t, self._prediction_type)
        return (alpha_t * x_pred - xz) / torch.square(sigma_t)

    def velocity(self, xz, time, stochastic_solver=False, conditions=None, training=False, **kwargs):
        # diffusion_model.py:501-559
        t = torch.as_tensor(time, dtype=xz.dtype)
        log_snr_t = self.noise_schedule.get_log_snr(t, training=training)
        while log_snr_t.ndim < xz.ndim:
            log_snr_t = log_snr_t.unsqueeze(-1)
        log_snr_t = torch.broadcast_to(log_snr_t, xz.shape[:-1] + (1,))
        score = self.score(xz, log_snr_t, conditions, training)
        f, g2 = self.noise_schedule.get_drift(log_snr_t, x=xz, training=training)
        out = f - g2 * score if stochastic_solver else f - 0.5 * g2 * score
        if not training:
            out = _mask(out, kwargs.get("target_mask", None))
        return out

    def diffusion_term(self, xz, time, training=False, **kwargs):
        # diffusion_model.py:561-595
        log_snr_t = self.noise_schedule.get_log_snr(torch.as_tensor(time, dtype=xz.dtype), training=training)
        while log_snr_t.ndim < xz.ndim:
            log_snr_t = log_snr_t.unsqueeze(-1)
        log_snr_t = torch.broadcast_to(log_snr_t, xz.shape[:-1] + (1,))
        g = torch.sqrt(self.noise_schedule.get_drift(log_snr_t))
        if not training:
            g = _mask(g, kwargs.get("target_mask", None))
        return g

    def _velocity_trace(self, xz, time, conditions=None, **kwargs):
        # diffusion_model.py `_velocity_trace` -> utils/jacobian/jacobian_trace.py:10-55 (exact trace, D reverse passes)
        with torch.enable_grad():
            x = xz.detach().clone().requires_grad_(True)
            v = self.velocity(x, time=time, stochastic_solver=False, conditions=conditions, **kwargs)
            trace = torch.zeros(x.shape[:-1], dtype=x.dtype)
            for d in range(x.shape[-1]):
                (g,) = torch.autograd.grad(v[..., d].sum(), x, retain_graph=True)
                trace = trace + g[..., d]
        return v.detach(), trace.detach().unsqueeze(-1)

    _IK = ("start_time", "stop_time", "method", "steps", "min_steps", "max_steps", "atol", "rtol", "info")

    def _density_integrate(self, x, conditions, ik):
        def deltas(time, xz, trace):
            v, tr = self._velocity_trace(xz, time=time, conditions=conditions)
            return {"xz": v, "trace": tr}
        state = {"xz": x, "trace": torch.zeros(x.shape[:-1] + (1,), dtype=x.dtype)}
        state = integrate(deltas, state, **ik)
        return state["xz"], state["trace"].squeeze(-1)

    def forward(self, x, conditions=None, density=False, **kwargs):
        # diffusion_model.py:719-780
        ik = {"start_time": 0.0, "stop_time": 1.0} | self.integrate_kwargs | kwargs
        if ik["method"] in _STOCHASTIC:
            ik["method"] = "tsit5"  # :733-739
        ik = {k: v for k, v in ik.items() if k in self._IK}
        if density:
            z, trace = self._density_integrate(x, conditions, ik)
            return z, self.base_distribution.log_prob(z) + trace

        def deltas(time, xz):
            return {"xz": self.velocity(xz, time=time, stochastic_solver=False, conditions=conditions)}
        return integrate(deltas, {"xz": x}, **ik)["xz"]

    def inverse(self, z, conditions=None, density=False, **kwargs):
        # diffusion_model.py:782-869, deterministic branch :857-866; density branch :804-826
        kwargs.pop("seed", None)
        ik = {"start_time": 1.0, "stop_time": 0.0} | self.integrate_kwargs | kwargs
        ik = {k: v for k, v in ik.items() if k in self._IK}
        if kwargs.get("target_mask") is not None:  # :793-798: the fixed coordinates start at their targets
            tm = torch.broadcast_to(kwargs["target_mask"], z.shape)
            z = _mask(z, tm, torch.broadcast_to(kwargs["targets_fixed"], z.shape))
        if density:
            if ik["method"] in _STOCHASTIC:
                ik["method"] = "tsit5"
            x, trace = self._density_integrate(z, conditions, ik)
            return x, self.base_distribution.log_prob(z) - trace
        if ik["method"] in _STOCHASTIC:
            # SDE branch, diffusion_model.py:830-856
            from .integrate import integrate_stochastic

            def sde_drift(time, xz):
                return {"xz": self.velocity(xz, time=time, stochastic_solver=True, conditions=conditions,
                                            target_mask=kwargs.get("target_mask"))}

            def sde_diffusion(time, xz):
                return {"xz": self.diffusion_term(xz, time=time, target_mask=kwargs.get("target_mask"))}
            extra = {k: kwargs[k] for k in ("noise_fn", "generator", "e_rel", "e_abs", "r", "adapt_safety", "corrector_steps",
                                            "step_size_factor") if k in kwargs}
            score_fn = None
            if "corrector_steps" in kwargs or ik["method"] == "langevin":  # :845-849
                def score_fn(time, xz):
                    lam = self.noise_schedule.get_log_snr(torch.as_tensor(time, dtype=torch.float32), training=False)
                    lam = torch.broadcast_to(lam.reshape((1,) * xz.ndim), xz.shape[:-1] + (1,))
                    return {"xz": self.score(xz, lam, conditions=conditions)}
            return integrate_stochastic(sde_drift, sde_diffusion, {"xz": z}, score_fn=score_fn,
                                        noise_schedule=self.noise_schedule, **ik, **extra)["xz"]

        def deltas(time, xz):
            return {"xz": self.velocity(xz, time=time, stochastic_solver=False, conditions=conditions,
                                        target_mask=kwargs.get("target_mask"))}
        return integrate(deltas, {"xz": z}, **ik)["xz"]

    def compute_metrics(self, x, conditions=None, sample_weight=None, stage="training", fixed=None, generator=None,
                        dropout_masks=None, cond_keep=None, target_keep=None):
        # diffusion_model.py:142-232; `fixed=(u0, eps_t)` bypasses the RNG for parity runs, `cond_keep` (B,1) /
        # `target_keep` (B,D) are fixed 0 / 1 masks in place of the draws of :155-156 / :177
        training = stage == "training"
        B = x.shape[0]
        if conditions is not None and self.drop_cond_prob > 0:
            conditions = (cond_keep * conditions) if cond_keep is not None else \
                randomly_mask_along_axis(conditions, self.drop_cond_prob, generator)
        if fixed is None:
            u0 = torch.rand((1,), generator=generator, dtype=x.dtype)
            eps_t = None
        else:
            u0, eps_t = fixed
        i = torch.arange(0, B, dtype=x.dtype)
        t = (u0 + i / float(B)) % 1
        log_snr_t = self.noise_schedule.get_log_snr(t, training=stage in ("training", "validation"))
        while log_snr_t.ndim < x.ndim:
            log_snr_t = log_snr_t.unsqueeze(-1)
        alpha_t, sigma_t = self.noise_schedule.get_alpha_sigma(log_snr_t)
        w = self.noise_schedule.get_weights_for_snr(log_snr_t)
        if eps_t is None:
            eps_t = torch.randn(x.shape, generator=generator, dtype=x.dtype)
        diffused_x = alpha_t * x + sigma_t * eps_t
        mask_x = target_keep if target_keep is not None else random_mask(x.shape, self.drop_target_prob, generator)  # :177
        if not isinstance(mask_x, float):
            diffused_x = mask_x * diffused_x + (1 - mask_x) * x  # :178
        pred = self._subnet_pred(diffused_x, self._transform_log_snr(log_snr_t), conditions, training, dropout_masks)
        x_pred = self.convert_prediction_to_x(pred, diffused_x, alpha_t, sigma_t, log_snr_t, self._prediction_type)
        if self._loss_type == "noise":
            noise_pred = (diffused_x - alpha_t * x_pred) / sigma_t
            per_row = torch.mean(mask_x * (noise_pred - eps_t) ** 2, dim=-1)
        elif self._loss_type == "velocity":
            velocity_pred = (alpha_t * diffused_x - x_pred) / sigma_t
            v_t = alpha_t * eps_t - sigma_t * x
            per_row = torch.mean(mask_x * (velocity_pred - v_t) ** 2, dim=-1)
        else:
            sd = getattr(self.noise_schedule, "sigma_data", 1.0)
            x1 = torch.sqrt(torch.exp(-log_snr_t) + sd ** 2) / (torch.exp(-log_snr_t / 2) * sd)
            x2 = (sd * alpha_t) / (torch.exp(-log_snr_t / 2) * torch.sqrt(torch.exp(-log_snr_t) + sd ** 2))
            per_row = torch.mean(mask_x * (x1 * x_pred - x2 * diffused_x - (x1 * x - x2 * diffused_x)) ** 2, dim=-1)
        # Reference quirk kept on purpose (diffusion_model.py:212): weights_for_snr is (B,1) (log_snr_t was
        # right-expanded at :165) while ops.mean(..., axis=-1) is (B,), so the product broadcasts to (B,B) and
        # weighted_mean averages the outer product, i.e. mean(w) * mean(per-row mse).
        loss = w * per_row
        loss = torch.mean(loss * sample_weight if sample_weight is not None else loss)
        return {"loss": loss}
