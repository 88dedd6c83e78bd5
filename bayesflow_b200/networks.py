"""Host-side mirror of the reference's inference-network surface (torch ``nn.Module``s, no Keras):

  InferenceNetwork  <- bayesflow/networks/inference/inference_network.py:12-138
  TimeMLP           <- bayesflow/networks/subnets/mlp/time_mlp.py:14-216           (weights + config only)
  FlowMatching      <- bayesflow/networks/inference/flow_matching/flow_matching.py:27-337
  DiffusionModel    <- bayesflow/networks/inference/diffusion/diffusion_model.py:35-869 (probability-flow ODE path)
  DiagonalNormal    <- bayesflow/distributions/diagonal_normal.py:17-105

Same constructor keywords, method names, kwargs merge order and error types as the reference.  All arithmetic on
the hot path runs in libbfk (sm_100a CUDA) through the C ABI in include/bfk.h; there is no eager / CPU path.
Parameters are stored in the reference's variable order with (in, out) kernels so that ``state_dict()`` values
can be exchanged with a Keras ``FlowMatching`` via get_weights()/set_weights().
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
from collections.abc import Sequence

import numpy as np
import torch
from torch import nn

from . import _lib
from .integrate import integrate as _integrate

# networks/defaults.py:13-26, 50, 52
TIME_MLP_DEFAULTS = {
    "widths": (256, 256, 256, 256, 256), "time_embedding_dim": 32, "fourier_scale": 30.0, "activation": "mish",
    "kernel_initializer": "he_normal", "residual": True, "dropout": 0.05, "norm": "layer", "merge": "concat",
    "film_use_gamma": False,
}
FLOW_MATCHING_INTEGRATE_DEFAULTS = {"method": "tsit5", "steps": "adaptive"}
DIFFUSION_INTEGRATE_DEFAULTS = {"method": "two_step_adaptive", "steps": "adaptive"}
DETERMINISTIC_METHODS = ["euler", "rk45", "tsit5"]
STOCHASTIC_METHODS = ["euler_maruyama", "sea", "shark", "two_step_adaptive", "langevin"]


def _he_normal(fan_in, fan_out, generator=None):
    std = math.sqrt(2.0 / fan_in) / 0.87962566103423978
    w = torch.empty(fan_in, fan_out)
    nn.init.trunc_normal_(w, 0.0, std, -2 * std, 2 * std, generator=generator)
    return w


def _glorot_uniform(fan_in, fan_out, generator=None):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(fan_in, fan_out, generator=generator) * 2 - 1) * lim


class _Dense(nn.Module):
    """Parameter holder for a keras.layers.Dense: kernel (in,out), bias (out)."""

    def __init__(self, fan_in, units, init="he_normal"):
        super().__init__()
        self.kernel = nn.Parameter(_he_normal(fan_in, units) if init == "he_normal" else _glorot_uniform(fan_in, units))
        self.bias = nn.Parameter(torch.zeros(units))


class DiagonalNormal(nn.Module):
    """distributions/diagonal_normal.py:17-105 (fixed mean / std)."""

    def __init__(self, mean=0.0, std=1.0, **kwargs):
        super().__init__()
        self.mean, self.std, self.dim, self.built = mean, std, None, False

    def build(self, input_shape):
        if self.built:
            return
        self.dim = int(input_shape[-1])
        self.register_buffer("_mean", torch.broadcast_to(torch.as_tensor(self.mean, dtype=torch.float32), (self.dim,)).clone())
        self.register_buffer("_std", torch.broadcast_to(torch.as_tensor(self.std, dtype=torch.float32), (self.dim,)).clone())
        self.built = True

    def log_prob(self, samples, *, normalize=True):
        result = -0.5 * torch.sum((samples - self._mean) ** 2 / self._std ** 2, dim=-1)
        if normalize:
            result = result + (-0.5 * self.dim * math.log(2.0 * math.pi) - torch.sum(torch.log(self._std)))
        return result

    def sample(self, batch_shape, seed=None, row_offset: int = 0):
        """mean + std * z with z written in HBM by `bfk_normal_fill` (Philox4x32-10 keyed by `seed`; the counter is the
        GLOBAL element index, so rows [row_offset, row_offset + rows) equal the same rows of one big call)."""
        if isinstance(batch_shape, int):
            batch_shape = (batch_shape,)
        dev = self._mean.device
        if dev.type != "cuda":
            raise RuntimeError("bayesflow_b200 generates base samples on CUDA only (no CPU fallback); move the network "
                               "to a CUDA device")
        if isinstance(seed, torch.Generator):
            seed = int(torch.randint(0, 2 ** 62, (1,), generator=seed, device=seed.device).item())
        elif seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        out = torch.empty(tuple(batch_shape) + (self.dim,), device=dev, dtype=torch.float32)
        rows = out.numel() // self.dim
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().bfk_normal_fill(_lib.ptr(out), rows, self.dim, int(row_offset), _lib.ptr(self._mean),
                                                  _lib.ptr(self._std), C.c_uint64(int(seed) & (2 ** 64 - 1)), None,
                                                  _lib.stream_ptr()), "bfk_normal_fill")
        return out


class _Block(nn.Module):
    def __init__(self, fan_in, width, T, film_use_gamma, norm, projector=False):
        super().__init__()
        self.dense = _Dense(fan_in, width)
        self.film = nn.Module()
        if film_use_gamma:
            self.film.to_gamma = _Dense(T, width)
        self.film.to_beta = _Dense(T, width)
        if norm == "layer":
            self.norm = nn.Module()
            self.norm.gamma = nn.Parameter(torch.ones(width))
            self.norm.beta = nn.Parameter(torch.zeros(width))
        elif norm == "rms":
            self.norm = nn.Module()
            self.norm.scale = nn.Parameter(torch.ones(width))
        if projector:  # residual projector for unequal in / out widths (conditional_dense_block.py:152-156)
            self.projector = _Dense(fan_in, width)


class TimeMLP(nn.Module):
    """Weights and configuration of the reference TimeMLP (time_mlp.py:54-183).  The arithmetic of ``call``
    (time_mlp.py:125-147) lives in the CUDA kernels; this class only owns parameters in reference order."""

    def __init__(self, widths: Sequence[int] = (256, 256), *, time_embedding_dim: int = 32, time_emb=None,
                 fourier_scale: float = 30.0, activation="mish", kernel_initializer="he_normal", residual: bool = True,
                 dropout=0.05, norm="layer", merge="concat", film_use_gamma: bool = False, **kwargs):
        super().__init__()
        if len(widths) == 0:
            raise ValueError("TimeMLP requires at least one hidden width.")
        if merge not in ("add", "concat"):
            raise ValueError(f"Unknown merge mode: {merge!r} (expected 'add' or 'concat').")
        if time_emb is not None:
            raise NotImplementedError("custom time_emb layers are outside the accelerated path")
        if not isinstance(activation, str) or activation not in _lib.ACT:
            raise ValueError(f"activation {activation!r} is not supported by the CUDA path; one of {sorted(_lib.ACT)}")
        if norm not in (None, "layer", "rms"):
            if isinstance(norm, str) and norm != "batch":
                raise ValueError(f"Unknown normalization strategy: {norm!r}.")
            raise NotImplementedError(f"norm={norm!r} is outside the accelerated path (layer / rms / None)")
        if kernel_initializer != "he_normal":
            raise NotImplementedError("only kernel_initializer='he_normal' is implemented")
        if time_embedding_dim != 1 and time_embedding_dim % 2 != 0:
            raise ValueError(f"Embedding dimension must be even, but is {time_embedding_dim}.")
        self.widths = tuple(int(w) for w in widths)
        self.time_embedding_dim, self.fourier_scale = int(time_embedding_dim), float(fourier_scale)
        self.activation, self.kernel_initializer, self.residual = activation, kernel_initializer, bool(residual)
        self.dropout, self.norm, self.merge, self.film_use_gamma = dropout, norm, merge, bool(film_use_gamma)
        self.built = False

    @property
    def T(self):
        return self.time_embedding_dim + 1 if self.time_embedding_dim > 1 else 1

    def build(self, input_shape):
        if self.built:
            return
        x_shape, t_shape, conditions_shape = input_shape
        D, W0 = int(x_shape[-1]), self.widths[0]
        if self.time_embedding_dim > 1:
            self.time_emb = nn.Module()
            self.time_emb.frequencies = nn.Parameter(torch.randn(self.time_embedding_dim // 2) * 0.05)
        self.x_proj = _Dense(D, W0)
        self.c_proj = self.merge_proj = None
        if conditions_shape is not None:
            Cdim = int(conditions_shape[-1])
            self.c_proj = _Dense(Cdim, W0)
            self.merge_proj = _Dense(2 * W0 if self.merge == "concat" else W0, W0)
        blocks, prev = [], W0
        for w in self.widths:
            # unequal widths: zero-padded to one common width inside the kernels, with a residual projector where a
            # block changes the width (served by the layer-wise training / density kernels, see _RunTE)
            if w != prev and (w % 8 or prev % 8):
                raise NotImplementedError("unequal hidden widths must be multiples of 8 on the accelerated path")
            blocks.append(_Block(prev, w, self.T, self.film_use_gamma, self.norm, projector=bool(self.residual and w != prev)))
            prev = w
        self.blocks = nn.ModuleList(blocks)
        self.built = True

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[0][:-1]) + (self.widths[-1],)

    def get_config(self):
        return {"widths": self.widths, "time_embedding_dim": self.time_embedding_dim, "time_emb": None,
                "fourier_scale": self.fourier_scale, "activation": self.activation,
                "kernel_initializer": self.kernel_initializer, "residual": self.residual, "dropout": self.dropout,
                "norm": self.norm, "merge": self.merge, "film_use_gamma": self.film_use_gamma}


def find_network(arg, **kwargs):
    """utils/dispatch/find_network.py:5-36 restricted to what the accelerated path covers."""
    if isinstance(arg, str):
        if arg == "time_mlp":
            return TimeMLP(**kwargs)
        raise ValueError(f"Unknown or unsupported network name for the accelerated path: {arg!r}")
    if isinstance(arg, type):
        if issubclass(arg, TimeMLP):
            return arg(**kwargs)
        raise TypeError(f"Cannot build a subnet from {arg!r}; the accelerated path needs a TimeMLP")
    if isinstance(arg, TimeMLP):
        return arg
    raise TypeError(f"Cannot infer network from {arg!r}.")


# ======================================================================================================================
# CUDA engine: owns the packed weights and turns python calls into C-ABI calls
# ======================================================================================================================
class _Engine:
    def __init__(self, net: "InferenceNetwork"):
        self.net = net
        self._packed = None
        self._key = None
        self.cfg = None
        self.layerwise = False
        self.train_engine_sampler = False
        self.kernel_events = None  # set to a list to have CUDA events recorded around the sampler kernel
        self.launches = 0          # kernels launched so far (bench.py reports it as gpu_launches)

    # -- configuration / weights -------------------------------------------------------------------------------------
    def _make_cfg(self):
        sub, net = self.net.subnet, self.net
        cfg = _lib.NetConfig()
        cfg.D, cfg.C = net._D, (net._C or 0)
        cfg.W, cfg.L, cfg.T = sub.widths[0], len(sub.widths), sub.T
        if any(w != sub.widths[0] for w in sub.widths):
            # unequal hidden widths: W = the common padded width, block_width[l] = the true widths (include/bfk.h)
            cfg.W = (max(sub.widths) + 63) // 64 * 64
            for l, w in enumerate(sub.widths):
                cfg.block_width[l] = int(w)
        cfg.activation = _lib.ACT[sub.activation]
        cfg.merge = _lib.MERGE[sub.merge] if net._C else 0
        cfg.norm = _lib.NORM[sub.norm]
        cfg.residual, cfg.film_use_gamma = int(sub.residual), int(sub.film_use_gamma)
        cfg.precision = _lib.PREC[net.precision]
        cfg.fourier_scale = sub.fourier_scale
        return cfg

    def prepare(self):
        net, sub = self.net, self.net.subnet
        params = list(net.parameters())
        key = (net.precision,) + tuple((p.data_ptr(), p._version) for p in params)
        if self._packed is not None and key == self._key:
            return
        L = _lib.lib()
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("bayesflow_b200: move the network to a CUDA device first (no CPU fallback)")
        self.cfg = cfg = self._make_cfg()
        self.train_engine_sampler = any(cfg.block_width[l] for l in range(cfg.L))
        if self.train_engine_sampler:
            # unequal hidden widths: every velocity evaluation is a forward pass of the training kernels (_RunTE)
            self.layerwise = False
            self._packed, self._key = torch.empty(1, dtype=torch.uint8, device=dev), key
            return
        self.layerwise = not L.bfk_fused_supported(C.byref(cfg))
        if self.layerwise:
            # W > 256 or D > 8: the layer-wise tcgen05 path (bfk_sample_layerwise) reads the fp32 weights directly
            from .training import _bind, fill_weights
            _bind(L)
            L.bfk_layerwise_workspace_bytes.restype = C.c_size_t
            L.bfk_layerwise_workspace_bytes.argtypes = [C.POINTER(_lib.NetConfig), C.c_int64, C.c_int64, C.c_int64]
            if L.bfk_layerwise_workspace_bytes(C.byref(cfg), 128, 1, 1) == 0:
                why = L.bfk_last_error_string().decode()
                if L.bfk_train_workspace_bytes(C.byref(cfg), 128, 0) != 0:
                    # e.g. film_use_gamma or norm="rms" with W > 256 / D > 8: no sampler kernel, but the training forward
                    # has it -> host-driven sampling over bfk_train_step (_RunTE), a coverage path
                    self.train_engine_sampler, self.layerwise = True, False
                    self._packed, self._key = torch.empty(1, dtype=torch.uint8, device=dev), key
                    return
                raise NotImplementedError(f"config D={cfg.D} W={cfg.W} L={cfg.L}: neither the fused sampler (D<=8, W<=256, "
                                          f"W%32==0) nor the layer-wise path covers it: " + why)
            self.w, self._keep_w = fill_weights(net, lambda p: p.detach())
            self._packed, self._key = torch.empty(1, dtype=torch.uint8, device=dev), key
            self._ws = None
            return
        w = _lib.Weights()
        keep = []

        def P(t):
            t = t.detach().to(torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        if sub.time_embedding_dim > 1:
            w.freq = P(sub.time_emb.frequencies)
        w.x_k, w.x_b = P(sub.x_proj.kernel), P(sub.x_proj.bias)
        if net._C:
            w.c_k, w.c_b = P(sub.c_proj.kernel), P(sub.c_proj.bias)
            w.m_k, w.m_b = P(sub.merge_proj.kernel), P(sub.merge_proj.bias)
        for i, b in enumerate(sub.blocks):
            w.dense_k[i], w.dense_b[i] = P(b.dense.kernel), P(b.dense.bias)
            if sub.film_use_gamma:
                w.film_g_k[i], w.film_g_b[i] = P(b.film.to_gamma.kernel), P(b.film.to_gamma.bias)
            w.film_b_k[i], w.film_b_b[i] = P(b.film.to_beta.kernel), P(b.film.to_beta.bias)
            if sub.norm == "layer":
                w.norm_g[i], w.norm_b[i] = P(b.norm.gamma), P(b.norm.beta)
            elif sub.norm == "rms":
                w.norm_g[i] = P(b.norm.scale)
        w.out_k, w.out_b = P(net.output_projector.kernel), P(net.output_projector.bias)
        nbytes = L.bfk_packed_bytes(C.byref(cfg))
        packed = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _lib.check(L.bfk_pack(C.byref(cfg), C.byref(w), _lib.ptr(packed), _lib.stream_ptr()), "bfk_pack")
        self._packed, self._key = packed, key
        self._tstride = int(L.bfk_table_stride(C.byref(cfg)))

    # -- per-call prologues --------------------------------------------------------------------------------------------
    def cond_tables(self, conditions, B, repeats=None):
        """conditions -> (pre_add, post_add, rows_per_group).  Rows of a dataset share their condition vector
        (approximators/helpers/samplers.py:51-63), so the conditions half of TimeMLP is computed once per dataset."""
        cond, repeats = self.group_conditions(conditions, B, repeats)
        if cond is None:
            return None, None, 1
        G = cond.shape[0]
        W = self.cfg.W
        post = torch.empty(G, W, device=cond.device, dtype=torch.float32)
        pre = torch.empty(G, W, device=cond.device, dtype=torch.float32) if self.net.subnet.merge == "add" else None
        _lib.check(_lib.lib().bfk_cond_prologue(C.byref(self.cfg), _lib.ptr(self._packed), _lib.ptr(cond), G,
                                                _lib.ptr(pre), _lib.ptr(post), _lib.stream_ptr()), "bfk_cond_prologue")
        self.launches += 1
        return pre, post, repeats

    def group_conditions(self, conditions, B, repeats=None):
        """-> ((G, C) conditions, rows per dataset): consecutive rows sharing one condition vector are one dataset."""
        net = self.net
        if conditions is None or not net._C:
            return None, 1
        cond = conditions.detach().to(torch.float32)
        if cond.ndim != 2 or cond.shape[-1] != net._C:
            raise ValueError(f"conditions must have shape (rows, {net._C}), got {tuple(cond.shape)}")
        if repeats is None:
            if cond.shape[0] != B:
                raise ValueError(f"conditions has {cond.shape[0]} rows, expected {B}")
            repeats = 1
            if net.detect_repeats and B > 1:
                # the reference's Sampler hands over each dataset's conditions repeated num_samples times: recover
                # (datasets, repeats) with two streaming launches (no sort / unique pass over the (N*S, C) tensor)
                cond = cond.contiguous()
                if not cond.is_cuda:
                    raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback); got conditions on "
                                       + str(cond.device))
                out2 = torch.empty(2, device=cond.device, dtype=torch.int64)
                with torch.cuda.device(cond.device):
                    _lib.check(_lib.lib().bfk_detect_repeats(_lib.ptr(cond), B, net._C, C.c_void_p(out2.data_ptr()),
                                                             _lib.stream_ptr()),
                               "bfk_detect_repeats")
                self.launches += 2
                R, bad = (int(v) for v in out2.tolist())
                if not bad and 1 < R <= B and B % R == 0:
                    cond, repeats = cond[::R], R
        elif cond.shape[0] * repeats != B:
            raise ValueError(f"conditions rows ({cond.shape[0]}) x repeats ({repeats}) != batch ({B})")
        return cond.contiguous(), repeats

    def eval_table(self, times: np.ndarray):
        dev = self._packed.device
        t = torch.from_numpy(np.ascontiguousarray(times, dtype=np.float32)).to(dev, non_blocking=True)
        table = torch.empty(t.numel(), self._tstride, device=dev, dtype=torch.float32)
        sched = self.net._schedule_struct()
        _lib.check(_lib.lib().bfk_eval_table(C.byref(self.cfg), _lib.ptr(self._packed), C.byref(sched), _lib.ptr(t),
                                             t.numel(), _lib.ptr(table), _lib.stream_ptr()), "bfk_eval_table")
        self.launches += 1
        return table


class _Run:
    """One sampling call: the engine plus the per-call hoisted tables."""

    def __init__(self, engine, pre, post, rpg, target_mask):
        self.e, self.pre, self.post, self.rpg = engine, pre, post, rpg
        self.tmask = None
        if target_mask is not None:
            tm = torch.as_tensor(target_mask, dtype=torch.float32, device=engine._packed.device)
            if tm.numel() != engine.cfg.D:
                raise NotImplementedError("target_mask must be feature-level, shape (D,), on the accelerated path")
            self.tmask = tm.reshape(-1).contiguous()
        self.info = {}

    def _check_x(self, x):
        if not x.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        if x.ndim != 2 or x.shape[-1] != self.e.cfg.D:
            raise ValueError(f"expected (rows, {self.e.cfg.D}) state, got {tuple(x.shape)}")
        return x.detach().to(torch.float32).contiguous()

    def velocity(self, x, time: float):
        e = self.e
        x = self._check_x(x)
        table = e.eval_table(np.asarray([time], dtype=np.float32))
        out = torch.empty_like(x)
        _lib.check(_lib.lib().bfk_velocity(C.byref(e.cfg), _lib.ptr(e._packed), _lib.ptr(table), _lib.ptr(self.pre),
                                           _lib.ptr(self.post), self.rpg, _lib.ptr(self.tmask), _lib.ptr(x),
                                           _lib.ptr(out), x.shape[0], _lib.stream_ptr()), "bfk_velocity")
        e.launches += 1
        return out

    def run_fixed(self, x, times, hs, method):
        e = self.e
        x = self._check_x(x)
        table = e.eval_table(times)
        h = torch.from_numpy(np.ascontiguousarray(hs, dtype=np.float32)).to(x.device, non_blocking=True)
        out = torch.empty_like(x)
        ev = None
        if e.kernel_events is not None:  # bench.py: CUDA events around the dominant kernel, on the launching stream
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        _lib.check(_lib.lib().bfk_sample_fixed(C.byref(e.cfg), _lib.ptr(e._packed), _lib.ptr(table), _lib.ptr(h),
                                               _lib.METHOD[method], h.numel(), _lib.ptr(self.pre), _lib.ptr(self.post),
                                               self.rpg, _lib.ptr(self.tmask), _lib.ptr(x), _lib.ptr(out), x.shape[0],
                                               _lib.stream_ptr()), "bfk_sample_fixed")
        if ev is not None:
            ev[1].record()
            e.kernel_events.append(ev)
        e.launches += 1
        self.info = {"steps": int(h.numel()), "evals": int(len(times)), "rejected": 0}
        return out

    def run_adaptive(self, x, start_time, stop_time, min_steps, max_steps, method, atol=1e-6, rtol=1e-4):
        """integrate_adaptive (integrate.py:291-406).  The step-size controller runs on the device (bfk_adaptive_plan);
        the host only enqueues trial steps and polls the `done` flag every few trials."""
        e, L, K = self.e, _lib.lib(), _lib.CTRL
        if method == "euler":
            raise ValueError("Adaptive step sizing is not supported for the 'euler' method.")  # integrate.py:304-306
        if method not in ("rk45", "tsit5"):
            if isinstance(method, str):
                raise ValueError(f"Unknown integration method name: {method!r}")
            raise TypeError(f"Invalid integration method: {method!r}")
        if max_steps <= min_steps:
            raise ValueError("Maximum number of steps must be greater than minimum number of steps.")
        x = self._check_x(x)
        B, m, st = x.shape[0], _lib.METHOD[method], _lib.stream_ptr
        ybuf = torch.empty((2,) + tuple(x.shape), device=x.device, dtype=torch.float32)
        kbuf = torch.empty_like(ybuf)
        ybuf[0].copy_(x)
        # FSAL slope k1 = fn(start_time, state), integrate.py:323
        table1 = e.eval_table(np.asarray([start_time], dtype=np.float32))
        _lib.check(L.bfk_velocity(C.byref(e.cfg), _lib.ptr(e._packed), _lib.ptr(table1), _lib.ptr(self.pre),
                                  _lib.ptr(self.post), self.rpg, _lib.ptr(self.tmask), _lib.ptr(ybuf[0]),
                                  _lib.ptr(kbuf[0]), B, st()), "bfk_velocity")
        ctrl = torch.zeros(K["floats"], device=x.device, dtype=torch.float32)
        _lib.check(L.bfk_adaptive_init(_lib.ptr(ctrl), float(start_time), float(stop_time), int(min_steps),
                                       int(max_steps), float(atol), float(rtol), st()), "bfk_adaptive_init")
        table = torch.empty(6, e._tstride, device=x.device, dtype=torch.float32)
        times, acc = ctrl[K["times"]:K["times"] + 6], ctrl[K["acc"]:K["acc"] + 4]
        sched = e.net._schedule_struct()
        group = getattr(e.net, "step_control_group", None)
        e.launches += 3

        def trial():
            _lib.check(L.bfk_eval_table(C.byref(e.cfg), _lib.ptr(e._packed), C.byref(sched), _lib.ptr(times), 6,
                                        _lib.ptr(table), st()), "bfk_eval_table")
            _lib.check(L.bfk_adaptive_trial(C.byref(e.cfg), _lib.ptr(e._packed), _lib.ptr(table), _lib.ptr(ctrl), m,
                                            _lib.ptr(self.pre), _lib.ptr(self.post), self.rpg, _lib.ptr(self.tmask),
                                            _lib.ptr(ybuf), _lib.ptr(kbuf), B, st()), "bfk_adaptive_trial")
            if group is not None:  # the reference's GLOBAL step control over sharded rows
                torch.distributed.all_reduce(acc, op=torch.distributed.ReduceOp.MAX, group=group)
            e.launches += 3

        n, next_check, cap = 0, int(min_steps), 8 * int(max_steps) + 64
        while True:
            _lib.check(L.bfk_adaptive_plan(_lib.ptr(ctrl), m, st()), "bfk_adaptive_plan")
            trial()
            n += 1
            if n >= next_check:
                if float(ctrl[K["done"]]) != 0.0:  # device -> host sync, once every 8 trials after min_steps
                    break
                next_check = n + 8
            if n > cap:
                raise RuntimeError("adaptive integration did not terminate")
        _lib.check(L.bfk_adaptive_plan_final(_lib.ptr(ctrl), m, st()), "bfk_adaptive_plan_final")
        trial()
        _lib.check(L.bfk_adaptive_commit(_lib.ptr(ctrl), st()), "bfk_adaptive_commit")
        c = ctrl.cpu()
        if float(c[K["all_nan"]]) != 0.0:
            raise RuntimeError(f"All values are NaNs in state during integration at {float(c[K['time']])}.")
        self.info = {"steps": int(c[K["step"]]), "rejected": int(c[K["rejected"]]), "trials": int(c[K["trials"]]),
                     "evals": 1 + 6 * int(c[K["trials"]]), "final_time": float(c[K["time"]])}
        return ybuf[int(c[K["sel"]])].clone()


class _RunLW:
    """One sampling call on the layer-wise path (bfk_sample_layerwise): same interface as _Run."""

    # rows per chunk = 16 x 4736: at W = 1024 a chunk is 592 x 4 output tiles per layer = a whole number of waves of the
    # 148 persistent CTAs, and enough tiles per CTA to hide the pipeline fill / last-epilogue tail (r01ae sweep: 439 us
    # per evaluation per 8192 rows at 8192-row chunks, 300 us at 75776).  r02 sweep (1M rows, W = 1024, after the
    # setmaxnreg change): 18,944 rows (L2-resident images) 404 k draws/s, 75,776 426 k, 151,552 446 k, 200 k .. 500 k rows
    # 451-463 k: fewer, longer launches win over L2 residency (DRAM is not the bound), so 64 x 4736
    CHUNK_ROWS = 303104

    def __init__(self, engine, cond, rpg, target_mask):
        self.e, self.cond, self.rpg = engine, cond, rpg
        self.tmask = None
        if target_mask is not None:
            tm = torch.as_tensor(target_mask, dtype=torch.float32, device=engine._packed.device)
            if tm.numel() != engine.cfg.D:
                raise NotImplementedError("target_mask must be feature-level, shape (D,), on the accelerated path")
            self.tmask = tm.reshape(-1).contiguous()
        self.info = {}

    _check_x = _Run._check_x

    def _call(self, x, times, hs, method, velocity_only):
        import os
        e, L = self.e, _lib.lib()
        dev = x.device
        t = torch.from_numpy(np.ascontiguousarray(times, dtype=np.float32)).to(dev, non_blocking=True)
        h = None if hs is None else torch.from_numpy(np.ascontiguousarray(hs, dtype=np.float32)).to(dev, non_blocking=True)
        n_steps = 1 if hs is None else int(h.numel())
        B = x.shape[0]
        chunk = int(os.environ.get("BFK_LW_CHUNK_ROWS", self.CHUNK_ROWS))
        chunk = min(chunk, (B + 127) // 128 * 128)
        G = 0 if self.cond is None else self.cond.shape[0]
        L.bfk_layerwise_workspace_bytes.restype = C.c_size_t
        L.bfk_layerwise_workspace_bytes.argtypes = [C.POINTER(_lib.NetConfig), C.c_int64, C.c_int64, C.c_int64]
        need = L.bfk_layerwise_workspace_bytes(C.byref(e.cfg), chunk, int(t.numel()), G)
        if need == 0:
            raise NotImplementedError("bfk_layerwise_workspace_bytes: " + L.bfk_last_error_string().decode())
        if e._ws is None or e._ws.numel() < need:
            e._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        out = torch.empty_like(x)
        sched = e.net._schedule_struct()
        L.bfk_sample_layerwise.argtypes = [C.POINTER(_lib.NetConfig), C.POINTER(_lib.Weights), C.POINTER(_lib.Schedule),
                                           C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int64, C.c_int64,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p,
                                           C.c_size_t, C.c_void_p]
        ev = None
        if e.kernel_events is not None:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        _lib.check(L.bfk_sample_layerwise(C.byref(e.cfg), C.byref(e.w), C.byref(sched), _lib.ptr(t), _lib.ptr(h),
                                          _lib.METHOD[method], n_steps, _lib.ptr(self.cond), G, self.rpg, _lib.ptr(self.tmask),
                                          _lib.ptr(x), _lib.ptr(out), B, int(velocity_only), chunk, _lib.ptr(e._ws),
                                          e._ws.numel(), _lib.stream_ptr()), "bfk_sample_layerwise")
        if ev is not None:
            ev[1].record()
            e.kernel_events.append(ev)
        cfg = e.cfg
        # combine + x_proj (+ merge_proj) + L fused blocks + output projector (+ finish); the small stats memsets are not
        # counted as kernels
        per_eval = 1 + (2 if cfg.C else 1) + cfg.L + 1 + (1 if (sched.model == 1 or self.tmask is not None) else 0)
        if ((cfg.D + 31) // 32 * 32) % 64:  # x_proj K (D padded to 32) not a multiple of 64: GEMM + activation kernels instead of the fused stage
            per_eval += (2 if cfg.C else 1)
        n_chunks = (B + chunk - 1) // chunk
        e.launches += cfg.L + 4 + n_chunks * (int(t.numel()) * per_eval + (0 if velocity_only else n_steps))
        self.info = {"steps": n_steps, "evals": int(t.numel()), "rejected": 0, "path": "layerwise", "chunk_rows": chunk}
        return out

    def velocity(self, x, time: float):
        return self._call(self._check_x(x), np.asarray([time], dtype=np.float32), None, "euler", True)

    def run_fixed(self, x, times, hs, method):
        return self._call(self._check_x(x), times, hs, method, False)

    def run_adaptive(self, x, start_time, stop_time, min_steps, max_steps, method, atol=1e-6, rtol=1e-4):
        """integrate_adaptive (integrate.py:291-406) on the layer-wise path: every velocity evaluation is one
        bfk_sample_layerwise(velocity_only) pass, the Runge-Kutta combinations are bfk_rk_combine, and the host runs the
        reference's global step controller in fp32 (one (err, max|state|) read per trial step)."""
        from .density import _integrate_adaptive
        x = self._check_x(x)
        evals = [0]

        def fn(time, xz):
            evals[0] += 1
            return (self.velocity(xz, float(np.float32(time))),)

        info = {}
        (out,) = _integrate_adaptive(fn, (x,), float(start_time), float(stop_time), int(min_steps), int(max_steps), method,
                                     info, atol, rtol, group=getattr(self.e.net, "step_control_group", None))
        self.info = {"steps": int(info["steps"]), "rejected": int(info["rejected"]), "evals": evals[0], "path": "layerwise"}
        return out


class _RunTE:
    """Sampling for networks with unequal hidden widths (residual projector): every velocity evaluation is one forward
    pass of the layer-wise TRAINING kernels (`bfk_train_step`, forward only, zero-padded common width), the Runge-Kutta
    combinations are `bfk_rk_combine`, and the host runs the reference's integrators (utils/integrate.py:31-406) - the
    machinery of the density path without the Jacobian trace."""

    def __init__(self, net, conditions, B, repeats, target_mask):
        from .training import train_engine
        self.net, self.eng = net, train_engine(net)
        self.cond = None
        if net._C:
            cond = conditions.detach().to(torch.float32)
            if repeats:
                cond = cond.repeat_interleave(int(repeats), dim=0)
            if cond.ndim != 2 or cond.shape != (B, net._C):
                raise ValueError(f"conditions must have shape ({B}, {net._C}), got {tuple(cond.shape)}")
            self.cond = cond.contiguous()
        self.affine = net._velocity_trace_fn(self.cond, B, {}).affine  # DiffusionModel: v = cZ x + cP pred at net-time
        self.tmask = None
        if target_mask is not None:
            tm = torch.as_tensor(target_mask, dtype=torch.float32, device=self.eng.device)
            if tm.numel() != net._D:
                raise NotImplementedError("target_mask must be feature-level, shape (D,), on the accelerated path")
            self.tmask = tm.reshape(1, -1).contiguous()
        self.info, self.evals = {}, 0

    def _check_x(self, x):
        if not x.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        if x.ndim != 2 or x.shape[-1] != self.net._D:
            raise ValueError(f"expected (rows, {self.net._D}) state, got {tuple(x.shape)}")
        return x.detach().to(torch.float32).contiguous()

    def velocity(self, x, time: float):
        x = self._check_x(x)
        cZ, cP, t_net = self.affine(float(time))
        tvec = torch.full((x.shape[0],), float(np.float32(t_net)), device=x.device, dtype=torch.float32)
        _, pred = self.eng.step(x, tvec, self.cond, None, backward=False, want_pred=True, want_loss=False)
        self.evals += 1
        if cZ != 0.0 or cP != 1.0:
            pred = float(np.float32(cZ)) * x + float(np.float32(cP)) * pred
        return pred if self.tmask is None else pred * self.tmask

    def run_fixed(self, x, times, hs, method):
        from .density import _step
        x = self._check_x(x)
        per = len(times) // max(len(hs), 1)
        state = (x,)
        for i, h in enumerate(hs):
            state, _, _, _ = _step(lambda t, xz: (self.velocity(xz, float(np.float32(t))),), state, float(times[i * per]),
                                   float(h), method)
        self.info = {"steps": len(hs), "evals": self.evals, "rejected": 0, "path": "train-engine"}
        return state[0]

    def run_adaptive(self, x, start_time, stop_time, min_steps, max_steps, method, atol=1e-6, rtol=1e-4):
        from .density import _integrate_adaptive
        x = self._check_x(x)
        info = {}
        (out,) = _integrate_adaptive(lambda t, xz: (self.velocity(xz, float(np.float32(t))),), (x,), float(start_time),
                                     float(stop_time), int(min_steps), int(max_steps), method, info, atol, rtol,
                                     group=getattr(self.net, "step_control_group", None))
        self.info = {"steps": int(info["steps"]), "rejected": int(info["rejected"]), "evals": self.evals,
                     "path": "train-engine"}
        return out


def _as_shape(batch_shape, kwargs):
    # utils/decorators.py:102-114 allow_batch_size
    if batch_shape is None and "batch_size" in kwargs:
        batch_shape = kwargs.pop("batch_size")
    if isinstance(batch_shape, (int, np.integer)):
        batch_shape = (int(batch_shape),)
    return tuple(batch_shape)


class InferenceNetwork(nn.Module):
    """inference_network.py:12-138."""

    _SUBNET_MASK_KEYS = {"attention_mask", "mask"}

    def __init__(self, base_distribution="normal", precision: str = "fp16", detect_repeats: bool = True, **kwargs):
        super().__init__()
        if isinstance(base_distribution, str):
            if base_distribution != "normal":
                raise NotImplementedError("only base_distribution='normal' is on the accelerated path")
            base_distribution = DiagonalNormal()
        self.base_distribution = base_distribution
        if precision not in _lib.PREC:
            raise ValueError(f"precision must be one of {sorted(_lib.PREC)}")
        self.precision = precision  # tensor-core operand type (fp32 accumulate); B200-specific knob
        self.detect_repeats = detect_repeats
        self.built = False
        self._metrics = []
        self._engine = _Engine(self)

    def _device_guard(self, *tensors):
        """The C ABI launches on the CURRENT device's stream: make the tensors' device current for the call, so a
        network on cuda:1 works while cuda:0 is current (one process driving several GPUs)."""
        for t in tensors:
            if torch.is_tensor(t) and t.is_cuda:
                return torch.cuda.device(t.device)
        for p in self.parameters():
            if p.is_cuda:
                return torch.cuda.device(p.device)
            break
        return contextlib.nullcontext()

    def call(self, xz, conditions=None, inverse=False, density=False, training=False, **kwargs):
        with self._device_guard(xz):
            if inverse:
                return self._inverse(xz, conditions=conditions, density=density, training=training, **kwargs)
            return self._forward(xz, conditions=conditions, density=density, training=training, **kwargs)

    forward = call

    def _forward(self, x, conditions=None, density=False, training=False, **kwargs):
        raise NotImplementedError

    def _inverse(self, z, conditions=None, density=False, training=False, **kwargs):
        raise NotImplementedError

    def sample(self, batch_shape=None, conditions=None, seed=None, **kwargs):
        # inference_network.py:109-125
        batch_shape = _as_shape(batch_shape, kwargs)
        if "compute_prior_score" in kwargs:
            raise NotImplementedError("compositional sampling is out of scope (SURVEY.md §2 row 2)")
        row_offset = kwargs.pop("row_offset", 0)
        if getattr(self, "_wants_noise_offset", False):  # samplers that draw more noise after z (ConsistencyModel)
            kwargs["noise_row_offset"] = row_offset
        samples = self.base_distribution.sample(batch_shape, seed=seed, row_offset=row_offset)
        flat = samples.reshape(-1, samples.shape[-1])
        out = self(flat, conditions=conditions, inverse=True, density=False, seed=seed, **kwargs)
        return out.reshape(samples.shape)

    def log_prob(self, samples, conditions=None, **kwargs):
        _, log_density = self(samples, conditions=conditions, inverse=False, density=True, **kwargs)
        return log_density

    def _loss_and_grads(self, x, conditions=None, sample_weight=None, stage="training", seed=None, world=1, backward=True,
                        **kwargs):
        raise NotImplementedError

    def compute_metrics(self, x, conditions=None, sample_weight=None, stage="training", **kwargs):
        """{"loss": 0-d tensor}.  With grad mode on and stage == "training" the kernels also produce every parameter
        gradient; ``loss.backward()`` then fills ``p.grad`` exactly like the reference's autograd path."""
        from .training import _LossWithGrads, train_engine
        backward = torch.is_grad_enabled() and stage == "training"
        extra = {k: kwargs[k] for k in ("fixed", "row_offset", "ot_seed", "noise_seed", "cond_keep", "target_keep")
                 if k in kwargs}
        with self._device_guard(x[1] if isinstance(x, (tuple, list)) else x):
            loss = self._loss_and_grads(x, conditions, sample_weight, stage=stage, seed=kwargs.get("seed"),
                                        backward=backward, **extra)
        if backward:
            eng = train_engine(self)
            loss = _LossWithGrads.apply(loss, eng, *eng.params)
        return {"loss": loss}

    def _graphable_training(self):
        return False

    @staticmethod
    def _keep_mask(rows, cols, drop_prob, seed, row_offset, device):
        """0 / 1 keep mask with P(keep) = 1 - drop_prob (utils/tensor_utils.py:445-470 `random_mask`: uniform > drop_prob),
        drawn as z > Phi^-1(drop_prob) from the Philox normal stream keyed by (seed, GLOBAL row): data-parallel ranks
        mask disjoint rows of one batch."""
        from statistics import NormalDist
        if drop_prob >= 1.0:
            return torch.zeros(rows, cols, device=device, dtype=torch.float32)
        z = torch.empty(rows, cols, device=device, dtype=torch.float32)
        with torch.cuda.device(device):
            _lib.check(_lib.lib().bfk_normal_fill(_lib.ptr(z), rows, cols, int(row_offset), None, None,
                                                  C.c_uint64(int(seed) & (2 ** 64 - 1)), None, _lib.stream_ptr()),
                       "bfk_normal_fill")
        return (z > NormalDist().inv_cdf(float(drop_prob))).to(torch.float32)

    def _dropout_masks(self, B, D, conditions, noise_seed, row_offset, kwargs, device):
        """-> (conditions with dropped rows zeroed, (B,D) target keep mask or None).  `cond_keep` (B,1) / `target_keep`
        (B,D) in kwargs replace the draws (parity runs)."""
        keep_t = kwargs.get("target_keep")
        if self.drop_cond_prob > 0 and conditions is not None:
            keep_c = kwargs.get("cond_keep")
            if keep_c is None:
                keep_c = self._keep_mask(B, 1, self.drop_cond_prob, int(noise_seed) ^ 0x6A09E667F3BCC908, row_offset, device)
            conditions = conditions.to(torch.float32) * torch.as_tensor(keep_c, device=device, dtype=torch.float32).reshape(B, 1)
        if self.drop_target_prob > 0 and keep_t is None:
            keep_t = self._keep_mask(B, D, self.drop_target_prob, int(noise_seed) ^ 0x3C6EF372FE94F82B, row_offset, device)
        if keep_t is not None:
            keep_t = torch.as_tensor(keep_t, device=device, dtype=torch.float32).reshape(B, D).contiguous()
        return conditions, keep_t

    def _dropout_seed(self, seed):
        if seed is None:
            self._drop_calls = getattr(self, "_drop_calls", 0) + 1
            seed = (torch.initial_seed() * 0x9E3779B1 + self._drop_calls) & (2 ** 63 - 1)
        return seed

    @property
    def variables(self):
        return list(self.parameters())

    # -- shared build of subnet + projector (flow_matching.py:180-199, diffusion_model.py build) -----------------------
    def _build_common(self, xz_shape, conditions_shape):
        self._D = int(xz_shape[-1])
        self._C = int(conditions_shape[-1]) if conditions_shape is not None else None
        self.base_distribution.build(xz_shape)
        time_shape = (xz_shape[0], 1)
        self.subnet.build((xz_shape, time_shape, conditions_shape))
        self.output_projector = _Dense(self.subnet.widths[-1], self._D, init="glorot_uniform")

    def _run(self, conditions, B, kwargs) -> _Run:
        if not self.built:
            raise RuntimeError("network is not built; call build(xz_shape, conditions_shape) first")
        if (conditions is None) != (self._C is None):
            raise ValueError("network was built " + ("without" if self._C is None else "with") + " conditions")
        self._engine.prepare()
        if self.unconditional_mode and conditions is not None:
            conditions = torch.zeros_like(conditions)
        if self._engine.train_engine_sampler:
            return _RunTE(self, conditions, B, kwargs.get("conditions_repeats"), kwargs.get("target_mask"))
        if self._engine.layerwise:
            cond, rpg = self._engine.group_conditions(conditions, B, kwargs.get("conditions_repeats"))
            return _RunLW(self._engine, cond, rpg, kwargs.get("target_mask"))
        pre, post, rpg = self._engine.cond_tables(conditions, B, kwargs.get("conditions_repeats"))
        return _Run(self._engine, pre, post, rpg, kwargs.get("target_mask"))

    @staticmethod
    def _apply_targets(x, kwargs):
        target_mask, targets_fixed = kwargs.get("target_mask"), kwargs.get("targets_fixed")
        if target_mask is not None:
            tm = torch.broadcast_to(torch.as_tensor(target_mask, dtype=x.dtype, device=x.device), x.shape)
            tf = torch.broadcast_to(torch.as_tensor(targets_fixed, dtype=x.dtype, device=x.device), x.shape)
            x = tm * x + (1 - tm) * tf
        return x

    _INTEGRATE_KEYS = ("start_time", "stop_time", "min_steps", "max_steps", "steps", "method", "atol", "rtol")

    def _velocity_trace_fn(self, conditions, B, kwargs):
        from .density import VelocityTrace
        return VelocityTrace(self, conditions, B, max_steps=kwargs.get("max_steps"), target_mask=kwargs.get("target_mask"),
                             seed=kwargs.get("seed"))

    def _integrate_density(self, x, conditions, integrate_kwargs, kwargs):
        """integrate({"xz", "trace"}) with the Jacobian trace of the velocity (flow_matching.py:272-284,
        diffusion_model.py:755-771) -> (xz_end (B,D), trace_end (B,))."""
        from .density import integrate_density
        if not self.built:
            raise RuntimeError("network is not built; call build(xz_shape, conditions_shape) first")
        if (conditions is None) != (self._C is None):
            raise ValueError("network was built " + ("without" if self._C is None else "with") + " conditions")
        if not x.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        x = x.detach().to(torch.float32).contiguous()
        if x.ndim != 2 or x.shape[-1] != self._D:
            raise ValueError(f"expected (rows, {self._D}) inputs, got {tuple(x.shape)}")
        fn = self._velocity_trace_fn(conditions, x.shape[0], kwargs)
        ik = {k: v for k, v in integrate_kwargs.items() if k in self._INTEGRATE_KEYS}
        info = {}
        state = (x, torch.zeros(x.shape[0], 1, device=x.device, dtype=torch.float32))
        # rows sharded over ranks (parallel.log_prob_sharded): the adaptive controller's maxima are all-reduced
        xz, trace = integrate_density(fn, state, info=info, group=getattr(self, "step_control_group", None), **ik)
        info["evals"] = fn.evals
        self.last_integration_info = info
        return xz, trace.squeeze(-1)

    def _integrate_call(self, run, x, integrate_kwargs):
        ik = {k: v for k, v in integrate_kwargs.items() if k in self._INTEGRATE_KEYS}  # filter_kwargs, dict_utils.py:52-63
        out = _integrate(run, x, **ik)
        self.last_integration_info = run.info
        return out


class FlowMatching(InferenceNetwork):
    """flow_matching.py:27-337.  Extra keywords (`precision`, `detect_repeats`) are B200-specific knobs."""

    def __init__(self, subnet="time_mlp", base_distribution="normal", use_optimal_transport: bool = False, loss_fn="mse",
                 integrate_kwargs=None, optimal_transport_kwargs=None, subnet_kwargs=None,
                 time_power_law_alpha: float = 0.0, drop_cond_prob: float = 0.0, drop_target_prob: float = 0.0, **kwargs):
        super().__init__(base_distribution, **kwargs)
        self.use_optimal_transport = use_optimal_transport
        self.integrate_kwargs = FLOW_MATCHING_INTEGRATE_DEFAULTS | (integrate_kwargs or {})
        from .optimal_transport import OPTIMAL_TRANSPORT_DEFAULTS
        self.optimal_transport_kwargs = OPTIMAL_TRANSPORT_DEFAULTS | (optimal_transport_kwargs or {})  # flow_matching.py:113
        if loss_fn not in ("mse", "mae"):
            raise NotImplementedError("loss_fn must be 'mse' or 'mae' on the accelerated path")
        self.loss_fn = loss_fn
        self.time_power_law_alpha = float(time_power_law_alpha)
        if self.time_power_law_alpha <= -1.0:
            raise ValueError("'time_power_law_alpha' must be greater than -1.0.")
        subnet_kwargs = subnet_kwargs or {}
        if subnet == "time_mlp":
            subnet_kwargs = TIME_MLP_DEFAULTS | subnet_kwargs
        self.subnet = find_network(subnet, **subnet_kwargs)
        self.output_projector = None
        self.drop_cond_prob = drop_cond_prob
        self.unconditional_mode = False
        self.drop_target_prob = drop_target_prob

    def build(self, xz_shape, conditions_shape=None):
        if self.built:
            return
        self._build_common(xz_shape, conditions_shape)
        self.built = True

    def _schedule_struct(self):
        s = _lib.Schedule()
        s.model = 0
        return s

    def get_config(self):
        return {"subnet": self.subnet.get_config(), "base_distribution": "normal",
                "use_optimal_transport": self.use_optimal_transport, "loss_fn": self.loss_fn,
                "integrate_kwargs": self.integrate_kwargs, "optimal_transport_kwargs": self.optimal_transport_kwargs,
                "time_power_law_alpha": self.time_power_law_alpha, "drop_cond_prob": self.drop_cond_prob,
                "drop_target_prob": self.drop_target_prob}

    def velocity(self, xz, time, conditions=None, training=False, **kwargs):
        # flow_matching.py:220-236
        if training:
            raise NotImplementedError("training-mode velocity goes through compute_metrics")
        if torch.is_tensor(time):
            if time.numel() != 1 and not bool((time == time.reshape(-1)[0]).all()):
                raise NotImplementedError("per-row time is only used by the training path (compute_metrics)")
            time = float(time.reshape(-1)[0])
        with self._device_guard(xz):
            run = self._run(conditions, xz.shape[0], kwargs)
            return run.velocity(xz, float(np.float32(time)))

    # -- training (flow_matching.py:132-178) ------------------------------------------------------------------------------
    def _graphable_training(self):
        """the training step has no host-side decisions: Trainer(cuda_graph=True) may capture it"""
        bd = self.base_distribution
        standard = (isinstance(bd, DiagonalNormal) and not torch.is_tensor(bd.mean) and not torch.is_tensor(bd.std)
                    and float(bd.mean) == 0.0 and float(bd.std) == 1.0)
        ragged = any(w != self.subnet.widths[0] for w in self.subnet.widths)  # padded staging launches around the step
        return standard and not ragged and not self.use_optimal_transport and self.drop_cond_prob == 0 and self.drop_target_prob == 0

    def _loss_and_grads(self, x, conditions=None, sample_weight=None, stage="training", seed=None, world=1, backward=True,
                        row_offset=0, noise_seed=None, seed_dev=None, loss_into_slot=False, **kwargs):
        """`row_offset`: global index of this shard's first row; `noise_seed`: keys x0 / t of the step (Philox streams that
        are functions of the GLOBAL row, so data-parallel ranks draw disjoint rows of ONE batch); `seed_dev`: two device
        words that override the seeds (CUDA-graph replay)."""
        from .training import train_engine
        if (conditions is None) != (self._C is None):
            raise ValueError("network was built " + ("without" if self._C is None else "with") + " conditions")
        eng = train_engine(self)
        if isinstance(x, Sequence) and not torch.is_tensor(x):
            x0, x1, t, x, target = x  # bypasses all RNG (flow_matching.py:140-141)
        else:
            x1 = x
            if not x1.is_cuda:
                raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
            if noise_seed is None:
                noise_seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            if self._graphable_training():
                # flow_matching.py:143-163 in one launch: x0, t = u^(1+alpha), x and the target velocity
                x1 = eng._f32(x1)
                x, target = torch.empty_like(x1), torch.empty_like(x1)
                t = torch.empty(x1.shape[0], device=x1.device, dtype=torch.float32)
                with torch.cuda.device(x1.device):
                    _lib.check(eng.L.bfk_fm_train_batch(_lib.ptr(x1), x1.shape[0], x1.shape[1], int(row_offset),
                                                        C.c_uint64(int(noise_seed) & (2 ** 64 - 1)), _lib.ptr(seed_dev),
                                                        float(self.time_power_law_alpha), None, _lib.ptr(t), _lib.ptr(x),
                                                        _lib.ptr(target), _lib.stream_ptr()), "bfk_fm_train_batch")
                eng.launches += 1
            else:
                x0 = self.base_distribution.sample(tuple(x1.shape[:-1]), seed=noise_seed, row_offset=row_offset)
                if self.use_optimal_transport:
                    # flow_matching.py:146-157: x1 (and its conditions) are re-drawn along the transport plan, x0 is kept
                    from .optimal_transport import optimal_transport
                    x0, x1, conditions = optimal_transport(x0, x1, conditions=conditions, seed=kwargs.get("ot_seed"),
                                                           **self.optimal_transport_kwargs)
                # u by GLOBAL row: Phi(z) of a second Philox normal stream
                zu = DiagonalNormal().to(x1.device)
                zu.build((None, 1))
                zu = zu.to(x1.device).sample((x0.shape[0],), seed=(int(noise_seed) ^ 0x2545F4914F6CDD1D) & (2 ** 63 - 1),
                                             row_offset=row_offset).reshape(-1)
                u = 0.5 * (1.0 + torch.erf(zu * (1.0 / math.sqrt(2.0))))
                t = u ** (1 + self.time_power_law_alpha)
                x, target = eng.interpolate(x0, x1, t)
        keep_t = None
        if self.drop_cond_prob > 0 or self.drop_target_prob > 0 or "target_keep" in kwargs:
            # flow_matching.py:167-175: conditions of dropped rows zeroed (classifier-free guidance); dropped target entries
            # see the clean x1 and leave the loss (the loss kernel multiplies the residual by the mask)
            if noise_seed is None:
                noise_seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            x = eng._f32(x)
            conditions, keep_t = self._dropout_masks(x.shape[0], x.shape[1], conditions, noise_seed, row_offset, kwargs,
                                                     x.device)
            if keep_t is not None:
                x = keep_t * x + (1.0 - keep_t) * eng._f32(x1)
        training = stage == "training"
        rate = float(self.subnet.dropout or 0.0) if training else 0.0
        seed = self._dropout_seed(seed)
        loss, _ = eng.step(x, t, conditions, target, row_weight=sample_weight, loss=self.loss_fn, dropout=rate, seed=seed,
                           global_batch=x.shape[0] * world, backward=backward, row_offset=row_offset, seed_dev=seed_dev,
                           loss_into_slot=loss_into_slot, elem_mask=keep_t)
        return loss

    def _forward(self, x, conditions=None, density=False, training=False, **kwargs):
        # flow_matching.py:254-294
        integrate_kwargs = self.integrate_kwargs | kwargs
        x = self._apply_targets(x, kwargs)
        if density:
            # flow_matching.py:272-284
            z, trace = self._integrate_density(x, conditions, integrate_kwargs | {"start_time": 1.0, "stop_time": 0.0}, kwargs)
            return z, self.base_distribution.log_prob(z) + trace
        run = self._run(conditions, x.shape[0], kwargs)
        return self._integrate_call(run, x, integrate_kwargs | {"start_time": 1.0, "stop_time": 0.0})

    def _inverse(self, z, conditions=None, density=False, training=False, **kwargs):
        # flow_matching.py:296-337
        kwargs.pop("seed", None)
        integrate_kwargs = self.integrate_kwargs | kwargs
        z = self._apply_targets(z, kwargs)
        if density:
            # flow_matching.py:316-326
            x, trace = self._integrate_density(z, conditions, integrate_kwargs | {"start_time": 0.0, "stop_time": 1.0}, kwargs)
            return x, self.base_distribution.log_prob(z) - trace
        run = self._run(conditions, z.shape[0], kwargs)
        return self._integrate_call(run, z, integrate_kwargs | {"start_time": 0.0, "stop_time": 1.0})


class EDMNoiseSchedule:
    """schedules/edm_noise_schedule.py:21-59 (parameters only; the scalar algebra runs in bfk_eval_table)."""
    name = "edm_noise_schedule"

    def __init__(self, sigma_data=1.0, sigma_min=1e-4, sigma_max=80.0, variance_type="preserving"):
        self.sigma_data, self.sigma_min, self.sigma_max, self.rho = sigma_data, sigma_min, sigma_max, 7
        self._variance_type = variance_type
        self.log_snr_min, self.log_snr_max = -2 * math.log(sigma_max), -2 * math.log(sigma_min)

    def validate(self):
        if self.log_snr_min >= self.log_snr_max:
            raise ValueError("min_log_snr must be less than max_log_snr.")

    # Training-time schedule algebra on (B,) device vectors (float64, then cast): O(B) plumbing next to the O(B*W*W)
    # kernels.  edm_noise_schedule.py:61-113, noise_schedule.py:100-119.
    def get_log_snr(self, t, training):
        if training:
            loc, scale = 2 * 1.2, 2 * 1.2  # -2*p_mean, 2*p_std
            snr = loc + scale * torch.erfinv(2 * (1 - t) - 1) * math.sqrt(2)
            return torch.clip(snr, self.log_snr_min - 1, self.log_snr_max + 1)
        smin, smax = self.sigma_min ** (1 / self.rho), self.sigma_max ** (1 / self.rho)
        return -2 * self.rho * torch.log(smax + (1 - t) * (smin - smax))

    def get_alpha_sigma(self, log_snr_t):
        if self._variance_type == "preserving":
            return torch.sqrt(torch.sigmoid(log_snr_t)), torch.sqrt(torch.sigmoid(-log_snr_t))
        return torch.ones_like(log_snr_t), torch.sqrt(torch.exp(-log_snr_t))

    def get_weights_for_snr(self, log_snr_t):
        return 1 + torch.exp(-log_snr_t) / self.sigma_data ** 2


class CosineNoiseSchedule:
    """schedules/cosine_noise_schedule.py:26-55."""
    name = "cosine_noise_schedule"

    def __init__(self, min_log_snr=-15.0, max_log_snr=15.0, shift=0.0, weighting="sigmoid"):
        self._variance_type, self._shift, self._weighting = "preserving", shift, weighting
        self.log_snr_min, self.log_snr_max = float(min_log_snr), float(max_log_snr)
        self._t_min = 2 / math.pi * math.atan(math.exp((2 * shift - self.log_snr_max) * 0.5))
        self._t_max = 2 / math.pi * math.atan(math.exp((2 * shift - self.log_snr_min) * 0.5))

    def validate(self):
        if self.log_snr_min >= self.log_snr_max:
            raise ValueError("min_log_snr must be less than max_log_snr.")

    # cosine_noise_schedule.py:57-82
    def get_log_snr(self, t, training):
        tt = self._t_min + (self._t_max - self._t_min) * t
        return -2 * torch.log(torch.tan(math.pi * tt * 0.5)) + 2 * self._shift

    def get_alpha_sigma(self, log_snr_t):
        return torch.sqrt(torch.sigmoid(log_snr_t)), torch.sqrt(torch.sigmoid(-log_snr_t))

    def get_weights_for_snr(self, log_snr_t):
        if self._weighting == "sigmoid":
            return torch.sigmoid(-log_snr_t + 2)
        if self._weighting == "likelihood_weighting":
            t = 2 / math.pi * torch.arctan(torch.exp((2 * self._shift - log_snr_t) * 0.5))
            tt = self._t_min + (self._t_max - self._t_min) * t
            dsnr_dt = -(2 * math.pi) / torch.sin(math.pi * tt) * (self._t_max - self._t_min)
            beta = -(torch.exp(-log_snr_t) / (1 + torch.exp(-log_snr_t))) * dsnr_dt
            return beta / torch.sigmoid(-log_snr_t)
        return torch.ones_like(log_snr_t)


def find_noise_schedule(arg, **kwargs):
    # networks/inference/diffusion/dispatch.py:6-35
    if isinstance(arg, str):
        if arg == "edm":
            return EDMNoiseSchedule(**kwargs)
        if arg == "cosine":
            return CosineNoiseSchedule(**kwargs)
        raise ValueError(f"Unsupported noise schedule name: '{arg}'.")
    if isinstance(arg, type):
        return arg(**kwargs)
    if isinstance(arg, (EDMNoiseSchedule, CosineNoiseSchedule)):
        return arg
    raise TypeError(f"Unknown noise schedule type: {type(arg)}.")


class DiffusionModel(InferenceNetwork):
    """diffusion_model.py:35-869, probability-flow ODE sampling path (deterministic methods)."""

    _wants_noise_offset = True  # the reverse-SDE increments are keyed by the GLOBAL row (stochastic.py)

    def __init__(self, *, subnet="time_mlp", noise_schedule="edm", prediction_type="F", loss_type="noise",
                 subnet_kwargs=None, schedule_kwargs=None, integrate_kwargs=None, drop_cond_prob: float = 0.0,
                 drop_target_prob: float = 0.0, **kwargs):
        super().__init__(base_distribution="normal", **kwargs)
        if prediction_type not in ["noise", "velocity", "F", "x", "score", "potential"]:
            raise ValueError(f"Unknown prediction type: {prediction_type}")
        if loss_type not in ["noise", "velocity", "F"]:
            raise ValueError(f"Unknown loss type: {loss_type}")
        if prediction_type == "potential":
            raise NotImplementedError("prediction_type='potential' is out of scope (SURVEY.md §2 row 2)")
        self._prediction_type, self._loss_type = prediction_type, loss_type
        self.noise_schedule = find_noise_schedule(noise_schedule, **(schedule_kwargs or {}))
        self.noise_schedule.validate()
        self.integrate_kwargs = DIFFUSION_INTEGRATE_DEFAULTS | (integrate_kwargs or {})
        subnet_kwargs = subnet_kwargs or {}
        if subnet == "time_mlp":
            subnet_kwargs = TIME_MLP_DEFAULTS | subnet_kwargs
        self.subnet = find_network(subnet, **subnet_kwargs)
        self.output_projector = None
        self.drop_cond_prob, self.drop_target_prob = drop_cond_prob, drop_target_prob
        self.unconditional_mode = False

    def build(self, xz_shape, conditions_shape=None):
        if self.built:
            return
        self._build_common(xz_shape, conditions_shape)
        self.built = True

    def _schedule_struct(self):
        ns, s = self.noise_schedule, _lib.Schedule()
        s.model, s.prediction = 1, _lib.PRED[self._prediction_type]
        s.exploding = int(ns._variance_type == "exploding")
        s.log_snr_min, s.log_snr_max = ns.log_snr_min, ns.log_snr_max
        if isinstance(ns, EDMNoiseSchedule):
            s.schedule = 0
            s.sigma_data, s.sigma_min, s.sigma_max, s.rho = ns.sigma_data, ns.sigma_min, ns.sigma_max, ns.rho
        else:
            s.schedule, s.sigma_data = 1, 1.0
            s.shift, s.t_min, s.t_max = ns._shift, ns._t_min, ns._t_max
        return s

    # -- training (diffusion_model.py:142-232) ---------------------------------------------------------------------------
    def _x_pred_coefs(self, alpha, sigma, log_snr):
        """convert_prediction_to_x (diffusion_model.py:351-405) is affine in (pred, z): x_pred = cP*pred + cZ*z."""
        pt = self._prediction_type
        if pt == "velocity":
            return -sigma, alpha
        if pt == "noise":
            return -sigma / alpha, 1 / alpha
        if pt == "F":
            sd = getattr(self.noise_schedule, "sigma_data", 1.0)
            e = torch.exp(-log_snr)
            return torch.exp(-log_snr / 2) * sd / torch.sqrt(e + sd ** 2), (sd ** 2 * alpha) / (e + sd ** 2)
        if pt == "x":
            return torch.ones_like(alpha), torch.zeros_like(alpha)
        return sigma ** 2 / alpha, 1 / alpha  # score

    def _loss_and_grads(self, x, conditions=None, sample_weight=None, stage="training", seed=None, world=1, backward=True,
                        fixed=None, row_offset=0, noise_seed=None, seed_dev=None, loss_into_slot=False, **kwargs):
        """One DM training step through bfk_train_step.  Every loss_type is `w * mean_d (a*pred - target)^2` with a
        per-row scalar `a` and a per-element target assembled here from the schedule; the network forward/backward,
        the loss reduction and all parameter gradients are the C-ABI kernels.  `fixed=(u0, eps_t)` bypasses the RNG
        (parity runs); `row_offset` is this rank's first GLOBAL row (low-discrepancy t uses the global index,
        diffusion_model.py:159-161)."""
        from .training import train_engine
        if (conditions is None) != (self._C is None):
            raise ValueError("network was built " + ("without" if self._C is None else "with") + " conditions")
        if not torch.is_tensor(x) or not x.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        eng = train_engine(self)
        ns, dev, f64 = self.noise_schedule, x.device, torch.float64
        B = x.shape[0]
        Bg = B * world
        if fixed is None:
            if noise_seed is None:
                noise_seed = int(torch.randint(0, 2 ** 62, (1,)).item())
                if world > 1:  # every rank must key the same global streams
                    sd = torch.tensor([noise_seed], device=dev, dtype=torch.int64)
                    torch.distributed.broadcast(sd, src=0)
                    noise_seed = int(sd.item())
            # u0: one uniform for the whole (global) batch; eps: Philox normals of the GLOBAL rows (bfk_normal_fill)
            u0 = torch.rand((1,), device="cpu", dtype=torch.float32,
                            generator=torch.Generator().manual_seed(int(noise_seed) & (2 ** 62 - 1))).to(dev)
            eps = self.base_distribution.sample((B,), seed=(int(noise_seed) ^ 0x2545F4914F6CDD1D) & (2 ** 63 - 1),
                                                row_offset=row_offset)
        else:
            u0, eps = (torch.as_tensor(v, device=dev, dtype=torch.float32) for v in fixed)
        sched_training = stage in ("training", "validation")
        # fp32 t exactly as the reference forms it, then the schedule in double
        tg = (u0 + torch.arange(0, Bg, device=dev, dtype=torch.float32) / float(Bg)) % 1
        lam_g = ns.get_log_snr(tg.to(f64), training=sched_training)
        w_mean = ns.get_weights_for_snr(lam_g).mean()  # the reference's (B,1)*(B,) broadcast: mean(w) * mean(per-row)
        lam = lam_g[row_offset:row_offset + B].unsqueeze(-1)
        alpha, sigma = ns.get_alpha_sigma(lam)
        x64, eps64 = x.to(f64), eps.to(f64)
        z = (alpha * x64 + sigma * eps64).to(torch.float32)
        keep_t = None
        if self.drop_cond_prob > 0 or self.drop_target_prob > 0 or "target_keep" in kwargs:
            # diffusion_model.py:155-156, 177-178: dropped rows lose their conditions; dropped target entries see the clean x
            # and leave the loss (the loss kernel multiplies the residual by the mask, :209-224)
            if noise_seed is None:
                noise_seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            conditions, keep_t = self._dropout_masks(B, x.shape[1], conditions, noise_seed, row_offset, kwargs, dev)
            if keep_t is not None:
                z = keep_t * z + (1.0 - keep_t) * x.to(torch.float32)
        z64 = z.to(f64)
        cP, cZ = self._x_pred_coefs(alpha, sigma, lam)
        if self._loss_type == "noise":
            a, target = -alpha * cP / sigma, eps64 - ((1 - alpha * cZ) / sigma) * z64
        elif self._loss_type == "velocity":
            a, target = -cP / sigma, (alpha * eps64 - sigma * x64) - ((alpha - cZ) / sigma) * z64
        else:
            sd = getattr(ns, "sigma_data", 1.0)
            x1 = torch.sqrt(torch.exp(-lam) + sd ** 2) / (torch.exp(-lam / 2) * sd)
            a, target = x1 * cP, x1 * (x64 - cZ * z64)
        t_net = 2 * ((lam - ns.log_snr_min) / (ns.log_snr_max - ns.log_snr_min)) - 1  # _transform_log_snr, :711-717
        row_w = w_mean.to(torch.float32).expand(B)
        if sample_weight is not None:
            row_w = row_w * sample_weight.reshape(-1).to(torch.float32)
        rate = float(self.subnet.dropout or 0.0) if stage == "training" else 0.0
        # The cotangent 2*w*a*resid enters the backward as an fp16 image; some (prediction, loss) pairs have |a| ~ 600 at
        # the noisy end, so it is pre-scaled by a power of two (bounded on a fixed time grid, no device sync) and every
        # gradient un-scaled by grad_scale - exact in floating point.
        row_w, cot_scale = row_w * self._cotangent_scale(stage), self._cotangent_scale(stage)
        loss, _ = eng.step(z, t_net.reshape(-1).to(torch.float32), conditions, target.to(torch.float32),
                           pred_scale=a.reshape(-1).to(torch.float32), row_weight=row_w.contiguous(), loss="mse",
                           dropout=rate, seed=self._dropout_seed(seed), global_batch=Bg, backward=backward,
                           cotangent_scale=cot_scale, row_offset=row_offset, loss_into_slot=loss_into_slot, elem_mask=keep_t)
        return loss

    def _cotangent_scale(self, stage):
        key = (stage in ("training", "validation"), self._loss_type, self._prediction_type)
        cache = self.__dict__.setdefault("_cot_scale_cache", {})
        if key not in cache:
            ns = self.noise_schedule
            t = (torch.arange(0, 4096, dtype=torch.float64) + 0.5) / 4096
            lam = ns.get_log_snr(t, training=key[0])
            alpha, sigma = ns.get_alpha_sigma(lam)
            cP, _ = self._x_pred_coefs(alpha, sigma, lam)
            if self._loss_type == "noise":
                a = alpha * cP / sigma
            elif self._loss_type == "velocity":
                a = cP / sigma
            else:
                sd = getattr(ns, "sigma_data", 1.0)
                a = torch.sqrt(torch.exp(-lam) + sd ** 2) / (torch.exp(-lam / 2) * sd) * cP
            bound = float(ns.get_weights_for_snr(lam).mean() * a.abs().max())
            cache[key] = 2.0 ** -max(0, math.ceil(math.log2(max(bound, 1e-30))))
        return cache[key]

    def velocity(self, xz, time, stochastic_solver=False, conditions=None, training=False, **kwargs):
        # diffusion_model.py:501-559
        if stochastic_solver or training:
            raise NotImplementedError("only the probability-flow ODE velocity is on the accelerated path")
        if torch.is_tensor(time):
            time = float(time.reshape(-1)[0])
        run = self._run(conditions, xz.shape[0], kwargs)
        return run.velocity(xz, float(np.float32(time)))

    def _velocity_trace_fn(self, conditions, B, kwargs):
        fn = super()._velocity_trace_fn(conditions, B, kwargs)
        L, sched = _lib.lib(), self._schedule_struct()
        L.bfk_schedule_coefs.argtypes = [C.POINTER(_lib.Schedule), C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                         C.POINTER(C.c_double)]

        def affine(time):
            cz, cp, nt = C.c_double(), C.c_double(), C.c_double()
            _lib.check(L.bfk_schedule_coefs(C.byref(sched), float(np.float32(time)), C.byref(cz), C.byref(cp), C.byref(nt)),
                       "bfk_schedule_coefs")
            return cz.value, cp.value, nt.value
        fn.affine = affine
        return fn

    def _forward(self, x, conditions=None, density=False, training=False, **kwargs):
        # diffusion_model.py:719-780
        integrate_kwargs = {"start_time": 0.0, "stop_time": 1.0}
        integrate_kwargs |= self.integrate_kwargs
        integrate_kwargs |= kwargs
        if integrate_kwargs["method"] in STOCHASTIC_METHODS:
            # "Stochastic methods are not supported for density evaluation. Falling back to tsit5 ODE solver." (:733-739)
            integrate_kwargs["method"] = "tsit5"
        x = self._apply_targets(x, kwargs)
        if density:
            z, trace = self._integrate_density(x, conditions, integrate_kwargs, kwargs)
            return z, self.base_distribution.log_prob(z) + trace
        run = self._run(conditions, x.shape[0], kwargs)
        return self._integrate_call(run, x, integrate_kwargs)

    def _inverse(self, z, conditions=None, density=False, training=False, **kwargs):
        # diffusion_model.py:782-869
        seed = kwargs.pop("seed", None)
        noise_row_offset = int(kwargs.pop("noise_row_offset", 0))  # this shard's first GLOBAL row (SDE increments)
        integrate_kwargs = {"start_time": 1.0, "stop_time": 0.0}
        integrate_kwargs |= self.integrate_kwargs
        integrate_kwargs |= kwargs
        z = self._apply_targets(z, kwargs)
        if density:
            # diffusion_model.py:804-826 (stochastic methods fall back to tsit5 for the density)
            if integrate_kwargs["method"] in STOCHASTIC_METHODS:
                integrate_kwargs["method"] = "tsit5"
            x, trace = self._integrate_density(z, conditions, integrate_kwargs, kwargs)
            return x, self.base_distribution.log_prob(z) - trace
        if integrate_kwargs["method"] in STOCHASTIC_METHODS:
            # diffusion_model.py:830-856: reverse SDE (the network default is two_step_adaptive, defaults.py:50)
            from .stochastic import ReverseSDE, integrate_stochastic
            if not z.is_cuda:
                raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
            run = self._run(conditions, z.shape[0], kwargs)
            sde = ReverseSDE(self, run, kwargs.get("target_mask"))
            if kwargs.get("target_mask") is not None and (integrate_kwargs.get("corrector_steps") or
                                                           integrate_kwargs["method"] == "langevin"):
                # the reference's score_fn ignores the target mask (diffusion_model.py:845-849): the Langevin updates use
                # the unmasked velocity
                sde.score_run = self._run(conditions, z.shape[0], {k: v for k, v in kwargs.items() if k != "target_mask"})
            keys = ("start_time", "stop_time", "steps", "method", "min_steps", "max_steps", "corrector_steps", "e_rel",
                    "e_abs", "r", "adapt_safety", "noise_fn", "step_size_factor")
            info = {}
            out = integrate_stochastic(sde, run._check_x(z), seed=seed, info=info, row_offset=noise_row_offset,
                                       **{k: v for k, v in integrate_kwargs.items() if k in keys})
            self.last_integration_info = info
            return out
        run = self._run(conditions, z.shape[0], kwargs)
        return self._integrate_call(run, z, integrate_kwargs)
