"""Running-moment standardization on the edges of the sampling / training path.

`Standardize` mirrors bayesflow/networks/helpers/standardization/standardize.py:11-170 and `Standardization`
bayesflow/networks/helpers/standardization/standardization.py:11-88 (same names, arguments and error behaviour).  The
location-scale transform and the moment update run in libbfk (`bfk_standardize`, `bfk_batch_moments`,
`bfk_moments_merge`, include/bfk.h); there is no CPU path — tensors must live on a CUDA device."""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist
from torch import nn

from . import _lib


def _flatten(tree):
    """keras.tree.flatten order: dict values by sorted key, sequences in order, leaves as they are."""
    if isinstance(tree, dict):
        out = []
        for k in sorted(tree):
            out += _flatten(tree[k])
        return out
    if isinstance(tree, (list, tuple)) and not (len(tree) > 0 and all(isinstance(v, int) or v is None for v in tree)):
        out = []
        for v in tree:
            out += _flatten(v)
        return out
    return [tree]


def _pack_like(tree, flat):
    it = iter(flat)

    def rec(t):
        if isinstance(t, dict):
            vals = {k: rec(t[k]) for k in sorted(t)}
            return {k: vals[k] for k in t}
        if isinstance(t, (list, tuple)):
            return type(t)(rec(v) for v in t)
        return next(it)

    return rec(tree)


class Standardize(nn.Module):
    """standardize.py:11-170.  Buffers per flattened input: moving_mean (F), moving_m2 (F), count ()."""

    def __init__(self, group=None, **kwargs):
        super().__init__()
        self.moving_mean = None
        self.moving_m2 = None
        self.count = None
        self.built = False
        # data-parallel training: every rank folds the batch moments of ALL ranks (rank order) into its running totals,
        # so the buffers stay identical without a broadcast
        self.group = group

    def build(self, input_shape):
        shapes = _flatten(input_shape) if isinstance(input_shape, dict) else [tuple(input_shape)]
        self.moving_mean, self.moving_m2, self.count = [], [], []
        for i, shape in enumerate(shapes):
            F = int(shape[-1])
            for name, store, init in (("moving_mean", self.moving_mean, torch.zeros(F)),
                                      ("moving_m2", self.moving_m2, torch.zeros(F)), ("count", self.count, torch.zeros(()))):
                self.register_buffer(f"{name}_{i}", init)
                store.append(f"{name}_{i}")
        self.built = True

    def _buf(self, store, index) -> torch.Tensor:
        return getattr(self, store[index])

    def moving_std(self, index: int) -> torch.Tensor:
        """standardize.py:27-37: 1 where M2 == 0, even if count > 0."""
        m2, count = self._buf(self.moving_m2, index), self._buf(self.count, index)
        return torch.where(m2 > 0, torch.sqrt(m2 / count), torch.ones_like(m2))

    def call(self, x, stage: str = "inference", forward: bool = True, log_det_jac: bool = False,
             transformation_type: str = "location_scale"):
        if (forward or log_det_jac) and transformation_type != "location_scale":
            raise ValueError('Non-default transformation (i.e. transformation_type != "location_scale") '
                             "is not supported for forward or log_det_jac.")
        if not self.built:
            raise RuntimeError("Standardize is not built; call build(input_shape) first")
        flattened = _flatten(x) if isinstance(x, (dict, list, tuple)) else [x]
        outputs, ldjs = [], []
        for idx, val in enumerate(flattened):
            if stage == "training":
                self._update_moments(val, idx)
            if forward or transformation_type == "location_scale":
                out = self._location_scale(val, idx, forward)
            else:
                # point-estimate transforms (covariance-like heads; not on the sampling path): F-sized device ops
                std = self.moving_std(idx).reshape((1,) * (val.ndim - 1) + (-1,))
                if transformation_type == "both_sides_scale":
                    out = val * std * std.movedim(-1, -2)
                elif transformation_type == "left_side_scale":
                    out = val * std.movedim(-1, -2)
                elif transformation_type == "right_side_scale_inverse":
                    out = val / std
                elif transformation_type == "identity":
                    out = val
                elif isinstance(transformation_type, str):
                    raise NotImplementedError(f"No implementation for transformation type {transformation_type!r}")
                else:
                    raise TypeError(f"Unexpected type for transformation_type {transformation_type!r} "
                                    f"of type {type(transformation_type)!r}")
            outputs.append(out)
            if log_det_jac:
                ldj = torch.sum(torch.log(torch.abs(self.moving_std(idx))), dim=-1)
                ldj = ldj.expand(val.shape[:-1]).clone() if val.ndim > 1 else ldj
                ldjs.append(-ldj if forward else ldj)
        packed = _pack_like(x, outputs) if isinstance(x, (dict, list, tuple)) else outputs[0]
        if log_det_jac:
            return packed, (_pack_like(x, ldjs) if isinstance(x, (dict, list, tuple)) else ldjs[0])
        return packed

    forward = call

    # ---- kernels --------------------------------------------------------------------------------------------------
    def _check(self, x, index):
        mean = self._buf(self.moving_mean, index)
        if not torch.is_tensor(x) or not x.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        if not mean.is_cuda:
            raise RuntimeError("Standardize buffers are not on a CUDA device; call .cuda() after build()")
        if x.shape[-1] != mean.shape[0]:
            raise ValueError(f"expected last axis {mean.shape[0]}, got {tuple(x.shape)}")
        return x.detach().to(torch.float32).contiguous()

    def _location_scale(self, x, index, forward: bool, out=None):
        xc = self._check(x, index)
        F = xc.shape[-1]
        rows = xc.numel() // F
        out = torch.empty_like(xc) if out is None else out
        _lib.check(_lib.lib().bfk_standardize(_lib.ptr(xc), _lib.ptr(out), rows, F, _lib.ptr(self._buf(self.moving_mean, index)),
                                              _lib.ptr(self._buf(self.moving_m2, index)),
                                              _lib.ptr(self._buf(self.count, index)), 1 if forward else 0,
                                              _lib.stream_ptr()), "bfk_standardize")
        return out

    def batch_moments(self, x, index: int = 0) -> torch.Tensor:
        """(2F+1) record {mean[F], m2[F], count} of this batch, all axes but the last reduced (standardize.py:150-157)."""
        xc = self._check(x, index)
        F = xc.shape[-1]
        rows = xc.numel() // F
        L = _lib.lib()
        ws = torch.empty(max(1, L.bfk_moments_workspace_floats(rows, F)), device=xc.device, dtype=torch.float32)
        stats = torch.empty(2 * F + 1, device=xc.device, dtype=torch.float32)
        _lib.check(L.bfk_batch_moments(_lib.ptr(xc), rows, F, _lib.ptr(ws), _lib.ptr(stats), _lib.stream_ptr()),
                   "bfk_batch_moments")
        return stats

    def _update_moments(self, x, index: int):
        stats = self.batch_moments(x, index)
        parts = stats.unsqueeze(0)
        if self.group is not None and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            world = dist.get_world_size(self.group)
            parts = torch.empty(world, stats.numel(), device=stats.device, dtype=torch.float32)
            dist.all_gather_into_tensor(parts, stats, group=self.group)
        F = stats.numel() // 2
        _lib.check(_lib.lib().bfk_moments_merge(_lib.ptr(self._buf(self.moving_mean, index)),
                                                _lib.ptr(self._buf(self.moving_m2, index)),
                                                _lib.ptr(self._buf(self.count, index)), _lib.ptr(parts.contiguous()),
                                                parts.shape[0], F, _lib.stream_ptr()), "bfk_moments_merge")


class Standardization(nn.Module):
    """standardization.py:11-88: one `Standardize` per named approximator variable."""

    ALL = ("inference_variables", "summary_variables", "inference_conditions")

    def __init__(self, standardize=None, **kwargs):
        super().__init__()
        if isinstance(standardize, str) and standardize != "all":
            self.standardize = [standardize]
        else:
            self.standardize = standardize or []
        self.standardize_layers = None
        if self.standardize != "all":
            self.standardize_layers = nn.ModuleDict({var: Standardize() for var in self.standardize})
        self.built = False

    def maybe_standardize(self, x, key: str, stage: str = "inference", forward: bool = True, log_det_jac: bool = False,
                          transformation_type: str = "location_scale"):
        if x is None:
            return (None, 0.0) if log_det_jac else None
        if key not in self.standardize:
            return (x, 0.0) if log_det_jac else x
        return self.standardize_layers[key](x, forward=forward, stage=stage, log_det_jac=log_det_jac,
                                            transformation_type=transformation_type)

    def build(self, data_shapes: dict) -> None:
        if self.standardize == "all":
            self.standardize = [var for var in self.ALL if var in data_shapes]
            self.standardize_layers = nn.ModuleDict({var: Standardize() for var in self.standardize})
        for var, layer in self.standardize_layers.items():
            layer.build(data_shapes[var])
        self.built = True

    def get_config(self):
        return {"standardize": self.standardize}
