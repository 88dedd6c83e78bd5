"""A test-only stand-in for the `keras` package (torch backend), just wide enough to run the reference's own hot-path
sources UNMODIFIED (see ref_loader.py).  TEST INFRASTRUCTURE: never imported by bayesflow_b200/ or by bench.py.

Keras (>= 3.12, `pyproject.toml:28` of the reference, not installable here) semantics this file ASSUMES — every entry
is a one-liner below and is the complete list of what the reference-backed pin takes on trust:

  layers.Dense                y = matmul(x, kernel[in, out]) + bias; kernel_initializer default glorot_uniform, bias zeros
  layers.LayerNormalization   last axis, epsilon = 1e-3, biased variance, gamma * (x - mean) * rsqrt(var + eps) + beta
  layers.RMSNormalization     last axis, epsilon = 1e-6 (keras default None -> backend epsilon... keras 3.12 uses 1e-6
                              through ops.rms_normalization), x * rsqrt(mean(x^2) + eps) * scale
  layers.Dropout              inverted dropout when training is truthy, identity otherwise
  layers.Activation           applies the function; layers.Identity returns its input
  activations                 mish = x * tanh(softplus(x)); silu = swish = x * sigmoid(x); gelu(approximate=True) (the
                              keras 3 default is the tanh form); relu; tanh; elu; softplus; sigmoid; linear
  losses.get("mse"/"mae")     mean over the LAST axis of (y_true - y_pred)^2 / |y_true - y_pred|
  initializers                he_normal = truncated normal(+-2 sigma), sigma = sqrt(2 / fan_in) / 0.87962566103423978;
                              random_normal = N(0, 0.05^2); glorot_uniform; zeros; ones
  Layer                       `built` becomes True after build() returns (keras wraps build); __call__ builds from the
                              input shapes if needed and forwards `training` only to call() signatures that take it
  ops.*                       the torch backend's eager semantics: python scalars become float32 / int tensors,
                              cond / while_loop / fori_loop are python control flow, dtype() returns a string
  random.*                    torch generators (streams differ from keras: parity tests feed fixed noise)
"""
from __future__ import annotations

import inspect
import math
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

_DTYPES = {"float32": torch.float32, "float64": torch.float64, "float16": torch.float16, "bfloat16": torch.bfloat16,
           "int32": torch.int32, "int64": torch.int64, "bool": torch.bool, "uint8": torch.uint8}
_DTYPE_NAMES = {v: k for k, v in _DTYPES.items()}


def _to_dtype(d):
    if d is None or isinstance(d, torch.dtype):
        return d
    return _DTYPES[str(d)]


def _t(x, dtype=None):
    """torch backend convert_to_tensor: python floats -> float32, python ints -> int32 (floatx = float32)."""
    dtype = _to_dtype(dtype)
    if isinstance(x, torch.Tensor):
        return x if dtype is None or x.dtype == dtype else x.to(dtype)
    if isinstance(x, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(x))
        if dtype is None and t.dtype == torch.float64:
            return t  # keras keeps numpy dtypes
        return t if dtype is None else t.to(dtype)
    if isinstance(x, bool):
        return torch.tensor(x, dtype=dtype or torch.bool)
    if isinstance(x, int):
        return torch.tensor(x, dtype=dtype or torch.int32)
    if isinstance(x, float):
        return torch.tensor(x, dtype=dtype or torch.float32)
    if isinstance(x, (list, tuple)):
        if len(x) and all(isinstance(v, torch.Tensor) for v in x):
            return torch.stack(list(x)) if dtype is None else torch.stack(list(x)).to(dtype)
        t = torch.tensor(x)
        if dtype is None and t.dtype == torch.float64:
            t = t.to(torch.float32)
        return t if dtype is None else t.to(dtype)
    return torch.as_tensor(x, dtype=dtype)


def _pair(a, b):
    """binary-op operand promotion of the torch backend (python scalars adopt the tensor operand's dtype; a python
    float against an integer tensor promotes to float32)."""
    ta, tb = isinstance(a, torch.Tensor), isinstance(b, torch.Tensor)
    if ta and tb:
        return a, b
    if not ta and not tb:
        return _t(a), _t(b)
    ten, sc = (a, b) if ta else (b, a)
    if isinstance(sc, float) and not ten.dtype.is_floating_point:
        ten = ten.to(torch.float32)
    sc = _t(sc, ten.dtype) if not isinstance(sc, (np.ndarray, list, tuple)) else _t(sc)
    return (ten, sc) if ta else (sc, ten)


# ------------------------------------------------------------------------------------------------------------ ops
class _Ops(types.ModuleType):
    pass


ops = _Ops("keras.ops")


def _axis_kw(axis, keepdims):
    kw = {}
    if axis is not None:
        kw["dim"] = axis
    if keepdims:
        kw["keepdim"] = True
    return kw


def _reduce(fn):
    def red(x, axis=None, keepdims=False):
        x = _t(x)
        return fn(x, **_axis_kw(axis, keepdims))
    return red


def _max(x, axis=None, keepdims=False, initial=None):
    x = _t(x)
    return x.amax(**_axis_kw(axis, keepdims)) if axis is not None else x.max()


def _min(x, axis=None, keepdims=False, initial=None):
    x = _t(x)
    return x.amin(**_axis_kw(axis, keepdims)) if axis is not None else x.min()


def _norm(x, ord=None, axis=None, keepdims=False):
    return torch.linalg.vector_norm(_t(x), ord=2 if ord is None else ord, dim=axis, keepdim=keepdims)


def _cond(pred, true_fn, false_fn):
    return true_fn() if bool(pred) else false_fn()


def _while_loop(cond, body, loop_vars, maximum_iterations=None):
    is_seq = isinstance(loop_vars, (list, tuple))
    loop_vars = tuple(loop_vars) if is_seq else (loop_vars,)
    it = 0
    while bool(cond(*loop_vars)) and (maximum_iterations is None or it < maximum_iterations):
        loop_vars = body(*loop_vars)
        loop_vars = tuple(loop_vars) if isinstance(loop_vars, (list, tuple)) else (loop_vars,)
        it += 1
    return loop_vars if is_seq else loop_vars[0]


def _fori_loop(lower, upper, body_fun, init_val):
    val = init_val
    for i in range(lower, upper):
        val = body_fun(i, val)
    return val


def _bin(fn):
    def f(a, b):
        a, b = _pair(a, b)
        return fn(a, b)
    return f


def _clip(x, x_min, x_max):
    x = _t(x)
    return torch.clip(x, _t(x_min, x.dtype) if not isinstance(x_min, torch.Tensor) else x_min,
                      _t(x_max, x.dtype) if not isinstance(x_max, torch.Tensor) else x_max)


def _where(c, a=None, b=None):
    c = _t(c)
    if a is None:
        return torch.where(c)
    a, b = _pair(a, b)
    return torch.where(c.bool(), a, b)


def _cast(x, dtype):
    return _t(x).to(_to_dtype(dtype))


def _shape(x):
    return tuple(x.shape)


def _dtype(x):
    return _DTYPE_NAMES.get(x.dtype, str(x.dtype)) if isinstance(x, torch.Tensor) else str(np.asarray(x).dtype)


def _top_k(x, k, sorted=True):
    return torch.topk(_t(x), k, sorted=sorted)


for _name, _fn in {
    "shape": _shape, "dtype": _dtype, "ndim": lambda x: _t(x).ndim, "size": lambda x: _t(x).numel(),
    "convert_to_tensor": lambda x, dtype=None, sparse=None, ragged=None: _t(x, dtype),
    "convert_to_numpy": lambda x: _t(x).detach().cpu().numpy(), "is_tensor": lambda x: isinstance(x, torch.Tensor),
    "cast": _cast, "copy": lambda x: _t(x).clone(),
    "abs": lambda x: torch.abs(_t(x)), "sign": lambda x: torch.sign(_t(x)), "sqrt": lambda x: torch.sqrt(_t(x)),
    "exp": lambda x: torch.exp(_t(x)), "expm1": lambda x: torch.expm1(_t(x)), "log": lambda x: torch.log(_t(x)),
    "log1p": lambda x: torch.log1p(_t(x)), "square": lambda x: torch.square(_t(x)), "sin": lambda x: torch.sin(_t(x)),
    "cos": lambda x: torch.cos(_t(x)), "tan": lambda x: torch.tan(_t(x)), "arctan": lambda x: torch.arctan(_t(x)),
    "tanh": lambda x: torch.tanh(_t(x)), "sigmoid": lambda x: torch.sigmoid(_t(x)), "softplus": lambda x: F.softplus(_t(x)),
    "erf": lambda x: torch.erf(_t(x)), "erfinv": lambda x: torch.erfinv(_t(x)), "isnan": lambda x: torch.isnan(_t(x)),
    "isinf": lambda x: torch.isinf(_t(x)), "isfinite": lambda x: torch.isfinite(_t(x)),
    "power": _bin(torch.pow), "maximum": _bin(torch.maximum), "minimum": _bin(torch.minimum), "less": _bin(torch.lt),
    "less_equal": _bin(torch.le), "greater": _bin(torch.gt), "greater_equal": _bin(torch.ge), "equal": _bin(torch.eq),
    "add": _bin(torch.add), "subtract": _bin(torch.sub), "multiply": _bin(torch.mul), "divide": _bin(torch.div),
    "matmul": lambda a, b: torch.matmul(_t(a), _t(b)),
    "logical_and": lambda a, b: torch.logical_and(_t(a), _t(b)), "logical_or": lambda a, b: torch.logical_or(_t(a), _t(b)),
    "logical_not": lambda a: torch.logical_not(_t(a)),
    "all": _reduce(torch.all), "any": _reduce(torch.any), "sum": _reduce(torch.sum), "mean": _reduce(torch.mean),
    "max": _max, "min": _min, "norm": _norm, "clip": _clip, "where": _where,
    "zeros": lambda shape, dtype=None: torch.zeros(tuple(shape), dtype=_to_dtype(dtype) or torch.float32),
    "ones": lambda shape, dtype=None: torch.ones(tuple(shape), dtype=_to_dtype(dtype) or torch.float32),
    "full": lambda shape, v, dtype=None: torch.full(tuple(shape), v, dtype=_to_dtype(dtype) or torch.float32),
    "zeros_like": lambda x, dtype=None: torch.zeros_like(_t(x), dtype=_to_dtype(dtype)),
    "ones_like": lambda x, dtype=None: torch.ones_like(_t(x), dtype=_to_dtype(dtype)),
    "eye": lambda n, m=None, dtype=None: torch.eye(n, m or n, dtype=_to_dtype(dtype) or torch.float32),
    "arange": lambda *a, dtype=None: torch.arange(*a, dtype=_to_dtype(dtype)),
    "stack": lambda xs, axis=0: torch.stack([_t(x) for x in xs], dim=axis),
    "concatenate": lambda xs, axis=0: torch.cat([_t(x) for x in xs], dim=axis),
    "split": lambda x, n, axis=0: torch.tensor_split(_t(x), n, dim=axis),
    "reshape": lambda x, shape: torch.reshape(_t(x), tuple(shape)),
    "expand_dims": lambda x, axis: torch.unsqueeze(_t(x), axis) if isinstance(axis, int) else
    _expand_dims_multi(_t(x), axis),
    "squeeze": lambda x, axis=None: torch.squeeze(_t(x)) if axis is None else torch.squeeze(_t(x), dim=axis),
    "broadcast_to": lambda x, shape: torch.broadcast_to(_t(x), tuple(shape)),
    "tile": lambda x, reps: torch.tile(_t(x), tuple(reps)),
    "repeat": lambda x, repeats, axis=None: torch.repeat_interleave(_t(x), repeats, dim=axis),
    "flip": lambda x, axis=None: torch.flip(_t(x), dims=[axis] if isinstance(axis, int) else list(axis)),
    "linspace": lambda start, stop, num=50, dtype=None: torch.linspace(float(start), float(stop), int(num),
                                                                        dtype=_to_dtype(dtype) or torch.float32),
    "full_like": lambda x, v, dtype=None: torch.full_like(_t(x), float(v), dtype=_to_dtype(dtype)),
    "take": lambda x, idx, axis=None: torch.take(_t(x), _t(idx).long()) if axis is None else
    torch.index_select(_t(x), axis, _t(idx).long().reshape(-1)),
    "take_along_axis": lambda x, idx, axis=None: torch.take_along_dim(_t(x), _t(idx).long(), dim=axis),
    "top_k": _top_k, "triu": lambda x, k=0: torch.triu(_t(x), k), "tril": lambda x, k=0: torch.tril(_t(x), k),
    "trace": lambda x, offset=0, axis1=0, axis2=1: torch.diagonal(_t(x), offset, axis1, axis2).sum(-1),
    "inv": lambda x: torch.linalg.inv(_t(x)), "solve": lambda a, b: torch.linalg.solve(_t(a), _t(b)),
    "expm": lambda x: torch.linalg.matrix_exp(_t(x)),
    "transpose": lambda x, axes=None: _t(x).permute(*axes) if axes is not None else _t(x).T,
    "logsumexp": lambda x, axis=None, keepdims=False: torch.logsumexp(_t(x), dim=axis if axis is not None else tuple(range(_t(x).ndim)), keepdim=keepdims),
    "diagonal": lambda x, offset=0, axis1=0, axis2=1: torch.diagonal(_t(x), offset, axis1, axis2),
    "log10": lambda x: torch.log10(_t(x)), "floor": lambda x: torch.floor(_t(x)), "round": lambda x: torch.round(_t(x)),
    "cond": _cond, "while_loop": _while_loop, "fori_loop": _fori_loop,
    "stop_gradient": lambda x: _t(x).detach(),
}.items():
    setattr(ops, _name, _fn)


def _expand_dims_multi(x, axes):
    for a in sorted(a if a >= 0 else a + x.ndim + len(axes) for a in axes):
        x = x.unsqueeze(a)
    return x


def _scatter_update(x, indices, updates):
    x = _t(x).clone()
    idx = _t(indices).long()
    x[tuple(idx.t())] = _t(updates, x.dtype)
    return x


ops.scatter_update = _scatter_update


# ------------------------------------------------------------------------------------------------------------ random
class SeedGenerator:
    def __init__(self, seed=None, name=None, **kwargs):
        self.generator = torch.Generator()
        if seed is None:
            self.generator.seed()
        else:
            self.generator.manual_seed(int(seed))


def _gen(seed):
    if isinstance(seed, SeedGenerator):
        return seed.generator
    if isinstance(seed, int):
        return torch.Generator().manual_seed(seed)
    return None


random = types.ModuleType("keras.random")
random.SeedGenerator = SeedGenerator
random.normal = lambda shape, mean=0.0, stddev=1.0, dtype=None, seed=None: (
    mean + stddev * torch.randn(tuple(shape), dtype=_to_dtype(dtype) or torch.float32, generator=_gen(seed)))
random.uniform = lambda shape, minval=0.0, maxval=1.0, dtype=None, seed=None: (
    minval + (maxval - minval) * torch.rand(tuple(shape), dtype=_to_dtype(dtype) or torch.float32, generator=_gen(seed)))
random.categorical = lambda logits, num_samples, dtype="int32", seed=None: torch.multinomial(
    torch.softmax(_t(logits).double(), dim=-1), num_samples, replacement=True, generator=_gen(seed))
random.shuffle = lambda x, axis=0, seed=None: torch.index_select(
    _t(x), axis, torch.randperm(_t(x).shape[axis], generator=_gen(seed)))


# ------------------------------------------------------------------------------------------------------------ activations
def _gelu(x, approximate=True):
    return F.gelu(x, approximate="tanh" if approximate else "none")


_ACTIVATIONS = {
    "mish": lambda x: x * torch.tanh(F.softplus(x)), "silu": F.silu, "swish": F.silu, "gelu": _gelu, "relu": F.relu,
    "tanh": torch.tanh, "elu": F.elu, "softplus": F.softplus, "sigmoid": torch.sigmoid, "linear": lambda x: x,
}
activations = types.ModuleType("keras.activations")
for _k, _v in _ACTIVATIONS.items():
    setattr(activations, _k, _v)


def _activation_get(identifier):
    if identifier is None:
        return _ACTIVATIONS["linear"]
    if isinstance(identifier, str):
        return _ACTIVATIONS[identifier]
    return identifier  # a callable or a Layer


activations.get = _activation_get


# ------------------------------------------------------------------------------------------------------------ initializers
class Initializer:
    def __call__(self, shape, dtype=None):
        raise NotImplementedError


def _fans(shape):
    return (shape[0], shape[0]) if len(shape) == 1 else (shape[-2], shape[-1])


class _FnInit(Initializer):
    def __init__(self, fn):
        self.fn = fn

    def __call__(self, shape, dtype=None):
        return self.fn(tuple(shape)).to(_to_dtype(dtype) or torch.float32)


def _he_normal(shape):
    std = math.sqrt(2.0 / _fans(shape)[0]) / 0.87962566103423978
    w = torch.empty(shape)
    torch.nn.init.trunc_normal_(w, 0.0, std, -2 * std, 2 * std)
    return w


def _glorot_uniform(shape):
    fi, fo = _fans(shape)
    lim = math.sqrt(6.0 / (fi + fo))
    return (torch.rand(shape) * 2 - 1) * lim


_INITS = {"he_normal": _he_normal, "glorot_uniform": _glorot_uniform, "zeros": torch.zeros, "ones": torch.ones,
          "random_normal": lambda s: torch.randn(s) * 0.05}
initializers = types.ModuleType("keras.initializers")
initializers.Initializer = Initializer


def _init_get(identifier):
    if isinstance(identifier, Initializer):
        return identifier
    if isinstance(identifier, str):
        return _FnInit(_INITS[identifier])
    if isinstance(identifier, torch.Tensor):
        return _FnInit(lambda s, v=identifier: v.clone().reshape(s))
    if callable(identifier):
        return identifier
    raise ValueError(identifier)


initializers.get = _init_get


# ------------------------------------------------------------------------------------------------------------ losses
class Loss:
    pass


losses = types.ModuleType("keras.losses")
losses.Loss = Loss
_LOSSES = {
    "mse": lambda y_true, y_pred: torch.mean(torch.square(_t(y_true) - _t(y_pred)), dim=-1),
    "mae": lambda y_true, y_pred: torch.mean(torch.abs(_t(y_true) - _t(y_pred)), dim=-1),
}
_LOSSES["mean_squared_error"] = _LOSSES["mse"]
_LOSSES["mean_absolute_error"] = _LOSSES["mae"]
losses.get = lambda identifier: _LOSSES[identifier] if isinstance(identifier, str) else identifier


# ------------------------------------------------------------------------------------------------------------ Layer
class Layer:
    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        if "build" in cls.__dict__:
            user_build = cls.__dict__["build"]

            def build(self, *a, __b=user_build, **k):
                r = __b(self, *a, **k)
                self.built = True  # keras wraps build() the same way (Layer.__new__ -> build_wrapper)
                return r
            build.__wrapped__ = user_build
            cls.build = build

    def __init__(self, *, name=None, dtype=None, trainable=True, **kwargs):
        self.name = name or type(self).__name__.lower()
        self.built = False
        self._own_weights = []
        object.__setattr__(self, "trainable", trainable)
        self.dtype = "float32"

    def add_weight(self, shape=None, initializer=None, dtype=None, trainable=True, name=None, **kwargs):
        init = _init_get(initializer if initializer is not None else "glorot_uniform")
        w = torch.nn.Parameter(init(tuple(shape), dtype=dtype).detach().clone(), requires_grad=bool(trainable))
        w._keras_name = name
        self._own_weights.append(w)
        return w

    def build(self, *args, **kwargs):
        self.built = True

    def compute_output_shape(self, input_shape, *a, **k):
        return input_shape

    def call(self, *args, **kwargs):
        raise NotImplementedError

    def __call__(self, *args, **kwargs):
        if not self.built:
            shapes = tuple(_shape_tree(a) for a in args)
            try:
                self.build(*shapes)
            except TypeError:
                self.build(*shapes, **{k + "_shape": _shape_tree(v) for k, v in kwargs.items()
                                       if isinstance(v, torch.Tensor)})
            self.built = True
        if "training" in kwargs:
            params = inspect.signature(self.call).parameters
            if "training" not in params and not any(p.kind == p.VAR_KEYWORD for p in params.values()):
                kwargs.pop("training")
        return self.call(*args, **kwargs)

    def _sublayers(self):
        for v in self.__dict__.values():
            if isinstance(v, Layer):
                yield v
            elif isinstance(v, (list, tuple)):
                for u in v:
                    if isinstance(u, Layer):
                        yield u

    @property
    def weights(self):
        out = list(self._own_weights)
        seen = {id(w) for w in out}
        for sub in self._sublayers():
            for w in sub.weights:
                if id(w) not in seen:
                    seen.add(id(w))
                    out.append(w)
        return out

    variables = weights

    @property
    def trainable_weights(self):
        return [w for w in self.weights if w.requires_grad]

    def get_config(self):
        return {"name": self.name, "trainable": self.trainable, "dtype": self.dtype}

    @classmethod
    def from_config(cls, config):
        return cls(**config)


def _shape_tree(x):
    if isinstance(x, torch.Tensor):
        return tuple(x.shape)
    if isinstance(x, (list, tuple)):
        return tuple(_shape_tree(v) for v in x)
    return None


class Model(Layer):
    pass


class Sequential(Model):
    def __init__(self, layers=None, **kw):
        super().__init__(**kw)
        self.layers = list(layers or [])

    def call(self, x, **kw):
        for layer in self.layers:
            x = layer(x)
        return x


class Dense(Layer):
    def __init__(self, units, activation=None, use_bias=True, kernel_initializer="glorot_uniform",
                 bias_initializer="zeros", **kw):
        super().__init__(**kw)
        self.units, self.use_bias = int(units), use_bias
        self.kernel_initializer, self.bias_initializer = kernel_initializer, bias_initializer
        self.activation = _activation_get(activation)
        self.kernel = self.bias = None

    def build(self, input_shape):
        if self.built:
            return
        self.kernel = self.add_weight(shape=(input_shape[-1], self.units), initializer=self.kernel_initializer, name="kernel")
        if self.use_bias:
            self.bias = self.add_weight(shape=(self.units,), initializer=self.bias_initializer, name="bias")

    def call(self, x):
        y = torch.matmul(x, self.kernel)
        if self.bias is not None:
            y = y + self.bias
        return self.activation(y)

    def compute_output_shape(self, input_shape):
        return tuple(input_shape[:-1]) + (self.units,)


class LayerNormalization(Layer):
    def __init__(self, axis=-1, epsilon=1e-3, center=True, scale=True, **kw):
        super().__init__(**kw)
        self.axis, self.epsilon = axis, epsilon
        self.gamma = self.beta = None

    def build(self, input_shape):
        if self.built:
            return
        self.gamma = self.add_weight(shape=(input_shape[-1],), initializer="ones", name="gamma")
        self.beta = self.add_weight(shape=(input_shape[-1],), initializer="zeros", name="beta")

    def call(self, x, training=None):
        mean = x.mean(dim=-1, keepdim=True)
        var = (x - mean).square().mean(dim=-1, keepdim=True)
        return (x - mean) * torch.rsqrt(var + self.epsilon) * self.gamma + self.beta


class RMSNormalization(Layer):
    def __init__(self, axis=-1, epsilon=1e-6, **kw):
        super().__init__(**kw)
        self.axis, self.epsilon = axis, epsilon
        self.scale = None

    def build(self, input_shape):
        if self.built:
            return
        self.scale = self.add_weight(shape=(input_shape[-1],), initializer="ones", name="scale")

    def call(self, x, training=None):
        return x * torch.rsqrt(x.square().mean(dim=-1, keepdim=True) + self.epsilon) * self.scale


class BatchNormalization(Layer):
    def __init__(self, **kw):
        super().__init__(**kw)

    def call(self, x, training=None):
        raise NotImplementedError("BatchNormalization is outside the pinned path")


class Dropout(Layer):
    def __init__(self, rate, seed=None, **kw):
        super().__init__(**kw)
        self.rate = rate

    def call(self, x, training=None):
        return F.dropout(x, self.rate, training=True) if training and self.rate > 0 else x


class Activation(Layer):
    def __init__(self, activation, **kw):
        super().__init__(**kw)
        self.activation = _activation_get(activation)

    def call(self, x):
        return self.activation(x)


class Identity(Layer):
    def call(self, x):
        return x


layers = types.ModuleType("keras.layers")
for _c in (Layer, Dense, LayerNormalization, RMSNormalization, BatchNormalization, Dropout, Activation, Identity):
    setattr(layers, _c.__name__, _c)


# ------------------------------------------------------------------------------------------------------------ saving / misc
_REGISTRY = {}


def _register(package="Custom", name=None):
    def deco(obj):
        _REGISTRY[f"{package}>{name or obj.__name__}"] = obj
        return obj
    return deco


saving = types.ModuleType("keras.saving")
saving.register_keras_serializable = _register
saving.get_registered_object = lambda name, custom_objects=None, module_objects=None: (
    (custom_objects or {}).get(name) or _REGISTRY.get(name) or (module_objects or {}).get(name))
saving.get_registered_name = lambda obj: next((k for k, v in _REGISTRY.items() if v is obj), getattr(obj, "__name__", None))
saving.serialize_keras_object = lambda obj: obj      # configs are only passed through, never written, in the pin tests
saving.deserialize_keras_object = lambda config, custom_objects=None, safe_mode=True, **kw: config

backend = types.ModuleType("keras.backend")
backend.backend = lambda: "torch"
backend.floatx = lambda: "float32"
backend.epsilon = lambda: 1e-7


def _map_structure(fn, *structs):
    s0 = structs[0]
    if isinstance(s0, dict):
        return {k: _map_structure(fn, *(s[k] for s in structs)) for k in s0}
    if isinstance(s0, (list, tuple)) and not isinstance(s0, torch.Size):
        return type(s0)(_map_structure(fn, *vs) for vs in zip(*structs))
    return fn(*structs)


def _flatten(s):
    if isinstance(s, dict):
        return [v for k in sorted(s) for v in _flatten(s[k])]
    if isinstance(s, (list, tuple)) and not isinstance(s, torch.Size):
        return [v for u in s for v in _flatten(u)]
    return [s]


tree = types.ModuleType("keras.tree")
tree.map_structure = _map_structure
tree.flatten = _flatten

utils = types.ModuleType("keras.utils")
utils.is_keras_tensor = lambda x: False


def build_module():
    keras = types.ModuleType("keras")
    keras.__version__ = "3.12-stub"
    keras.__path__ = []  # mark as a package so that `from keras import ops` and `import keras.saving` work
    keras.__bayesflow_b200_stub__ = True
    for name, mod in (("ops", ops), ("random", random), ("activations", activations), ("initializers", initializers),
                      ("losses", losses), ("layers", layers), ("saving", saving), ("backend", backend), ("tree", tree),
                      ("utils", utils)):
        setattr(keras, name, mod)
    keras.Layer, keras.Model, keras.Sequential = Layer, Model, Sequential
    keras.Initializer, keras.Loss = Initializer, Loss
    keras.Variable = torch.nn.Parameter
    keras.KerasTensor = type("KerasTensor", (), {})
    return keras


def install():
    """Put the stand-in into sys.modules as `keras` (refuses to shadow a real keras install)."""
    if "keras" in sys.modules and not getattr(sys.modules["keras"], "__bayesflow_b200_stub__", False):
        raise RuntimeError("a real keras is imported; the stand-in is for images without keras only")
    if "keras" not in sys.modules:
        try:
            import importlib.util
            if importlib.util.find_spec("keras") is not None:
                raise RuntimeError("a real keras is installed; use it instead of the stand-in")
        except (ImportError, ValueError):
            pass
        keras = build_module()
        sys.modules["keras"] = keras
        for sub in ("ops", "random", "activations", "initializers", "losses", "layers", "saving", "backend", "tree", "utils"):
            sys.modules[f"keras.{sub}"] = getattr(keras, sub)
    return sys.modules["keras"]
