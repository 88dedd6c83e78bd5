"""Training side of the hot path: forward + loss + backward of the velocity network on the layer-wise tcgen05 path
(``bfk_train_step``), the optimizer (``bfk_adam_step``) and the data-parallel gradient exchange.

  TrainEngine.loss_and_grads  <- FlowMatching.compute_metrics + autograd backward
                                 flow_matching.py:132-178, torch_approximator.py:88-99
  Adam                        <- keras.optimizers.Adam(lr, clipnorm=1.5) / AdamW as BasicWorkflow builds them
                                 workflows/basic_workflow.py:514-526
  Trainer.train_step          <- TorchApproximator.train_step (zero_grad, compute_metrics, backward, optimizer step)
                                 approximators/backend_approximators/torch_approximator.py:62-113

Parameters live in ONE flat fp32 buffer (the nn.Parameters are views into it) and so do the gradients: the
data-parallel exchange is a single NCCL all-reduce and the optimizer a single kernel over all variables.
No CPU or eager fallback: everything below raises unless libbfk.so is loaded and the tensors are CUDA tensors.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib


class TrainArgs(C.Structure):
    _fields_ = [("x", C.c_void_p), ("t", C.c_void_p), ("cond", C.c_void_p), ("target", C.c_void_p),
                ("pred_scale", C.c_void_p), ("row_weight", C.c_void_p), ("B", C.c_int64), ("loss", C.c_int32),
                ("dw_swap", C.c_int32), ("dropout", C.c_float), ("loss_scale", C.c_float), ("grad_scale", C.c_float),
                ("seed", C.c_uint64), ("huber_c", C.c_float), ("seed_dev", C.c_void_p), ("row_offset", C.c_int64),
                ("elem_mask", C.c_void_p)]


def _bind(L):
    if getattr(L, "_train_bound", False):
        return L
    L.bfk_train_workspace_bytes.restype = C.c_size_t
    L.bfk_train_workspace_bytes.argtypes = [C.POINTER(_lib.NetConfig), C.c_int64, C.c_int32]
    L.bfk_train_step.argtypes = [C.POINTER(_lib.NetConfig), C.POINTER(_lib.Weights), C.POINTER(_lib.Weights),
                                 C.POINTER(TrainArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    L.bfk_fm_interpolate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
    L.bfk_adam_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p]
    L.bfk_adam_step_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                    C.c_float, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p]
    L.bfk_fm_train_batch.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_uint64, C.c_void_p, C.c_float,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.bfk_train_vjp.argtypes = [C.POINTER(_lib.NetConfig), C.POINTER(_lib.Weights), C.POINTER(TrainArgs), C.c_void_p,
                                C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    L.bfk_rk_combine.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_float, C.c_int64, C.c_void_p,
                                 C.c_void_p]
    L.bfk_train_dropout_mask.argtypes = [C.c_uint64, C.c_void_p, C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_float,
                                         C.c_void_p, C.c_void_p]
    L._train_bound = True
    return L


def dropout_mask(seed, layer, rows, width, rate, row_offset=0, device="cuda"):
    """The keep scales (0 or 1/(1-rate)) `bfk_train_step` applies in block `layer` for global rows
    row_offset .. row_offset+rows-1 (`bfk_train_dropout_mask`): a counter-based function of (seed, layer, row, column)."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("dropout_mask runs on a CUDA device (there is no CPU path)")
    L = _bind(_lib.lib())
    out = torch.empty(rows, width, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        _lib.check(L.bfk_train_dropout_mask(int(seed), None, int(layer), int(row_offset), int(rows), int(width),
                                            float(rate), _lib.ptr(out), _lib.stream_ptr()), "bfk_train_dropout_mask")
    return out


def fill_weights(net, get):
    """bfk_weights struct whose pointers are ``get(parameter)`` (the parameter itself or its gradient view)."""
    sub, w, keep = net.subnet, _lib.Weights(), []

    def P(p):
        t = get(p)
        if t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda:
            raise RuntimeError("training needs contiguous fp32 CUDA parameters (no CPU fallback)")
        keep.append(t)
        return t.data_ptr()

    if sub.time_embedding_dim > 1:
        w.freq = P(sub.time_emb.frequencies)
    w.x_k, w.x_b = P(sub.x_proj.kernel), P(sub.x_proj.bias)
    if net._C:
        w.c_k, w.c_b = P(sub.c_proj.kernel), P(sub.c_proj.bias)
        w.m_k, w.m_b = P(sub.merge_proj.kernel), P(sub.merge_proj.bias)
    for i, b in enumerate(sub.blocks):
        if hasattr(b, "projector"):
            w.proj_k[i], w.proj_b[i] = P(b.projector.kernel), P(b.projector.bias)
        w.dense_k[i], w.dense_b[i] = P(b.dense.kernel), P(b.dense.bias)
        w.film_b_k[i], w.film_b_b[i] = P(b.film.to_beta.kernel), P(b.film.to_beta.bias)
        if sub.film_use_gamma:
            w.film_g_k[i], w.film_g_b[i] = P(b.film.to_gamma.kernel), P(b.film.to_gamma.bias)
        if sub.norm == "layer":
            w.norm_g[i], w.norm_b[i] = P(b.norm.gamma), P(b.norm.beta)
        elif sub.norm == "rms":
            w.norm_g[i] = P(b.norm.scale)
    w.out_k, w.out_b = P(net.output_projector.kernel), P(net.output_projector.bias)
    return w, keep


class TrainEngine:
    """Flat parameter / gradient storage + the C-ABI training step for one built inference network."""

    def __init__(self, net):
        if not net.built:
            raise RuntimeError("build the network before training it")
        params = list(net.parameters())
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("bayesflow_b200: move the network to a CUDA device first (no CPU fallback)")
        self.net, self.params, self.device = net, params, dev
        self.L = _bind(_lib.lib())
        sizes = [p.numel() for p in params]
        offs = [0]
        for n in sizes:
            offs.append(offs[-1] + n)
        self.flat = torch.empty(offs[-1], device=dev, dtype=torch.float32)
        # gradients + ONE more float: the loss rides in the same buffer, so the data-parallel exchange is one all-reduce
        self.grad_ext = torch.zeros(offs[-1] + 1, device=dev, dtype=torch.float32)
        self.grad = self.grad_ext[:offs[-1]]
        self.loss_slot = self.grad_ext[offs[-1]:].view(())
        self.version = 0  # bumped whenever step() overwrites the gradient buffer (see _LossWithGrads)
        self._gviews = {}
        for p, a, b in zip(params, offs[:-1], offs[1:]):
            self.flat[a:b].copy_(p.data.reshape(-1).to(torch.float32))
            p.data = self.flat[a:b].view(p.shape)
            self._gviews[p] = self.grad[a:b].view(p.shape)
        self.seg_offsets = torch.tensor(offs, device=dev, dtype=torch.int64)
        self.offsets = offs[:-1]
        self.n_seg = len(sizes)
        self.cfg = net._engine._make_cfg()
        if net.precision != "fp16":
            raise NotImplementedError("the training path uses fp16 operands (bf16 cannot meet the 2e-3 gradient tolerance)")
        self.ragged = any(self.cfg.block_width[l] for l in range(self.cfg.L))
        if self.ragged:
            self._build_padded_arena()
        else:
            self.w, self._keep_w = fill_weights(net, lambda p: p.data)
            self.g, self._keep_g = fill_weights(net, lambda p: self._gviews[p])
        self._ws, self._ws_key = None, None
        self.launches = 0

    # -- unequal hidden widths: the kernels see every weight zero-padded to the common width cfg.W ------------------------
    def _build_padded_arena(self):
        """Padded twins of the parameters and of their gradients (include/bfk.h `block_width`).  `_pad_params` refreshes
        the twins before a step (`bfk_pad2d`), `_unpad_grads` copies the valid gradient regions back (`bfk_unpad2d`)."""
        net, sub, Wm = self.net, self.net.subnet, int(self.cfg.W)
        jobs, shapes = [], {}

        def padded_shape(p, rows_pad=None):
            if p.ndim == 1:
                return (1, p.shape[0]), (1, Wm)
            return tuple(p.shape), (rows_pad if rows_pad is not None else p.shape[0], Wm)

        def add(p, rows_pad=None, halves=1):
            (R, Cc), (Rp, Cp) = padded_shape(p, rows_pad)
            shapes[p] = (R, Cc, Rp, Cp, halves)

        W0 = sub.widths[0]
        add(sub.x_proj.kernel), add(sub.x_proj.bias)
        if net._C:
            add(sub.c_proj.kernel), add(sub.c_proj.bias)
            concat = sub.merge == "concat"
            add(sub.merge_proj.kernel, rows_pad=(2 * Wm if concat else Wm), halves=2 if concat else 1)
            add(sub.merge_proj.bias)
        for b in sub.blocks:
            add(b.dense.kernel, rows_pad=Wm), add(b.dense.bias)
            add(b.film.to_beta.kernel), add(b.film.to_beta.bias)
            if sub.film_use_gamma:
                add(b.film.to_gamma.kernel), add(b.film.to_gamma.bias)
            if sub.norm == "layer":
                add(b.norm.gamma), add(b.norm.beta)
            elif sub.norm == "rms":
                add(b.norm.scale)
            if hasattr(b, "projector"):
                add(b.projector.kernel, rows_pad=Wm), add(b.projector.bias)
        # the output projector reads W rows (the padded last block output); its D columns stay as they are
        pk = net.output_projector.kernel
        shapes[pk] = (pk.shape[0], pk.shape[1], Wm, pk.shape[1], 1)
        total = sum(Rp * Cp for (_, _, Rp, Cp, _) in shapes.values())
        self.pflat = torch.zeros(total, device=self.device, dtype=torch.float32)
        self.pgrad = torch.zeros(total, device=self.device, dtype=torch.float32)
        self._pviews, self._pgviews, off = {}, {}, 0
        for p, (R, Cc, Rp, Cp, halves) in shapes.items():
            self._pviews[p] = self.pflat[off:off + Rp * Cp]
            self._pgviews[p] = self.pgrad[off:off + Rp * Cp]
            if halves == 2:  # concat merge kernel: the h0 rows and the condition rows are padded separately
                h, hp = R // 2, Rp // 2
                jobs.append((p, 0, h, Cc, off, hp, Cp))
                jobs.append((p, h * Cc, h, Cc, off + hp * Cp, hp, Cp))
            else:
                jobs.append((p, 0, R, Cc, off, Rp, Cp))
            off += Rp * Cp
        self._pad_jobs = jobs
        self.w, self._keep_w = fill_weights(net, lambda p: self._pviews[p] if p in self._pviews else p.data)
        self.g, self._keep_g = fill_weights(net, lambda p: self._pgviews[p] if p in self._pgviews else self._gviews[p])

    def _pad_params(self):
        st = _lib.stream_ptr()
        for p, src_off, R, Cc, dst_off, Rp, Cp in self._pad_jobs:
            _lib.check(self.L.bfk_pad2d(C.c_void_p(p.data.data_ptr() + 4 * src_off), R, Cc,
                                        C.c_void_p(self.pflat.data_ptr() + 4 * dst_off), Rp, Cp, st), "bfk_pad2d")

    def _unpad_grads(self):
        st = _lib.stream_ptr()
        for p, src_off, R, Cc, dst_off, Rp, Cp in self._pad_jobs:
            _lib.check(self.L.bfk_unpad2d(C.c_void_p(self.pgrad.data_ptr() + 4 * dst_off), Cp,
                                          C.c_void_p(self._gviews[p].data_ptr() + 4 * src_off), R, Cc, st), "bfk_unpad2d")

    def grad_of(self, p):
        return self._gviews[p]

    def _workspace(self, B, backward):
        key = (int(B), bool(backward))
        if self._ws is None or self._ws_key != key:
            n = self.L.bfk_train_workspace_bytes(C.byref(self.cfg), int(B), int(backward))
            if n == 0:
                raise NotImplementedError("bfk_train_workspace_bytes: " + self.L.bfk_last_error_string().decode())
            if self._ws is None or self._ws.numel() < n:
                self._ws = torch.empty(n, device=self.device, dtype=torch.uint8)
            self._ws_key = key
        return self._ws

    @staticmethod
    def _f32(t, shape=None):
        if t is None:
            return None
        if not t.is_cuda:
            raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback)")
        t = t.detach().to(torch.float32).contiguous()
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return t

    def step(self, x, t, conditions, target, *, pred_scale=None, row_weight=None, loss="mse", dropout=0.0, seed=0,
             global_batch=None, backward=True, zero_grad=True, want_pred=False, want_loss=True, _dw_swap=0,
             cotangent_scale=1.0, huber_c=0.0, row_offset=0, seed_dev=None, loss_into_slot=False, elem_mask=None):
        """One forward (+ loss, + backward).  Gradients are accumulated into ``self.grad`` (zeroed first unless
        ``zero_grad=False``) scaled by 1/(global_batch*D).  Returns (loss 0-d device tensor or None, pred or None)."""
        net = self.net
        D = net._D
        x = self._f32(x)
        B = x.shape[0]
        if x.ndim != 2 or x.shape[1] != D:
            raise ValueError(f"expected (rows, {D}) inputs, got {tuple(x.shape)}")
        t = self._f32(t).reshape(-1)
        if t.numel() != B:
            raise ValueError("time must have one entry per row")
        cond = self._f32(conditions, (B, net._C)) if net._C else None
        if net._C and cond is None:
            raise ValueError("network was built with conditions")
        target = self._f32(target, (B, D)) if target is not None else None
        pred_scale = self._f32(pred_scale, (B,)) if pred_scale is not None else None
        row_weight = self._f32(row_weight.reshape(-1), (B,)) if row_weight is not None else None
        gb = int(global_batch or B)
        a = TrainArgs()
        a.x, a.t, a.cond, a.target = x.data_ptr(), t.data_ptr(), (cond.data_ptr() if cond is not None else None), \
            (target.data_ptr() if target is not None else None)
        a.pred_scale = pred_scale.data_ptr() if pred_scale is not None else None
        a.row_weight = row_weight.data_ptr() if row_weight is not None else None
        a.B, a.loss, a.dw_swap = B, {"mse": 0, "mae": 1, "pseudo_huber": 2}[loss], int(_dw_swap)
        a.dropout, a.seed = float(dropout), int(seed) & (2 ** 64 - 1)
        a.huber_c = float(huber_c)
        a.row_offset = int(row_offset)
        a.seed_dev = seed_dev.data_ptr() if seed_dev is not None else None
        elem_mask = self._f32(elem_mask, (B, D)) if elem_mask is not None else None  # kept alive until the launch below
        a.elem_mask = elem_mask.data_ptr() if elem_mask is not None else None
        # `cotangent_scale` (a power of two already folded into row_weight by the caller) keeps the fp16 cotangent image
        # in range; loss and gradients are un-scaled here
        a.loss_scale = 1.0 / (gb * (D if loss in ("mse", "mae", "pseudo_huber") else 1)) / cotangent_scale
        a.grad_scale = 1.0 / (gb * D) / cotangent_scale
        ws = self._workspace(B, backward)
        if self.ragged:
            self._pad_params()
            if backward:
                if not zero_grad:
                    raise NotImplementedError("gradient accumulation over several steps is not built for unequal hidden widths")
                self.pgrad.zero_()
        if backward and zero_grad:
            self.grad_ext.zero_()
            self.version += 1
        if loss_into_slot:
            if not (backward and zero_grad):
                self.loss_slot.zero_()
            loss_t = self.loss_slot
        else:
            loss_t = torch.zeros((), device=self.device, dtype=torch.float32) if want_loss else None
        pred = torch.empty(B, D, device=self.device, dtype=torch.float32) if want_pred else None
        rc = self.L.bfk_train_step(C.byref(self.cfg), C.byref(self.w), C.byref(self.g) if backward else None, C.byref(a),
                                   _lib.ptr(loss_t), _lib.ptr(pred), _lib.ptr(ws), ws.numel(), _lib.stream_ptr())
        _lib.check(rc, "bfk_train_step")
        if self.ragged and backward:
            self._unpad_grads()
        self.launches += self.launches_per_step(backward)
        return loss_t, pred

    # -- density path (density.py): forward that keeps its activations, then reverse passes w.r.t. the input ----------
    def forward_keep(self, x, t, conditions):
        """Forward only (no dropout); activations stay in the backward-sized workspace.  -> (pred, args, workspace)."""
        net, D = self.net, self.net._D
        x = self._f32(x)
        B = x.shape[0]
        if x.ndim != 2 or x.shape[1] != D:
            raise ValueError(f"expected (rows, {D}) inputs, got {tuple(x.shape)}")
        t = self._f32(t).reshape(-1)
        cond = self._f32(conditions, (B, net._C)) if net._C else None
        a = TrainArgs()
        a.x, a.t, a.cond, a.B = x.data_ptr(), t.data_ptr(), (cond.data_ptr() if cond is not None else None), B
        a.dropout, a.loss_scale, a.grad_scale = 0.0, 1.0, 1.0
        a._keep = (x, t, cond)
        ws = self._workspace(B, True)
        if self.ragged:
            self._pad_params()
        pred = torch.empty(B, D, device=self.device, dtype=torch.float32)
        _lib.check(self.L.bfk_train_step(C.byref(self.cfg), C.byref(self.w), None, C.byref(a), None, _lib.ptr(pred),
                                         _lib.ptr(ws), ws.numel(), _lib.stream_ptr()), "bfk_train_step")
        self.launches += self.launches_per_step(False) - 1
        return pred, a, ws

    def vjp(self, args, ws, *, col=-1, dout=None, trace=None, want_dx=False):
        """One reverse pass over the activations of the last forward_keep (bfk_train_vjp)."""
        dx = torch.empty(args.B, self.net._D, device=self.device, dtype=torch.float32) if want_dx else None
        dout = self._f32(dout, (args.B, self.net._D)) if dout is not None else None
        _lib.check(self.L.bfk_train_vjp(C.byref(self.cfg), C.byref(self.w), C.byref(args), _lib.ptr(dout), int(col),
                                        _lib.ptr(dx), _lib.ptr(trace), _lib.ptr(ws), ws.numel(), _lib.stream_ptr()),
                   "bfk_train_vjp")
        Lb, merge = self.cfg.L, 1 if self.net._C else 0
        self.launches += 3 + 3 * Lb + (4 if merge else 1) + 1
        return dx

    def launches_per_step(self, backward=True):
        """Kernels bfk_train_step enqueues (counted from csrc/bfk_train_host.cuh)."""
        Lb, merge = self.cfg.L, 1 if self.net._C else 0
        # one batched weight-pack launch (pack_batch_kernel) replaces the 2 L + 2 forward and 2 L + 2 backward pack launches
        # per block: Dense GEMM + block kernel; the FiLM shifts of all blocks are ONE GEMM (N = L * W) on the side stream
        fwd = 1 + 2 + 2 + (3 if merge else 1) + 2 * Lb + 1 + 1 + 1
        if not backward:
            return fwd
        bwd = 2 + Lb * 5 + (4 if merge else 1) + 3
        return fwd + bwd

    def interpolate(self, x0, x1, t):
        x0, x1, t = self._f32(x0), self._f32(x1), self._f32(t).reshape(-1)
        x, target = torch.empty_like(x1), torch.empty_like(x1)
        _lib.check(self.L.bfk_fm_interpolate(_lib.ptr(x0), _lib.ptr(x1), _lib.ptr(t), x1.shape[0], x1.shape[1], _lib.ptr(x),
                                             _lib.ptr(target), _lib.stream_ptr()), "bfk_fm_interpolate")
        self.launches += 1
        return x, target


class _LossWithGrads(torch.autograd.Function):
    """Makes ``compute_metrics(...)["loss"].backward()`` work like the reference's torch path: the kernels already
    produced every parameter gradient; backward() hands them to autograd."""

    @staticmethod
    def forward(ctx, loss, eng, *params):
        ctx.eng = eng
        ctx.n = len(params)
        ctx.version = eng.version
        return loss.clone()

    @staticmethod
    def backward(ctx, grad_out):
        eng = ctx.eng
        if ctx.version != eng.version:
            raise RuntimeError("bayesflow_b200: another compute_metrics call overwrote this loss' gradients before "
                               "backward() ran; call backward() right after each compute_metrics (the engine keeps ONE "
                               "gradient buffer)")
        return (None, None) + tuple(eng.grad_of(p) * grad_out for p in eng.params)


def train_engine(net) -> TrainEngine:
    eng = getattr(net, "_train_engine", None)
    if eng is None or any(p.data_ptr() != eng.flat.data_ptr() + 4 * a for p, a in zip(eng.params, eng.offsets)):
        eng = TrainEngine(net)
        object.__setattr__(net, "_train_engine", eng)
    return eng


class Adam:
    """keras.optimizers.Adam / AdamW semantics (Keras 3 `adam.py`: m += (g-m)(1-b1); v += (g^2-v)(1-b2);
    p -= lr*sqrt(1-b2^t)/(1-b1^t) * m/(sqrt(v)+eps)); `clipnorm` clips every variable's gradient separately, as Keras
    does; `learning_rate` may be a float or a callable of the (0-based) step (e.g. cosine_decay)."""

    def __init__(self, net, learning_rate=5e-4, beta_1=0.9, beta_2=0.999, epsilon=1e-7, clipnorm=None, weight_decay=0.0):
        self.eng = train_engine(net)
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
        self.clipnorm, self.weight_decay = clipnorm, weight_decay
        self.m = torch.zeros_like(self.eng.flat)
        self.v = torch.zeros_like(self.eng.flat)
        self._sq = torch.zeros(self.eng.n_seg, device=self.eng.device, dtype=torch.float32)
        self.iterations = 0

    def hyper(self):
        """(lr_t, weight_decay * lr) of the NEXT step"""
        t = self.iterations + 1
        lr = self.learning_rate(self.iterations) if callable(self.learning_rate) else self.learning_rate
        return lr * math.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t), float(self.weight_decay) * lr

    def step_dev(self, hyper_dev):
        """the same update with (lr_t, wd*lr) read from `hyper_dev` (2 device floats): capturable in a CUDA graph;
        the caller advances `iterations`"""
        e = self.eng
        _lib.check(e.L.bfk_adam_step_dev(_lib.ptr(e.flat), _lib.ptr(e.grad), _lib.ptr(self.m), _lib.ptr(self.v),
                                         C.c_void_p(e.seg_offsets.data_ptr()), e.n_seg, _lib.ptr(self._sq),
                                         float(self.clipnorm or 0.0), _lib.ptr(hyper_dev), self.beta_1, self.beta_2,
                                         self.epsilon, _lib.stream_ptr()), "bfk_adam_step_dev")
        e.launches += 2 if self.clipnorm else 1

    def step(self):
        e = self.eng
        t = self.iterations + 1
        lr = self.learning_rate(self.iterations) if callable(self.learning_rate) else self.learning_rate
        lr_t = lr * math.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t)
        _lib.check(e.L.bfk_adam_step(_lib.ptr(e.flat), _lib.ptr(e.grad), _lib.ptr(self.m), _lib.ptr(self.v),
                                     C.c_void_p(e.seg_offsets.data_ptr()), e.n_seg, _lib.ptr(self._sq),
                                     float(self.clipnorm or 0.0), lr_t, self.beta_1, self.beta_2, self.epsilon,
                                     float(self.weight_decay) * lr, _lib.stream_ptr()), "bfk_adam_step")
        e.launches += 2 if self.clipnorm else 1
        self.iterations = t
        e.net._engine._key = None  # the kernel wrote the parameters in place: the sampler must re-pack its weight blob


def cosine_decay(initial_learning_rate, decay_steps, alpha=0.0):
    """keras.optimizers.schedules.CosineDecay without warmup (basic_workflow.py:514-520)."""
    def lr(step):
        s = min(step, decay_steps)
        return initial_learning_rate * ((1 - alpha) * 0.5 * (1 + math.cos(math.pi * s / decay_steps)) + alpha)
    return lr


class Trainer:
    """TorchApproximator.train_step for the accelerated network (torch_approximator.py:62-113): zero grads,
    compute_metrics, backward, (all-reduce), optimizer step - without autograd in the loop.

    cuda_graph=True captures the whole step (noise, forward, loss, backward, the NCCL all-reduce, Adam) once per batch
    shape and replays it: ~85 kernel launches become one graph launch.  What changes from step to step (the noise / dropout
    seed, lr_t of the schedule) is read by the kernels from a small device buffer that the host refreshes before each
    replay.  Graph mode covers FlowMatching on tensor input with the default sampling (no optimal transport, no
    sample_weight); anything else takes the eager path."""

    def __init__(self, net, optimizer=None, process_group=None, world_size=None, cuda_graph=False, seed=None):
        self.net, self.eng = net, train_engine(net)
        self.opt = optimizer or Adam(net, 5e-4, clipnorm=1.5)
        self.group = process_group
        if world_size is None:
            world_size = torch.distributed.get_world_size(process_group) if torch.distributed.is_initialized() else 1
        self.world = world_size
        self.rank = torch.distributed.get_rank(process_group) if self.world > 1 else 0
        self.cuda_graph = bool(cuda_graph)
        # one noise seed per step, identical on every rank: rank r draws rows [r*B, (r+1)*B) of the GLOBAL batch
        self._seed0 = int(seed if seed is not None else torch.initial_seed()) & (2 ** 62 - 1)
        self._step_no = 0
        self._graphs = {}

    def _step_seed(self):
        return (self._seed0 * 0x9E3779B1 + 0x632BE59BD9B4E019 * (self._step_no + 1)) & (2 ** 63 - 1)

    def train_step(self, x, conditions=None, sample_weight=None, seed=None):
        """x: this rank's shard of the batch (or the (x0, x1, t, x, target) tuple).  Returns the GLOBAL mean loss as a
        0-d device tensor (after the all-reduce when world_size > 1)."""
        with torch.cuda.device(self.eng.device):
            return self._train_step(x, conditions, sample_weight, seed)

    def _train_step(self, x, conditions, sample_weight, seed):
        if self.cuda_graph and torch.is_tensor(x) and sample_weight is None and seed is None \
                and self.net._graphable_training():
            return self._train_step_graph(x, conditions)
        net, eng = self.net, self.eng
        extra = {}
        if torch.is_tensor(x):
            # global row index of this rank's first row: noise, dropout masks and DiffusionModel's low-discrepancy times
            # (diffusion_model.py:159-161) are functions of the GLOBAL row
            extra["row_offset"] = self.rank * x.shape[0]
            extra["noise_seed"] = self._step_seed()
            if seed is None:
                seed = self._step_seed()  # dropout masks: keyed by the step seed and the GLOBAL row (as in graph replay)
        self._step_no += 1
        loss = net._loss_and_grads(x, conditions, sample_weight, stage="training", seed=seed, world=self.world,
                                   loss_into_slot=True, **extra)
        if self.world > 1:
            # ONE bucket: the gradients and the loss (last element) ride the same all-reduce (SUM); the scaling by
            # 1/B_global happened on device
            torch.distributed.all_reduce(eng.grad_ext, op=torch.distributed.ReduceOp.SUM, group=self.group)
        self.opt.step()
        return loss

    # -- CUDA-graph replay ------------------------------------------------------------------------------------------------
    def _train_step_graph(self, x, conditions):
        net, eng, opt = self.net, self.eng, self.opt
        key = (tuple(x.shape), None if conditions is None else tuple(conditions.shape))
        g = self._graphs.get(key)
        if g is None:
            dev = eng.device
            g = {"x": torch.empty_like(x, dtype=torch.float32),
                 "c": None if conditions is None else torch.empty_like(conditions, dtype=torch.float32),
                 # device words the kernels read: [0:2] seed (lo, hi), [2:4] lr_t, wd*lr; one copy a step from a RING of
                 # pinned host rows - the copy is asynchronous, so a row is rewritten only after the copy that read it has
                 # run (one shared row let a host that runs ahead of the GPU hand many steps the same seed / lr)
                 "params": torch.zeros(4, device=dev, dtype=torch.float32),
                 "host": torch.zeros(self.HOST_RING, 4, dtype=torch.float32).pin_memory(),
                 "host_ev": [None] * self.HOST_RING, "slot": 0}
            g["seed_dev"], g["hyper_dev"] = g["params"][0:2], g["params"][2:4]

            def body():
                loss = net._loss_and_grads(g["x"], g["c"], None, stage="training", world=self.world, loss_into_slot=True,
                                           row_offset=self.rank * x.shape[0], noise_seed=0, seed_dev=g["seed_dev"])
                if self.world > 1:
                    torch.distributed.all_reduce(eng.grad_ext, op=torch.distributed.ReduceOp.SUM, group=self.group)
                opt.step_dev(g["hyper_dev"])
                return loss

            g["x"].copy_(x)
            if conditions is not None:
                g["c"].copy_(conditions)
            self._refresh(g)
            # warm-up on a side stream (workspace allocation, function attributes, NCCL), optimizer state restored after
            saved = (eng.flat.clone(), opt.m.clone(), opt.v.clone())
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(2):
                    body()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            eng.flat.copy_(saved[0])
            opt.m.copy_(saved[1])
            opt.v.copy_(saved[2])
            graph = torch.cuda.CUDAGraph()
            l0 = eng.launches
            with torch.cuda.graph(graph):
                g["loss"] = body()
            g["graph"], g["launches"] = graph, eng.launches - l0  # kernels one replay executes
            eng.launches = l0
            self._graphs[key] = g
        g["x"].copy_(x, non_blocking=True)
        if conditions is not None:
            g["c"].copy_(conditions, non_blocking=True)
        self._refresh(g)
        g["graph"].replay()
        eng.launches += g["launches"]
        self._step_no += 1
        opt.iterations += 1
        net._engine._key = None  # parameters changed in place: the sampler must re-pack its weight blob
        return g["loss"]

    HOST_RING = 32  # steps the host may run ahead of the GPU before `_refresh` waits for a pinned row to be free

    def _refresh(self, g):
        seed = self._step_seed()
        lr_t, wd_lr = self.opt.hyper()
        slot = g["slot"]
        g["slot"] = (slot + 1) % self.HOST_RING
        if g["host_ev"][slot] is not None:
            g["host_ev"][slot].synchronize()  # the copy that last read this row has completed
        row = g["host"][slot]
        words = row.view(torch.int32)
        for i, w in enumerate((seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)):
            words[i] = w - (1 << 32) if w >= (1 << 31) else w
        row[2], row[3] = lr_t, wd_lr
        g["params"].copy_(row, non_blocking=True)
        if g["host_ev"][slot] is None:
            g["host_ev"][slot] = torch.cuda.Event()
        g["host_ev"][slot].record()
