"""ctypes binding of libbfk.so (C ABI in include/bfk.h).  There is NO fallback: if the library is missing or a call
fails, the caller gets an exception."""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbfk.so")

BFK_MAX_BLOCKS = 16
ACT = {"mish": 0, "silu": 1, "swish": 1, "relu": 2, "gelu": 3, "tanh": 4}
MERGE = {None: 0, "concat": 1, "add": 2}
NORM = {None: 0, "layer": 1, "rms": 2}
PREC = {"fp16": 0, "bf16": 1}
METHOD = {"euler": 0, "rk45": 1, "tsit5": 2}
PRED = {"velocity": 0, "noise": 1, "F": 2, "x": 3, "score": 4}
SCHED = {"edm": 0, "cosine": 1}
# include/bfk.h BFK_CTRL_*
CTRL = {"time": 0, "step_size": 1, "h": 2, "step": 3, "rejected": 4, "sel": 11, "done": 12, "all_nan": 14, "trials": 15,
        "acc": 20, "times": 24, "floats": 32}

_fp = C.POINTER(C.c_float)


class NetConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("D", "C", "W", "L", "T", "activation", "merge", "norm", "residual", "film_use_gamma", "precision")] + \
               [("fourier_scale", C.c_float), ("block_width", C.c_int32 * BFK_MAX_BLOCKS)]


class Weights(C.Structure):
    _fields_ = [("freq", C.c_void_p), ("x_k", C.c_void_p), ("x_b", C.c_void_p), ("c_k", C.c_void_p), ("c_b", C.c_void_p),
                ("m_k", C.c_void_p), ("m_b", C.c_void_p)] + \
               [(n, C.c_void_p * BFK_MAX_BLOCKS) for n in
                ("dense_k", "dense_b", "film_g_k", "film_g_b", "film_b_k", "film_b_b", "norm_g", "norm_b")] + \
               [("out_k", C.c_void_p), ("out_b", C.c_void_p)] + \
               [(n, C.c_void_p * BFK_MAX_BLOCKS) for n in ("proj_k", "proj_b")]


class Schedule(C.Structure):
    _fields_ = [("model", C.c_int32), ("prediction", C.c_int32), ("schedule", C.c_int32), ("exploding", C.c_int32)] + \
               [(n, C.c_double) for n in ("sigma_data", "sigma_min", "sigma_max", "rho", "log_snr_min", "log_snr_max",
                                          "shift", "t_min", "t_max")]


_lib = None


def lib():
    """Load libbfk.so or raise.  Never substitutes another implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"bayesflow_b200: {LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU or eager fallback.")
    L = C.CDLL(LIB_PATH)
    L.bfk_version.restype = C.c_int
    L.bfk_last_error_string.restype = C.c_char_p
    L.bfk_packed_bytes.restype = C.c_size_t
    L.bfk_packed_bytes.argtypes = [C.POINTER(NetConfig)]
    L.bfk_table_stride.restype = C.c_size_t
    L.bfk_table_stride.argtypes = [C.POINTER(NetConfig)]
    L.bfk_fused_supported.argtypes = [C.POINTER(NetConfig)]
    L.bfk_pack.argtypes = [C.POINTER(NetConfig), C.POINTER(Weights), C.c_void_p, C.c_void_p]
    L.bfk_cond_prologue.argtypes = [C.POINTER(NetConfig), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                    C.c_void_p]
    L.bfk_eval_table.argtypes = [C.POINTER(NetConfig), C.c_void_p, C.POINTER(Schedule), C.c_void_p, C.c_int64,
                                 C.c_void_p, C.c_void_p]
    L.bfk_velocity.argtypes = [C.POINTER(NetConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.bfk_sample_fixed.argtypes = [C.POINTER(NetConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                   C.c_void_p]
    L.bfk_adaptive_init.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int64, C.c_int64, C.c_double, C.c_double,
                                    C.c_void_p]
    L.bfk_adaptive_plan.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.bfk_adaptive_plan_final.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.bfk_adaptive_commit.argtypes = [C.c_void_p, C.c_void_p]
    L.bfk_adaptive_trial.argtypes = [C.POINTER(NetConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                     C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.bfk_packed_operand_bytes.restype = C.c_size_t
    L.bfk_packed_operand_bytes.argtypes = [C.c_int64, C.c_int64, C.c_int32]
    L.bfk_pack_kmajor.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                  C.c_void_p]
    L.bfk_pack_kmajor_t.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                    C.c_void_p, C.c_void_p]
    L.bfk_gemm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                           C.c_int32, C.c_int32, C.c_void_p, C.c_float, C.c_int32, C.c_void_p]
    L.bfk_detect_repeats.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    L.bfk_normal_fill.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_uint64,
                                  C.c_void_p, C.c_void_p]
    L.bfk_schedule_alpha_sigma.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.bfk_sde_langevin_update.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_double,
                                          C.c_double, C.c_double, C.c_int32, C.c_void_p, C.c_void_p]
    L.bfk_pad2d.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    L.bfk_unpad2d.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    L.bfk_standardize.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_int32, C.c_void_p]
    L.bfk_lincomb3.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_float, C.c_void_p, C.c_float, C.c_void_p, C.c_int64,
                               C.c_void_p]
    L.bfk_moments_workspace_floats.restype = C.c_size_t
    L.bfk_moments_workspace_floats.argtypes = [C.c_int64, C.c_int32]
    L.bfk_batch_moments.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    L.bfk_batch_moments_static.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    L.bfk_static_scratch_floats.restype = C.c_size_t
    L.bfk_static_scratch_floats.argtypes = []
    L.bfk_moments_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    L.bfk_ot_cost.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    L.bfk_ot_ratio.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]
    L.bfk_ot_scale.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int64, C.c_void_p]
    L.bfk_ot_work_floats.restype = C.c_size_t
    L.bfk_ot_work_floats.argtypes = [C.c_int64, C.c_int64]
    L.bfk_ot_init.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_float, C.c_int64, C.c_void_p]
    L.bfk_ot_log_sinkhorn_step.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
    L.bfk_ot_plain_mode.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.bfk_ot_sample_plan.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                     C.c_uint64, C.c_void_p, C.c_void_p]
    L.bfk_ot_sample.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_uint64, C.c_void_p, C.c_void_p]
    _lib = L
    return L


def check(rc: int, what: str):
    if rc == 0:
        return
    msg = lib().bfk_last_error_string().decode()
    if rc < 0:
        raise ValueError(f"{what}: {msg} (code {rc})")
    raise RuntimeError(f"{what}: CUDA error {rc}: {msg}")


def ptr(t: torch.Tensor | None):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("bayesflow_b200 runs on CUDA tensors only (no CPU fallback); got a tensor on " + str(t.device))
    if t.dtype != torch.float32 and t.dtype != torch.uint8:
        raise TypeError(f"expected float32, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError("tensor must be contiguous")
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
