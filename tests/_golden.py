"""Loader for tests/golden/*.pt — outputs of the reference's OWN classes (tools/make_golden.py, run in the build
container on the keras stand-in).  The default-network case stores no weights: they are rebuilt from
oracle.init_params and checked against the stored SHA-256."""
import hashlib
import os

import torch

import oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FM_CASES = ["fm_cfg2_default", "fm_small", "fm_small_silu_add", "fm_small_projector"]
DM_CASES = ["dm_small_edm_F", "dm_small_cosine_velocity"]
_PKEYS = ("merge", "film_use_gamma", "norm", "residual", "time_embedding_dim")


def params_digest(params):
    h = hashlib.sha256()
    for k, v in params.items():
        h.update(k.encode())
        h.update(v.contiguous().numpy().tobytes())
    return h.hexdigest()


def load(name):
    g = torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    if "params" not in g:
        sk = g.get("subnet_kwargs", {})
        p = oracle.perturb_params(oracle.init_params(g["D"], g["C"], g["widths"], **{k: v for k, v in sk.items() if k in _PKEYS}))
        g["params"] = p
    g["params_ok"] = params_digest(g["params"]) == g["params_sha256"]
    return g


def subnet_kwargs(g):
    return {"dropout": 0.0, "widths": tuple(g["widths"])} | g.get("subnet_kwargs", {})
