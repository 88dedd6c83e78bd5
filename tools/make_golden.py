"""Writes tests/golden/*.pt: outputs of the REFERENCE's own classes (loaded by path, unmodified, on the keras stand-in
— tests/_keras_stub/) on fixed seeded inputs.  Run in the build container (the only place /root/reference exists):

    python tools/make_golden.py

The `-m gpu` tests (tests/test_gpu_golden.py) compare the CUDA path with these files on the GPU box, where the
reference does not exist; tests/test_golden_oracle.py compares oracle/ with them anywhere."""
import hashlib
import logging
import os
import re
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "_keras_stub"))
import ref_loader  # noqa: E402

import oracle  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
_PKEYS = ("merge", "film_use_gamma", "norm", "residual", "time_embedding_dim")


def params_digest(params):
    h = hashlib.sha256()
    for k, v in params.items():
        h.update(k.encode())
        h.update(v.contiguous().numpy().tobytes())
    return h.hexdigest()


def build(R, D, C, widths, dm=False, model_kwargs=None, **sk):
    params = oracle.perturb_params(oracle.init_params(D, C, widths, **{k: v for k, v in sk.items() if k in _PKEYS}))
    kw = {"dropout": 0.0, "widths": widths} | sk
    cls = R.DiffusionModel if dm else R.FlowMatching
    ref = cls(subnet_kwargs=kw, **(model_kwargs or {}))
    ref.build((None, D), (None, C) if C else None)
    ref_loader.load_params(ref, params)
    return ref, params


def adaptive_with_counts(fn):
    found = {}

    class H(logging.Handler):
        def emit(self, record):
            m = re.search(r"after ([0-9.]+) steps with ([0-9.]+) rejected", record.getMessage())
            if m:
                found["steps"], found["rejected"] = float(m.group(1)), float(m.group(2))
    lg = logging.getLogger("bayesflow")
    h = H(level=logging.DEBUG)
    old = lg.level
    lg.addHandler(h)
    lg.setLevel(logging.DEBUG)
    try:
        out = fn()
    finally:
        lg.removeHandler(h)
        lg.setLevel(old)
    return out, found


def grads_by_name(ref, params):
    alias = {"norm": "norm_layer", "frequencies": "w"}
    out = {}
    for name in params:
        obj = ref
        parts = name.split(".")
        for part in parts[:-1]:
            part = alias.get(part, part)
            obj = obj[int(part)] if part.isdigit() else getattr(obj, part)
        out[name] = getattr(obj, alias.get(parts[-1], parts[-1])).grad.clone()
    return out


def fm_case(R, name, D, C, widths, B, rows_per_group, store_params, **sk):
    ref, params = build(R, D, C, widths, **sk)
    g = torch.Generator().manual_seed(20260105)
    z = torch.randn(B, D, generator=g)
    G = (B + rows_per_group - 1) // rows_per_group
    c = torch.randn(G, C, generator=g).repeat_interleave(rows_per_group, dim=0)[:B].contiguous() if C else None
    out = {"kind": "fm", "D": D, "C": C, "widths": widths, "subnet_kwargs": sk, "rows_per_group": rows_per_group,
           "z": z, "conditions": c, "params_sha256": params_digest(params)}
    if store_params:
        out["params"] = params
    with torch.no_grad():
        for t in (0.0, 0.37, 1.0):
            out[f"velocity_t{t}"] = ref.velocity(z, t, conditions=c)
        out["euler20"] = ref(z, conditions=c, inverse=True, method="euler", steps=20)
        out["rk45_5"] = ref(z, conditions=c, inverse=True, method="rk45", steps=5)
        out["tsit5_5"] = ref(z, conditions=c, inverse=True, method="tsit5", steps=5)
        out["forward_euler10"] = ref(z, conditions=c, inverse=False, method="euler", steps=10)
        x, cnt = adaptive_with_counts(lambda: ref(z, conditions=c, inverse=True, method="tsit5", steps="adaptive"))
        out["tsit5_adaptive"], out["tsit5_adaptive_counts"] = x, cnt
        x, cnt = adaptive_with_counts(lambda: ref(z, conditions=c, inverse=True, method="rk45", steps="adaptive",
                                                  min_steps=3, max_steps=40))
        out["rk45_adaptive_3_40"], out["rk45_adaptive_3_40_counts"] = x, cnt
    # training: tuple input (flow_matching.py:140-141), loss + gradients from torch autograd on the reference graph
    x1, x0, t = torch.randn(B, D, generator=g), torch.randn(B, D, generator=g), torch.rand(B, 1, generator=g)
    tup = (x0, x1, t, t * x1 + (1 - t) * x0, x1 - x0)
    for w in ref.subnet.weights + ref.output_projector.weights:
        w.grad = None
    loss = ref.compute_metrics(tup, conditions=c, stage="training")["loss"]
    loss.backward()
    out["train_x0"], out["train_x1"], out["train_t"] = x0, x1, t
    out["train_loss"] = loss.detach()
    gr = grads_by_name(ref, params)
    out["train_grads"] = gr if store_params else {k: v for k, v in gr.items() if v.numel() <= 512}
    out["train_grad_norms"] = {k: float(v.norm()) for k, v in gr.items()}
    if D <= 4 and store_params:
        xs = z[:16]
        cs = c[:16] if C else None
        out["log_prob_euler8"] = ref.log_prob(xs, conditions=cs, method="euler", steps=8).detach()
    torch.save(out, os.path.join(OUT, name + ".pt"))
    print(name, "ok", {k: tuple(v.shape) for k, v in out.items() if isinstance(v, torch.Tensor)})


def dm_case(R, name, D, C, widths, B, rows_per_group, schedule, ptype):
    mk = {"noise_schedule": schedule, "prediction_type": ptype}
    ref, params = build(R, D, C, widths, dm=True, model_kwargs=mk)
    g = torch.Generator().manual_seed(20260106)
    z = torch.randn(B, D, generator=g)
    G = (B + rows_per_group - 1) // rows_per_group
    c = torch.randn(G, C, generator=g).repeat_interleave(rows_per_group, dim=0)[:B].contiguous()
    out = {"kind": "dm", "D": D, "C": C, "widths": widths, "model_kwargs": mk, "rows_per_group": rows_per_group,
           "z": z, "conditions": c, "params": params, "params_sha256": params_digest(params)}
    with torch.no_grad():
        for t in (0.9, 0.5, 0.05):
            out[f"velocity_t{t}"] = ref.velocity(z, time=t, stochastic_solver=False, conditions=c)
        out["euler50"] = ref(z, conditions=c, inverse=True, method="euler", steps=50)
        out["rk45_10"] = ref(z, conditions=c, inverse=True, method="rk45", steps=10)
    import keras
    x = torch.randn(B, D, generator=g)
    u0, eps = torch.rand((1,), generator=g), torch.randn(B, D, generator=g)
    old_u, old_n = keras.random.uniform, keras.random.normal
    keras.random.uniform = lambda shape, *a, **k: u0.reshape(tuple(shape)).clone()
    keras.random.normal = lambda shape, *a, **k: eps.reshape(tuple(shape)).clone()
    try:
        for w in ref.subnet.weights + ref.output_projector.weights:
            w.grad = None
        loss = ref.compute_metrics(x, conditions=c, stage="training")["loss"]
        loss.backward()
    finally:
        keras.random.uniform, keras.random.normal = old_u, old_n
    out["train_x"], out["train_u0"], out["train_eps"] = x, u0, eps
    out["train_loss"] = loss.detach()
    out["train_grads"] = grads_by_name(ref, params)
    torch.save(out, os.path.join(OUT, name + ".pt"))
    print(name, "ok")


def main():
    os.makedirs(OUT, exist_ok=True)
    R = ref_loader.reference()
    torch.manual_seed(0)
    # the default (cfg2) network: weights are NOT stored (2 MB) — rebuilt from oracle.init_params(seed), digest checked
    fm_case(R, "fm_cfg2_default", 2, 32, (256,) * 5, 300, 100, store_params=False)
    fm_case(R, "fm_small", 2, 6, (64,) * 3, 200, 50, store_params=True)
    fm_case(R, "fm_small_silu_add", 3, 7, (64,) * 2, 130, 13, store_params=True, activation="silu", merge="add")
    fm_case(R, "fm_small_projector", 2, 6, (64, 96, 32), 150, 50, store_params=True)
    dm_case(R, "dm_small_edm_F", 4, 20, (64,) * 3, 200, 50, "edm", "F")
    dm_case(R, "dm_small_cosine_velocity", 4, 20, (64,) * 3, 200, 50, "cosine", "velocity")


if __name__ == "__main__":
    main()
