"""Throughput of the §8f rank-4 paths on one B200 (CUDA events, 3 warm-ups, median of 5): ConsistencyModel multistep
sampling at the cfg2 shape, ConsistencyModel training step at the cfg5 batch, log-Sinkhorn pairing at n = 1024 / 4096.
Prints one JSON line per measurement."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bayesflow_b200 as bf  # noqa: E402


def med(fn, iters=5, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[iters // 2]


def main():
    torch.manual_seed(0)
    D, Cd, N, S = 2, 32, 1000, 1000
    cm = bf.ConsistencyModel(1000, subnet_kwargs={"dropout": 0.0})
    cm.build((None, D), (None, Cd))
    cm.cuda()
    cond = torch.randn(N, Cd, device="cuda")
    smp = bf.Sampler()
    for steps in (2, 11):
        ms = med(lambda: smp.sample(cm, S, conditions=cond, seed=1, steps=steps))
        print(json.dumps({"path": "ConsistencyModel Sampler.sample", "workload": f"cfg2 shape 1000x1000 draws, {steps} steps",
                          "ms": round(ms, 3), "draws_per_s": round(N * S / ms * 1e3)}), flush=True)
    B = 65536
    cmt = bf.ConsistencyModel(1000, subnet_kwargs={"dropout": 0.05})
    cmt.build((None, D), (None, Cd))
    cmt.cuda()
    x, c = torch.randn(B, D, device="cuda"), torch.randn(B, Cd, device="cuda")

    def train():
        loss = cmt.compute_metrics(x, conditions=c, stage="training")["loss"]
        loss.backward()
    ms = med(train)
    print(json.dumps({"path": "ConsistencyModel compute_metrics + backward", "workload": f"batch {B}, 256x5 subnet",
                      "ms": round(ms, 3), "samples_per_s": round(B / ms * 1e3)}), flush=True)
    for n in (1024, 4096):
        x0, x1, cc = torch.randn(n, D, device="cuda"), torch.randn(n, D, device="cuda"), torch.randn(n, Cd, device="cuda")
        ms = med(lambda: bf.optimal_transport(x0, x1, conditions=cc, seed=3, **bf.optimal_transport.__globals__["OPTIMAL_TRANSPORT_DEFAULTS"]))
        info = dict(bf.log_sinkhorn_plan.last_info)
        ms_u = med(lambda: bf.optimal_transport(x0, x1, seed=3, method="log_sinkhorn", regularization=0.1, max_steps=100))
        info_u = dict(bf.log_sinkhorn_plan.last_info)
        print(json.dumps({"path": "optimal_transport (FlowMatching defaults, conditional)", "n": n, "ms": round(ms, 3),
                          "sinkhorn_steps": info["steps"], "w": info["w"], "unconditional_ms": round(ms_u, 3),
                          "unconditional_steps": info_u["steps"],
                          "matrix_sweep_gbs": round(info_u["steps"] * 2 * n * n * 4 / (ms_u * 1e-3) / 1e9, 1)}), flush=True)
    fm = bf.FlowMatching(use_optimal_transport=True, subnet_kwargs={"dropout": 0.05})
    fm.build((None, D), (None, Cd))
    fm.cuda()
    xb, cb = torch.randn(4096, D, device="cuda"), torch.randn(4096, Cd, device="cuda")

    def train_ot():
        fm.compute_metrics(xb, conditions=cb, stage="training")["loss"].backward()
    ms = med(train_ot)
    fm.use_optimal_transport = False
    ms0 = med(train_ot)
    print(json.dumps({"path": "FlowMatching compute_metrics + backward, batch 4096", "with_ot_ms": round(ms, 3),
                      "without_ot_ms": round(ms0, 3)}), flush=True)


if __name__ == "__main__":
    main()
