#!/usr/bin/env python
"""bench.py — posterior draws/sec of the FlowMatching / DiffusionModel ODE sampling hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]

One "step" = one pass of the hot path over one batch: N datasets x S draws pushed through the whole ODE loop
(`InferenceNetwork._inverse` -> `integrate`).  Default workload = BASELINE.json configs[1]: SIR FlowMatching, fixed
32-dim summary conditions, 1k datasets x 1k draws, Euler 100 steps.  Under torchrun every rank runs the same per-GPU
workload on its own datasets (weak scaling, no data-path collective: rows are independent).

Keys of the JSON line: see DESIGN.md "Measurement".  `--impl reference` times the reference's CPU algorithm (the oracle
port; the reference itself needs Keras, which is absent here and on the GPU box) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# name -> model, D, C, widths, datasets N, draws S, method, ODE steps
WORKLOADS = {
    "cfg1_twomoons_fm_euler100": dict(model="fm", D=2, C=2, widths=(256,) * 5, N=512, S=1000, method="euler", ode_steps=100),
    "cfg2_sir_fm_euler100": dict(model="fm", D=2, C=32, widths=(256,) * 5, N=1000, S=1000, method="euler", ode_steps=100),
    "cfg2_sir_fm_rk45x100": dict(model="fm", D=2, C=32, widths=(256,) * 5, N=1000, S=1000, method="rk45", ode_steps=100),
    "cfg3_lv_dm_euler100": dict(model="dm", D=4, C=20, widths=(256,) * 5, N=1000, S=4000, method="euler", ode_steps=100),
    "cfg3_lv_dm_rk45x100": dict(model="dm", D=4, C=20, widths=(256,) * 5, N=1000, S=4000, method="rk45", ode_steps=100),
    "cfg4_wide_fm_euler100": dict(model="fm", D=64, C=32, widths=(1024,) * 8, N=1000, S=1000, method="euler", ode_steps=100),
}
# training workloads (BASELINE.json configs[4]): GLOBAL batch, split evenly over the ranks (strong scaling)
TRAIN_WORKLOADS = {
    "cfg5_fm_train": dict(model="fm", D=2, C=32, widths=(256,) * 5, batch=65536, dropout=0.05),
    "cfg5_fm_train_wide": dict(model="fm", D=64, C=32, widths=(1024,) * 8, batch=65536, dropout=0.05),
}
DEFAULT_WORKLOAD = "cfg2_sir_fm_euler100"
EVALS_PER_STEP = {"euler": 1, "rk45": 6, "tsit5": 6}


def workload_config(name, w, n_gpus, scaling="strong"):
    cfg = {"workload": name, "network": "FlowMatching" if w["model"] == "fm" else "DiffusionModel(edm,F)",
           "subnet": f"time_mlp widths={w['widths'][0]}x{len(w['widths'])}", "D": w["D"], "C": w["C"],
           "draws_per_dataset": w["S"], "solver": f"{w['method']} fixed {w['ode_steps']} steps",
           "l2": "state+conditions stream once per step; weights are L2-resident by design; 256 MB L2 flush between steps"}
    if scaling == "strong":
        cfg["datasets_total"] = w["N"]
        cfg["sharding"] = (f"the {w['N']} datasets of the stated problem split over {n_gpus} GPU(s) by "
                           "bayesflow_b200.parallel.sample_sharded (dataset_shard), no data-path collective")
    else:
        cfg["datasets_per_gpu"] = w["N"]
        cfg["sharding"] = f"{w['N']} datasets per GPU on {n_gpus} GPU(s), no collective"
    return cfg


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def count_flops_per_eval(D, C, widths, T=33):
    """2*in*out per dense layer exactly as the reference executes them (SURVEY.md §8d formula) — the NAIVE count: the
    per-dataset conditions GEMMs and the per-row FiLM GEMMs are included although the kernel hoists them."""
    W0 = widths[0]
    f = 2 * D * W0
    if C:
        f += 2 * C * W0 + 2 * (2 * W0) * W0
    prev = W0
    for w in widths:
        f += 2 * prev * w + 2 * T * w + (2 * prev * w if prev != w else 0)
        prev = w
    return f + 2 * prev * D


def flops_per_draw(w):
    return count_flops_per_eval(w["D"], w["C"], w["widths"]) * EVALS_PER_STEP[w["method"]] * w["ode_steps"]


def measured_peaks():
    """MEASURED_PEAKS.json (driver-written) else the fallback B200_PROFILING.md states."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"tflops": float(p.get("bf16_tflops_sustained") or p["bf16_tflops"]), "hbm_gbs": float(p["hbm_gbs"]),
                "source": "MEASURED_PEAKS.json bf16_tflops_sustained (fp16 operands run at the bf16 rate)"}
    return {"tflops": 1366.1, "hbm_gbs": 6548.8, "source": "fallback: B200_PROFILING.md sustained bf16 figure"}


def ncu_traffic(name):
    """dram bytes per launch of the sampler kernel from the committed `ncu --set full` capture (profiles/), or None."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(path):
        return json.load(open(path)).get(name)
    return None


def rows_total_fn(world, N, S, steps):
    return world * N * S * steps


def measure_edge(dev, peak_gbs):
    """HBM roofline of the HBM-bound edge kernels (bfk_standardize, bfk_batch_moments, bfk_normal_fill) on 1 GiB inputs
    (larger than L2), CUDA events on the launching stream, median of 10 after 3 warm-ups."""
    import ctypes as C
    import torch
    import bayesflow_b200 as bf
    from bayesflow_b200 import _lib
    F, rows = 32, 8 * 1024 * 1024
    x = torch.randn(rows, F, device=dev)
    out = torch.empty_like(x)
    st = bf.Standardize()
    st.build((None, F))
    st.to(dev)
    st(x[:4096], stage="training")
    nbytes = x.numel() * 4

    def med(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
        for a, b in ev:
            a.record()
            fn()
            b.record()
        torch.cuda.synchronize()
        return sorted(a.elapsed_time(b) for a, b in ev)[5] * 1e-3

    def fill():
        _lib.check(_lib.lib().bfk_normal_fill(_lib.ptr(out), rows, F, 0, None, None, C.c_uint64(7), None, _lib.stream_ptr()),
                   "bfk_normal_fill")
    res = {}
    for name, fn, traffic in (("standardize_forward", lambda: st._location_scale(x, 0, True, out=out), 2 * nbytes),
                              ("standardize_inverse_inplace", lambda: st._location_scale(out, 0, False, out=out), 2 * nbytes),
                              ("batch_moments", lambda: st.batch_moments(x, 0), nbytes),
                              ("normal_fill", fill, nbytes)):
        t = med(fn)
        res[name] = {"bound": "hbm", "achieved": traffic / t / 1e9, "peak": peak_gbs, "unit": "GB/s",
                     "frac": traffic / t / 1e9 / peak_gbs, "algorithmic_bytes": traffic, "ms": t * 1e3}
    res["normal_fill"]["note"] = "issue-bound (Philox4x32-10 + Box-Muller: ~30 instructions per 4 bytes), not HBM-bound"
    return res


# --------------------------------------------------------------------------------------------------- CPU arm
def make_oracle(w, state_dict):
    import oracle
    kw = {"dropout": 0.0, "widths": w["widths"]}
    cls = oracle.OracleFlowMatching if w["model"] == "fm" else oracle.OracleDiffusionModel
    return cls(state_dict, w["D"], subnet_kwargs=kw)


def cpu_rate(w, orc, n_datasets, threads):
    """draws/s of the reference's CPU algorithm (oracle port: same ATen ops the Keras/torch backend issues, fp32,
    autograd off as bayesflow/__init__.py:80) on `n_datasets` datasets of the SAME workload."""
    import torch
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(3456)
    S = w["S"]
    z = torch.randn(n_datasets * S, w["D"], generator=g)
    c = torch.randn(n_datasets, w["C"], generator=g).repeat_interleave(S, dim=0) if w["C"] else None  # samplers.py:51-63
    t0 = time.perf_counter()
    with torch.no_grad():
        out = orc.inverse(z, c, method=w["method"], steps=w["ode_steps"])
    dt = time.perf_counter() - t0
    assert bool(torch.isfinite(out).all())
    return n_datasets * S / dt, dt


def cpu_state_dict(w):
    import oracle
    return oracle.init_params(w["D"], w["C"] or None, w["widths"])


CPU_SAMPLE_DATASETS = 16   # the SAME bounded sample in `cpu_baseline` and in `--impl reference`: the port's rate depends on
                            # the working set (activations of n*S rows x 512 columns: 16 datasets = 33 MB per op), so the
                            # sample is fixed, not calibrated to a time budget


def cpu_sample_datasets(w):
    # cfg3 (4000 draws per dataset) and cfg4 (wide net) are ~4x / ~20x heavier per dataset
    per_ds = w["S"] * count_flops_per_eval(w["D"], w["C"], w["widths"])
    return max(1, min(CPU_SAMPLE_DATASETS, int(CPU_SAMPLE_DATASETS * 1000 * 1020416 / per_ds)))


def run_cpu_baseline(w, state_dict, budget_s=15.0):
    orc = make_oracle(w, state_dict)
    threads = os.cpu_count() or 1
    cpu_rate(w, orc, 1, threads)  # warm-up: one dataset
    n = cpu_sample_datasets(w)
    rate, dt = cpu_rate(w, orc, n, threads)
    return {"value": rate, "unit": "draws/s", "cores": threads, "kind": "port",
            "sample": f"{n} dataset(s) x {w['S']} draws of {w['method']}-{w['ode_steps']} ({dt:.1f} s); torch fp32 eager, "
                      f"torch.set_num_threads({threads})"}


def run_reference_arm(args, name, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sd = cpu_state_dict(w)
    orc = make_oracle(w, sd)
    threads = os.cpu_count() or 1
    n = cpu_sample_datasets(w)  # the same bounded sample as the `cpu_baseline` leg of the GPU arm
    for _ in range(max(args.warmup, 1)):
        cpu_rate(w, orc, 1, threads)
    tot_rows, tot_t = 0, 0.0
    for _ in range(args.steps):
        r, dt = cpu_rate(w, orc, n, threads)
        tot_rows += n * w["S"]
        tot_t += dt
    value = tot_rows / tot_t
    line = {"impl": "reference", "metric": "posterior draws/sec", "value": value, "unit": "draws/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(name, w, args.gpus, args.scaling),
            "cpu_baseline": {"value": value, "unit": "draws/s", "cores": threads, "kind": "port",
                             "sample": f"each step = {n} dataset(s) x {w['S']} draws of the workload on the host cores; the "
                                       "reference needs Keras (absent), so this is the oracle port of its torch-CPU path"},
            "e2e": {"value": value, "unit": "draws/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------- GPU arm
def run_ours(args, name, w):
    import torch
    import torch.distributed as dist
    import bayesflow_b200 as bf
    from bayesflow_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()  # fail loudly if libbfk.so is missing

    # ---- network: random init of the named architecture (seed 1234), identical on every rank
    torch.manual_seed(1234)
    kw = {"dropout": 0.0, "widths": w["widths"]}
    net = (bf.FlowMatching if w["model"] == "fm" else bf.DiffusionModel)(subnet_kwargs=kw)
    net.build((None, w["D"]), (None, w["C"]) if w["C"] else None)
    state_cpu = {k: v.detach().clone() for k, v in net.state_dict().items() if "base_distribution" not in k}
    net.to(dev)

    # ---- synthetic inputs
    N, S, D, Cd = w["N"], w["S"], w["D"], w["C"]
    strong = args.scaling == "strong"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    ikw = dict(method=w["method"], steps=w["ode_steps"])
    if strong:
        # the STATED problem (N datasets x S draws) split over the ranks by dataset: every rank holds the (N, C) conditions
        # (identical on all ranks) and draws its datasets' rows through bayesflow_b200.parallel.sample_sharded, i.e.
        # InferenceNetwork.sample: z is generated in HBM (Philox keyed by the GLOBAL row) inside the timed call
        from bayesflow_b200.parallel import dataset_shard, sample_sharded
        lo, hi = dataset_shard(N, rank, world)
        n_local = hi - lo
        g = torch.Generator().manual_seed(2345)
        cond_host = torch.randn(N, Cd, generator=g).pin_memory() if Cd else None
        cond_dev = cond_host.to(dev) if Cd else None
        out_host = torch.empty(max(n_local, 1) * S, D).pin_memory()
        z_host = None
        if not Cd:
            raise SystemExit("bench.py: --scaling strong shards by dataset and needs a conditional workload")

        def step_resident():
            return sample_sharded(net, S, cond_dev, seed=1234, global_step_control=False, **ikw)

        def step_e2e():
            c = cond_host.to(dev, non_blocking=True)
            out = sample_sharded(net, S, c, seed=1234, global_step_control=False, **ikw)
            out_host[: n_local * S].copy_(out.reshape(-1, D), non_blocking=True)
            return out
        h2d_bytes = int(cond_host.numel() * 4)
        d2h_bytes = int(n_local * S * D * 4)
    else:
        # weak scaling: N datasets per GPU, fixed z fed through InferenceNetwork.call(inverse=True)
        n_local = N
        g = torch.Generator().manual_seed(2345 + rank)
        cond_host = torch.randn(N, Cd, generator=g).pin_memory() if Cd else None
        z_host = torch.randn(N * S, D, generator=g).pin_memory()
        out_host = torch.empty(N * S, D).pin_memory()
        cond_dev = cond_host.to(dev) if Cd else None
        z_dev = z_host.to(dev)

        def step_resident():
            # the reference-facing call (InferenceNetwork.call(inverse=True)); conditions stay (N, C) + repeat count instead
            # of the (N*S, C) tensor samplers.py:51-63 materialises
            return net(z_dev, conditions=cond_dev, conditions_repeats=S if Cd else None, inverse=True, **ikw)

        def step_e2e():
            c = cond_host.to(dev, non_blocking=True) if Cd else None
            z = z_host.to(dev, non_blocking=True)
            out = net(z, conditions=c, conditions_repeats=S if Cd else None, inverse=True, **ikw)
            out_host.copy_(out, non_blocking=True)
            return out
        h2d_bytes = int(z_host.numel() * 4 + (cond_host.numel() * 4 if Cd else 0))
        d2h_bytes = int(out_host.numel() * 4)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        eng = net._engine
        barrier()
        eng.kernel_events = []
        launches0 = eng.launches
        pairs = []
        t_wall = time.perf_counter()
        for _ in range(steps):
            flush.fill_(1)  # L2 flush between timed iterations (outside the event pair)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            pairs.append((a, b))
        barrier()
        wall = time.perf_counter() - t_wall
        total_ms = sum(a.elapsed_time(b) for a, b in pairs)
        kern_ms = [a.elapsed_time(b) for a, b in eng.kernel_events]
        launches = eng.launches - launches0
        eng.kernel_events = None
        if world > 1:
            t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        return total_ms, kern_ms, launches, wall

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    total_ms, kern_ms, launches, wall = timed(step_resident, args.steps, args.warmup)
    clk = clocks.stop() if rank == 0 else None
    e2e_ms, _, _, _ = timed(step_e2e, args.steps, max(1, args.warmup // 2))
    info = dict(net.last_integration_info)

    # the approximator-shaped call: Sampler.sample with HOST (N, C) conditions, z drawn in HBM by
    # bfk_normal_fill, draws mapped back to the data scale by bfk_standardize, result copied to pinned host memory
    sampler_api = None
    if Cd and world == 1 and not args.no_train:
        try:
            import bayesflow_b200 as bf
            stdz = bf.Standardize()
            stdz.build((None, D))
            stdz.to(dev)
            stdz(torch.randn(4096, D, device=dev) * 3 + 1, stage="training")
            smp = bf.Sampler()
            out3_host = torch.empty(N, S, D).pin_memory()

            def step_sampler():
                c = cond_host.to(dev, non_blocking=True)
                out = smp.sample(net, S, conditions=c, seed=1234, standardize=stdz, **ikw)
                out3_host.copy_(out, non_blocking=True)
                return out
            s_ms, _, s_launches, _ = timed(step_sampler, args.steps, max(1, args.warmup // 2))
            sampler_api = {"call": "Sampler.sample(net, num_samples, conditions=(N, C) host) -> (N, S, D) host",
                           "value": world * N * S * args.steps / (s_ms * 1e-3), "unit": "draws/s",
                           "ms_per_step": s_ms / args.steps, "h2d_bytes_per_step": int(cond_host.numel() * 4),
                           "d2h_bytes_per_step": int(out3_host.numel() * 4), "gpu_launches": s_launches + 2 * args.steps}
        except Exception as exc:  # the headline line must not depend on the secondary one
            sampler_api = {"unavailable": repr(exc)[:200]}

    rows_total = (N if strong else world * N) * S * args.steps
    value = rows_total / (total_ms * 1e-3)
    e2e_value = rows_total / (e2e_ms * 1e-3)
    fpd = flops_per_draw(w)
    peaks = measured_peaks()
    k_avg_ms = sum(kern_ms) / max(len(kern_ms), 1)
    achieved = (n_local * S * fpd) / (k_avg_ms * 1e-3) / 1e12 if kern_ms else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = run_cpu_baseline(w, state_cpu)
    train = None
    if not args.no_train:
        # second half of BASELINE.json's metric: FlowMatching train samples/sec (configs[4]), same run, same ranks
        del flush
        torch.cuda.empty_cache()
        tw = TRAIN_WORKLOADS["cfg5_fm_train"]
        train = measure_training(tw, dev, world, rank, max(args.steps, 5), 3, False)
        train["config"] = train_config("cfg5_fm_train", tw, world)
        train["metric"] = "FlowMatching train samples/sec"
    is_layerwise = bool(net._engine.layerwise)
    wide = None
    train_wide = None
    if not args.no_train and world > 1 and name == DEFAULT_WORKLOAD:
        # configs[3] and the wide training net at every N: same process group, same timing rules
        del net
        torch.cuda.empty_cache()
        try:
            wide = measure_sampling_simple(WORKLOADS["cfg4_wide_fm_euler100"], "cfg4_wide_fm_euler100", dev, world, rank, args.scaling)
        except Exception as exc:
            wide = {"unavailable": repr(exc)[:200]}
        try:
            torch.cuda.empty_cache()
            tww = TRAIN_WORKLOADS["cfg5_fm_train_wide"]
            train_wide = measure_training(tww, dev, world, rank, 5, 3, False)
            train_wide["config"] = train_config("cfg5_fm_train_wide", tww, world)
        except Exception as exc:
            train_wide = {"unavailable": repr(exc)[:200]}
    if not args.no_train and world == 1 and name == DEFAULT_WORKLOAD:
        # configs[3] (wide subnet, 1M draws) measured in the same run as a secondary line: a fresh process of this script
        # with the same timing rules (3 warm-ups, L2 flush, device events), 2 timed steps of ~2.6 s
        del net
        torch.cuda.empty_cache()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--workload", "cfg4_wide_fm_euler100", "--steps", "2",
                                "--warmup", "3", "--no-train", "--no-cpu-baseline"], capture_output=True, text=True, timeout=600)
            wide = json.loads(r.stdout.strip().splitlines()[-1])
            wide = {k: wide[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "warmup", "config", "e2e",
                                         "gpu_launches", "roofline", "clocks")}
        except Exception as exc:  # the headline line must not depend on the secondary one
            wide = {"unavailable": repr(exc)[:200]}
        try:  # the wide training net at N = 1 too, so that the scaling of cfg5_fm_train_wide has its base point
            torch.cuda.empty_cache()
            tww = TRAIN_WORKLOADS["cfg5_fm_train_wide"]
            train_wide = measure_training(tww, dev, world, rank, 5, 3, False)
            train_wide["config"] = train_config("cfg5_fm_train_wide", tww, world)
        except Exception as exc:
            train_wide = {"unavailable": repr(exc)[:200]}
    edge = None
    if rank == 0 and world == 1 and not args.no_train:
        try:
            torch.cuda.empty_cache()
            edge = measure_edge(dev, peaks["hbm_gbs"])
        except Exception as exc:
            edge = {"unavailable": repr(exc)[:200]}
    adaptive = None
    if rank == 0 and world == 1 and not args.no_train:
        try:
            torch.cuda.empty_cache()
            adaptive = measure_adaptive(dev)
        except Exception as exc:
            adaptive = {"unavailable": repr(exc)[:200]}
    if rank == 0:
        line = {"metric": "posterior draws/sec", "value": value, "unit": "draws/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
                "vs_baseline": None, "dtype": "f16 tensor-core operands, f32 accumulate/state", "data": "synthetic",
                "config": workload_config(name, w, world, args.scaling),
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": "draws/s", "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": d2h_bytes, "ms_per_step": e2e_ms / args.steps,
                        "call": ("bayesflow_b200.parallel.sample_sharded (InferenceNetwork.sample on this rank's datasets)"
                                 if strong else "InferenceNetwork.call(z, conditions, inverse=True)")},
                "gpu_launches": launches,
                "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                             "frac": (achieved / peaks["tflops"]) if achieved else None, "traffic": ncu_traffic(name),
                             "kernel": ("bfk::lw_block_gemm_kernel<0> (fused Dense+block epilogue, 83 % of the call) + x_proj/merge/out stages: all launches of one bfk_sample_layerwise call"
                                        if is_layerwise else "bfk::sampler_fold_kernel"), "kernel_ms": k_avg_ms, "kernel_share_of_step": k_avg_ms * args.steps / total_ms,
                             "flop_per_draw_naive": fpd, "peak_source": peaks["source"],
                             "hbm_gbs_algorithmic": (n_local * S * D * 8) / (k_avg_ms * 1e-3) / 1e9 if kern_ms else None},
                "cpu_baseline": cpu,
                "integration": info, "wall_s_timed_region": wall, "train": train, "wide_cfg4": wide,
                "cfg5_fm_train_wide": train_wide,
                "sampler_api": sampler_api, "edge_kernels": edge, "adaptive_cfg1": adaptive}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure_sampling_simple(w, name, dev, world, rank, scaling, steps=2, warmup=3):
    """Secondary sampling line inside an existing process group (cfg4 at N > 1): device-timed, max over ranks."""
    import torch
    import torch.distributed as dist
    import bayesflow_b200 as bf
    from bayesflow_b200.parallel import dataset_shard, sample_sharded
    torch.manual_seed(1234)
    net = bf.FlowMatching(subnet_kwargs={"dropout": 0.0, "widths": w["widths"]})
    net.build((None, w["D"]), (None, w["C"]))
    net.to(dev)
    N, S = w["N"], w["S"]
    g = torch.Generator().manual_seed(2345)
    cond = torch.randn(N, w["C"], generator=g).to(dev)
    ikw = dict(method=w["method"], steps=w["ode_steps"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    if scaling == "strong":
        fn = lambda: sample_sharded(net, S, cond, seed=1234, global_step_control=False, **ikw)  # noqa: E731
        rows = N * S
    else:
        fn = lambda: bf.Sampler().sample(net, S, conditions=cond, seed=1234 + rank, **ikw)  # noqa: E731
        rows = world * N * S
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    pairs = []
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        pairs.append((a, b))
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in pairs)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    fpd = flops_per_draw(w)
    peaks = measured_peaks()
    achieved = rows * steps * fpd / (ms * 1e-3) / 1e12 / world
    return {"metric": "posterior draws/sec", "value": rows * steps / (ms * 1e-3), "unit": "draws/s", "ms_per_step": ms / steps,
            "steps": steps, "warmup": warmup, "scaling": scaling, "config": workload_config(name, w, world, scaling),
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tflops"], "note": "per GPU, whole call (all launches of bfk_sample_layerwise)"}}


def measure_adaptive(dev, steps=3, warmup=3):
    """BASELINE.json configs[0]'s solver at configs[0]'s shape: FlowMatching.sample with the network's DEFAULT integrator
    (tsit5, steps="adaptive", utils/integrate.py:291-406; networks/defaults.py:44-48) on 512 x 1000 two-moons-shaped rows.
    The step controller is global (max error over all rows), so the evaluation count is data dependent and is reported."""
    import torch
    import bayesflow_b200 as bf
    w = WORKLOADS["cfg1_twomoons_fm_euler100"]
    torch.manual_seed(1234)
    net = bf.FlowMatching(subnet_kwargs={"dropout": 0.0, "widths": w["widths"]})
    net.build((None, w["D"]), (None, w["C"]))
    net.to(dev)
    N, S = w["N"], w["S"]
    cond = torch.randn(N, w["C"], generator=torch.Generator().manual_seed(2345)).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    fn = lambda: bf.Sampler().sample(net, S, conditions=cond, seed=1234)  # noqa: E731  (no integrate kwargs: the default)
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    pairs = []
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        pairs.append((a, b))
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in pairs) / steps
    info = dict(getattr(net, "last_integration_info", {}) or {})
    return {"metric": "posterior draws/sec", "value": N * S / (ms * 1e-3), "unit": "draws/s", "ms_per_step": ms, "steps": steps,
            "warmup": warmup, "config": {"workload": "cfg1_twomoons_fm_tsit5_adaptive", "D": w["D"], "C": w["C"],
                                         "datasets": N, "draws_per_dataset": S,
                                         "solver": "tsit5 steps='adaptive' (the network default), global step control"},
            "integration": {k: info.get(k) for k in ("method", "steps", "rejected", "evals") if k in info}}


# --------------------------------------------------------------------------------------------------- training arms
def train_config(name, w, world):
    return {"workload": name, "network": "FlowMatching", "subnet": f"time_mlp widths={w['widths'][0]}x{len(w['widths'])}",
            "D": w["D"], "C": w["C"], "global_batch": w["batch"], "per_gpu_batch": w["batch"] // world,
            "dropout": w["dropout"], "step": "compute_metrics (x0~N, t~U on device) + backward + gradient all-reduce + "
            "Adam(clipnorm=1.5)", "parallelism": f"dp{world}",
            "l2": "activations of one step (>1 GB) exceed L2; 256 MB L2 flush between steps"}


def train_flops_per_sample(w):
    return 3 * count_flops_per_eval(w["D"], w["C"], w["widths"])  # forward + dX + dW, naive count


def cpu_train_rate(w, n_rows, threads, steps=1):
    """samples/s of the reference's CPU training step (oracle port + torch autograd + torch Adam)."""
    import torch
    import oracle
    torch.set_num_threads(threads)
    params = oracle.init_params(w["D"], w["C"] or None, w["widths"])
    ps = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    orc = oracle.OracleFlowMatching(ps, w["D"], subnet_kwargs={"dropout": w["dropout"], "widths": w["widths"]})
    opt = torch.optim.Adam(list(ps.values()), 5e-4)
    g = torch.Generator().manual_seed(4567)
    x, c = torch.randn(n_rows, w["D"], generator=g), torch.randn(n_rows, w["C"], generator=g)
    t0 = time.perf_counter()
    for _ in range(steps):
        opt.zero_grad()
        loss = orc.compute_metrics(x, conditions=c, stage="training")["loss"]
        loss.backward()
        torch.nn.utils.clip_grad_norm_(list(ps.values()), 1.5)
        opt.step()
    dt = time.perf_counter() - t0
    return steps * n_rows / dt, dt


def run_train_reference_arm(args, name, w):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = os.cpu_count() or 1
    rate, _ = cpu_train_rate(w, 2048, threads)
    n = int(max(2048, min(w["batch"], 10.0 * rate)) // 1024 * 1024)
    for _ in range(max(args.warmup - 1, 0)):
        cpu_train_rate(w, 2048, threads)
    rate, dt = cpu_train_rate(w, n, threads, steps=args.steps)
    line = {"impl": "reference", "metric": "FlowMatching train samples/sec", "value": rate, "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": train_config(name, w, args.gpus),
            "cpu_baseline": {"value": rate, "unit": "samples/s", "cores": threads, "kind": "port",
                             "sample": f"each step = one optimizer step on {n} rows (of the 65536-row batch), oracle port + torch "
                                       "autograd + torch Adam on the host cores"},
            "e2e": {"value": rate, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def measure_training(w, dev, world, rank, steps, warmup, want_cpu):
    """-> dict with the training numbers (device-timed, max over ranks)."""
    import torch
    import torch.distributed as dist
    import bayesflow_b200 as bf
    from bayesflow_b200.training import Adam, Trainer, cosine_decay

    torch.manual_seed(1234)
    net = bf.FlowMatching(subnet_kwargs={"dropout": w["dropout"], "widths": w["widths"]})
    net.build((None, w["D"]), (None, w["C"]))
    net.to(dev)
    tr = Trainer(net, Adam(net, cosine_decay(5e-4, 10000), clipnorm=1.5), world_size=world,
                 cuda_graph=os.environ.get("BFK_TRAIN_GRAPH", "1") != "0")
    Bl = w["batch"] // world
    g = torch.Generator().manual_seed(4567 + rank)
    x_host = torch.randn(Bl, w["D"], generator=g).pin_memory()
    c_host = torch.randn(Bl, w["C"], generator=g).pin_memory()
    x_dev, c_dev = x_host.to(dev), c_host.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    def step_resident():
        return tr.train_step(x_dev, c_dev)

    def step_e2e():
        loss = tr.train_step(x_host.to(dev, non_blocking=True), c_host.to(dev, non_blocking=True))
        loss_host.copy_(loss, non_blocking=True)
        return loss

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k, wu):
        for _ in range(wu):
            fn()
        barrier()
        l0 = tr.eng.launches
        pairs = []
        for _ in range(k):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            last = fn()
            b.record()
            pairs.append((a, b))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in pairs)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, tr.eng.launches - l0, float(last)

    ms, launches, loss = timed(step_resident, steps, warmup)
    e2e_ms, _, _ = timed(step_e2e, steps, max(1, warmup // 2))
    fps = train_flops_per_sample(w)
    peaks = measured_peaks()
    achieved = w["batch"] / world * fps / (ms / steps * 1e-3) / 1e12
    out = {"value": w["batch"] * steps / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / steps,
           "e2e": {"value": w["batch"] * steps / (e2e_ms * 1e-3), "unit": "samples/s", "ms_per_step": e2e_ms / steps,
                   "h2d_bytes_per_step": int((x_host.numel() + c_host.numel()) * 4), "d2h_bytes_per_step": 4},
           "gpu_launches": launches, "loss_last": loss,
           "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                        "frac": achieved / peaks["tflops"], "traffic": None, "kernel": "whole step (bfk::gemm_kernel + "
                        "bfk::gemm_dw_kernel + block kernels), per GPU", "flop_per_sample_naive": fps,
                        "peak_source": peaks["source"]}}
    if want_cpu:
        threads = os.cpu_count() or 1
        rate, dt = cpu_train_rate(w, 8192, threads)
        out["cpu_baseline"] = {"value": rate, "unit": "samples/s", "cores": threads, "kind": "port",
                               "sample": f"one optimizer step on 8192 rows ({dt:.1f} s): oracle port + torch autograd + torch Adam"}
    return out


def run_train_ours(args, name, w):
    import torch
    import torch.distributed as dist
    from bayesflow_b200 import _lib
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    m = measure_training(w, dev, world, rank, args.steps, args.warmup, rank == 0 and world == 1 and not args.no_cpu_baseline)
    clk = clocks.stop() if rank == 0 else None
    if rank == 0:
        line = {"metric": "FlowMatching train samples/sec", "value": m["value"], "unit": "samples/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": m["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f16 tensor-core operands, f32 accumulate/master weights",
                "data": "synthetic", "config": train_config(name, w, world), "clocks": clk, "e2e": m["e2e"],
                "gpu_launches": m["gpu_launches"], "roofline": m["roofline"], "cpu_baseline": m.get("cpu_baseline"),
                "loss_last": m["loss_last"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS) + sorted(TRAIN_WORKLOADS))
    ap.add_argument("--no-train", action="store_true", help="skip the short training measurement added to the default line")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default): the stated N x S problem is split over the ranks by dataset (sample_sharded); "
                         "weak: N datasets per GPU")
    args = ap.parse_args()
    if args.workload in TRAIN_WORKLOADS:
        w = TRAIN_WORKLOADS[args.workload]
        (run_train_reference_arm if args.impl == "reference" else run_train_ours)(args, args.workload, w)
        return
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, args.workload, w)
    else:
        if args.warmup < 3:
            print("bench.py: note: fewer than 3 warm-up steps requested", file=sys.stderr)
        run_ours(args, args.workload, w)


if __name__ == "__main__":
    # stdout carries exactly ONE JSON line: libraries that write to fd 1 (NCCL prints its version banner there) are sent to
    # stderr, and the line is written to the saved descriptor
    sys.stdout.flush()
    _real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(_real_stdout, "w")
    main()
