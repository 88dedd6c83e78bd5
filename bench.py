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
}
DEFAULT_WORKLOAD = "cfg2_sir_fm_euler100"
EVALS_PER_STEP = {"euler": 1, "rk45": 6, "tsit5": 6}


def flops_per_draw(w):
    import oracle
    return oracle.count_flops_per_eval(w["D"], w["C"], w["widths"]) * EVALS_PER_STEP[w["method"]] * w["ode_steps"]


def workload_config(name, w, n_gpus):
    return {"workload": name, "network": "FlowMatching" if w["model"] == "fm" else "DiffusionModel(edm,F)",
            "subnet": f"time_mlp widths={w['widths'][0]}x{len(w['widths'])}", "D": w["D"], "C": w["C"],
            "datasets_per_gpu": w["N"], "draws_per_dataset": w["S"], "solver": f"{w['method']} fixed {w['ode_steps']} steps",
            "sharding": f"rows split by dataset over {n_gpus} GPU(s), no collective",
            "l2": "state+conditions stream once per step; weights (1 MB) are L2-resident by design; 256 MB L2 flush between steps"}


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
