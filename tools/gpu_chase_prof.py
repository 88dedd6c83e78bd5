"""Timing builds only (libbfk built with -DBFK_PROFILE): one cfg2 Euler launch of the chase kernel, then the per-warp
cycle accounting (bfk_debug_read_prof).  usage: python tools/gpu_chase_prof.py [steps] [datasets]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bayesflow_b200 as bf
from bayesflow_b200 import _lib

if os.environ.get("BFK_LIB"):  # timing-experiment variant built by tools/build_variants.sh
    _lib.LIB_PATH = os.path.abspath(os.environ["BFK_LIB"])


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    S = 1000
    torch.manual_seed(0)
    net = bf.FlowMatching(subnet_kwargs={"dropout": 0.0})
    net.build((None, 2), (None, 32))
    net.cuda()
    z = torch.randn(N * S, 2, device="cuda")
    c = torch.randn(N, 32, device="cuda")
    for it in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        x = net(z, conditions=c, conditions_repeats=S, inverse=True, method="euler", steps=steps)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"euler-{steps} {N}x{S}: {ms:.2f} ms  -> {ms / steps * 100:.1f} ms per 100 evals, {N * S / (ms / steps * 100) / 1e3:.2f} M draws/s @100", flush=True)
    L = _lib.lib()
    if not hasattr(L, "bfk_debug_read_prof"):
        return
    buf = np.zeros(148 * 18 * 8, dtype=np.int64)
    L.bfk_debug_read_prof.argtypes = [C.c_void_p]
    L.bfk_debug_read_prof(buf.ctypes.data)
    p = buf.reshape(148, 18, 8).astype(np.float64)
    rows_per_cta = 128 if os.environ.get("BFK_SAMPLER") == "chase" else 256
    n_tiles = (N * S + rows_per_cta - 1) // rows_per_cta
    per_cta_tiles = n_tiles / 148.0
    phases = per_cta_tiles * steps  # evaluations per CTA
    epi = p[:, :16, :].mean(axis=(0, 1)) / phases
    names = ["ode+tail", "film+bar", "x_proj", "wait_acc", "epi_merge", "blk_tmemwait", "blk_math", "blk_announce"]
    if os.environ.get("BFK_SAMPLER") != "chase":
        names = ["ode+tail", "wait_film", "x_proj", "wait_acc", "epi_merge", "epi_block", "stats", "announce"]
    print("epilogue warps, cycles per evaluation (mean over CTAs and warps):")
    for n, v in zip(names, epi):
        print(f"  {n:10s} {v:9.0f}")
    print(f"  total      {epi.sum():9.0f}")
    iss = p[:, 16, :].mean(axis=0) / phases
    print("issuer per evaluation: issue %.0f  wait_slab %.0f  wait_wfull %.0f" % (iss[0], iss[1], iss[2]))
    pr = p[:, 17, :].mean(axis=0) / phases
    print("producer per evaluation: issue %.0f  wait_empty %.0f  film %.0f" % (pr[0], pr[1], pr[2]))
    # spread across column groups
    for cg in range(4):
        e = p[:, cg * 4:(cg + 1) * 4, :].mean(axis=(0, 1)) / phases
        print(f"  cg{cg}: " + " ".join(f"{n}={v:.0f}" for n, v in zip(names, e)))


if __name__ == "__main__":
    main()
