"""Achieved HBM GB/s of the approximator-edge kernels (bfk_standardize, bfk_batch_moments, bfk_normal_fill) on inputs
larger than L2, timed with CUDA events on the launching stream.  Prints one JSON line per kernel.
usage (GPU box): python tools/gpu_edge_bench.py [rows] [F]"""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bayesflow_b200 as bf  # noqa: E402
from bayesflow_b200 import _lib  # noqa: E402


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2] * 1e-3


def main():
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 16 * 1024 * 1024
    peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    peak = peaks["hbm_gbs"]
    L = _lib.lib()
    for F in ([int(sys.argv[2])] if len(sys.argv) > 2 else [2, 32, 64]):
        r = rows * 32 // max(F, 2) if F < 32 else rows * 32 // F
        x = torch.randn(r, F, device="cuda")
        out = torch.empty_like(x)
        st = bf.Standardize()
        st.build((None, F))
        st.cuda()
        st(x[:4096], stage="training")
        nbytes = x.numel() * 4
        for name, fn, traffic in (
            ("standardize_fwd", lambda: st._location_scale(x, 0, True, out=out), 2 * nbytes),
            ("standardize_inv_inplace", lambda: st._location_scale(out, 0, False, out=out), 2 * nbytes),
            ("batch_moments", lambda: st.batch_moments(x, 0), nbytes),
            ("torch_copy", lambda: out.copy_(x), 2 * nbytes),
        ):
            t = timed(fn)
            print(json.dumps({"kernel": name, "rows": r, "F": F, "ms": round(t * 1e3, 4), "algorithmic_bytes": traffic,
                              "achieved_gbs": round(traffic / t / 1e9, 1), "peak_gbs": peak,
                              "frac": round(traffic / t / 1e9 / peak, 3)}), flush=True)
        dist = bf.DiagonalNormal()
        dist.build((None, F))
        dist.cuda()
        zbuf = torch.empty(r, F, device="cuda")

        def fill():
            _lib.check(L.bfk_normal_fill(_lib.ptr(zbuf), r, F, 0, None, None, C.c_uint64(7), None, _lib.stream_ptr()), "fill")
        t = timed(fill)
        t2 = timed(lambda: zbuf.normal_())
        print(json.dumps({"kernel": "normal_fill", "rows": r, "F": F, "ms": round(t * 1e3, 4), "algorithmic_bytes": nbytes,
                          "achieved_gbs": round(nbytes / t / 1e9, 1), "peak_gbs": peak, "frac": round(nbytes / t / 1e9 / peak, 3),
                          "torch_normal_ms": round(t2 * 1e3, 4)}), flush=True)
        del x, out, zbuf


if __name__ == "__main__":
    main()
