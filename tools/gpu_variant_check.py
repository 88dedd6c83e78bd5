"""Timing + parity of one libbfk build variant (BFK_LIB=build/libbfk_<name>.so, tools/build_variants.sh):
cfg2-shaped Euler launch timed with CUDA events, and rel-L2 of Euler-100 draws against the CPU oracle (B = 3000)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

from bayesflow_b200 import _lib

if os.environ.get("BFK_LIB"):
    _lib.LIB_PATH = os.path.abspath(os.environ["BFK_LIB"])
from _util import make_pair, rel_l2, grouped_conditions  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    N, S = int(sys.argv[2]) if len(sys.argv) > 2 else 1000, 1000
    act = sys.argv[3] if len(sys.argv) > 3 else "mish"
    torch.manual_seed(0)
    net, orc = make_pair(2, 32, (256,) * 5, activation=act)
    z = torch.randn(N * S, 2, device="cuda")
    c = torch.randn(N, 32, device="cuda")
    best = 1e9
    for it in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        net(z, conditions=c, conditions_repeats=S, inverse=True, method="euler", steps=steps)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    B = 3000
    zz, cc = torch.randn(B, 2), grouped_conditions(B, 32, 1000)
    ref = orc.inverse(zz, cc, method="euler", steps=100)
    out = net(zz.cuda(), conditions=cc.cuda(), inverse=True, method="euler", steps=100)
    print(f"{act} euler-{steps} {N}x{S}: {best / steps * 100:.1f} ms per 100 evals, {N * S / (best / steps * 100) / 1e3:.2f} M draws/s"
          f"   rel-L2 vs oracle (euler-100, B={B}) = {rel_l2(out, ref):.3e}", flush=True)


if __name__ == "__main__":
    main()
