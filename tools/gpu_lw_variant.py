"""cfg4-shaped timing + parity of one libbfk variant (BFK_LIB=...): 1024x8 subnet, D=64, Euler, layer-wise sampler."""
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
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    N, S = int(sys.argv[2]) if len(sys.argv) > 2 else 1000, 1000
    torch.manual_seed(0)
    net, orc = make_pair(64, 32, (1024,) * 8)
    z = torch.randn(N * S, 64, device="cuda")
    c = torch.randn(N, 32, device="cuda")
    best = 1e9
    for it in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        net(z, conditions=c, conditions_repeats=S, inverse=True, method="euler", steps=steps)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    B = 300
    zz, cc = torch.randn(B, 64), grouped_conditions(B, 32, 100)
    ref = orc.inverse(zz, cc, method="euler", steps=10)
    out = net(zz.cuda(), conditions=cc.cuda(), inverse=True, method="euler", steps=10)
    print(f"cfg4 euler-{steps} {N}x{S}: {best / steps * 100:.1f} ms per 100 evals, {N * S / (best / steps * 100):.1f} k draws/s"
          f"   rel-L2 vs oracle (euler-10, B={B}) = {rel_l2(out, ref):.3e}", flush=True)


if __name__ == "__main__":
    main()
