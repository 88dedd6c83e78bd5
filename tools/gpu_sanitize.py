"""CI-sized shapes of every hand-rolled mbarrier/TMEM kernel, to be run under compute-sanitizer
(tools/gpu_sanitize.sh).  Prints the parity of each stage so a sanitizer-clean run is also a correct run."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from _util import make_pair, rel_l2, grouped_conditions  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    torch.manual_seed(0)
    if which in ("all", "fused"):
        # fused on-chip sampler (sampler_*_kernel): velocity, Euler, one adaptive trial sequence
        net, orc = make_pair(2, 32)
        B = 300
        z = torch.randn(B, 2)
        c = grouped_conditions(B, 32, 100)
        v = net.velocity(z.cuda(), 0.37, conditions=c.cuda()).cpu()
        print("fused velocity", rel_l2(v, orc.velocity(z, 0.37, c)), flush=True)
        x = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=3).cpu()
        print("fused euler-3", rel_l2(x, orc.inverse(z, c, method="euler", steps=3)), flush=True)
        x = net(z.cuda(), conditions=c.cuda(), inverse=True, method="rk45", steps=2).cpu()
        print("fused rk45-2", rel_l2(x, orc.inverse(z, c, method="rk45", steps=2)), flush=True)
        x = net(z.cuda(), conditions=c.cuda(), inverse=True, method="tsit5", steps="adaptive", min_steps=2, max_steps=6).cpu()
        print("fused tsit5-adaptive", rel_l2(x, orc.inverse(z, c, method="tsit5", steps="adaptive", min_steps=2, max_steps=6)), flush=True)
        net, orc = make_pair(4, 20, widths=(128,) * 2, norm="rms", residual=False)  # generic kernel
        z = torch.randn(200, 4)
        c = grouped_conditions(200, 20, 50)
        x = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=2).cpu()
        print("generic euler-2", rel_l2(x, orc.inverse(z, c, method="euler", steps=2)), flush=True)
    if which in ("all", "lw"):
        # layer-wise sampler (lw_block_gemm_kernel)
        net, orc = make_pair(16, 8, widths=(512,) * 2)
        z = torch.randn(200, 16)
        c = grouped_conditions(200, 8, 50)
        x = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=2).cpu()
        print("layerwise euler-2", rel_l2(x, orc.inverse(z, c, method="euler", steps=2)), flush=True)
    if which in ("all", "train"):
        # training step (gemm_persist_kernel, gemm_dw_kernel, block fwd/bwd)
        net, orc = make_pair(2, 32, widths=(128,) * 2)
        B = 256
        g = torch.Generator().manual_seed(5)
        x1, x0, t = torch.randn(B, 2, generator=g), torch.randn(B, 2, generator=g), torch.rand(B, 1, generator=g)
        c = torch.randn(B, 32, generator=g)
        tup = (x0, x1, t, t * x1 + (1 - t) * x0, x1 - x0)
        m = net.compute_metrics(tuple(v.cuda() for v in tup), conditions=c.cuda(), stage="training")
        m["loss"].backward()
        ref = orc.compute_metrics(tup, conditions=c)
        print("train loss", float(m["loss"]), float(ref["loss"]), flush=True)
    torch.cuda.synchronize()
    print("sanitize driver done", flush=True)


if __name__ == "__main__":
    main()
