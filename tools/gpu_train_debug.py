"""First-contact check of the training path: per-tensor gradient errors against oracle autograd (prints, no asserts)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch

from _util import make_pair, rel_l2
from bayesflow_b200.training import train_engine


def oracle_grads(orc, tup, c, w=None):
    ps = {k: v.clone().requires_grad_(True) for k, v in orc.p.items()}
    old = orc.p, orc.subnet.p
    orc.p = orc.subnet.p = ps
    try:
        loss = orc.compute_metrics(tup, conditions=c, sample_weight=w, stage="training")["loss"]
        loss.backward()
    finally:
        orc.p, orc.subnet.p = old
    return float(loss), {k: v.grad for k, v in ps.items()}


def run(D, C, widths, B, swap, **sk):
    net, orc = make_pair(D, C, widths, **sk)
    g = torch.Generator().manual_seed(5678)
    x1, x0, t = torch.randn(B, D, generator=g), torch.randn(B, D, generator=g), torch.rand(B, 1, generator=g)
    c = torch.randn(B, C, generator=g) if C else None
    x, target = t * x1 + (1 - t) * x0, x1 - x0
    ref_loss, ref = oracle_grads(orc, (x0, x1, t, x, target), c)
    eng = train_engine(net)
    loss, pred = eng.step(x.cuda(), t.cuda(), c.cuda() if C else None, target.cuda(), want_pred=True, _dw_swap=swap)
    torch.cuda.synchronize()
    pref = orc.velocity(x, t, c)
    print(f"D={D} C={C} W={widths} B={B} swap={swap} {sk}: pred rel={rel_l2(pred, pref):.2e} loss {float(loss):.6f} vs {ref_loss:.6f}")
    worst = 0
    for name, p in net.named_parameters():
        if name not in ref or ref[name] is None:
            continue
        e = rel_l2(eng.grad_of(p), ref[name])
        worst = max(worst, e)
        if e > 1e-3 or "--all" in sys.argv:
            print(f"    {name:45s} rel={e:.2e}  |ref|={float(ref[name].norm()):.3e}")
    print(f"    worst gradient rel-L2 = {worst:.2e}", flush=True)


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    run(2, 32, (256,) * 5, 1000, 0)
    run(2, 32, (256,) * 5, 1000, 1)
    run(2, 32, (256,) * 5, 4097, 0)
    run(4, None, (128,) * 3, 777, 0)
    run(3, 7, (128,) * 2, 300, 0, merge="add")
    run(2, 6, (128,) * 2, 300, 0, activation="silu", norm=None)
    run(64, 32, (1024,) * 2, 512, 0)
    # throughput probe
    import bayesflow_b200 as bf
    from bayesflow_b200.training import Trainer, Adam
    net = bf.FlowMatching()
    net.build((None, 2), (None, 32))
    net.cuda()
    tr = Trainer(net, Adam(net, 5e-4, clipnorm=1.5), world_size=1)
    B = 65536
    x, c = torch.randn(B, 2, device="cuda"), torch.randn(B, 32, device="cuda")
    for i in range(3):
        l = tr.train_step(x, c)
    torch.cuda.synchronize()
    t0 = time.time()
    n = 10
    for i in range(n):
        l = tr.train_step(x, c)
    torch.cuda.synchronize()
    dt = (time.time() - t0) / n
    print(f"train step B={B}: {dt*1e3:.2f} ms -> {B/dt/1e6:.2f} M samples/s, loss {float(l):.4f}")
