"""Small cfg4-shaped run for ncu: 2 chunks x 3 Euler steps of the layer-wise sampler."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bayesflow_b200 as bf

torch.manual_seed(0)
net = bf.FlowMatching(subnet_kwargs={"widths": (1024,) * 8, "dropout": 0.0})
net.build((None, 64), (None, 32))
net.cuda()
B = int(os.environ.get("B", 16384))
z = torch.randn(B, 64, device="cuda")
c = torch.randn(B // 1024, 32, device="cuda")
steps = int(os.environ.get("STEPS", 3))
for it in range(2):
    torch.cuda.synchronize(); t0 = time.time()
    out = net(z, conditions=c, conditions_repeats=1024, inverse=True, method="euler", steps=steps)
    torch.cuda.synchronize(); dt = time.time() - t0
    print(f"B={B} euler-{steps}: {dt*1e3:.1f} ms -> {dt/steps/ (B/8192)*1e6:.0f} us per eval per 8192-row chunk", flush=True)
