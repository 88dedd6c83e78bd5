import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from _util import make_pair, grouped_conditions
import oracle.integrate as oi
import bayesflow_b200.stochastic as st
torch.manual_seed(0)
net, orc = make_pair(3, 6, (128,) * 2, dm=True, model_kwargs={"noise_schedule": "cosine"})
B = 400
z, c = torch.randn(B, 3), grouped_conditions(B, 6, 100)
g = torch.Generator().manual_seed(9)
Z = [torch.randn(B, 3, generator=g) for _ in range(1100)]
oi.TRACE, st.TRACE = [], []
ref = orc.inverse(z, c, method="two_step_adaptive", steps="adaptive", noise_fn=lambda i, s: {"xz": Z[i]})
out = net(z.cuda(), conditions=c.cuda(), inverse=True, method="two_step_adaptive", steps="adaptive", noise_fn=lambda i, shape: Z[i])
print(len(oi.TRACE), len(st.TRACE), float(ref.abs().max()), float(out.abs().max()))
for i, (a, b) in enumerate(zip(oi.TRACE, st.TRACE)):
    flag = "" if (a[3] == b[3] and abs(a[2] - b[2]) < 1e-2 * max(1, abs(a[2]))) else "  <<<"
    if flag or i < 5 or i > len(st.TRACE) - 8:
        print(i, a, b, flag)
