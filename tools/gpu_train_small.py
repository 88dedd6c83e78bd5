"""Per-GPU share of the N = 8 training step on ONE GPU: the 256x5 FlowMatching step at 65536 / 8 = 8192 rows (no
all-reduce), CUDA-graph trainer and eager, device-timed.  usage: python tools/gpu_train_small.py [rows] [steps] [graph]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bayesflow_b200 as bf  # noqa: E402
from bayesflow_b200.training import Adam, Trainer, cosine_decay  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
modes = [bool(int(sys.argv[3]))] if len(sys.argv) > 3 else [True, False]
dev = torch.device("cuda:0")
for graph in modes:
    torch.manual_seed(1234)
    net = bf.FlowMatching(subnet_kwargs={"dropout": 0.05, "widths": (256,) * 5})
    net.build((None, 2), (None, 32))
    net.to(dev)
    tr = Trainer(net, Adam(net, cosine_decay(5e-4, 10000), clipnorm=1.5), world_size=1, cuda_graph=graph)
    x, c = torch.randn(rows, 2, device=dev), torch.randn(rows, 32, device=dev)
    for _ in range(5):
        tr.train_step(x, c)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = tr.eng.launches
    a.record()
    for _ in range(steps):
        loss = tr.train_step(x, c)
    b.record()
    torch.cuda.synchronize()
    print(f"rows {rows} graph {int(graph)}: {a.elapsed_time(b) / steps * 1e3:.1f} us / step, "
          f"{(tr.eng.launches - l0) / steps:.0f} launches / step, loss {float(loss):.4f}", flush=True)
