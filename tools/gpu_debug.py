"""First-contact GPU check: prints the error of every stage against the oracle instead of asserting."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bayesflow_b200 as bf
import oracle


def make(D, C, widths=(256,) * 5, dm=False, **sk):
    params = oracle.perturb_params(oracle.init_params(D, C, widths, **{k: v for k, v in sk.items() if k in ("merge", "film_use_gamma", "norm", "residual")}))
    kw = {"dropout": 0.0, "widths": widths} | sk
    cls, ocls = (bf.DiffusionModel, oracle.OracleDiffusionModel) if dm else (bf.FlowMatching, oracle.OracleFlowMatching)
    net = cls(subnet_kwargs=kw)
    net.build((None, D), (None, C) if C else None)
    missing = net.load_state_dict(params, strict=False)
    assert not missing.unexpected_keys, missing
    net.cuda()
    return net, ocls(params, D, subnet_kwargs=kw)


def rel(a, b):
    return float((a - b).norm() / b.norm())


def main():
    torch.manual_seed(0)
    print(torch.cuda.get_device_name(0), flush=True)
    for (D, C, widths, B) in [(2, 32, (256,) * 5, 128), (2, 32, (256,) * 5, 1000), (2, None, (256,) * 5, 300),
                              (4, 20, (128,) * 3, 777), (2, 2, (64,) * 2, 257)]:
        net, orc = make(D, C, widths)
        z = torch.randn(B, D)
        c = torch.randn(B, C) if C else None
        if C and B >= 300:
            c = torch.randn((B + 99) // 100, C).repeat_interleave(100, dim=0)[:B]
        v_ref = orc.velocity(z, 0.37, c)
        t0 = time.time()
        v = net.velocity(z.cuda(), 0.37, conditions=c.cuda() if C else None).cpu()
        torch.cuda.synchronize()
        print(f"D={D} C={C} W={widths[0]}x{len(widths)} B={B}: velocity rel={rel(v, v_ref):.3e}  ({time.time()-t0:.2f}s)", flush=True)
        if rel(v, v_ref) > 1e-2:
            print("  v[:3]", v[:3].tolist(), "\n  ref ", v_ref[:3].tolist(), flush=True)
        for method, steps in [("euler", 20), ("rk45", 5), ("tsit5", 5)]:
            x_ref = orc.inverse(z, c, method=method, steps=steps)
            x = net(z.cuda(), conditions=c.cuda() if C else None, inverse=True, method=method, steps=steps).cpu()
            print(f"   {method}-{steps}: rel={rel(x, x_ref):.3e}", flush=True)
    # DM
    net, orc = make(4, 20, dm=True)
    z = torch.randn(500, 4); c = torch.randn(5, 20).repeat_interleave(100, dim=0)
    for t in (0.9, 0.5, 0.05):
        print(f"DM velocity t={t}: rel={rel(net.velocity(z.cuda(), t, conditions=c.cuda()).cpu(), orc.velocity(z, t, conditions=c)):.3e}", flush=True)
    x_ref = orc.inverse(z, c, method="euler", steps=50)
    x = net(z.cuda(), conditions=c.cuda(), inverse=True, method="euler", steps=50).cpu()
    print(f"DM euler-50: rel={rel(x, x_ref):.3e}", flush=True)
    # throughput probe
    net, orc = make(2, 32)
    N, S = 1000, 1000
    z = torch.randn(N * S, 2, device="cuda"); c = torch.randn(N, 32, device="cuda").repeat_interleave(S, dim=0)
    for it in range(3):
        torch.cuda.synchronize(); t0 = time.time()
        x = net(z, conditions=c, inverse=True, method="euler", steps=100)
        torch.cuda.synchronize(); dt = time.time() - t0
        print(f"cfg2 Euler-100 1M draws: {dt*1e3:.1f} ms -> {N*S/dt/1e6:.2f} M draws/s, {N*S/dt*102.04e6/1e12:.1f} TFLOP/s naive", flush=True)
    print("mean/std", x.mean(0).tolist(), x.std(0).tolist())


if __name__ == "__main__":
    main()
