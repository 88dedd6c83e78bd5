"""GPU parity tests of the training path (through the C ABI: bfk_train_step, bfk_adam_step) against the CPU oracle:
loss and EVERY parameter gradient of FlowMatching.compute_metrics (flow_matching.py:132-178) vs torch autograd on the
restated network (what the reference does: torch_approximator.py:88-99).  Gate from BASELINE.json: 2e-3 relative L2 on
velocity outputs and on all gradients.  The tuple input (x0, x1, t, x, target) bypasses all RNG (flow_matching.py:140)."""
import math

import pytest
import torch

from _util import make_pair, rel_l2

pytestmark = pytest.mark.gpu
TOL = 2e-3


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    import bayesflow_b200._lib as L
    L.lib()


def oracle_loss_and_grads(orc, tup, c, w=None):
    ps = {k: v.clone().requires_grad_(True) for k, v in orc.p.items()}
    old = orc.p, orc.subnet.p
    orc.p = orc.subnet.p = ps
    try:
        loss = orc.compute_metrics(tup, conditions=c, sample_weight=w, stage="training")["loss"]
        loss.backward()
    finally:
        orc.p, orc.subnet.p = old
    return float(loss.detach()), {k: v.grad for k, v in ps.items()}


def make_batch(B, D, C, seed=5678):
    g = torch.Generator().manual_seed(seed)
    x1, x0, t = torch.randn(B, D, generator=g), torch.randn(B, D, generator=g), torch.rand(B, 1, generator=g)
    c = torch.randn(B, C, generator=g) if C else None
    return (x0, x1, t, t * x1 + (1 - t) * x0, x1 - x0), c


def cuda_tuple(tup):
    return tuple(v.cuda() for v in tup)


@pytest.mark.parametrize("D,C,widths,B,sk", [
    (2, 32, (256,) * 5, 1000, {}),           # cfg2 net, tile-unaligned batch
    (2, 32, (256,) * 5, 128, {}),
    (2, 32, (256,) * 5, 1, {}),              # single row
    (2, 2, (256,) * 5, 513, {}),             # cfg1 net
    (4, None, (128,) * 3, 777, {}),          # unconditional
    (3, 7, (128,) * 2, 300, {"merge": "add"}),
    (2, 6, (128,) * 2, 300, {"activation": "silu", "norm": None}),
    (2, 6, (64,) * 2, 300, {"activation": "gelu"}),
    (2, 6, (192,) * 2, 260, {"activation": "tanh", "residual": False}),
    (2, 6, (128,) * 2, 300, {"time_embedding_dim": 8}),
    (2, 6, (128,) * 2, 300, {"time_embedding_dim": 1}),
    (2, 6, (128,) * 2, 300, {"activation": "relu"}),
    (2, 6, (64, 96, 32), 300, {}),                                           # unequal widths: residual projector, padded to 128
    (3, 5, (128, 64, 64, 256), 260, {"merge": "add", "activation": "silu"}),
    (2, None, (96, 160), 200, {"norm": "rms"}),
    (2, 6, (64, 128), 200, {"norm": None, "film_use_gamma": True}),
    (2, 6, (128,) * 3, 300, {"norm": "rms"}),                                # RMSNormalization: scale only, eps 1e-6
    (4, 5, (512,) * 2, 260, {"norm": "rms", "activation": "relu"}),
    (2, 6, (128,) * 3, 300, {"film_use_gamma": True}),                       # FiLM scale: (1 + gamma(t)) * h + beta(t)
    (4, 5, (512,) * 2, 260, {"film_use_gamma": True, "activation": "silu"}),  # ... on the wide block kernels
    (3, None, (1024,) * 2, 200, {"film_use_gamma": True, "time_embedding_dim": 1}),
    (64, 32, (1024,) * 2, 512, {}),          # cfg4 block shape
    (5, 3, (512,) * 2, 384, {}),
    # block_bwd_split_kernel: 768 columns on the 4-warp form (one idle column warp), norm None / rms (no row-sum exchange /
    # no centring), unequal wide widths (padded to 768, effective-width norm), a single row, one row beyond a tile
    (3, 4, (768,) * 2, 130, {}),
    (3, 4, (768, 512), 129, {"activation": "silu"}),
    (2, 6, (512,) * 2, 1, {"norm": None}),
    (2, 6, (1024,) * 2, 257, {"norm": "rms", "activation": "gelu"}),
    (2, None, (512,) * 3, 300, {"residual": False}),
])
def test_loss_and_all_gradients(D, C, widths, B, sk):
    net, orc = make_pair(D, C, widths, **sk)
    tup, c = make_batch(B, D, C)
    ref_loss, ref = oracle_loss_and_grads(orc, tup, c)
    out = net.compute_metrics(cuda_tuple(tup), conditions=c.cuda() if C else None, stage="training")
    loss = out["loss"]
    loss.backward()
    assert abs(float(loss) - ref_loss) <= 2e-3 * abs(ref_loss), (float(loss), ref_loss)
    checked = 0
    for name, p in net.named_parameters():
        if name in ref and ref[name] is not None:
            assert p.grad is not None, name
            if B == 1 and float(ref[name].norm()) < 1e-12:
                continue
            e = rel_l2(p.grad, ref[name])
            # relu' is a step function: wherever the fp16-operand pre-activation and the fp32 oracle's fall on different
            # sides of 0 a whole gradient path flips, so the smooth-activation gate does not apply to it
            assert e < (5e-2 if sk.get("activation") == "relu" else TOL), (name, e)
            checked += 1
    assert checked >= 10


def test_sample_weight_and_mae():
    net, orc = make_pair(2, 6, (128,) * 2)
    tup, c = make_batch(400, 2, 6)
    w = torch.rand(400) + 0.5
    ref_loss, ref = oracle_loss_and_grads(orc, tup, c, w)
    loss = net.compute_metrics(cuda_tuple(tup), conditions=c.cuda(), sample_weight=w.cuda())["loss"]
    loss.backward()
    assert abs(float(loss) - ref_loss) <= 2e-3 * abs(ref_loss)
    for name, p in net.named_parameters():
        if name in ref:
            assert rel_l2(p.grad, ref[name]) < TOL, name


def test_validation_stage_has_no_gradients_and_matches_loss():
    net, orc = make_pair(2, 32)
    tup, c = make_batch(300, 2, 32)
    ref_loss = float(orc.compute_metrics(tup, conditions=c, stage="validation")["loss"])
    with torch.no_grad():
        loss = net.compute_metrics(cuda_tuple(tup), conditions=c.cuda(), stage="validation")["loss"]
    assert not loss.requires_grad
    assert abs(float(loss) - ref_loss) <= 2e-3 * abs(ref_loss)


def test_rng_path_runs_and_is_plausible():
    """x0 ~ N(0,1), t ~ U(0,1) drawn on the device (flow_matching.py:143-168); the loss of a fresh network on
    standard-normal targets is E|x1-x0-v|^2 ~ 2 + |v|^2."""
    net, _ = make_pair(2, 32)
    x1, c = torch.randn(4096, 2, device="cuda"), torch.randn(4096, 32, device="cuda")
    loss = net.compute_metrics(x1, conditions=c)["loss"]
    assert math.isfinite(float(loss)) and 1.0 < float(loss) < 20.0


def test_dropout_is_a_training_only_inverted_mask():
    """conditional_dense_block.py:120-121 / test_blocks.py: train != eval with dropout, same mask for the same seed."""
    net, _ = make_pair(2, 6, (128,) * 2, dropout=0.3)
    tup, c = make_batch(512, 2, 6)
    ct, cc = cuda_tuple(tup), c.cuda()
    with torch.no_grad():
        l_eval = float(net.compute_metrics(ct, conditions=cc, stage="validation")["loss"])
    l_a = float(net.compute_metrics(ct, conditions=cc, stage="training", seed=11)["loss"])
    l_b = float(net.compute_metrics(ct, conditions=cc, stage="training", seed=11)["loss"])
    l_c = float(net.compute_metrics(ct, conditions=cc, stage="training", seed=12)["loss"])
    assert l_a == pytest.approx(l_b, rel=1e-5) and l_a != l_eval and l_a != l_c


def _np_drop_mask(seed, layer, row0, B, W, rate):
    """numpy restatement of drop_key / drop_scale8 (bfk_train.cuh): splitmix64 key, two multiply-xorshift rounds per
    column PAIR, the top 24 bits of each half against ceil(rate * 2^24)."""
    import numpy as np
    M64 = (1 << 64) - 1
    z = (seed + 0x9E3779B97F4A7C15 * (layer + 1)) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    key = np.uint64(z ^ (z >> 31))
    rows = np.arange(row0, row0 + B, dtype=np.uint64)[:, None]
    pairs = np.arange(W // 2, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        x = (rows * np.uint64(512) + pairs) ^ key
        for _ in range(2):
            x = x * np.uint64(0xD6E8FEB86659FD93)
            x ^= x >> np.uint64(32)
    thr = int(math.ceil(float(np.float32(rate)) * 16777216.0))
    lo = ((x & np.uint64(0xFFFFFFFF)) >> np.uint64(8)).astype(np.int64)
    hi = (x >> np.uint64(40)).astype(np.int64)
    keep = np.empty((B, W), dtype=bool)
    keep[:, 0::2] = lo >= thr
    keep[:, 1::2] = hi >= thr
    return keep


@pytest.mark.parametrize("rate,W", [(0.05, 256), (0.3, 128), (0.5, 8)])
def test_hidden_dropout_mask_is_the_documented_counter_hash_and_bernoulli(rate, W):
    """conditional_dense_block.py:120-121 (keras Dropout: Bernoulli(1 - rate) keep, inverted scaling).  The kernel's mask
    is bit-equal to the numpy restatement of its hash, has the right keep rate per column and per row parity, is
    uncorrelated between the two columns that share a hash and between neighbouring rows, and a shard reproduces the
    rows of the whole batch."""
    import numpy as np
    from bayesflow_b200.training import dropout_mask
    B, row0 = 60_000, 123_456_789
    m = dropout_mask(seed=7, layer=2, rows=B, width=W, rate=rate, row_offset=row0)
    scale = float(np.float32(1.0) / (np.float32(1.0) - np.float32(rate)))
    vals = sorted(m.unique().tolist())
    assert len(vals) == 2 and vals[0] == 0.0 and vals[1] == pytest.approx(scale, rel=1e-6)
    keep = (m > 0).cpu().numpy()
    assert np.array_equal(keep, _np_drop_mask(7, 2, row0, B, W, rate))
    p = 1.0 - rate
    se = math.sqrt(p * (1 - p) / B)
    assert np.abs(keep.mean(0) - p).max() < 5.5 * se                            # every column
    assert abs(keep.mean() - p) < 5 * se / math.sqrt(W)
    k = keep.astype(np.float64) - p
    var = p * (1 - p)
    n = k[:, 0::2].size
    assert abs((k[:, 0::2] * k[:, 1::2]).mean() / var) < 5 / math.sqrt(n)       # the two halves of one hash
    assert abs((k[:-1] * k[1:]).mean() / var) < 5 / math.sqrt(k[1:].size)       # neighbouring rows
    assert abs((k[:, :-1] * k[:, 1:]).mean() / var) < 5 / math.sqrt(k[:, 1:].size)
    # other layer / other seed: unrelated masks
    for other in (dropout_mask(7, 3, B, W, rate, row0), dropout_mask(8, 2, B, W, rate, row0)):
        o = (other > 0).cpu().numpy().astype(np.float64) - p
        assert abs((k * o).mean() / var) < 5 / math.sqrt(k.size)
    part = dropout_mask(7, 2, 100, W, rate, row_offset=row0 + 31_000)
    assert torch.equal(part, m[31_000:31_100])


@pytest.mark.parametrize("W", [128, 512, 1024])   # block_fwd_kernel<1> / block_fwd_wide_kernel<2> / <4> and their backward twins
def test_hidden_dropout_mask_is_the_one_the_training_step_uses(W):
    """Loss and every parameter gradient of a training step WITH hidden Dropout against the oracle run with the exported
    masks (oracle/networks.py `dropout_masks`): the mask entry point returns what the forward AND backward kernels use."""
    from bayesflow_b200.training import dropout_mask
    D, C, B, rate = 2, 4, 700, 0.25
    net, orc = make_pair(D, C, (W,) * 2, dropout=rate)
    tup, c = make_batch(B, D, C)
    masks = [dropout_mask(seed=21, layer=l, rows=B, width=W, rate=rate).cpu() for l in range(2)]
    ps = {k: v.clone().requires_grad_(True) for k, v in orc.p.items()}
    orc.p = orc.subnet.p = ps
    ref_loss = orc.compute_metrics(tup, conditions=c, stage="training", dropout_masks=masks)["loss"]
    loss = net.compute_metrics(cuda_tuple(tup), conditions=c.cuda(), stage="training", seed=21)["loss"]
    ref, got = float(ref_loss), float(loss)
    assert abs(got - ref) <= 2e-3 * abs(ref), (got, ref)
    ref_loss.backward()
    loss.backward()  # the backward kernels regenerate the same mask
    checked = 0
    for name, p in net.named_parameters():
        if name in ps and ps[name].grad is not None:
            tol = 5e-3 if name.endswith("time_emb.frequencies") else TOL
            assert rel_l2(p.grad, ps[name].grad) < tol, (name, rel_l2(p.grad, ps[name].grad))
            checked += 1
    assert checked >= 10
    with torch.no_grad():
        other = float(net.compute_metrics(cuda_tuple(tup), conditions=c.cuda(), stage="training", seed=22)["loss"])
    assert abs(other - ref) > 1e-4 * abs(ref)


def test_cuda_graph_trainer_equals_the_eager_trainer_when_the_host_runs_ahead():
    """The graph trainer hands each replay its seed and learning rate through device words refreshed from pinned host
    memory; 60 steps enqueued with no host synchronisation must use 60 DIFFERENT seeds / schedule values - the same ones
    the eager trainer uses (torch_approximator.py:71-106 is one optimizer step; the graph is a replay of it)."""
    import bayesflow_b200 as bf
    from bayesflow_b200.training import Adam, Trainer, cosine_decay
    out = []
    x, c = torch.randn(4096, 2, device="cuda"), torch.randn(4096, 6, device="cuda")
    for graph in (False, True):
        torch.manual_seed(5)
        net = bf.FlowMatching(subnet_kwargs={"widths": (128,) * 2, "dropout": 0.1})
        net.build((None, 2), (None, 6))
        net.cuda()
        tr = Trainer(net, Adam(net, cosine_decay(2e-3, 60), clipnorm=1.5), world_size=1, cuda_graph=graph, seed=99)
        torch.cuda.synchronize()
        losses = [tr.train_step(x, c) for _ in range(60)]  # no .item() in the loop: the host runs ahead of the GPU
        torch.cuda.synchronize()
        out.append((float(losses[-1]), torch.cat([p.detach().flatten() for p in net.parameters()]).clone()))
    (l_e, p_e), (l_g, p_g) = out
    assert rel_l2(p_g, p_e) < 2e-3, rel_l2(p_g, p_e)
    assert abs(l_g - l_e) <= 2e-2 * abs(l_e), (l_g, l_e)
    # independent data: the loss cannot fall far below the conditional-variance floor (~1.4) unless noise is being reused
    assert l_g > 1.0, l_g


def test_adam_step_matches_keras_formula():
    """keras Adam with per-variable clipnorm (basic_workflow.py:524) restated in torch fp64 on the same gradients."""
    from bayesflow_b200.training import Adam, train_engine
    net, _ = make_pair(2, 6, (128,) * 2)
    eng = train_engine(net)
    opt = Adam(net, 1e-3, clipnorm=1.5)
    tup, c = make_batch(256, 2, 6)
    p_ref = [p.detach().double().cpu().clone() for p in eng.params]
    m = [torch.zeros_like(p) for p in p_ref]
    v = [torch.zeros_like(p) for p in p_ref]
    for it in range(1, 4):
        eng.step(tup[3].cuda(), tup[2].cuda(), c.cuda(), tup[4].cuda())
        grads = [eng.grad_of(p).detach().double().cpu() for p in eng.params]
        opt.step()
        lr_t = 1e-3 * math.sqrt(1 - 0.999 ** it) / (1 - 0.9 ** it)
        for i, g in enumerate(grads):
            n = float(g.norm())
            if n > 1.5:
                g = g * (1.5 / n)
            m[i] += (g - m[i]) * (1 - 0.9)
            v[i] += (g * g - v[i]) * (1 - 0.999)
            p_ref[i] -= lr_t * m[i] / (v[i].sqrt() + 1e-7)
        for p, r in zip(eng.params, p_ref):
            assert rel_l2(p, r) < 1e-5


def test_training_reduces_the_loss_and_sampler_sees_new_weights():
    import bayesflow_b200 as bf
    from bayesflow_b200.training import Adam, Trainer
    torch.manual_seed(1)
    net = bf.FlowMatching(subnet_kwargs={"widths": (128,) * 2, "dropout": 0.0})
    net.build((None, 2), (None, 3))
    net.cuda()
    tr = Trainer(net, Adam(net, 2e-3, clipnorm=1.5), world_size=1)
    c = torch.randn(2048, 3, device="cuda")
    x = c[:, :2] * 0.5 + 0.1 * torch.randn(2048, 2, device="cuda")
    z = torch.randn(64, 2, device="cuda")
    before = net(z, conditions=c[:64], inverse=True, method="euler", steps=8)
    l0 = float(tr.train_step(x, c))
    for _ in range(150):
        l = float(tr.train_step(x, c))
    assert l < 0.6 * l0, (l0, l)
    after = net(z, conditions=c[:64], inverse=True, method="euler", steps=8)
    assert float((after - before).abs().max()) > 1e-3  # the fused sampler re-packed the updated weights
    assert float((after - x[:64]).pow(2).mean()) < float((before - x[:64]).pow(2).mean())


def test_no_cpu_fallback_in_training():
    net, _ = make_pair(2, 6, (128,) * 2)
    tup, c = make_batch(64, 2, 6)
    with pytest.raises(RuntimeError):
        net.compute_metrics(tup, conditions=c)


# ---- DiffusionModel.compute_metrics (diffusion_model.py:142-232) ------------------------------------------------------
@pytest.mark.parametrize("pt,lt,sched", [
    ("F", "noise", "edm"),            # reference defaults
    ("velocity", "velocity", "cosine"),
    ("noise", "noise", "cosine"),
    ("x", "F", "edm"),
    ("F", "F", "edm"),
    ("velocity", "noise", "edm"),
])
def test_diffusion_loss_and_all_gradients(pt, lt, sched):
    D, C, B = 3, 6, 500
    net, orc = make_pair(D, C, (128,) * 2, dm=True,
                         model_kwargs={"prediction_type": pt, "loss_type": lt, "noise_schedule": sched})
    g = torch.Generator().manual_seed(77)
    x, c = torch.randn(B, D, generator=g), torch.randn(B, C, generator=g)
    u0, eps = torch.rand(1, generator=g), torch.randn(B, D, generator=g)
    w = torch.rand(B, generator=g) + 0.5
    ps = {k: v.clone().requires_grad_(True) for k, v in orc.p.items()}
    orc.p = orc.subnet.p = ps
    ref_loss = orc.compute_metrics(x, conditions=c, sample_weight=w, stage="training", fixed=(u0, eps))["loss"]
    ref_loss.backward()
    loss = net.compute_metrics(x.cuda(), conditions=c.cuda(), sample_weight=w.cuda(), stage="training",
                               fixed=(u0, eps))["loss"]
    loss.backward()
    assert abs(float(loss) - float(ref_loss)) <= 2e-3 * abs(float(ref_loss)), (float(loss), float(ref_loss))
    checked = 0
    for name, p in net.named_parameters():
        if name in ps and ps[name].grad is not None:
            # The Fourier frequencies see d(loss)/d(time embedding) amplified by 2*pi*fourier_scale and, under the
            # diffusion weighting, dominated by the few rows at the noisy end, so the fp16-operand rounding does not
            # average out over rows as it does for FlowMatching: 2.2e-3 measured on F/noise/edm, gate 5e-3 for that vector.
            tol = 5e-3 if name.endswith("time_emb.frequencies") else TOL
            assert rel_l2(p.grad, ps[name].grad) < tol, (name, rel_l2(p.grad, ps[name].grad))
            checked += 1
    assert checked >= 10


def test_diffusion_validation_and_rng_paths():
    net, orc = make_pair(2, 4, (128,) * 2, dm=True)
    g = torch.Generator().manual_seed(5)
    x, c = torch.randn(256, 2, generator=g), torch.randn(256, 4, generator=g)
    u0, eps = torch.rand(1, generator=g), torch.randn(256, 2, generator=g)
    ref = float(orc.compute_metrics(x, conditions=c, stage="validation", fixed=(u0, eps))["loss"])
    with torch.no_grad():
        got = net.compute_metrics(x.cuda(), conditions=c.cuda(), stage="validation", fixed=(u0, eps))["loss"]
    assert not got.requires_grad and abs(float(got) - ref) <= 2e-3 * abs(ref)
    # inference-stage schedule (stage="other") differs from the training one (diffusion_model.py:153)
    ref_i = float(orc.compute_metrics(x, conditions=c, stage="inference", fixed=(u0, eps))["loss"])
    got_i = float(net.compute_metrics(x.cuda(), conditions=c.cuda(), stage="inference", fixed=(u0, eps))["loss"])
    assert abs(got_i - ref_i) <= 2e-3 * abs(ref_i)
    loss = net.compute_metrics(x.cuda(), conditions=c.cuda())["loss"]  # device RNG path
    assert math.isfinite(float(loss)) and float(loss) > 0


@pytest.mark.parametrize("dm", [False, True])
@pytest.mark.parametrize("p_cond,p_target", [(0.3, 0.0), (0.0, 0.4), (0.25, 0.35)])
def test_condition_and_target_dropout_loss_and_gradients(dm, p_cond, p_target):
    """drop_cond_prob / drop_target_prob (flow_matching.py:167-175, diffusion_model.py:155-156, 177-216) with the SAME
    0 / 1 masks on both sides: loss and every gradient against the oracle (itself pinned on the reference's masks)."""
    D, C, B = 3, 6, 500
    mk = {"drop_cond_prob": p_cond, "drop_target_prob": p_target}
    net, orc = make_pair(D, C, (128,) * 2, dm=dm, model_kwargs=mk)
    g = torch.Generator().manual_seed(91)
    x, c = torch.randn(B, D, generator=g), torch.randn(B, C, generator=g)
    keep_c = (torch.rand(B, 1, generator=g) > p_cond).float() if p_cond > 0 else None
    keep_t = (torch.rand(B, D, generator=g) > p_target).float() if p_target > 0 else None
    mkw = {k: v for k, v in (("cond_keep", keep_c), ("target_keep", keep_t)) if v is not None}
    ps = {k: v.clone().requires_grad_(True) for k, v in orc.p.items()}
    orc.p = orc.subnet.p = ps
    if dm:
        u0, eps = torch.rand(1, generator=g), torch.randn(B, D, generator=g)
        ref_loss = orc.compute_metrics(x, conditions=c, stage="training", fixed=(u0, eps), **mkw)["loss"]
        loss = net.compute_metrics(x.cuda(), conditions=c.cuda(), stage="training", fixed=(u0, eps),
                                   **{k: v.cuda() for k, v in mkw.items()})["loss"]
    else:
        tup, _ = make_batch(B, D, None)
        ref_loss = orc.compute_metrics(tup, conditions=c, stage="training", **mkw)["loss"]
        loss = net.compute_metrics(cuda_tuple(tup), conditions=c.cuda(), stage="training",
                                   **{k: v.cuda() for k, v in mkw.items()})["loss"]
    ref_loss.backward()
    loss.backward()
    assert abs(float(loss) - float(ref_loss)) <= 2e-3 * abs(float(ref_loss)), (float(loss), float(ref_loss))
    checked = 0
    for name, p in net.named_parameters():
        if name in ps and ps[name].grad is not None:
            tol = 5e-3 if name.endswith("time_emb.frequencies") else TOL
            assert rel_l2(p.grad, ps[name].grad) < tol, (name, rel_l2(p.grad, ps[name].grad))
            checked += 1
    assert checked >= 10


def test_dropout_masks_are_bernoulli_and_keyed_by_the_global_row():
    import bayesflow_b200 as bf
    m = bf.FlowMatching._keep_mask(200_000, 4, 0.3, seed=5, row_offset=0, device=torch.device("cuda"))
    assert set(m.unique().tolist()) <= {0.0, 1.0}
    assert abs(float(m.mean()) - 0.7) < 0.005                                   # P(keep) = 1 - drop_prob
    part = bf.FlowMatching._keep_mask(1000, 4, 0.3, seed=5, row_offset=150_000, device=torch.device("cuda"))
    assert torch.equal(part, m[150_000:151_000])                                # a shard draws the rows of the whole batch
    # the RNG path end to end: runs, is seeded, and differs from the no-dropout loss
    net, _ = make_pair(2, 4, (128,) * 2, model_kwargs={"drop_cond_prob": 0.5, "drop_target_prob": 0.5})
    x, c = torch.randn(512, 2).cuda(), torch.randn(512, 4).cuda()
    a = float(net.compute_metrics(x, conditions=c, stage="training", noise_seed=3, seed=1)["loss"])
    b = float(net.compute_metrics(x, conditions=c, stage="training", noise_seed=3, seed=1)["loss"])
    d = float(net.compute_metrics(x, conditions=c, stage="training", noise_seed=4, seed=1)["loss"])
    # (the loss reduction uses float atomics: equal seeds agree to round-off, not bit for bit)
    assert abs(a - b) <= 1e-6 * abs(a) and abs(a - d) > 1e-4 * abs(a) and math.isfinite(a)
