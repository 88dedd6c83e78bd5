"""GPU parity tests (through the C ABI) for the approximator edge of the sampling path:
bfk_normal_fill, bfk_standardize, bfk_batch_moments / bfk_moments_merge, and the `Sampler` that passes (N, C)
conditions without materialising the repeat.  Checker: oracle/edge.py."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle
from _util import make_pair, rel_l2

pytestmark = pytest.mark.gpu


def _fill(rows, D, seed, row_offset=0, mean=None, std=None, raw=False, misalign=0):
    from bayesflow_b200 import _lib
    buf = torch.full((rows * D + misalign,), float("nan"), device="cuda")
    out = buf[misalign:]
    words = torch.zeros(rows * D, dtype=torch.int32, device="cuda") if raw else None
    m = None if mean is None else torch.tensor(mean, dtype=torch.float32, device="cuda")
    s = None if std is None else torch.tensor(std, dtype=torch.float32, device="cuda")
    _lib.check(_lib.lib().bfk_normal_fill(C.c_void_p(out.data_ptr()), rows, D, row_offset, _lib.ptr(m), _lib.ptr(s),
                                          C.c_uint64(seed), None if words is None else C.c_void_p(words.data_ptr()),
                                          _lib.stream_ptr()), "bfk_normal_fill")
    torch.cuda.synchronize()
    z = out.cpu().numpy().reshape(rows, D)
    return (z, words.cpu().numpy().view(np.uint32)) if raw else z


@pytest.mark.parametrize("rows,D,off", [(1, 1, 0), (5, 3, 0), (1000, 2, 0), (333, 7, 41), (64, 64, 3), (4097, 4, 1 << 29)])
def test_normal_fill_words_bit_exact_and_normals_match_oracle(rows, D, off):
    seed = 0x1234_5678_9ABC_DEF0
    z, words = _fill(rows, D, seed, off, raw=True)
    assert np.array_equal(words, oracle.philox_words(rows * D, seed, off * D))  # integer work: bit-exact
    ref = oracle.normal_stream(rows, D, seed, off)
    assert not np.isnan(z).any()
    np.testing.assert_allclose(z, ref, atol=2e-5, rtol=1e-5)  # fp32 log/sincos vs float64


def test_normal_fill_known_answer_counter_zero():
    # Random123 KAT: counter 0, key 0 -> 6627e8d5 e169c58d bc57ac4c 9b00dbd8
    _, words = _fill(1, 4, 0, raw=True)
    assert [int(w) for w in words] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]


def test_normal_fill_shards_equal_the_single_stream_and_misaligned_output():
    full = _fill(4000, 3, seed=99)
    lo = _fill(1333, 3, seed=99, row_offset=0)
    hi = _fill(4000 - 1333, 3, seed=99, row_offset=1333, misalign=1)
    assert np.array_equal(np.concatenate([lo, hi]), full)
    scaled = _fill(4000, 3, seed=99, mean=[1.0, -2.0, 0.5], std=[2.0, 1.0, 0.25])
    np.testing.assert_allclose(scaled, full * np.array([2.0, 1.0, 0.25]) + np.array([1.0, -2.0, 0.5]), atol=1e-6)


def test_normal_fill_moments_at_full_size():
    # 4M values: mean / std / kurtosis / tail mass of a standard normal (statistical parity, cf. test_integrate.py:96-222)
    z = _fill(1_000_000, 4, seed=5)
    assert abs(z.mean()) < 2e-3 and abs(z.std() - 1) < 2e-3
    assert abs((z.astype(np.float64) ** 4).mean() - 3.0) < 0.02
    assert 0.0024 < (np.abs(z) > 3).mean() < 0.0030
    assert abs(np.corrcoef(z[:, 0], z[:, 1])[0, 1]) < 3e-3 and abs(np.corrcoef(z[:-1, 0], z[1:, 0])[0, 1]) < 3e-3


@pytest.mark.parametrize("shape", [(8, 4), (1, 3), (4, 20, 5), (1000, 2), (4099, 33), (513, 64), (70, 300), (3, 1024)])
def test_standardize_training_matches_oracle(shape):
    import bayesflow_b200 as bf
    g = torch.Generator().manual_seed(sum(shape))
    F = shape[-1]
    scale, shift = torch.rand(F, generator=g) * 5 + 0.1, torch.randn(F, generator=g) * 10
    layer = bf.Standardize()
    layer.build(shape)
    layer.cuda()
    ref = oracle.OracleStandardize(F, dtype=torch.float64)
    for _ in range(3):  # running totals over three batches (standardize.py:159-170)
        x = torch.randn(*shape, generator=g) * scale + shift
        out = layer(x.cuda(), stage="training")
        want = ref(x.double(), stage="training")
        np.testing.assert_allclose(layer._buf(layer.moving_mean, 0).cpu(), ref.moving_mean, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(layer._buf(layer.moving_m2, 0).cpu(), ref.moving_m2, rtol=2e-4, atol=1e-5)
        assert float(layer._buf(layer.count, 0)) == float(ref.count)
        if ref.count > 1:
            np.testing.assert_allclose(out.cpu(), want, rtol=1e-4, atol=2e-4)
    x = torch.randn(*shape, generator=g) * scale + shift
    z = layer(x.cuda(), stage="inference", forward=True)
    back, ldj = layer(z, stage="inference", forward=False, log_det_jac=True)
    np.testing.assert_allclose(back.cpu(), x, rtol=1e-5, atol=1e-4)  # test_standardization.py:72-89
    _, ldj_ref = ref(x.double(), forward=False, log_det_jac=True)
    assert ldj.shape == x.shape[:-1]
    np.testing.assert_allclose(ldj.cpu(), ldj_ref, rtol=1e-5, atol=1e-5)


def test_standardize_reference_tests_on_the_kernels():
    import bayesflow_b200 as bf
    # test_standardization.py:11-30
    x = torch.randn(8, 4).cuda()
    std = bf.Standardization(standardize="all")
    std.build(dict(inference_variables=x.shape))
    std.cuda()
    layer = std.standardize_layers["inference_variables"]
    out = layer(x, stage="training").cpu().numpy()
    np.testing.assert_allclose(layer._buf(layer.moving_mean, 0).cpu(), x.cpu().numpy().mean(0), atol=1e-5)
    np.testing.assert_allclose(out.std(0), 1.0, atol=1e-5)
    # :33-53 constant batch -> std 1, output 0
    const = bf.Standardization(standardize="a")
    const.build(dict(a=(8, 4)))
    const.cuda()
    out = const.maybe_standardize(torch.ones(8, 4).cuda(), key="a", stage="training").cpu().numpy()
    np.testing.assert_allclose(out, 0.0, atol=1e-5)
    # keys that are not standardized pass through (standardization.py:44-52)
    y = torch.randn(3, 2).cuda()
    assert const.maybe_standardize(y, key="b") is y and const.maybe_standardize(None, key="a", log_det_jac=True) == (None, 0.0)
    with pytest.raises(ValueError):
        layer(x, forward=True, transformation_type="left_side_scale")
    with pytest.raises(RuntimeError):
        layer(x.cpu())  # no CPU fallback
    # nested input: one set of running totals per leaf, sorted by key (standardize.py:39-48, 86-90)
    nested = bf.Standardize()
    nested.build({"b": (4, 3), "a": (2, 3, 5)})
    nested.cuda()
    data = {"b": torch.randn(4, 3).cuda() * 2 + 1, "a": torch.randn(2, 3, 5).cuda() - 4}
    nested(data, stage="training")
    rec = nested(nested(data, stage="inference"), stage="inference", forward=False)
    assert nested._buf(nested.moving_mean, 0).shape == (5,) and float(nested._buf(nested.count, 0)) == 6.0
    for k in data:
        np.testing.assert_allclose(rec[k].cpu(), data[k].cpu(), atol=1e-4)


def test_moments_merge_of_rank_records_equals_one_big_batch():
    import bayesflow_b200 as bf
    from bayesflow_b200 import _lib
    g = torch.Generator().manual_seed(4)
    x = (torch.randn(6000, 10, generator=g) * 3 + 7).cuda()
    whole, parts = bf.Standardize(), bf.Standardize()
    for m in (whole, parts):
        m.build((None, 10))
        m.cuda()
    whole._update_moments(x, 0)
    recs = torch.stack([parts.batch_moments(x[lo:lo + 1500], 0) for lo in range(0, 6000, 1500)])  # 4 "ranks"
    _lib.check(_lib.lib().bfk_moments_merge(_lib.ptr(parts._buf(parts.moving_mean, 0)), _lib.ptr(parts._buf(parts.moving_m2, 0)),
                                            _lib.ptr(parts._buf(parts.count, 0)), _lib.ptr(recs.contiguous()), 4, 10,
                                            _lib.stream_ptr()), "merge")
    np.testing.assert_allclose(parts._buf(parts.moving_mean, 0).cpu(), whole._buf(whole.moving_mean, 0).cpu(), rtol=1e-6)
    np.testing.assert_allclose(parts._buf(parts.moving_m2, 0).cpu(), whole._buf(whole.moving_m2, 0).cpu(), rtol=1e-5)
    assert float(parts._buf(parts.count, 0)) == 6000.0
    # bit-reproducible: fixed merge order, no atomics
    again = bf.Standardize()
    again.build((None, 10))
    again.cuda()
    again._update_moments(x, 0)
    assert torch.equal(again._buf(again.moving_m2, 0), whole._buf(whole.moving_m2, 0))


@pytest.mark.parametrize("D,Cd,widths", [(2, 32, (256,) * 5), (16, 8, (512,) * 2)])
def test_sampler_without_materialised_conditions_matches_oracle(D, Cd, widths):
    """Sampler.sample((N, C), num_samples) == oracle integration of the SAME z on the repeated conditions
    (samplers.py:51-63, 119-148); fused (cfg2 shape) and layer-wise engines."""
    import bayesflow_b200 as bf
    net, orc = make_pair(D, Cd, widths=widths)
    N, S, seed = 5, 64, 321
    cond = torch.randn(N, Cd)
    draws = bf.Sampler().sample(net, S, conditions=cond.cuda(), seed=seed, method="euler", steps=8)
    assert draws.shape == (N, S, D)
    z = torch.from_numpy(oracle.normal_stream(N * S, D, seed)).float()
    ref = orc.inverse(z, oracle.repeat_and_flatten_conditions(cond, S), method="euler", steps=8)
    assert rel_l2(draws.reshape(N * S, D), ref) < 2e-3
    # mini-batching over datasets draws the same noise rows (row_offset), so the result does not depend on batch_size
    split = bf.Sampler().sample(net, S, conditions=cond.cuda(), seed=seed, batch_size=2, method="euler", steps=8)
    assert rel_l2(split, draws) < 1e-6
    # inverse standardization fused behind the draws (continuous_approximator's sample path)
    st = bf.Standardize()
    st.build((None, D))
    st.cuda()
    theta = (torch.randn(500, D) * 4 + 2).cuda()
    st(theta, stage="training")
    scaled = bf.Sampler().sample(net, S, conditions=cond.cuda(), seed=seed, standardize=st, method="euler", steps=8)
    want = draws * st.moving_std(0) + st._buf(st.moving_mean, 0)
    assert rel_l2(scaled, want) < 1e-6
