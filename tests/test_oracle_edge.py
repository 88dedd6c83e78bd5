"""CPU tests for the approximator-edge oracle (oracle/edge.py): Philox known answers, and the reference's own
Standardize tests (tests/test_networks/test_standardization.py) restated against the oracle."""
import numpy as np
import pytest
import torch

import oracle


# Random123 kat_vectors, philox4x32 10 rounds: counter | key | expected
PHILOX_KAT = [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


@pytest.mark.parametrize("ctr,key,want", PHILOX_KAT)
def test_philox_known_answers(ctr, key, want):
    got = oracle.philox4x32_10(np.array([ctr], dtype=np.uint32), np.array(key, dtype=np.uint32))[0]
    assert tuple(int(v) for v in got) == want


def test_philox_words_offsets_are_slices_of_one_stream():
    full = oracle.philox_words(64, seed=1234567890123)
    for off, n in [(0, 64), (1, 7), (4, 8), (13, 30), (63, 1)]:
        assert np.array_equal(oracle.philox_words(n, 1234567890123, off), full[off:off + n])
    assert not np.array_equal(full, oracle.philox_words(64, seed=1234567890124))


def test_normal_stream_moments_and_sharding():
    z = oracle.normal_stream(50000, 4, seed=11)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1.0) < 0.01
    # excess kurtosis and tails of a standard normal
    assert abs((z ** 4).mean() - 3.0) < 0.1 and 0.0015 < (np.abs(z) > 3).mean() < 0.004
    part = oracle.normal_stream(100, 4, seed=11, row_offset=777)
    assert np.array_equal(part, z[777:877])
    shifted = oracle.normal_stream(100, 4, seed=11, row_offset=777, mean=[1, 2, 3, 4], std=[1, 2, 1, 0.5])
    assert np.allclose(shifted, part * np.array([1, 2, 1, 0.5]) + np.array([1, 2, 3, 4]))


# ---- reference tests for Standardize, restated (test_standardization.py) ------------------------------------------------
def test_forward_standardization_training():
    # test_standardization.py:11-30
    x = torch.randn(8, 4)
    layer = oracle.OracleStandardize(4)
    out = layer(x, stage="training")
    np.testing.assert_allclose(layer.moving_mean.numpy(), x.numpy().mean(0), atol=1e-5)
    assert out.shape == x.shape and not torch.isnan(out).any()
    np.testing.assert_allclose(out.numpy().std(0), 1.0, atol=1e-5)


def test_forward_standardization_training_constant_batch():
    # test_standardization.py:33-53: M2 == 0 -> std 1, output 0
    x = torch.ones(8, 4)
    layer = oracle.OracleStandardize(4)
    out = layer(x, stage="training")
    np.testing.assert_allclose(layer.moving_mean.numpy(), 1.0, atol=1e-5)
    np.testing.assert_allclose(out.numpy(), 0.0, atol=1e-5)


def test_inverse_standardization_ldj():
    # test_standardization.py:56-69
    x = torch.randn(1, 3)
    layer = oracle.OracleStandardize(3)
    layer(x, stage="training", forward=True)
    inv, ldj = layer(x, stage="inference", forward=False, log_det_jac=True)
    assert inv.shape == x.shape and ldj.shape == x.shape[:-1]


def test_consistency_forward_inverse():
    # test_standardization.py:72-89
    x = torch.randn(4, 20, 5)
    layer = oracle.OracleStandardize(5)
    layer(x, stage="training", forward=True)
    rec = layer(layer(x, stage="inference", forward=True), stage="inference", forward=False)
    np.testing.assert_allclose(x.numpy(), rec.numpy(), atol=1e-4)


def test_two_training_passes_equal_whole_data_moments():
    # test_standardization.py:121-170 (nested_accuracy_forward): after two passes the combined data has mean 0 / std 1
    a1, a2 = torch.randn(2, 3, 5) * 3 + 1, torch.randn(3, 3, 5) * 3 + 1
    layer = oracle.OracleStandardize(5)
    layer(a1, stage="training")
    layer(a2, stage="training")
    out = layer(torch.cat([a1, a2], 0), stage="inference").numpy()
    np.testing.assert_allclose(out.mean((0, 1)), 0.0, atol=1e-4)
    np.testing.assert_allclose(out.std((0, 1)), 1.0, atol=1e-4)
    assert float(layer.count) == 15.0


def test_non_default_transformation_is_rejected_forward():
    # standardize.py:79-83
    layer = oracle.OracleStandardize(3)
    with pytest.raises(ValueError):
        layer(torch.randn(2, 3), forward=True, transformation_type="both_sides_scale")


def test_repeat_and_flatten_is_dataset_major():
    # samplers.py:51-69
    c = torch.arange(6.0).reshape(3, 2)
    flat = oracle.repeat_and_flatten_conditions(c, 4)
    assert flat.shape == (12, 2) and torch.equal(flat[4:8], c[1].expand(4, 2))
    assert torch.equal(oracle.unflatten_samples(flat, 4)[:, 0], c)


# ---- ConsistencyModel oracle anchors (closed-form properties of the parameterisation) ---------------------------------
def test_consistency_boundary_condition_and_discretisation_end_points():
    D, Cd = 3, 4
    params = oracle.perturb_params(oracle.init_params(D, Cd, (64, 64)))
    cm = oracle.OracleConsistencyModel(params, D, total_steps=100, subnet_kwargs={"widths": (64, 64), "dropout": 0.0})
    x, c = torch.randn(7, D), torch.randn(7, Cd)
    # f(x, eps) = x: c_skip(eps) = 1, c_out(eps) = 0 (Song et al. 2023 eq. 5; consistency_model.py:336-342)
    assert torch.allclose(cm.consistency_function(x, cm.eps, c), x, atol=1e-6)
    dt = cm.discretize_time(11, dtype=torch.float64)
    assert abs(float(dt[0]) - cm.eps) < 1e-12 and abs(float(dt[-1]) - cm.max_time) < 1e-9 and bool((dt[1:] > dt[:-1]).all())
    # N(k) doubles from s0 + 1 to s1 + 1 (consistency_model.py:153-162, :201-203)
    assert cm.schedule_discretization(0) == 11.0 and cm.schedule_discretization(100) == 151.0
    assert sorted({int(cm.schedule_discretization(s)) for s in range(100)}) == [11, 21, 41, 81, 151]
    with pytest.raises(ValueError):
        oracle.OracleConsistencyModel(params, D, total_steps=5)


def test_consistency_loss_and_one_step_sampling_are_finite():
    D = 2
    params = oracle.perturb_params(oracle.init_params(D, None, (64, 64)))
    cm = oracle.OracleConsistencyModel(params, D, total_steps=100, subnet_kwargs={"widths": (64, 64), "dropout": 0.0})
    x = torch.randn(16, D)
    idx, noise = torch.randint(0, 9, (16,)), torch.randn(16, D)
    loss = cm.compute_metrics(x, stage="validation", fixed=(idx, noise))["loss"]
    assert float(loss) > 0 and torch.isfinite(loss)
    # two-step sampling (steps=1 divides 0 by 0 in _discretize_time, in the reference too): c_skip(T) ~ 1.6e-4, so the draw is c_out(T) * f, not the T-scaled input
    out = cm.inverse(torch.randn(64, D), steps=2)
    assert out.shape == (64, D) and bool(torch.isfinite(out).all())


# ---- optimal transport oracle, pinned by the reference's closed-form tests (test_optimal_transport.py) ------------------
def test_log_sinkhorn_plan_correct_marginals():
    # test_optimal_transport.py:215-224
    from oracle.optimal_transport import log_sinkhorn_plan
    x1, x2 = torch.randn(10, 2), torch.randn(20, 2)
    lp = log_sinkhorn_plan(x1, x2, atol=1e-5, max_steps=1000)
    assert torch.allclose(torch.logsumexp(lp, 0), torch.full((20,), -np.log(20.0)), atol=1e-3)
    assert torch.allclose(torch.logsumexp(lp, 1), torch.full((10,), -np.log(10.0)), atol=1e-3)


def test_log_sinkhorn_plan_matches_analytical_result():
    # test_optimal_transport.py:249-263: identical points -> outer product of the marginals
    from oracle.optimal_transport import log_sinkhorn_plan
    plan = torch.exp(log_sinkhorn_plan(torch.ones(16), torch.ones(64), regularization=0.1))
    np.testing.assert_allclose(plan.numpy(), np.full((16, 64), 1 / (16 * 64)), rtol=1e-4)


def test_assignment_of_a_shuffled_copy_is_recovered():
    # test_optimal_transport.py:98-116
    from oracle.optimal_transport import optimal_transport
    y = torch.randn(16, 2)
    p = torch.randperm(16)
    _, _, _, a = optimal_transport(y[p], y, regularization=0.01, seed=0, max_steps=10_000, return_assignments=True)
    assert int((a == p).sum()) >= 14


def test_conditional_weight_search_hits_the_ratio():
    # ot_utils.py:95-164: mean((M + w C) <= diag(M)) ~ condition_ratio
    from oracle.optimal_transport import squared_euclidean, cosine_distance, search_for_conditional_weight
    x0, x1, c = torch.randn(64, 2), torch.randn(64, 2), torch.randn(64, 5)
    M, Cm = squared_euclidean(x0, x1), cosine_distance(c, c)
    cost, w = search_for_conditional_weight(M, Cm, 0.1)
    ratio = float((cost <= M.diagonal().unsqueeze(1)).float().mean())
    assert w > 0 and abs(ratio - 0.1) < 0.02


@pytest.mark.parametrize("s", [0.3, 0.8])
def test_partial_ot_leaves_unmatched_mass(s):
    # test_optimal_transport.py:289-327
    from oracle.optimal_transport import log_sinkhorn_plan
    x, y = torch.randn(20, 2), torch.randn(20, 2)
    plan = torch.exp(log_sinkhorn_plan(x, y, regularization=0.1, max_steps=10_000, partial_factor=s))
    assert plan.shape == (20, 20)
    assert bool(((plan.sum(1) - s / 20).abs() < 0.05).all()) and bool(((plan.sum(0) - s / 20).abs() < 0.05).all())
    assert abs(float(plan.sum()) - s) < 1e-3
