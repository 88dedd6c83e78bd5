"""CPU oracle for the BayesFlow FlowMatching / DiffusionModel hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``bayesflow_b200/`` imports this package; it is used by
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``
as the checker and the timed CPU baseline, never as the product path.

It is a plain-PyTorch FP32 restatement of the reference (bayesflow v2.0.11, /root/reference) — the
reference itself cannot be imported here because Keras is not installed.  Each function cites
the reference file:line it follows.

Parity status: the INTEGRATOR is pinned against the reference's own closed-form tests
(tests/test_utils/test_integrate.py:15-25, 47-93; restated in tests/test_oracle_integrate.py).  The NETWORK
numerics (Keras Dense / LayerNormalization / mish semantics) have no golden vectors in the reference:
"parity unpinned" at the Keras boundary; the Keras semantics assumed are listed in oracle/networks.py.
"""
from .integrate import integrate, integrate_fixed, integrate_adaptive, integrate_scheduled  # noqa: F401
from .networks import (  # noqa: F401
    OracleTimeMLP, OracleFlowMatching, OracleDiffusionModel, EDMNoiseSchedule, CosineNoiseSchedule,
    DiagonalNormal, init_params, perturb_params, count_flops_per_eval,
)
from .edge import (  # noqa: F401
    OracleStandardize, philox4x32_10, philox_words, normal_stream, repeat_and_flatten_conditions, unflatten_samples,
)
from .consistency import OracleConsistencyModel, OracleStableConsistencyModel  # noqa: F401
