"""bayesflow_b200 — B200-native (sm_100a) FlowMatching / DiffusionModel hot path behind the BayesFlow
``InferenceNetwork`` surface.  See DESIGN.md."""
from . import networks  # noqa: F401
from .networks import (  # noqa: F401
    FlowMatching, DiffusionModel, InferenceNetwork, TimeMLP, DiagonalNormal, EDMNoiseSchedule, CosineNoiseSchedule,
)
from .integrate import integrate  # noqa: F401
from .standardization import Standardize, Standardization  # noqa: F401
from .samplers import Sampler  # noqa: F401
from .consistency import ConsistencyModel, StableConsistencyModel  # noqa: F401
from .optimal_transport import optimal_transport, log_sinkhorn, log_sinkhorn_plan  # noqa: F401
from .parallel import dataset_shard, sample_sharded, log_prob_sharded, gather_rows  # noqa: F401

__version__ = "0.1.0"
