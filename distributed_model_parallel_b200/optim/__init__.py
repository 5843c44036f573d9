from .fused_sgd import FlatFusedSGD  # noqa: F401
from .warmup import LinearWarmup  # noqa: F401
