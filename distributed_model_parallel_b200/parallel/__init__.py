from . import comm  # noqa: F401
from .distributed import DistributedDataParallel, FusedSGD  # noqa: F401
from .reducer import Reducer, compute_bucket_assignment_by_size  # noqa: F401
