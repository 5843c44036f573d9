"""Parallelism wrappers with torch's call shapes: ``DataParallel(module, device_ids)``, ``DistributedDataParallel(module)``,
``SyncBatchNorm`` / ``convert_sync_batchnorm``, the scatter / gather / replicate / parallel_apply building blocks, and the
``pipeline`` sub-package for layer-split model parallelism."""
from . import comm  # noqa: F401
from .data_parallel import DataParallel, data_parallel  # noqa: F401
from .distributed import DistributedDataParallel, FusedSGD  # noqa: F401
from .parallel_apply import parallel_apply  # noqa: F401
from .reducer import Reducer, compute_bucket_assignment_by_size  # noqa: F401
from .replicate import replicate  # noqa: F401
from .scatter_gather import gather, scatter, scatter_kwargs  # noqa: F401
from .sync_batchnorm import SyncBatchNorm, convert_sync_batchnorm  # noqa: F401
