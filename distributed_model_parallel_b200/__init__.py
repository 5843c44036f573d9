"""distributed_model_parallel_b200 -- a Blackwell-native data/model-parallel training library.

Public surface (same call shapes as the torch machinery the reference runs/studies):

* ``DataParallel(module, device_ids)``            single-process, thread-per-replica      (parallel/data_parallel.py)
* ``DistributedDataParallel(module, ...)``        multi-process, bucketed C++ Reducer     (parallel/distributed.py)
* ``pipeline`` (partition / p2p / schedules)      layer-split model parallel, micro-batched (parallel/pipeline/)
* ``SyncBatchNorm``, ``FusedSGD``, model zoo, synthetic data, trainers.
"""
__version__ = "0.1.0"

from . import models, ops  # noqa: F401
