"""distributed_model_parallel_b200 -- a Blackwell-native data/model-parallel training library.

Public surface (same call shapes as the torch machinery the reference runs/studies):

* ``DataParallel(module, device_ids)``            single-process, thread-per-replica      (parallel/data_parallel.py)
* ``DistributedDataParallel(module, ...)``        multi-process, bucketed C++ Reducer     (parallel/distributed.py)
* ``pipeline`` (partition / p2p / schedules)      layer-split model parallel, micro-batched (parallel/pipeline/)
* ``SyncBatchNorm``, ``FusedSGD``, model zoo, synthetic data, trainers.
"""
__version__ = "0.1.0"

from . import models, ops  # noqa: F401

_LAZY = {
    "DataParallel": ("parallel", "DataParallel"), "data_parallel": ("parallel", "data_parallel"),
    "DistributedDataParallel": ("parallel", "DistributedDataParallel"), "FusedSGD": ("parallel", "FusedSGD"),
    "SyncBatchNorm": ("parallel", "SyncBatchNorm"), "convert_sync_batchnorm": ("parallel", "convert_sync_batchnorm"),
    "FlatFusedSGD": ("optim", "FlatFusedSGD"), "LinearWarmup": ("optim", "LinearWarmup"), "TrainStep": ("train.engine", "TrainStep"),
}


def __getattr__(name):
    """``from distributed_model_parallel_b200 import DistributedDataParallel`` etc.; resolved on first use so that importing the
    package stays cheap (the parallel layer pulls in torch.distributed and the native runtime)."""
    import importlib

    if name in ("parallel", "pipeline", "optim", "train", "data", "utils"):
        return importlib.import_module("." + ("parallel.pipeline" if name == "pipeline" else name), __name__)
    if name in _LAZY:
        mod, attr = _LAZY[name]
        return getattr(importlib.import_module("." + mod, __name__), attr)
    raise AttributeError("module %r has no attribute %r" % (__name__, name))
