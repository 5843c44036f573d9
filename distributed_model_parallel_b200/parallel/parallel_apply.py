"""Thread-per-replica forward execution (the measured comparison point against multi-process DDP).

Behavioural model: torch/nn/parallel/parallel_apply.py:31-146 -- one Python thread per replica, each entering
its device, the caller's current stream on that device and the caller's grad / autocast modes; results are
collected in order and the first exception is re-raised with the replica's device in the message.
"""
from __future__ import annotations

import threading
from typing import Any, Dict, List, Optional, Sequence

import torch


class ReplicaError(RuntimeError):
    pass


def parallel_apply(modules: Sequence[torch.nn.Module], inputs: Sequence[Any], kwargs_tup: Optional[Sequence[Dict]] = None,
                   devices: Optional[Sequence] = None) -> List[Any]:
    assert len(modules) == len(inputs), "The number of modules %d is not equal to the number of inputs %d" % (len(modules), len(inputs))
    if kwargs_tup is None:
        kwargs_tup = ({},) * len(modules)
    assert len(kwargs_tup) == len(modules)
    if devices is None:
        devices = [None] * len(modules)
    devices = [torch.device(d) if d is not None and not isinstance(d, torch.device) else d for d in devices]
    streams = [torch.cuda.current_stream(d) if (d is not None and d.type == "cuda") else None for d in devices]
    lock = threading.Lock()
    results: Dict[int, Any] = {}
    grad_enabled = torch.is_grad_enabled()
    autocast_cuda = torch.is_autocast_enabled("cuda") if torch.cuda.is_available() else False
    autocast_dtype = torch.get_autocast_dtype("cuda") if torch.cuda.is_available() else None

    def _worker(i, module, inp, kwargs, device, stream):
        torch.set_grad_enabled(grad_enabled)
        try:
            if not isinstance(inp, (list, tuple)):
                inp = (inp,)
            if device is not None and device.type == "cuda":
                with torch.cuda.device(device), torch.cuda.stream(stream), \
                        torch.autocast("cuda", enabled=autocast_cuda, dtype=autocast_dtype):
                    out = module(*inp, **kwargs)
            else:
                out = module(*inp, **kwargs)
            with lock:
                results[i] = out
        except Exception as e:  # noqa: BLE001
            with lock:
                results[i] = ReplicaError("Caught %s in replica %d on device %s: %s" % (type(e).__name__, i, device, e))
                results[i].__cause__ = e

    if len(modules) > 1:
        threads = [threading.Thread(target=_worker, args=(i, m, inp, kw, d, s))
                   for i, (m, inp, kw, d, s) in enumerate(zip(modules, inputs, kwargs_tup, devices, streams))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    else:
        _worker(0, modules[0], inputs[0], kwargs_tup[0], devices[0], streams[0])

    outputs = []
    for i in range(len(inputs)):
        out = results[i]
        if isinstance(out, ReplicaError):
            raise out
        outputs.append(out)
    return outputs
