"""scatter / gather tree recursion and the DataParallel wrapper on CPU "devices" (container logic without GPUs)."""
import collections
import copy

import pytest
import torch
import torch.nn as nn

from distributed_model_parallel_b200.parallel import comm
from distributed_model_parallel_b200.parallel.data_parallel import DataParallel, data_parallel
from distributed_model_parallel_b200.parallel.parallel_apply import ReplicaError, parallel_apply
from distributed_model_parallel_b200.parallel.replicate import replicate
from distributed_model_parallel_b200.parallel.scatter_gather import gather, scatter, scatter_kwargs

CPU2 = [torch.device("cpu"), torch.device("cpu")]
Pair = collections.namedtuple("Pair", ["a", "b"])


def test_scatter_tree_recursion():
    x = torch.arange(12.).view(6, 2)
    out = scatter((x, [x, 7], {"k": x, "s": "str"}, Pair(x, x)), CPU2)
    assert len(out) == 2
    for i, part in enumerate(out):
        xs = x.chunk(2)[i]
        t, lst, dct, pair = part
        assert torch.equal(t, xs) and torch.equal(lst[0], xs) and lst[1] == 7
        assert torch.equal(dct["k"], xs) and dct["s"] == "str"
        assert isinstance(pair, Pair) and torch.equal(pair.b, xs)
    ins, kws = scatter_kwargs((x,), {}, CPU2)
    assert len(ins) == len(kws) == 2 and kws[0] == {}
    # fewer rows than devices -> fewer chunks
    assert len(scatter((torch.zeros(1, 3),), CPU2)) == 1


def test_gather_tree_and_scalars():
    a, b = torch.ones(2, 3), torch.zeros(4, 3)
    assert gather([a, b], "cpu").shape == (6, 3)
    g = gather([{"x": a, "y": (a, None)}, {"x": b, "y": (b, None)}], "cpu")
    assert g["x"].shape == (6, 3) and g["y"][1] is None
    with pytest.warns(UserWarning):
        s = gather([torch.tensor(1.0), torch.tensor(2.0)], "cpu")
    assert s.tolist() == [1.0, 2.0]
    p = gather([Pair(a, a), Pair(b, b)], "cpu")
    assert isinstance(p, Pair) and p.a.shape == (6, 3)


def test_scatter_gather_autograd_round_trip():
    x = torch.randn(6, 4, requires_grad=True)
    parts = scatter((x,), CPU2)
    y = gather([p[0] * (i + 1) for i, p in enumerate(parts)], "cpu")
    y.sum().backward()
    assert torch.equal(x.grad, torch.cat([torch.ones(3, 4), 2 * torch.ones(3, 4)]))


def test_coalesced_collectives_cpu():
    ts = [torch.randn(5, 3), torch.randn(7), torch.arange(4), torch.randn(100)]
    outs = comm.broadcast_coalesced(ts, CPU2, buffer_size=64)
    assert all(torch.equal(a, b) for a, b in zip(ts, outs[1]))
    assert outs[0][0] is ts[0]
    red = comm.reduce_add_coalesced([ts[:2], [t * 2 for t in ts[:2]]], "cpu", buffer_size=32)
    assert all(torch.allclose(r, 3 * t) for r, t in zip(red, ts[:2]))
    chunks = list(comm.take_tensors(ts, 40))
    assert sum(len(c) for c in chunks) == len(ts) and all(len({t.dtype for t in c}) == 1 for c in chunks)
    flat = comm.flatten_dense(ts[:2], pad_to=8)
    assert flat.numel() % 8 == 0
    assert all(torch.equal(a, b) for a, b in zip(comm.unflatten_dense(flat, ts[:2]), ts[:2]))


def test_replicate_shares_replica0_and_is_differentiable():
    m = nn.Sequential(nn.Linear(4, 4), nn.BatchNorm1d(4))
    reps = replicate(m, CPU2)
    assert len(reps) == 2
    assert reps[0][1].running_mean.data_ptr() == m[1].running_mean.data_ptr()      # BN stats persist through replica 0
    assert reps[1][0].weight.data_ptr() != m[0].weight.data_ptr()
    assert not isinstance(reps[1][0].weight, nn.Parameter)                           # non-leaf copies
    (reps[0](torch.randn(3, 4)).sum() + reps[1](torch.randn(3, 4)).sum()).backward()
    assert m[0].weight.grad is not None
    det = replicate(m, CPU2, detach=True)
    assert not det[1][0].weight.requires_grad


def test_parallel_apply_propagates_exceptions():
    class Boom(nn.Module):
        def forward(self, x):
            raise ValueError("bad replica")

    ok = nn.Identity()
    outs = parallel_apply([ok, ok], [(torch.ones(1),), (torch.zeros(1),)], None, CPU2)
    assert outs[0].item() == 1 and outs[1].item() == 0
    with pytest.raises(ReplicaError, match="replica 1"):
        parallel_apply([ok, Boom()], [(torch.ones(1),), (torch.ones(1),)], None, CPU2)


def test_data_parallel_matches_module_and_state_dict_prefix():
    torch.manual_seed(0)
    m = nn.Sequential(nn.Linear(8, 16), nn.ReLU(), nn.Linear(16, 4))
    r = copy.deepcopy(m)
    dp = DataParallel(m, device_ids=CPU2)
    x = torch.randn(6, 8)
    out = dp(x)
    out.sum().backward()
    o2 = r(x)
    o2.sum().backward()
    assert torch.allclose(out, o2, atol=1e-6)
    for a, b in zip(m.parameters(), r.parameters()):
        assert torch.allclose(a.grad, b.grad, atol=1e-6)
    assert list(dp.state_dict().keys()) == ["module." + k for k in r.state_dict().keys()]
    with torch.no_grad():
        assert torch.allclose(data_parallel(m, x, CPU2), o2, atol=1e-6)
    single = DataParallel(copy.deepcopy(r), device_ids=[torch.device("cpu")])
    assert torch.allclose(single(x), o2, atol=1e-6)
