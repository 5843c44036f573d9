"""DistributedDataParallel on CPU/gloo, world_size 2 (BASELINE.json config 1): parity with torch DDP."""
import copy

import pytest
import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from _dist_utils import run_gloo


def _small_net():
    from distributed_model_parallel_b200.models import ResNet18
    return ResNet18()


def _w_parity(rank, world):
    from distributed_model_parallel_b200.parallel import DistributedDataParallel as DDP

    torch.manual_seed(rank)  # different initial weights per rank: init sync must make them equal
    base = _small_net()
    ours = DDP(copy.deepcopy(base), bucket_cap_mb=5)
    ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base), bucket_cap_mb=5)
    o1 = torch.optim.SGD(ours.parameters(), lr=0.1, momentum=0.9)
    o2 = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
    for it in range(3):
        torch.manual_seed(100 + 10 * rank + it)
        x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 10, (4,))
        for m, o in ((ours, o1), (ref, o2)):
            o.zero_grad()
            F.cross_entropy(m(x), y).backward()
            o.step()
        for (n, p), q in zip(ours.module.named_parameters(), ref.module.parameters()):
            assert torch.allclose(p.grad, q.grad, atol=1e-6, rtol=1e-5), (it, n)
            assert torch.allclose(p, q, atol=1e-6, rtol=1e-5), (it, n)
    assert ours.reducer.stats["rebuilds"] == 1
    assert len(ours.reducer.buckets) > 1
    info = ours._get_ddp_logging_data()
    assert info["world_size"] == world and info["num_buckets"] == len(ours.reducer.buckets) and info["iterations"] == 3
    assert sum(info["bucket_sizes"]) >= info["total_parameter_bytes"] and info["rebuilds"] == 1
    assert all(k.startswith("module.") for k in ours.state_dict())
    # buffers (BN running stats) are re-broadcast from rank 0 at every forward, exactly like torch's wrapper
    for (n, b), c in zip(ours.module.named_buffers(), ref.module.buffers()):
        assert torch.allclose(b.float(), c.float(), atol=1e-6), n


def test_ddp_matches_torch_ddp_resnet18():
    run_gloo(_w_parity, 2)


class _Branchy(nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b, self.c = nn.Linear(8, 8), nn.Linear(8, 8), nn.Linear(8, 4)

    def forward(self, x, use_b=True):
        h = self.a(x)
        if use_b:
            h = self.b(h)
        return self.c(h)


def _w_features(rank, world):
    from distributed_model_parallel_b200.parallel import DistributedDataParallel as DDP

    torch.manual_seed(0)
    base = _Branchy()
    # --- no_sync accumulates locally, the next synced backward reduces the sum
    ours = DDP(copy.deepcopy(base))
    ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
    torch.manual_seed(rank + 1)
    xs = [torch.randn(3, 8) for _ in range(2)]
    for m in (ours, ref):
        with m.no_sync():
            m(xs[0]).sum().backward()
        m(xs[1]).sum().backward()
    for p, q in zip(ours.module.parameters(), ref.module.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-6)
    # --- unused parameters: error without the flag, zeros-contributing with it
    strict = DDP(copy.deepcopy(base))
    with pytest.raises(RuntimeError, match="no gradient"):
        strict(xs[0], use_b=False).sum().backward()
    lax = DDP(copy.deepcopy(base), find_unused_parameters=True)
    ref2 = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base), find_unused_parameters=True)
    for m in (lax, ref2):
        m(xs[0], use_b=False).sum().backward()
    for (n, p), q in zip(lax.module.named_parameters(), ref2.module.parameters()):
        if q.grad is None:
            assert p.grad is None, n  # unused on every rank: .grad is left alone, exactly like torch
        else:
            assert torch.allclose(p.grad, q.grad, atol=1e-6), n
    # --- communication hook replaces the built-in all-reduce
    calls = []

    def hook(state, bucket):
        calls.append(bucket.index())
        t = bucket.buffer().clone()
        fut = dist.all_reduce(t, async_op=True).get_future()
        return fut.then(lambda f: f.value()[0] / world)

    hooked = DDP(copy.deepcopy(base))
    hooked.register_comm_hook(None, hook)
    hooked(xs[1]).sum().backward()
    plain = DDP(copy.deepcopy(base))
    plain(xs[1]).sum().backward()
    assert calls
    for p, q in zip(hooked.module.parameters(), plain.module.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-6)
    # --- constructor validation
    frozen = copy.deepcopy(base)
    for p in frozen.parameters():
        p.requires_grad_(False)
    with pytest.raises(RuntimeError, match="doesn't have any parameter"):
        DDP(frozen)
    # --- gradient_as_bucket_view: grads alias the flat bucket
    view = DDP(copy.deepcopy(base), gradient_as_bucket_view=True)
    view(xs[0]).sum().backward()
    bk = view.reducer.buckets[0]
    assert bk.params[0].grad.data_ptr() == bk.views[0].data_ptr()


def test_ddp_no_sync_unused_hooks_validation():
    run_gloo(_w_features, 2)


def _w_join(rank, world):
    """Uneven inputs: rank r trains 2 + 2 r iterations inside join(); parity with torch DDP's join on the same data."""
    from distributed_model_parallel_b200.parallel import DistributedDataParallel as DDP

    torch.manual_seed(0)
    base = nn.Sequential(nn.Linear(8, 16), nn.ReLU(), nn.Linear(16, 4))
    ours = DDP(copy.deepcopy(base), bucket_cap_mb=0.0001, first_bucket_bytes=64)   # several buckets: the shadow must launch all of them, in order
    ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
    o1 = torch.optim.SGD(ours.parameters(), lr=0.05, momentum=0.9)
    o2 = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    iters = 2 + 2 * rank
    torch.manual_seed(10 + rank)
    data = [(torch.randn(5, 8), torch.randint(0, 4, (5,))) for _ in range(iters)]
    for m, o in ((ours, o1), (ref, o2)):
        with m.join():
            for x, y in data:
                o.zero_grad()
                F.cross_entropy(m(x), y).backward()
                o.step()
    assert len(ours.reducer.buckets) > 1
    # every rank ends with the parameters of the rank that trained longest, and they equal torch's
    for (n, p), q in zip(ours.module.named_parameters(), ref.module.parameters()):
        assert torch.allclose(p, q, atol=1e-6, rtol=1e-5), n
        gathered = [torch.empty_like(p) for _ in range(world)]
        dist.all_gather(gathered, p.detach().clone())
        assert all(torch.equal(gathered[0], g) for g in gathered), n
    # a second, even round still works after the join (no collective left half-done)
    x, y = data[0]
    F.cross_entropy(ours(x), y).backward()


def test_ddp_join_uneven_inputs_matches_torch():
    run_gloo(_w_join, 2)


def _w_bucket_sweep(rank, world):
    """Bucket-size sweep: whatever the bucket layout, the synchronised gradients are the same; a smaller
    batch on one rank (uneven last batch) is averaged per rank exactly as torch does."""
    from distributed_model_parallel_b200.parallel import DistributedDataParallel as DDP

    torch.manual_seed(0)
    base = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Flatten(), nn.Linear(8 * 8 * 8, 32), nn.ReLU(), nn.Linear(32, 10))
    torch.manual_seed(50 + rank)
    n = 6 if rank == 0 else 3                                   # uneven per-rank batch
    x, y = torch.randn(n, 3, 8, 8), torch.randint(0, 10, (n,))
    ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
    F.cross_entropy(ref(x), y).backward()
    want = [p.grad.clone() for p in ref.module.parameters()]
    layouts = set()
    for cap_mb, first in ((25.0, 1 << 20), (0.01, 1 << 20), (0.001, 256), (0.0001, 64)):
        m = DDP(copy.deepcopy(base), bucket_cap_mb=cap_mb, first_bucket_bytes=first)
        for _ in range(2):                                      # the second pass runs on the rebuilt (ready-order) buckets
            m.zero_grad(set_to_none=True)
            F.cross_entropy(m(x), y).backward()
            for p, g in zip(m.module.parameters(), want):
                assert torch.allclose(p.grad, g, atol=1e-6, rtol=1e-5), (cap_mb, first)
        layouts.add(len(m.reducer.buckets))
    assert len(layouts) >= 3, layouts                            # the sweep really produced different bucket counts


def test_ddp_bucket_size_sweep_and_uneven_batches():
    run_gloo(_w_bucket_sweep, 2)


def _tiny_cnn():
    return nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Conv2d(8, 8, 3, padding=1, stride=2), nn.ReLU(), nn.Flatten(),
                         nn.Linear(8 * 4 * 4, 10))


def _w_loss_curves(rank, world):
    """Loss-curve parity (SURVEY section 4 / the reference's de-facto integration test, Readme.md:283-294, at unit-test scale):
    20 SGD steps of the same model on the same global batches, (a) single process on the full batch, (b) our single-process
    DataParallel over two replicas, (c) our DistributedDataParallel over two ranks with the batch split -- the three loss
    curves coincide."""
    from distributed_model_parallel_b200.parallel import DataParallel
    from distributed_model_parallel_b200.parallel import DistributedDataParallel as DDP

    torch.manual_seed(0)
    base = _tiny_cnn()
    single, dp_m, ddp_m = copy.deepcopy(base), copy.deepcopy(base), copy.deepcopy(base)
    dp = DataParallel(dp_m, device_ids=[torch.device("cpu"), torch.device("cpu")])
    ddp = DDP(ddp_m, bucket_cap_mb=0.001, first_bucket_bytes=256)
    opts = [torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9) for m in (single, dp, ddp)]
    curves = [[], [], []]
    half = 4
    for step in range(20):
        torch.manual_seed(1000 + step)                          # the same global batch on every rank
        x, y = torch.randn(2 * half, 3, 8, 8), torch.randint(0, 10, (2 * half,))
        for i, (m, o) in enumerate(zip((single, dp, ddp), opts)):
            o.zero_grad()
            if i == 2:
                xs, ys = x[rank * half: (rank + 1) * half], y[rank * half: (rank + 1) * half]
                loss = F.cross_entropy(m(xs), ys)
                loss.backward()
                t = loss.detach().clone()
                dist.all_reduce(t)
                curves[i].append(float(t) / world)              # mean of the per-rank means = full-batch mean (equal halves)
            else:
                loss = F.cross_entropy(m(x), y)
                loss.backward()
                curves[i].append(float(loss))
            o.step()
    a, b, c = (torch.tensor(v) for v in curves)
    assert torch.allclose(a, b, atol=1e-5), (a - b).abs().max()
    assert torch.allclose(a, c, atol=1e-5), (a - c).abs().max()
    assert a[-1] < a[0]                                          # and it did learn something


def test_loss_curves_single_vs_dataparallel_vs_ddp():
    run_gloo(_w_loss_curves, 2)
