"""Whole models through the kernel path against the same models through the plain PyTorch restatement in fp32.

The reference's de-facto integration test is convergence parity between its two trainers (/root/reference/Readme.md:283-294);
offline, the equivalent is numerical parity of the kernel path with an fp32 oracle from the same seed: the loss, every
parameter's gradient (direction and size) after one backward, and the loss trajectory over three optimizer steps -- eagerly
and through the CUDA-graph step engine."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _build(name, **kw):
    from distributed_model_parallel_b200.models import build_model

    torch.manual_seed(1234)
    m = build_model(name, **kw).cuda()
    with torch.no_grad():  # non-trivial BatchNorm affine parameters so that their gradients are exercised
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.weight.uniform_(0.6, 1.4)
                mod.bias.normal_(0, 0.1)
    return m


def _batch(name, n):
    g = torch.Generator().manual_seed(5)
    if name.startswith("bert"):
        return torch.randint(0, 1000, (n, 64), generator=g).cuda(), torch.randint(0, 2, (n,), generator=g).cuda()
    hw = 64 if name == "resnet50" else 32
    classes = 100 if name == "resnet50" else 10
    return torch.randn(n, 3, hw, hw, generator=g).cuda(), torch.randint(0, classes, (n,), generator=g).cuda()


CASES = [
    # name, ctor kwargs, batch, min gradient cosine, max loss difference
    ("mobilenetv2", {}, 32, 0.97, 0.05),
    ("resnet18", {}, 32, 0.97, 0.05),
    ("resnet50", {"num_classes": 100}, 8, 0.95, 0.08),
    ("vgg16", {}, 16, 0.97, 0.05),
    ("bert_base", {"layers": 2, "vocab": 1000, "max_pos": 64}, 8, 0.97, 0.05),
]


@pytest.mark.parametrize("name,kw,batch,min_cos,max_dloss", CASES, ids=[c[0] for c in CASES])
def test_model_kernel_path_matches_fp32_reference(name, kw, batch, min_cos, max_dloss):
    from distributed_model_parallel_b200 import _native
    from distributed_model_parallel_b200.ops import functional as OF

    model = _build(name, **kw)
    ref, amp = copy.deepcopy(model), copy.deepcopy(model)
    x, y = _batch(name, batch)
    c0 = _native.launch_count()
    OF.path_counts.clear()
    loss = OF.cross_entropy(model(x), y)
    loss.backward()
    torch.cuda.synchronize()
    assert _native.launch_count() - c0 > 10, "the native kernels did not run"
    assert not any(k.endswith(":reference") for k in OF.path_counts), dict(OF.path_counts)
    with OF.use_reference_ops():
        loss_r = OF.cross_entropy(ref(x), y)
        loss_r.backward()
        # the yardstick for "bf16-level": the same network through torch's own bf16 autocast (cuDNN / cuBLAS kernels).  Per-channel
        # BatchNorm gradients in a randomly initialised network are sums with heavy cancellation and come out noisy in ANY bf16
        # implementation (cosines of 0.8 and below against fp32); the kernel path must be at least as close to fp32 as that.
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out_a = amp(x if x.dtype != torch.float32 else x.contiguous(memory_format=torch.channels_last))
        OF.cross_entropy(out_a.float(), y).backward()
    loss_a = float(OF.cross_entropy(out_a.float().detach(), y))
    assert abs(float(loss) - float(loss_r)) < max(max_dloss, 1.5 * abs(loss_a - float(loss_r)) + 0.02), (float(loss), float(loss_r), loss_a)
    trip = [(n, p.grad.float().flatten(), q.grad.float().flatten(), a.grad.float().flatten())
            for (n, p), (_, q), (_, a) in zip(model.named_parameters(), ref.named_parameters(), amp.named_parameters())]
    numel = sum(g.numel() for _, _, g, _ in trip)
    rms = (sum(float(g.pow(2).sum()) for _, _, g, _ in trip) / numel) ** 0.5   # typical size of one gradient element
    import math
    for n, gk, gr, ga in trip:
        assert torch.isfinite(gk).all(), n
        nk, nr, na = gk.norm().item(), gr.norm().item(), ga.norm().item()
        floor = 0.05 * rms * gr.numel() ** 0.5
        if nr < floor:
            # structurally (near-)zero gradient, e.g. a bias in front of a BatchNorm: only rounding noise is left
            assert nk < 4 * floor + 2 * na, "gradient of %s should be negligible: %.3g vs reference %.3g" % (n, nk, nr)
            continue
        if gr.numel() >= 16:
            cos_k = torch.dot(gk, gr).item() / (nk * nr + 1e-30)
            cos_a = torch.dot(ga, gr).item() / (na * nr + 1e-30)
            if cos_a > 0.9:  # a gradient that bf16 can resolve at all (others are checked through the whole-model cosine only)
                assert cos_k > min(min_cos, cos_a - 0.1), "gradient direction of %s: cosine %.4f (torch bf16 autocast: %.4f)" % (n, cos_k, cos_a)
        # 1-D (per-channel) gradients are sums with heavy cancellation: rounding noise adds to their norm in quadrature
        lim = 0.45 if gr.numel() <= 4096 else 0.22
        assert abs(math.log(nk / nr)) < max(lim, abs(math.log(max(na, 1e-30) / nr)) + 0.15), \
            "gradient norm of %s off by %.3f (torch bf16 autocast: %.3f)" % (n, nk / nr, na / nr)
    allk, allr, alla = (torch.cat([t[i] for t in trip]) for i in (1, 2, 3))
    cos_all = torch.dot(allk, allr).item() / (allk.norm().item() * allr.norm().item())
    cos_all_a = torch.dot(alla, allr).item() / (alla.norm().item() * allr.norm().item())
    assert cos_all > min(0.98, cos_all_a - 0.03), "whole-model gradient cosine %.4f (torch bf16 autocast: %.4f)" % (cos_all, cos_all_a)
    assert 0.85 < allk.norm().item() / allr.norm().item() < 1.18


@pytest.mark.parametrize("name,kw,batch", [(c[0], c[1], c[2]) for c in CASES[:2]], ids=[c[0] for c in CASES[:2]])
@pytest.mark.parametrize("graph", [False, True], ids=["eager", "graph"])
def test_three_steps_follow_the_fp32_trajectory(name, kw, batch, graph):
    """Three SGD steps on one batch: the kernel path (eager, and captured once and replayed by TrainStep) tracks the fp32
    reference's loss at every step and ends with parameters close to the reference's."""
    from distributed_model_parallel_b200.ops import functional as OF
    from distributed_model_parallel_b200.optim import FlatFusedSGD
    from distributed_model_parallel_b200.train.engine import TrainStep

    model = _build(name, **kw)
    ref = copy.deepcopy(model)
    x, y = _batch(name, batch)
    hp = dict(lr=0.05, momentum=0.9, weight_decay=1e-4)
    opt_r = torch.optim.SGD(ref.parameters(), **hp)
    losses_r = []
    with OF.use_reference_ops():
        for _ in range(3):
            opt_r.zero_grad()
            lr_ = OF.cross_entropy(ref(x), y)
            lr_.backward()
            opt_r.step()
            losses_r.append(float(lr_))
    step = TrainStep(model, FlatFusedSGD(model.parameters(), **hp), use_graph=graph)
    losses_k = []
    for _ in range(3):
        step(x, y)
        losses_k.append(step.read_stats()[0])
    assert (step.graph is not None) == graph
    for a, b in zip(losses_k, losses_r):
        assert abs(a - b) < 0.08 + 0.05 * abs(b), (losses_k, losses_r)
    assert losses_r[-1] < losses_r[0]  # the batch is being fitted, so the comparison is not of two flat lines
    num = den = 0.0
    for p, q in zip(model.parameters(), ref.parameters()):
        num += (p.detach().float() - q.detach().float()).pow(2).sum().item()
        den += q.detach().float().pow(2).sum().item()
    assert (num / den) ** 0.5 < 0.04, "parameters drifted %.4f from the fp32 run" % (num / den) ** 0.5
