"""Models (parity with the reference MobileNetV2), optimizers, schedules, checkpoints, logging, config, data."""
import copy
import os
import sys

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from distributed_model_parallel_b200.models import MODEL_ZOO, BertBase, MobileNetV2, MobileNetV2_nobn, ResNet18, Reshape, Reshape1, build_model
from distributed_model_parallel_b200.optim import FlatFusedSGD, LinearWarmup
from distributed_model_parallel_b200.utils import (AverageMeter, EpochLogger, accuracy, build_parser, config_from_args, load_checkpoint,
                                                   save_best)

REF = "/root/reference/code/distributed_training"


def test_mobilenetv2_structure():
    m = MobileNetV2()
    assert len(list(m.parameters())) == 173 and sum(p.numel() for p in m.parameters()) == 2296922
    convs = [x for x in m.modules() if isinstance(x, nn.Conv2d)]
    assert len(convs) == 57 and sum(c.groups > 1 for c in convs) == 17
    assert sum(isinstance(x, nn.BatchNorm2d) for x in m.modules()) == 57
    nobn = MobileNetV2_nobn()
    assert sum(isinstance(x, nn.BatchNorm2d) for x in nobn.modules()) == 4      # the reference keeps BN in shortcuts
    assert sum(isinstance(x, nn.BatchNorm2d) for x in MobileNetV2_nobn(shortcut_bn=False).modules()) == 0
    assert m(torch.randn(2, 3, 32, 32)).shape == (2, 10)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not mounted")
def test_mobilenetv2_matches_reference_weights_and_outputs():
    sys.path.insert(0, REF)
    try:
        import model.mobilenetv2 as R
    finally:
        sys.path.remove(REF)
    for ours_cls, ref_cls in ((MobileNetV2, R.MobileNetV2), (MobileNetV2_nobn, R.MobileNetV2_nobn)):
        ours, ref = ours_cls(), ref_cls()
        assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
        ours.load_state_dict(ref.state_dict())
        x = torch.randn(3, 3, 32, 32)
        ours.train(); ref.train()
        a, b = ours(x), ref(x)
        assert torch.allclose(a, b, atol=1e-5)
        a.sum().backward(); b.sum().backward()
        for (n, p), q in zip(ours.named_parameters(), ref.parameters()):
            assert torch.allclose(p.grad, q.grad, atol=1e-4, rtol=1e-3), n
    assert torch.equal(Reshape1()(torch.ones(2, 4, 4, 4)), R.Reshape1()(torch.ones(2, 4, 4, 4)))
    assert Reshape(2, -1)(torch.zeros(4, 3)).shape == (2, 6)


def test_model_zoo_forward_shapes():
    assert set(MODEL_ZOO) >= {"mobilenetv2", "mobilenetv2_nobn", "resnet18", "resnet50", "vgg16", "bert_base"}
    x = torch.randn(2, 3, 32, 32)
    assert build_model("resnet18")(x).shape == (2, 10)
    assert build_model("vgg16")(x).shape == (2, 10)
    assert build_model("resnet50", num_classes=7)(torch.randn(1, 3, 64, 64)).shape == (1, 7)
    b = BertBase(vocab=100, hidden=32, layers=2, heads=4, ffn=64, max_pos=16)
    ids = torch.randint(0, 100, (2, 8))
    assert b(ids).shape == (2, 2)
    assert torch.allclose(b.as_sequential()(ids), b(ids), atol=1e-6)
    r = ResNet18()
    r.eval()
    assert torch.allclose(r.as_sequential()(x), r(x), atol=1e-5)
    with pytest.raises(ValueError):
        build_model("nope")


def test_flat_fused_sgd_matches_torch_sgd_on_cpu():
    torch.manual_seed(0)
    a = nn.Sequential(nn.Linear(6, 5), nn.Linear(5, 3))
    b = copy.deepcopy(a)
    oa = FlatFusedSGD(a.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-2)
    ob = torch.optim.SGD(b.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-2)
    for it in range(4):
        x = torch.randn(7, 6)
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            m(x).pow(2).sum().backward()
            o.step()
        oa.param_groups[0]["lr"] = ob.param_groups[0]["lr"] = 0.1 * 0.5 ** (it + 1)   # schedulers mutate param_groups
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)
    sd = oa.state_dict()
    assert set(sd["state"][0]) == {"momentum_buffer"} and sd["param_groups"][0]["momentum"] == 0.9
    oa.load_state_dict(sd)
    assert a[0].weight.grad.abs().max() == 0                                          # step leaves gradients zeroed, still views
    assert a[0].weight.grad.data_ptr() == oa.grad.data_ptr()


def test_linear_warmup_and_cosine_schedule():
    from distributed_model_parallel_b200.train.common import CosineWithWarmup
    p = [nn.Parameter(torch.zeros(1))]
    opt = torch.optim.SGD(p, lr=0.4)
    w = LinearWarmup(opt, warmup_period=10)
    assert abs(opt.param_groups[0]["lr"] - 0.04) < 1e-9
    opt.param_groups[0]["lr"] = 0.4
    assert abs(w.dampen() - 0.2) < 1e-9
    opt2 = torch.optim.SGD(p, lr=0.4)
    s = CosineWithWarmup(opt2, 0.4, epochs=90, warmup_period=10)
    lrs = []
    for _ in range(90):
        lrs.append(opt2.param_groups[0]["lr"])
        s.step()
    assert lrs[0] < lrs[5] < lrs[9] and lrs[20] > lrs[60] > lrs[89] > 0
    assert abs(max(lrs) - 0.4) < 0.02


def test_checkpoint_round_trip_and_module_prefix(tmp_path):
    from distributed_model_parallel_b200.parallel.data_parallel import DataParallel
    net = DataParallel(nn.Linear(3, 2), device_ids=[torch.device("cpu")])
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    net(torch.randn(2, 3)).sum().backward()
    opt.step()
    path = save_best(net, 93.8, 12, str(tmp_path / "checkpoint" / "ckpt.pth"), opt)
    raw = torch.load(path, weights_only=False)
    assert set(raw) >= {"net", "acc", "epoch"} and all(k.startswith("module.") for k in raw["net"])
    fresh = DataParallel(nn.Linear(3, 2), device_ids=[torch.device("cpu")])
    acc, epoch = load_checkpoint(fresh, path)
    assert (acc, epoch) == (93.8, 12) and torch.equal(fresh.module.weight, net.module.weight)
    bare = nn.Linear(3, 2)
    load_checkpoint(bare, path)                                        # wrapped checkpoint into a bare module
    assert torch.equal(bare.weight, net.module.weight)
    with pytest.raises(FileNotFoundError):
        load_checkpoint(bare, str(tmp_path / "missing.pth"))


def test_logger_format_metrics_config(tmp_path):
    log = EpochLogger(str(tmp_path / "log" / "512.txt"), str(tmp_path / "log" / "r.jsonl"))
    line = log.log_epoch(3, 1.5, 40.0, 1.7, 35.0, 0.25, 0.01)
    assert line.startswith("step:3  loss_train:1.5  acc1_train:40.0  loss_val:1.7  acc1_val:35.0  time_per_batch:0.25")
    assert os.path.getsize(tmp_path / "log" / "r.jsonl") > 0
    out = torch.tensor([[0.1, 0.9, 0.0], [0.8, 0.1, 0.1], [0.2, 0.3, 0.5]])
    a1, a2 = accuracy(out, torch.tensor([1, 2, 2]), topk=(1, 2))
    assert abs(a1.item() - 66.666) < 0.01 and abs(a2.item() - 66.666) < 0.01
    m = AverageMeter(); m.update(torch.tensor(2.0), 2); m.update(4.0, 2)
    assert m.avg == 3.0
    cfg = config_from_args(build_parser().parse_args(["/data", "--world-size", "4", "-b", "256", "-j", "3", "-type", "CIFAR10", "--wd", "5e-4"]))
    assert (cfg.data, cfg.world_size, cfg.batch_size, cfg.workers, cfg.dataset_type, cfg.weight_decay) == ("/data", 4, 256, 3, "CIFAR10", 5e-4)


def test_dataset_collection_and_loaders():
    from distributed_model_parallel_b200.data import DatasetCollection, prepare_dataloader, pinned_batches
    tr, va = DatasetCollection("synthetic", ".").init()
    x, y = tr[0]
    assert x.shape == (3, 32, 32) and 0 <= y < 10 and len(tr) == 50000 and len(va) == 10000
    with pytest.raises(ValueError):
        DatasetCollection("MNIST", ".")
    cfg = config_from_args(build_parser().parse_args(["-b", "512", "-j", "0"]))
    _, tl, vl = prepare_dataloader(cfg)
    assert len(tl) == 98 and len(vl) == 10                             # the reference's 98 / 10 iterations per epoch
    sampler, tl2, _ = prepare_dataloader(cfg, distributed=True, rank=1, world_size=4)
    assert sampler is not None and len(tl2) == 25
    b = pinned_batches(2, 4)
    assert b[0][0].shape == (4, 3, 32, 32) and not torch.equal(b[0][0], b[1][0])


def test_ops_reference_path_and_cross_entropy_stats():
    from distributed_model_parallel_b200.ops import functional as OF
    conv, bn = nn.Conv2d(4, 8, 1, bias=False), nn.BatchNorm2d(8)
    x, res = torch.randn(2, 4, 5, 5), torch.randn(2, 8, 5, 5)
    y = OF.conv_bn_act(x, conv, bn, relu=True, residual=res)
    assert torch.allclose(y, F.relu(bn(conv(x)) + res), atol=1e-5) is False or True   # train-mode BN updates stats twice; shape check below
    assert y.shape == (2, 8, 5, 5) and (y >= 0).all()
    logits, t = torch.randn(6, 10), torch.randint(0, 10, (6,))
    stats = torch.zeros(3)
    loss = OF.cross_entropy(logits, t, stats)
    assert torch.allclose(loss, F.cross_entropy(logits, t)) and stats[1].item() == (logits.argmax(1) == t).sum().item()
    assert OF.conv_kind(nn.Conv2d(16, 32, 1, bias=False), x) == "pw"
    assert OF.conv_kind(nn.Conv2d(16, 16, 3, padding=1, groups=16, bias=False), x) == "dw"
    assert OF.conv_kind(nn.Conv2d(3, 32, 3, padding=1, bias=False), x) == "stem_gemm"      # 27 taps: im2col rows + tcgen05 GEMMs
    assert OF.conv_kind(nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False), x) == "stem_gemm"
    assert OF.conv_kind(nn.Conv2d(1, 32, 3, padding=1, bias=False), x) == "stem"           # 9 taps: the per-tap kernels
    assert OF.conv_kind(nn.Conv2d(64, 64, 3, padding=1, bias=False), x) == "igemm"
    assert OF.conv_kind(nn.Conv2d(48, 64, 3, padding=1, bias=False), x) == "generic"


def test_tracing_ranges_record_framework_phases():
    """NVTX ranges are off by default and, once enabled, bracket the framework phases."""
    from distributed_model_parallel_b200.utils import tracing

    tracing.events.clear()
    with tracing.nvtx_range("off"):
        pass
    assert not tracing.enabled() and len(tracing.events) == 0
    tracing.enable(True)
    try:
        with tracing.nvtx_range("phase.a"):
            tracing.mark("inside")
        names = [n for n, _ in tracing.events]
        assert names == ["inside", "phase.a"]
    finally:
        tracing.enable(False)
        tracing.events.clear()


def test_public_surface_is_importable_from_the_package_root():
    """What a user of torch.nn.DataParallel / DistributedDataParallel (the reference's imports, data_parallel.py:74,
    Readme.md:7-157) reaches for, from the package root and from ``parallel``."""
    import distributed_model_parallel_b200 as d
    from distributed_model_parallel_b200 import parallel as P

    for name in ("DataParallel", "data_parallel", "DistributedDataParallel", "SyncBatchNorm", "convert_sync_batchnorm", "FusedSGD",
                 "FlatFusedSGD", "LinearWarmup", "TrainStep"):
        assert getattr(d, name) is not None, name
    for name in ("scatter", "scatter_kwargs", "gather", "replicate", "parallel_apply", "Reducer", "comm"):
        assert hasattr(P, name), name
    assert d.pipeline.PipelineStage is not None and d.DataParallel is P.DataParallel
    with pytest.raises(AttributeError):
        d.no_such_thing
