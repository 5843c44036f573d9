"""End-to-end runs of the three trainers through the command-line front end, on CPU (gloo for the multi-process ones) with the
synthetic dataset: one or two tiny epochs each.  What is checked is the plumbing the reference scripts consist of
(/root/reference/code/distributed_training/data_parallel.py, model_parallel.py): argument handling, the epoch loop, the text log
line format, best-accuracy checkpointing and --resume, per-stage roles of the pipeline trainer."""
import json
import os
import socket

import torch

from distributed_model_parallel_b200 import cli


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _common(tmp_path, extra=()):
    return ["--dataset-type", "synthetic", "--model", "mobilenetv2", "--epochs", "1", "--steps-per-epoch", "2", "-b", "8",
            "--val-batch-size", "8", "-j", "0", "--lr", "0.05", "--warmup-period", "2", "--log-dir", str(tmp_path / "log"),
            "--checkpoint", str(tmp_path / "ckpt" / "ckpt.pth"), "--no-cuda-graph", *extra]


def test_dp_trainer_logs_checkpoints_and_resumes(tmp_path):
    assert cli.main(["dp", *_common(tmp_path)]) == 0
    logs = os.listdir(tmp_path / "log")
    text = [f for f in logs if f.endswith(".txt")]
    assert text, logs
    line = open(tmp_path / "log" / text[0]).read().strip().splitlines()[-1]
    assert line.startswith("step:0") and "loss_train:" in line and "acc1_val:" in line     # the reference's line format
    ck = torch.load(tmp_path / "ckpt" / "ckpt.pth", map_location="cpu", weights_only=False)
    assert {"net", "acc", "epoch"} <= set(ck) and ck["epoch"] == 0                        # data_parallel.py:146-150
    assert all(k.startswith("module.") for k in ck["net"])                               # DataParallel key space
    # --resume continues after the stored epoch
    assert cli.main(["dp", *_common(tmp_path, ["--resume", "--epochs", "2"])]) == 0
    lines = [ln for ln in open(tmp_path / "log" / text[0]).read().splitlines() if ln.startswith("step:")]
    assert [ln.split()[0] for ln in lines] == ["step:0", "step:1"]


def test_ddp_trainer_two_gloo_ranks(tmp_path):
    url = "tcp://127.0.0.1:%d" % _port()
    assert cli.main(["ddp", *_common(tmp_path, ["--world-size", "2", "--dist-backend", "gloo", "--comm-backend", "torch",
                                                 "--dist-url", url])]) == 0
    recs = [json.loads(ln) for ln in open(tmp_path / "log" / "ddp_rank0.jsonl")]
    assert recs and recs[-1]["epoch"] == 0 and recs[-1]["ddp"]["world_size"] == 2 and recs[-1]["ddp"]["iterations"] >= 2
    assert os.path.exists(tmp_path / "log" / "ddp_rank1.jsonl")                           # every rank keeps its own record
    assert os.path.exists(tmp_path / "ckpt" / "ckpt.pth")


def test_mp_trainer_two_stage_pipeline(tmp_path):
    url = "tcp://127.0.0.1:%d" % _port()
    assert cli.main(["mp", *_common(tmp_path, ["--world-size", "2", "--dist-backend", "gloo", "--dist-url", url, "--schedule", "1f1b",
                                                "--microbatches", "2", "-b", "32", "--val-batch-size", "64"])]) == 0
    logs = os.listdir(tmp_path / "log")
    text = [f for f in logs if f.endswith(".txt")]
    assert text, logs
    line = open(tmp_path / "log" / text[0]).read().strip().splitlines()[-1]
    assert line.startswith("step:0") and "time_per_batch:" in line                        # model_parallel.py:119-124
    # the stage that owns the loss logs real accuracies (reference: utils.py:65-78,98-109), not placeholders
    fields = dict(tok.split(":", 1) for tok in line.split() if ":" in tok)
    assert float(fields["acc1_val"]) > 0.0 and float(fields["acc1_train"]) > 0.0, line
    shard = torch.load(tmp_path / "ckpt" / "pipeline" / "stage0.pth", map_location="cpu", weights_only=False)
    assert shard["acc"] == float(fields["acc1_val"]) or abs(shard["acc"] - float(fields["acc1_val"])) < 1e-3
    # every stage wrote its shard; merged they are exactly the single-model state_dict
    from distributed_model_parallel_b200.models import MobileNetV2
    from distributed_model_parallel_b200.utils import merge_pipeline_shards

    merged = merge_pipeline_shards(str(tmp_path / "ckpt" / "pipeline"), 2)
    full = MobileNetV2()
    assert set(merged) == set(full.state_dict())
    full.load_state_dict(merged, strict=True)


def test_train_step_accepts_a_short_last_batch():
    """50000 % 512 = 336: the last batch of an epoch is short (drop_last=False, like the reference's loaders); the step engine
    must take it instead of failing on its static input buffers."""
    from distributed_model_parallel_b200.models import MobileNetV2
    from distributed_model_parallel_b200.train.engine import TrainStep

    torch.manual_seed(0)
    net = MobileNetV2()
    opt = torch.optim.SGD(net.parameters(), lr=0.01, momentum=0.9)
    step = TrainStep(net, opt, use_graph=False)
    g = torch.Generator().manual_seed(1)
    for n in (16, 10, 16):
        stats = step(torch.randn(n, 3, 32, 32, generator=g), torch.randint(0, 10, (n,), generator=g))
        assert torch.isfinite(stats[0])
    assert step.steps == 3 and step.eager_fallback_steps == 1
