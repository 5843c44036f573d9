"""Pipeline parallelism: partitioner, schedules (generated + simulated), transport metadata, and a 3-stage gloo run
whose gradients must equal the single-process full model."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F
from hypothesis import given, settings, strategies as st

from _dist_utils import run_gloo
from distributed_model_parallel_b200.parallel.pipeline import (TensorMeta, generate_schedule, layer_costs, partition_balanced,
                                                             reference_split, simulate_schedule, split_sequential)


@settings(max_examples=60, deadline=None)
@given(st.lists(st.floats(0.1, 100.0), min_size=1, max_size=30), st.integers(1, 8))
def test_partition_is_contiguous_and_optimal_vs_bruteforce_bound(costs, w):
    w = min(w, len(costs))
    b = partition_balanced(costs, w)
    assert b[0] == 0 and b[-1] == len(costs) and all(b[i] < b[i + 1] for i in range(w))
    worst = max(sum(costs[b[i]: b[i + 1]]) for i in range(w))
    assert worst >= max(max(costs), sum(costs) / w) - 1e-9          # lower bound
    # no single boundary move improves the bottleneck
    for i in range(1, w):
        for d in (-1, 1):
            c = list(b)
            c[i] += d
            if c[i - 1] < c[i] < c[i + 1]:
                assert max(sum(costs[c[j]: c[j + 1]]) for j in range(w)) >= worst - 1e-9


def test_reference_split_and_generic_split():
    from distributed_model_parallel_b200.models import MobileNetV2
    m = MobileNetV2()
    stages = reference_split(m)
    assert [len(s) for s in stages] == [4, 6, 6, 5]                   # stem+3 | 6 | 6 | 2 blocks + head conv + pool + linear
    x = torch.randn(2, 3, 32, 32)
    m.eval()
    y = x
    for s in stages:
        y = s(y)
    assert torch.allclose(y, m(x), atol=1e-5)
    dropped = reference_split(m, stem_relu=False)                     # the reference's stage 0 omits the stem ReLU
    assert not torch.allclose(dropped[0][0](x), stages[0][0](x), atol=1e-3)
    assert (stages[0][0](x) >= 0).all() and (dropped[0][0](x) < 0).any()
    for w in (2, 3, 5, 8):                                            # the reference only works for 4
        seq = m.as_sequential()
        st_ = split_sequential(seq, w, layer_costs(seq, "params"))
        assert sum(len(s) for s in st_) == len(seq)
    with pytest.raises(ValueError):
        partition_balanced([1.0, 2.0], 3)


@pytest.mark.parametrize("kind", ["gpipe", "1f1b"])
@pytest.mark.parametrize("w,m", [(2, 2), (4, 4), (4, 8), (8, 3), (3, 16)])
def test_schedules_are_complete_and_deadlock_free(kind, w, m):
    for s in range(w):
        order = generate_schedule(kind, w, s, m)
        assert sorted(o for o in order if o > 0) == list(range(1, m + 1))
        assert sorted(-o for o in order if o < 0) == list(range(1, m + 1))
        for i in range(1, m + 1):
            assert order.index(i) < order.index(-i)
    span, peak = simulate_schedule(kind, w, m, 1.0, 2.0)
    ideal = 3.0 * m
    assert span >= ideal and span <= ideal + 3.0 * (w - 1) + 1e-9      # bubble is at most (W-1) fwd+bwd slots
    if kind == "1f1b":
        assert all(peak[s] <= min(m, w - s) for s in range(w))         # bounded activation memory
    else:
        assert all(p == m for p in peak)


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 8), st.integers(1, 24), st.sampled_from(["gpipe", "1f1b"]))
def test_graph_slot_count_and_fifo_retirement(w, m, kind):
    """The CUDA-graph path of PipelineStage gives micro-batch i slot i % n_slots with n_slots = max_in_flight(order): that is
    only sound if (a) the in-flight depth never exceeds n_slots and (b) the occupant of a slot has retired (its backward has
    run) before the slot is filled again.  Checked against the native schedule generator for every stage."""
    from distributed_model_parallel_b200.parallel.pipeline import max_in_flight

    for s in range(w):
        order = generate_schedule(kind, w, s, m)
        n_slots = max_in_flight(order)
        assert n_slots == (m if kind == "gpipe" else min(m, w - s))
        occupant = {}
        for op in order:
            i = abs(op) - 1
            slot = i % n_slots
            if op > 0:
                assert slot not in occupant, (kind, w, s, m, order)
                occupant[slot] = i
            else:
                assert occupant.pop(slot) == i
        assert not occupant


def test_naive_schedule_is_one_batch_in_flight():
    assert generate_schedule("naive", 4, 2, 1) == [1, -1]
    with pytest.raises(ValueError):
        generate_schedule("naive", 4, 2, 3)
    span, peak = simulate_schedule("naive", 4, 1, 1.0, 2.0)
    assert span == 12.0 and peak == [1, 1, 1, 1]                      # fully serial: utilisation 1/W like the reference


def test_tensor_meta_round_trip():
    for t in (torch.zeros(5, 3, 8, 8), torch.zeros(4, 7, dtype=torch.bfloat16), torch.zeros(2, 3, 4, 4).contiguous(memory_format=torch.channels_last),
              torch.zeros(9, dtype=torch.int64)):
        m = TensorMeta.of(t)
        m2 = TensorMeta.unpack(m.pack())
        assert m == m2
        e = m2.empty("cpu")
        assert e.shape == t.shape and e.dtype == t.dtype and e.stride() == t.stride()
    assert TensorMeta.of(torch.zeros(8, 3)).with_batch(2).shape == (2, 3)


def _tiny_seq():
    torch.manual_seed(0)
    return nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.Conv2d(8, 8, 3, padding=1), nn.ReLU(), nn.Flatten(), nn.Linear(8 * 8 * 8, 16),
                         nn.ReLU(), nn.Linear(16, 10))


def _w_pipeline(rank, world, sched, loss_on, mb):
    from distributed_model_parallel_b200.parallel.pipeline import PipelineStage

    full = _tiny_seq()
    ref = copy.deepcopy(full)
    stages = split_sequential(full, world, layer_costs(full, "uniform"))
    torch.manual_seed(1)
    x, y = torch.randn(10, 3, 8, 8), torch.randint(0, 10, (10,))
    F.cross_entropy(ref(x), y).backward()
    params = list(stages[rank].parameters())
    opt = torch.optim.SGD(params, lr=0.1) if params else None
    st_ = PipelineStage(stages[rank], rank, world, optimizer=None, loss_fn=F.cross_entropy, schedule=sched, loss_on=loss_on, device="cpu")
    for rep in range(2):   # second step runs with cached metadata (no probe)
        for p in params:
            p.grad = None
        loss = st_.train_step(x if rank == 0 else None, y if rank == 0 else None, mb)
        owner = (rank == 0 and loss_on == "first") or (rank == world - 1 and loss_on == "last")
        assert (loss is not None) == owner
        if owner:
            assert abs(float(loss) - float(F.cross_entropy(ref(x), y))) < 1e-5
        refp = dict(ref.named_parameters())
        ids = {id(p): n for n, p in full.named_parameters()}
        for p in params:
            assert torch.allclose(p.grad, refp[ids[id(p)]].grad, atol=1e-5), (sched, rank, ids[id(p)])
    # ragged batch (the reference's 50000 % 512 case) flows without reconfiguration
    l2 = st_.train_step(x[:7] if rank == 0 else None, y[:7] if rank == 0 else None, mb)
    ev = st_.eval_step(x if rank == 0 else None, y if rank == 0 else None)
    if owner:
        assert ev[0] is not None and ev[1].shape == (10, 10) and l2 is not None


@pytest.mark.parametrize("sched,loss_on,mb", [("naive", "first", 1), ("gpipe", "last", 4), ("1f1b", "last", 4), ("gpipe", "first", 3)])
def test_pipeline_matches_full_model(sched, loss_on, mb):
    run_gloo(_w_pipeline, 3, sched, loss_on, mb)


def _w_reject(rank, world):
    from distributed_model_parallel_b200.parallel.pipeline import PipelineStage
    with pytest.raises(ValueError, match="1f1b"):
        PipelineStage(nn.Linear(2, 2), rank, world, schedule="1f1b", loss_on="first", device="cpu")


def test_1f1b_with_loss_on_first_is_rejected():
    run_gloo(_w_reject, 2)


def test_layer_costs_by_time_and_measured_boundaries_single_process():
    """Per-layer timing leaves the network input without a gradient (stem convolutions provide none on the kernel path) and
    restores BatchNorm buffers; measured_boundaries without a process group is the rank-0 measurement itself."""
    import torch.nn as nn
    from distributed_model_parallel_b200.parallel.pipeline import layer_costs, measured_boundaries

    torch.manual_seed(0)
    seq = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.BatchNorm2d(8), nn.ReLU(), nn.Conv2d(8, 8, 3, padding=1), nn.Flatten(),
                        nn.Linear(8 * 8 * 8, 4))
    x = torch.randn(2, 3, 8, 8)
    seen = []
    seq[0].register_forward_hook(lambda m, inp, out: seen.append(inp[0].requires_grad))
    costs = layer_costs(seq, "time", sample=x, repeats=1)
    assert len(costs) == len(seq) and all(c > 0 for c in costs)
    assert seen and not any(seen), "the network input must not require a gradient"
    assert all(p.grad is None for p in seq.parameters())
    before = [b.clone() for b in seq.buffers()]
    bounds = measured_boundaries(seq, 2, x, torch.device("cpu"), rank=0, mode="time")
    assert bounds[0] == 0 and bounds[-1] == len(seq) and len(bounds) == 3 and bounds[1] in range(1, len(seq))
    for a, b in zip(before, seq.buffers()):
        assert torch.equal(a, b), "train-mode timing passes leaked into BatchNorm statistics"


def test_bench_bubble_report_matches_the_schedule_simulator():
    """bench.py's pipeline line carries the schedule's ideal idle fraction (equal stages, backward = 2 x forward)."""
    import importlib.util
    import os

    spec = importlib.util.spec_from_file_location("dmp_bench", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert abs(bench.bubble_report("naive", 4, 1)["ideal_bubble_fraction"] - 0.75) < 1e-9        # one stage busy at a time
    r4, r8 = bench.bubble_report("1f1b", 4, 4), bench.bubble_report("1f1b", 4, 8)
    assert abs(r4["ideal_bubble_fraction"] - (1 - 12.0 / 21.0)) < 1e-9 and r4["peak_in_flight_per_stage"] == [4, 3, 2, 1]
    assert r8["ideal_bubble_fraction"] < r4["ideal_bubble_fraction"]
    assert bench.bubble_report("gpipe", 4, 8)["peak_in_flight_per_stage"] == [8, 8, 8, 8]     # GPipe holds every micro-batch
