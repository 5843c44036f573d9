"""Host-side Reducer logic (C++ runtime): bucket assignment and the ready/launch state machine."""
import sys

import pytest
import torch
from hypothesis import given, settings, strategies as st

from distributed_model_parallel_b200.parallel.reducer import NativeReducerState, compute_bucket_assignment_by_size

MiB = 1024 * 1024


def test_bucket_assignment_matches_torch():
    import torch.distributed as dist
    torch.manual_seed(0)
    for trial in range(5):
        sizes = torch.randint(1, 400000, (60,)).tolist()
        tensors = [torch.empty(s) for s in sizes]
        limits = [MiB, 5 * MiB]
        want, _ = dist._compute_bucket_assignment_by_size(tensors, limits)
        got = compute_bucket_assignment_by_size([s * 4 for s in sizes], limits)
        assert got == [list(b) for b in want]


def test_bucket_assignment_groups_by_key_and_orders_by_first_index():
    sizes = [100, 100, 100, 100, 100, 100]
    keys = [0, 1, 0, 1, 0, 1]
    b = compute_bucket_assignment_by_size(sizes, [250], keys)
    assert b == [[0, 2, 4], [1, 3, 5]]
    b = compute_bucket_assignment_by_size(sizes, [150, 1000], keys)
    assert b[0] == [0, 2] and b[1] == [1, 3]


@settings(max_examples=50, deadline=None)
@given(st.lists(st.integers(1, 10 ** 7), min_size=1, max_size=80), st.integers(1, 10 ** 7), st.integers(1, 10 ** 8))
def test_bucket_assignment_properties(sizes, first, cap):
    buckets = compute_bucket_assignment_by_size(sizes, [first, cap])
    flat = sorted(i for b in buckets for i in b)
    assert flat == list(range(len(sizes)))                     # partition
    assert all(b == sorted(b) for b in buckets)                  # order preserved inside a bucket
    assert [b[0] for b in buckets] == sorted(b[0] for b in buckets)
    for k, b in enumerate(buckets):                              # a bucket closes as soon as it reaches its limit
        lim = first if k == 0 else cap
        assert sum(sizes[i] for i in b[:-1]) < lim


def test_reducer_in_order_launch_and_errors():
    r = NativeReducerState([0, 0, 1, 1, 2], 3)
    r.prepare()
    assert r.mark_ready(4) == []          # bucket 2 complete but 0 and 1 have not launched
    assert r.mark_ready(2) == []
    assert r.mark_ready(0) == []
    assert r.mark_ready(1) == [0]         # bucket 0 launches
    assert r.mark_ready(3) == [1, 2]      # bucket 1 launches and releases the waiting bucket 2
    assert r.finalize() == []
    assert r.ready_order() == [4, 2, 0, 1, 3]
    r.prepare()
    r.mark_ready(0)
    with pytest.raises(RuntimeError, match="twice"):
        r.mark_ready(0)
    assert sorted(r.finalize()) == [1, 2, 3, 4]
    with pytest.raises(RuntimeError):
        r.mark_ready(1)                   # no iteration prepared


@settings(max_examples=30, deadline=None)
@given(st.integers(1, 40), st.integers(1, 8), st.randoms())
def test_reducer_launches_every_bucket_exactly_once_in_order(n, nb, rnd):
    nb = min(nb, n)
    bucket_of = [i % nb for i in range(n)]
    r = NativeReducerState(bucket_of, nb)
    r.prepare()
    order = list(range(n))
    rnd.shuffle(order)
    launched = []
    for p in order:
        launched += r.mark_ready(p, unused=rnd.random() < 0.2)
    assert launched == list(range(nb))
    assert r.finalize() == [] and r.num_launched() == nb


def test_comm_watchdog_polls_reports_once_and_stops():
    """Host watchdog: polls the probe, hands the first fault to the callback, then ends; stop() ends a healthy one."""
    import time

    from distributed_model_parallel_b200.parallel.watchdog import CommWatchdog

    calls, faults = [], []

    def probe():
        calls.append(1)
        if len(calls) == 3:
            raise RuntimeError("peer collective timed out waiting for rank 1")

    wd = CommWatchdog(probe, interval_s=0.02, on_fault=faults.append).start()
    t0 = time.time()
    while not faults and time.time() - t0 < 5:
        time.sleep(0.01)
    time.sleep(0.1)
    assert len(faults) == 1 and "rank 1" in str(faults[0]) and wd.fault is faults[0]
    assert len(calls) == 3 and wd.polls == 2                      # it stopped probing after the fault
    healthy = CommWatchdog(lambda: None, interval_s=0.01, on_fault=faults.append)
    with healthy:
        time.sleep(0.1)
    assert healthy.polls >= 2 and healthy.fault is None and not healthy._thread.is_alive()
