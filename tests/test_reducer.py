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


def test_native_hook_launches_plans_in_bucket_order_with_stream_fork_and_join():
    """The whole gradient-ready hook behind one C++ call (dmp_reducer_grad_ready): the kernel-library entry points are replaced
    by recording stubs, so the launch order, the arguments taken from each bucket's plan and the fork/join of the communication
    stream can be checked without a GPU."""
    import ctypes as C

    from distributed_model_parallel_b200.parallel.reducer import (PLAN_ALLREDUCE_FUSED, PLAN_CALLER, PLAN_NVLS_SGD, PLAN_SGD_FLAT,
                                                                  BucketPlan, LaunchFns)

    calls = []
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float
    FORK = C.CFUNCTYPE(i32, vp, vp, C.POINTER(vp), i32)
    JOIN = C.CFUNCTYPE(i32, vp, vp)
    ARF = C.CFUNCTYPE(i32, i32, i32, vp, vp, i32, i32, i64, f32, vp, vp, vp, vp, vp, i32, vp, vp, i32, i32, vp)
    NVS = C.CFUNCTYPE(i32, vp, vp, vp, vp, vp, vp, i32, vp, i32, i32, i64, f32, i32, vp)
    SGD = C.CFUNCTYPE(i32, vp, vp, vp, i32, vp, i64, vp, i32, i32, vp)
    fail = {"on": False}
    fork = FORK(lambda cur, comm, side, n: calls.append(("fork", cur, comm, [side[i] for i in range(n)])) or 0)
    join = JOIN(lambda cur, comm: calls.append(("join", cur, comm)) or 0)
    arf = ARF(lambda algo, dt, bufs, st, rank, W, n, scale, out, m, mom, sh, hp, nest, mp, sp_, za, mb, stream:
              (calls.append(("allreduce_fused", algo, bufs, rank, W, n, round(scale, 4), m, za, stream)) or (7 if fail["on"] else 0)))
    nvs = NVS(lambda g, m, s, master, mom, hp, nest, st, rank, W, n, scale, mb, stream: calls.append(("nvls_sgd", g, m, s, n, stream)) or 0)
    sgd = SGD(lambda master, sh, grad, dt, mom, n, hp, nest, zero, stream: calls.append(("sgd_flat", master, grad, n, zero, stream)) or 0)
    fns = LaunchFns()
    for name, cb in (("fork", fork), ("join", join), ("allreduce_fused", arf), ("allreduce_nvls_sgd", nvs), ("sgd_flat", sgd)):
        setattr(fns, name, C.cast(cb, C.c_void_p).value)

    def plan(kind, **kw):
        p = BucketPlan()
        p.kind = kind
        for k, v in kw.items():
            setattr(p, k, v)
        return p

    r = NativeReducerState([0, 0, 1, 2], 3)
    plans = [plan(PLAN_ALLREDUCE_FUSED, algo=1, bufs=0x1000, rank=1, world=4, numel=4096, scale=0.25, master=0x2000, zero_after=1),
             plan(PLAN_NVLS_SGD, mc_grad=0x10, mc_master=0x20, mc_shadow=0x30, numel=64),
             plan(PLAN_SGD_FLAT, master=0x40, grad=0x50, numel=8)]
    assert r.set_plans(fns, plans, comm_stream=0xC0, side_streams=[0xA1, 0xA2]) is True
    r.prepare()
    assert r.grad_ready(3, False, 0xBEEF) == [] and calls == []      # bucket 2 complete, but buckets launch in order
    assert r.grad_ready(2, False, 0xBEEF) == [] and calls == []
    assert r.grad_ready(0, False, 0xBEEF) == [] and calls == []
    assert r.grad_ready(1, False, 0xBEEF) == []                       # releases all three
    assert [c[0] for c in calls] == ["fork", "allreduce_fused", "fork", "nvls_sgd", "fork", "sgd_flat"]
    assert calls[0] == ("fork", 0xBEEF, 0xC0, [0xA1, 0xA2])
    assert calls[1] == ("allreduce_fused", 1, 0x1000, 1, 4, 4096, 0.25, 0x2000, 1, 0xC0)    # launched ON the communication stream
    assert calls[3] == ("nvls_sgd", 0x10, 0x20, 0x30, 64, 0xC0)
    assert calls[5] == ("sgd_flat", 0x40, 0x50, 8, 1, 0xC0)
    assert r.finish(0xBEEF) == [] and calls[-1] == ("join", 0xBEEF, 0xC0)
    assert r.native_launches() == 3
    # an iteration without launches does not join; a missing gradient is still reported
    del calls[:]
    r.prepare()
    r.grad_ready(0, False, 1)
    assert r.finish(1) == [1, 2, 3] and calls == []
    # a failing entry point surfaces as an error from the hook
    r.prepare()
    fail["on"] = True
    r.grad_ready(0, False, 1)
    with pytest.raises(RuntimeError):
        r.grad_ready(1, False, 1)
    fail["on"] = False
    # one caller-launched bucket (communication hook, torch.distributed back end) hands every launch back, in order
    r2 = NativeReducerState([0, 1], 2)
    assert r2.set_plans(fns, [plan(PLAN_ALLREDUCE_FUSED), plan(PLAN_CALLER)], 0xC0, []) is False
    r2.prepare()
    del calls[:]
    assert r2.grad_ready(1, False, 0) == []
    assert r2.grad_ready(0, False, 0) == [0, 1] and calls == []
