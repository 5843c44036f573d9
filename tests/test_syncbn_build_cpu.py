"""SyncBatchNorm on gloo vs BatchNorm over the concatenated batch; native build + header/ABI consistency."""
import copy
import os
import re
import subprocess

import torch
import torch.distributed as dist
import torch.nn as nn

from _dist_utils import run_gloo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _w_syncbn(rank, world):
    from distributed_model_parallel_b200.parallel.sync_batchnorm import SyncBatchNorm, convert_sync_batchnorm

    torch.manual_seed(0)
    net = nn.Sequential(nn.Conv2d(3, 6, 3, padding=1), nn.BatchNorm2d(6), nn.ReLU(), nn.Conv2d(6, 4, 1), nn.BatchNorm2d(4))
    ref = copy.deepcopy(net)
    sync = convert_sync_batchnorm(copy.deepcopy(net))
    assert sum(isinstance(m, SyncBatchNorm) for m in sync.modules()) == 2
    assert list(sync.state_dict().keys()) == list(net.state_dict().keys())
    torch.manual_seed(1)
    full = torch.randn(8, 3, 5, 5)
    g = torch.randn(8, 4, 5, 5)
    lo, hi = rank * 4, rank * 4 + 4
    out = sync(full[lo:hi])
    out.backward(g[lo:hi])
    out_ref = ref(full)
    out_ref.backward(g)
    assert torch.allclose(out, out_ref[lo:hi], atol=1e-5)
    for (n, p), q in zip(sync.named_parameters(), ref.parameters()):
        tot = p.grad.clone()
        dist.all_reduce(tot)                      # DDP would sum (then average) the per-rank parameter gradients
        assert torch.allclose(tot, q.grad, atol=1e-4), n
    assert torch.allclose(sync[1].running_mean, ref[1].running_mean, atol=1e-6)
    assert torch.allclose(sync[1].running_var, ref[1].running_var, atol=1e-5)
    sync.eval()
    ref.eval()
    assert torch.allclose(sync(full[lo:hi]), ref(full[lo:hi]), atol=1e-5)


def test_syncbn_matches_full_batch_bn():
    run_gloo(_w_syncbn, 2)


def test_native_libraries_build_and_export_the_declared_abi():
    from distributed_model_parallel_b200 import _build, _native

    klib, rlib = _build.build_all(verbose=False)
    assert klib.exists() and rlib.exists()
    assert _native.runtime().dmp_rt_version() >= 100
    header = open(os.path.join(ROOT, "include", "dmp_b200.h")).read()
    declared = set(re.findall(r"\b(dmp_[a-z0-9_]+)\s*\(", header))
    exported = set()
    for lib in (klib, rlib):
        out = subprocess.run(["nm", "-D", "--defined-only", str(lib)], capture_output=True, text=True).stdout
        exported |= set(re.findall(r" T (dmp_[a-z0-9_]+)", out))
    assert declared <= exported, "declared but not exported: %s" % sorted(declared - exported)
    assert exported <= declared, "exported but not declared in include/dmp_b200.h: %s" % sorted(exported - declared)
    # the kernel library really is sm_100a code with tensor-core / TMA instructions
    sass = subprocess.run(["cuobjdump", "-sass", str(klib)], capture_output=True, text=True).stdout
    assert "sm_100a" in sass or "SM100" in sass.upper()
    assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass


def test_graft_entry_build():
    import __graft_entry__ as g
    g.build()
