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


def _c_type_code(ctype: str) -> str:
    """Collapses a C parameter / return type of include/dmp_b200.h to the class the ctypes tables use."""
    t = ctype.replace("const", " ").strip()
    if "*" in t or "dmp_stream_t" in t:
        return "ptr"
    t = " ".join(t.split())
    return {"int": "i32", "long long": "i64", "unsigned long long": "u64", "float": "f32", "double": "f64", "void": "void",
            "uint32_t": "u32", "int64_t": "i64", "int32_t": "i32"}[t]


def test_ctypes_signatures_match_the_header():
    """Every declaration of include/dmp_b200.h against the argtypes / restype registered in _native.py (and by the modules that
    register runtime signatures): same arity, same scalar widths, pointers where the header has pointers.  A missing or narrower
    argtype makes ctypes pass 64-bit pointers as 32-bit ints -- silently, until a kernel faults."""
    import ctypes as C

    from distributed_model_parallel_b200 import _native
    from distributed_model_parallel_b200.parallel import reducer  # noqa: F401  (registers runtime signatures)
    from distributed_model_parallel_b200.parallel.pipeline import partition  # noqa: F401

    header = open(os.path.join(ROOT, "include", "dmp_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", " ", header, flags=re.S)
    decls = re.findall(r"([A-Za-z_][A-Za-z0-9_ \*]*?)\b(dmp_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", header)
    assert len(decls) > 60
    code_of = {C.c_void_p: "ptr", C.c_char_p: "ptr", C.c_int: "i32", C.c_longlong: "i64", C.c_ulonglong: "u64", C.c_float: "f32",
               C.c_double: "f64", C.c_uint: "u32", None: "void"}
    tables = dict(_native._KERNEL_SIGS)
    tables.update(_native._RUNTIME_SIGS)
    problems = []
    for ret, name, params in decls:
        if name not in tables:
            problems.append("%s: declared in the header but has no ctypes signature" % name)
            continue
        restype, argtypes = tables[name]
        want_args = [] if params.strip() in ("", "void") else [_c_type_code(re.sub(r"\b[A-Za-z_][A-Za-z0-9_]*\s*(\[\s*\])?\s*$", lambda m: "*" if m.group(1) else "", p.strip()) or p)
                                                              for p in params.split(",")]
        have_args = [code_of.get(a, "ptr" if hasattr(a, "contents") or getattr(a, "_type_", None) is not None and a not in code_of else "?")
                     for a in argtypes]
        if len(want_args) != len(have_args):
            problems.append("%s: header has %d parameters, ctypes table %d" % (name, len(want_args), len(have_args)))
            continue
        for i, (w, h) in enumerate(zip(want_args, have_args)):
            if w != h:
                problems.append("%s: parameter %d is %s in the header but %s in the ctypes table" % (name, i, w, h))
        if _c_type_code(ret) != code_of.get(restype, "?"):
            problems.append("%s: return type %s vs %s" % (name, _c_type_code(ret), code_of.get(restype, "?")))
    assert not problems, "\n".join(problems)
