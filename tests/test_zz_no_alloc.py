"""The allocation-free BatchNorm-statistics entry point, exercised through the C ABI.

``dmp_bn_stats_finalize_noalloc`` is for hosts that cannot give the library a scratch buffer because nothing in
the process may call cudaMalloc any more (all device memory was handed to the host's own arena at start-up, or
allocator calls are banned in the steady-state loop since they synchronise the device).  It keeps its
accumulator in a fixed 64 KiB array inside the loaded library, shared by all calls on a GPU, so the things
worth checking are: same numbers as the ordinary entry points (``dmp_bn_stats`` + ``dmp_bn_finalize`` with a
caller-provided accumulator), still the same numbers on a second call that finds the first call's leftovers in
the array, and a clean refusal above the size limit.

This file is named so that pytest collects it last: a ``-x`` run has then already executed every other test.
It does not depend on any of them and also passes alone in a fresh process.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bind():
    from distributed_model_parallel_b200 import _native

    lib = _native.kernels()
    vp = C.c_void_p
    fn = lib.dmp_bn_stats_finalize_noalloc
    fn.restype = C.c_int
    fn.argtypes = [vp, C.c_int, C.c_longlong, C.c_int, vp, vp, vp, vp, C.c_float, C.c_float, vp, vp, vp, vp, vp]
    lib.dmp_bn_noalloc_max_channels.restype = C.c_int
    lib.dmp_bn_noalloc_max_channels.argtypes = []
    return _native, lib, fn


def _close(a, b, rtol, atol, what):
    err = (a.double() - b.double()).abs()
    tol = atol + rtol * b.double().abs()
    assert bool((err <= tol).all()), "%s: max abs err %.3g" % (what, err.max().item())


def _run_both(nat, lib, fn, M, Cc, dtype, seed):
    torch.manual_seed(seed)
    x = (torch.randn(M, Cc, device="cuda") * 1.5 + 0.25).to(dtype)
    gamma = torch.rand(Cc, device="cuda") + 0.5
    beta = torch.randn(Cc, device="cuda") * 0.1
    code = 0 if dtype == torch.float32 else 1
    sp = torch.cuda.current_stream().cuda_stream
    eps, mom = 1e-5, 0.1

    def outputs():
        return [torch.empty(Cc, device="cuda") for _ in range(4)] + [torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")]

    # existing entry points, accumulator allocated by the caller
    mean0, invstd0, scale0, shift0, rm0, rv0 = outputs()
    accum = torch.empty(2 * Cc, device="cuda", dtype=torch.float64)
    nat.check(lib.dmp_bn_stats(x.data_ptr(), code, M, Cc, accum.data_ptr(), sp), "bn_stats")
    nat.check(lib.dmp_bn_finalize(accum.data_ptr(), float(M), gamma.data_ptr(), beta.data_ptr(), rm0.data_ptr(), rv0.data_ptr(),
                                  mom, eps, mean0.data_ptr(), invstd0.data_ptr(), scale0.data_ptr(), shift0.data_ptr(), Cc, sp),
              "bn_finalize")
    # the allocation-free entry point: only inputs and outputs are passed
    mean1, invstd1, scale1, shift1, rm1, rv1 = outputs()
    nat.check(fn(x.data_ptr(), code, M, Cc, gamma.data_ptr(), beta.data_ptr(), rm1.data_ptr(), rv1.data_ptr(), mom, eps,
                 mean1.data_ptr(), invstd1.data_ptr(), scale1.data_ptr(), shift1.data_ptr(), sp), "bn_stats_finalize_noalloc")
    torch.cuda.synchronize()
    for name, a, b in (("mean", mean1, mean0), ("invstd", invstd1, invstd0), ("scale", scale1, scale0),
                       ("shift", shift1, shift0), ("running_mean", rm1, rm0), ("running_var", rv1, rv0)):
        _close(a, b, 1e-4, 1e-5, "%s (M=%d C=%d %s)" % (name, M, Cc, dtype))
    # and both agree with a plain fp32 restatement
    _close(mean1, x.float().mean(0), 1e-4, 1e-4, "mean vs torch")
    _close(invstd1, torch.rsqrt(x.float().var(0, unbiased=False) + eps), 1e-3, 1e-4, "invstd vs torch")


def test_no_alloc_matches_existing_entry_point_twice_in_a_row():
    nat, lib, fn = _bind()
    # first call fills 2*1280 accumulator slots; the second, smaller and with other data, must not see them
    _run_both(nat, lib, fn, 128 * 4 * 4, 1280, torch.bfloat16, seed=1)
    _run_both(nat, lib, fn, 4096, 32, torch.bfloat16, seed=2)
    # a third call in fp32 with yet another channel count
    _run_both(nat, lib, fn, 1000, 24, torch.float32, seed=3)


def test_no_alloc_rejects_sizes_beyond_the_builtin_scratch():
    nat, lib, fn = _bind()
    cmax = lib.dmp_bn_noalloc_max_channels()
    assert cmax == 4096
    Cc = cmax + 8
    x = torch.zeros(8, Cc, device="cuda", dtype=torch.bfloat16)
    outs = [torch.empty(Cc, device="cuda") for _ in range(4)]
    rc = fn(x.data_ptr(), 1, 8, Cc, None, None, None, None, 0.1, 1e-5, outs[0].data_ptr(), outs[1].data_ptr(),
            outs[2].data_ptr(), outs[3].data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.dmp_last_error().decode()
    assert "exceeds" in msg and str(cmax) in msg
    torch.cuda.synchronize()
