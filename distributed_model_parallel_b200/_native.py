"""ctypes bindings for the two native libraries (C ABI declared in ``include/dmp_b200.h``).

``kernels()`` returns the CUDA kernel library, ``runtime()`` the host runtime.  Both are built on demand
(see ``_build.py``) and loaded once.  Every kernel entry point returns 0 on success; ``check`` turns a
non-zero code into a ``RuntimeError`` carrying ``dmp_last_error()``.  On a machine with a GPU a missing or
unbuildable library is a hard error: there is no silent fallback to PyTorch ops for CUDA tensors.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import _build

_lock = threading.Lock()
_klib = None
_rlib = None

vp = C.c_void_p
i32 = C.c_int
i64 = C.c_longlong
f32 = C.c_float
f64 = C.c_double

_KERNEL_SIGS = {
    # core
    "dmp_last_error": (C.c_char_p, []),
    "dmp_launch_count": (C.c_ulonglong, []),
    "dmp_reset_launch_count": (None, []),
    "dmp_version": (i32, []),
    "dmp_device_cc": (i32, []),
    "dmp_l2_flush": (i32, [vp, i64, vp]),
    "dmp_stream_fork": (i32, [vp, vp, vp, i32]),
    "dmp_stream_join": (i32, [vp, vp]),
    # gemm
    "dmp_gemm_bf16": (i32, [i32, vp, i64, vp, i64, vp, i64, i32, i32, i32, vp, i32, i32, vp]),
    "dmp_gemm_ref_f32out": (i32, [i32, vp, i64, vp, i64, vp, i64, i32, i32, i32, vp, i32, vp]),
    "dmp_conv_igemm_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_conv_igemm_wgrad": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_conv_igemm_fwd_strided": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_conv_igemm_dgrad_s2": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_conv_igemm_wgrad_strided": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_gemm_bf16_batched": (i32, [i32, vp, i64, i64, i64, vp, i64, i64, i64, vp, i64, i64, i64, i32, i32, i32, i32, i32, vp]),
    # pooling / transformer row-wise kernels (csrc/kernels/elementwise.cu)
    "dmp_maxpool_fwd": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_maxpool_bwd": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_layernorm_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, vp]),
    "dmp_layernorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, vp]),
    "dmp_gelu_fwd": (i32, [vp, vp, i64, vp]),
    "dmp_gelu_bwd": (i32, [vp, vp, vp, i64, vp]),
    "dmp_softmax_fwd": (i32, [vp, vp, i64, i32, f32, vp]),
    "dmp_softmax_bwd": (i32, [vp, vp, vp, i64, i32, f32, vp]),
    # optimizer
    "dmp_sgd_flat": (i32, [vp, vp, vp, i32, vp, i64, vp, i32, i32, vp]),
    "dmp_cast_flat": (i32, [vp, vp, i64, i32, vp]),
    # loss
    "dmp_xent_fwd_bwd": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, f32, vp]),
    # batch norm
    "dmp_bn_stats": (i32, [vp, i32, i64, i32, vp, vp]),
    "dmp_bn_finalize": (i32, [vp, f64, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, i32, vp]),
    "dmp_bn_apply": (i32, [vp, i32, vp, vp, vp, vp, i64, i32, i32, vp]),
    "dmp_bn_bwd_reduce": (i32, [vp, vp, vp, i32, vp, vp, i64, i32, i32, vp, vp]),
    "dmp_bn_bwd_dx": (i32, [vp, vp, vp, i32, vp, vp, vp, vp, f64, vp, vp, i64, i32, i32, vp]),
    "dmp_bn_bwd_params": (i32, [vp, vp, vp, i32, i32, vp]),
    "dmp_bn_fused_ws_bytes": (i64, [i32]),
    "dmp_gemm_bf16_colstats": (i32, [vp, i64, vp, i64, vp, i64, i32, i32, i32, vp, vp]),
    "dmp_dwconv3x3_fwd_colstats": (i32, [vp, vp, i32, vp, i32, i32, i32, i32, i32, vp, vp]),
    "dmp_bn_sums_bytes": (i64, [i32]),
    # conv -> bn -> relu chain fusions (BatchNorm reductions in the producers' epilogues, normalisation on load)
    "dmp_gemm_bf16_dgrad_bnbwd": (i32, [vp, i64, vp, i64, vp, i64, i32, i32, i32, vp, i64, vp, vp, vp, vp, i32, vp, vp]),
    "dmp_dwconv3x3_fwd_bnin_colstats": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp, vp, vp]),
    "dmp_dwconv3x3_dgrad_bnbwd": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]),
    "dmp_dwconv3x3_wgrad_bnin": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp]),
    "dmp_bn_bwd_dx_sums": (i32, [vp, vp, i32, vp, vp, vp, vp, vp, vp, i32, i64, i32, vp, vp]),
    "dmp_bn_apply_sums": (i32, [vp, i32, vp, vp, i64, i32, vp, vp, vp, vp, vp, f32, f32, vp, vp, vp, i32, vp]),
    "dmp_bn_fwd_fused": (i32, [vp, i32, vp, vp, i64, i32, vp, vp, vp, vp, vp, f32, f32, vp, vp, vp, i32, vp]),
    "dmp_bn_bwd_fused": (i32, [vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i64, i32, i32, vp, vp]),
    "dmp_bn_stats_finalize_noalloc": (i32, [vp, i32, i64, i32, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp]),
    "dmp_bn_noalloc_max_channels": (i32, []),
    # SyncBatchNorm with the cross-GPU combine inside the kernel (csrc/comm/syncbn.cu)
    "dmp_syncbn_ws_bytes": (i64, [i32]),
    "dmp_syncbn_fwd": (i32, [vp, vp, vp, i64, i32, vp, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, i32, i32, vp, i32, vp]),
    "dmp_syncbn_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i64, i32, i32, vp, vp, i32, i32, vp, vp]),
    # depthwise conv
    "dmp_dwconv3x3_fwd": (i32, [vp, vp, i32, vp, i32, i32, i32, i32, i32, vp]),
    "dmp_dwconv3x3_dgrad": (i32, [vp, vp, i32, vp, i32, i32, i32, i32, i32, vp]),
    "dmp_dwconv3x3_wgrad": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    # misc
    "dmp_stem_conv_fwd": (i32, [vp, i32, vp, i32, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_stem_conv_wgrad": (i32, [vp, vp, i32, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_stem_im2col": (i32, [vp, i32, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "dmp_pack_rows_bf16": (i32, [vp, i32, vp, i32, i32, i32, vp]),
    "dmp_pack_filter": (i32, [vp, i32, vp, i32, i32, i32, i32, i32, vp]),
    "dmp_avgpool_fwd": (i32, [vp, vp, i32, i32, i32, i32, vp]),
    "dmp_avgpool_bwd": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    "dmp_linear_small_fwd": (i32, [vp, vp, i32, vp, vp, i32, i32, i32, vp]),
    "dmp_linear_small_bwd": (i32, [vp, vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, vp]),
    "dmp_add_act": (i32, [vp, vp, vp, i64, i32, vp]),
    "dmp_relu_bwd": (i32, [vp, vp, vp, i64, vp]),
    # NVLS multicast objects (csrc/comm/multicast.cu)
    "dmp_mc_supported": (i32, [vp, i32]),
    "dmp_mc_create": (i32, [vp, i32, i64, vp, vp, vp, vp]),
    "dmp_mc_destroy": (i32, [vp]),
    "dmp_mc_xp_create": (i32, [i32, i32, i64, vp, vp, vp]),
    "dmp_mc_xp_import": (i32, [i32, i32, i32, i64, vp, vp]),
    "dmp_mc_xp_add_device": (i32, [vp]),
    "dmp_mc_xp_bind_map": (i32, [vp, vp, vp]),
    "dmp_mc_xp_destroy": (i32, [vp]),
    "dmp_allreduce_nvls": (i32, [vp, vp, i32, i32, i64, f32, i32, vp]),
    "dmp_allreduce_nvls_sgd": (i32, [vp, vp, vp, vp, vp, vp, i32, vp, i32, i32, i64, f32, i32, vp]),
    # peer-memory collectives and plumbing (csrc/comm/collectives.cu)
    "dmp_comm_state_bytes": (C.c_longlong, []),
    "dmp_comm_max_ranks": (C.c_int, []),
    "dmp_comm_max_blocks": (C.c_int, []),
    "dmp_allreduce_fused": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_float,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "dmp_broadcast": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_int,
                                C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "dmp_gather_peer": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "dmp_copy_peer": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]),
    "dmp_reduce_add_peer": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "dmp_reduce_add_peer_z": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "dmp_symm_malloc": (C.c_int, [C.c_longlong, C.POINTER(C.c_void_p)]),
    "dmp_symm_free": (C.c_int, [C.c_void_p]),
    "dmp_ipc_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "dmp_ipc_import": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "dmp_ipc_close": (C.c_int, [C.c_void_p]),
    "dmp_ipc_handle_bytes": (C.c_int, []),
    "dmp_enable_peer_access": (C.c_int, [C.c_int]),
    "dmp_comm_status": (C.c_int, [C.c_void_p, C.c_void_p]),
}

_RUNTIME_SIGS = {
    "dmp_rt_version": (i32, []),
}


def register_kernel_sigs(sigs: dict) -> None:
    """Other modules (comm, conv) add their entry points here before first load or after (applied lazily)."""
    _KERNEL_SIGS.update(sigs)
    if _klib is not None:
        _apply(_klib, sigs)


def register_runtime_sigs(sigs: dict) -> None:
    _RUNTIME_SIGS.update(sigs)
    if _rlib is not None:
        _apply(_rlib, sigs)


def _apply(lib, sigs):
    for name, (res, args) in sigs.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # stale library: force a rebuild next time
            raise RuntimeError("native library is missing symbol %s (stale build? run "
                               "`python -m distributed_model_parallel_b200._build --force`)" % name) from e
        fn.restype = res
        fn.argtypes = args


def kernels():
    """The CUDA kernel library (built for sm_100a). Raises if it cannot be built/loaded."""
    global _klib
    if _klib is None:
        with _lock:
            if _klib is None:
                path = _build.KERNEL_LIB
                if os.environ.get("DMP_B200_NO_BUILD") != "1":
                    path = _build.build_kernels(verbose=False)
                import torch  # noqa: F401  (loads the CUDA runtime this library shares with torch before dlopen)
                lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
                _apply(lib, _KERNEL_SIGS)
                _klib = lib
    return _klib


def runtime():
    """The host runtime library (pure C++, usable without a GPU)."""
    global _rlib
    if _rlib is None:
        with _lock:
            if _rlib is None:
                path = os.environ.get("DMP_B200_RT_LIB")  # e.g. an ASan/UBSan build, see tools/sanitize_runtime.sh
                if not path:
                    path = _build.RUNTIME_LIB
                    if os.environ.get("DMP_B200_NO_BUILD") != "1":
                        path = _build.build_runtime(verbose=False)
                lib = C.CDLL(str(path))
                _apply(lib, _RUNTIME_SIGS)
                _rlib = lib
    return _rlib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = kernels().dmp_last_error()
        raise RuntimeError("dmp_b200 %s failed (code %d): %s" % (what, rc, (msg or b"").decode(errors="replace")))


def ptr(t) -> int:
    """Device/host address of a tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def stream_ptr(stream=None) -> int:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def launch_count() -> int:
    return int(kernels().dmp_launch_count()) if _klib is not None else 0
