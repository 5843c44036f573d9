"""Numerics of every compute kernel against a plain PyTorch fp32 restatement of the same op.

The kernels are called through the C ABI (ctypes) exactly as the Python op layer calls them.
Tolerances are bf16-level where the kernel reads/writes bf16 and fp32-level otherwise.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _lib():
    from distributed_model_parallel_b200 import _native

    return _native


def _sp():
    return torch.cuda.current_stream().cuda_stream


def _close(a, b, rtol, atol, what="", max_bad_frac=0.0):
    """max_bad_frac > 0 tolerates isolated outliers: a ReLU mask taken from a bf16 activation can flip for elements
    whose pre-activation is within rounding distance of zero, which changes that one gradient by the whole dy."""
    a = a.float()
    b = b.float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = (err > tol).sum().item()
    assert bad <= max_bad_frac * a.numel(), "%s: %d/%d mismatches, max abs err %.4g (ref max %.4g)" % (
        what, bad, a.numel(), err.max().item(), b.abs().max().item())


GEMM_SHAPES = [
    # (M, N, K)
    (128, 128, 64), (256, 64, 128), (128, 32, 16), (1000, 96, 24), (4096, 144, 24), (2048, 1280, 320),
    (333, 200, 72), (8192, 16, 32), (512, 960, 160), (128, 256, 1024),
    # few output rows (weight-gradient shapes: M = Cout, long K): MN-major boxes that lie entirely past M / N are skipped
    (64, 96, 256), (24, 144, 512), (144, 24, 2048), (16, 96, 4096), (160, 320, 192),
]


@pytest.mark.parametrize("layout", [0, 1, 2])
@pytest.mark.parametrize("shape", GEMM_SHAPES)
def test_gemm_tcgen05(layout, shape):
    nat = _lib()
    lib = nat.kernels()
    M, N, K = shape
    if layout != 0 and N % 8:
        pytest.skip("N-contiguous B needs N % 8 == 0")
    if layout == 2 and M % 8:
        pytest.skip("M-contiguous A needs M % 8 == 0")
    if layout != 2 and K % 8:
        pytest.skip("K-contiguous A needs K % 8 == 0")
    torch.manual_seed(layout * 100 + M + N + K)
    dev = "cuda"
    a = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
    b = (torch.randn(N, K, device=dev) * 0.5).bfloat16()
    ref = a.float() @ b.float().t()
    if layout == 0:
        A, lda, B, ldb = a, K, b, K
    elif layout == 1:
        A, lda, B, ldb = a, K, b.t().contiguous(), N
    else:
        A, lda, B, ldb = a.t().contiguous(), M, b.t().contiguous(), N
    for out_mode in (0, 1, 2):
        if out_mode == 0:
            if N % 8:
                continue
            D = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        elif out_mode == 1:
            D = torch.empty(M, N, device=dev, dtype=torch.float32)
        else:
            D = torch.full((M, N), 1.0, device=dev, dtype=torch.float32)
        nat.check(lib.dmp_gemm_bf16(layout, A.data_ptr(), lda, B.data_ptr(), ldb, D.data_ptr(), N, M, N, K, None,
                                    out_mode, 0, _sp()), "gemm")
        torch.cuda.synchronize()
        want = ref + 1.0 if out_mode == 2 else ref
        tol = 2e-2 if out_mode == 0 else 2e-3
        _close(D, want, tol, tol * max(1.0, K ** 0.5 * 0.25), "gemm layout=%d mode=%d %s" % (layout, out_mode, shape))


def test_gemm_bias_and_ref():
    nat = _lib()
    lib = nat.kernels()
    M, N, K = 300, 768, 256
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    D = torch.empty(M, N, device="cuda")
    nat.check(lib.dmp_gemm_bf16(0, a.data_ptr(), K, b.data_ptr(), K, D.data_ptr(), N, M, N, K, bias.data_ptr(), 1, 0, _sp()))
    R = torch.empty(M, N, device="cuda")
    nat.check(lib.dmp_gemm_ref_f32out(0, a.data_ptr(), K, b.data_ptr(), K, R.data_ptr(), N, M, N, K, bias.data_ptr(), 0, _sp()))
    torch.cuda.synchronize()
    _close(D, a.float() @ b.float().t() + bias, 2e-3, 2e-2, "gemm+bias")
    _close(R, a.float() @ b.float().t() + bias, 2e-3, 2e-2, "gemm ref")


@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [1, 7, 1024, 2296922])
def test_sgd_flat(gdtype, n):
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(n)
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda").to(gdtype)
    m = torch.randn(n, device="cuda") * 0.1
    shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    lr, mu, wd, gs = 0.4, 0.9, 1e-4, 0.5
    hp = torch.tensor([lr, mu, wd, gs], device="cuda")
    p_ref, m_ref = p.clone(), m.clone()
    d = g.float() * gs + wd * p_ref
    m_ref = mu * m_ref + d
    p_ref = p_ref - lr * m_ref
    nat.check(lib.dmp_sgd_flat(p.data_ptr(), shadow.data_ptr(), g.data_ptr(), 0 if gdtype == torch.float32 else 1,
                               m.data_ptr(), n, hp.data_ptr(), 0, 1, _sp()))
    torch.cuda.synchronize()
    _close(p, p_ref, 1e-5, 1e-6, "sgd p")
    _close(m, m_ref, 1e-5, 1e-6, "sgd m")
    _close(shadow, p_ref.bfloat16(), 1e-2, 1e-3, "sgd shadow")
    assert g.float().abs().max().item() == 0.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,C", [(128, 10), (1000, 10), (37, 1000)])
def test_xent(dtype, B, C):
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(B + C)
    x = (torch.randn(B, C, device="cuda") * 3).to(dtype)
    t = torch.randint(0, C, (B,), device="cuda")
    dx = torch.empty_like(x)
    rows = torch.empty(3 * B, device="cuda")
    out = torch.empty(3, device="cuda")
    nat.check(lib.dmp_xent_fwd_bwd(x.data_ptr(), t.data_ptr(), dx.data_ptr(), rows.data_ptr(), out.data_ptr(), B, C,
                                   0 if dtype == torch.float32 else 1, 1.0 / B, _sp()))
    torch.cuda.synchronize()
    xr = x.float().clone().requires_grad_()
    loss = F.cross_entropy(xr, t)
    loss.backward()
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    _close(out[0:1], loss.detach().reshape(1), 1e-4, 1e-5, "xent loss")
    _close(dx, xr.grad, tol, tol * 0.1, "xent grad")
    top1 = (x.float().argmax(1) == t).sum().item()
    assert abs(out[1].item() - top1) <= (0 if dtype == torch.float32 else 2)
    k = min(5, C)
    top5 = (x.float().topk(k, 1).indices == t[:, None]).any(1).sum().item()
    assert abs(out[2].item() - top5) <= (0 if dtype == torch.float32 else 2)


@pytest.mark.parametrize("M,C", [(128 * 32 * 32, 32), (128 * 16 * 16, 144), (128 * 4 * 4, 1280), (1000, 24), (4096, 960)])
@pytest.mark.parametrize("relu,res", [(1, 0), (0, 1), (1, 1), (0, 0)])
def test_batchnorm_train(M, C, relu, res):
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(M + C)
    x = (torch.randn(M, C, device="cuda") * 2 + 0.5).bfloat16()
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda") * 0.1
    rm = torch.zeros(C, device="cuda")
    rv = torch.ones(C, device="cuda")
    resid = torch.randn(M, C, device="cuda").bfloat16() if res else None
    accum = torch.empty(2 * C, device="cuda", dtype=torch.float64)
    mean = torch.empty(C, device="cuda"); invstd = torch.empty_like(mean)
    scale = torch.empty_like(mean); shift = torch.empty_like(mean)
    y = torch.empty_like(x)
    eps, mom = 1e-5, 0.1
    sp = _sp()
    nat.check(lib.dmp_bn_stats(x.data_ptr(), 1, M, C, accum.data_ptr(), sp))
    nat.check(lib.dmp_bn_finalize(accum.data_ptr(), float(M), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                                  rv.data_ptr(), mom, eps, mean.data_ptr(), invstd.data_ptr(), scale.data_ptr(),
                                  shift.data_ptr(), C, sp))
    nat.check(lib.dmp_bn_apply(x.data_ptr(), 1, scale.data_ptr(), shift.data_ptr(), nat.ptr(resid), y.data_ptr(), M, C,
                               relu, sp))
    # reference (fp32 math on the same bf16 inputs)
    xr = x.float().clone().requires_grad_()
    g_r = gamma.clone().requires_grad_(); b_r = beta.clone().requires_grad_()
    rm_r = torch.zeros(C, device="cuda"); rv_r = torch.ones(C, device="cuda")
    rr = resid.float().clone().requires_grad_() if res else None
    z = F.batch_norm(xr, rm_r, rv_r, g_r, b_r, True, mom, eps)
    if res:
        z = z + rr
    yr = F.relu(z) if relu else z
    torch.cuda.synchronize()
    _close(mean, x.float().mean(0), 1e-4, 1e-4, "bn mean")
    _close(rm, rm_r, 1e-4, 1e-5, "bn running_mean")
    _close(rv, rv_r, 1e-3, 1e-4, "bn running_var")
    _close(y, yr, 2e-2, 2e-2, "bn y")
    # backward
    dy = torch.randn(M, C, device="cuda").bfloat16()
    yr.backward(dy.float())
    acc2 = torch.empty(2 * C, device="cuda", dtype=torch.float64)
    dx = torch.empty_like(x); dres = torch.empty_like(x) if res else None
    dgamma = torch.empty(C, device="cuda"); dbeta = torch.empty(C, device="cuda")
    # the kernel masks with its own bf16 output y (what the op layer saves)
    nat.check(lib.dmp_bn_bwd_reduce(dy.data_ptr(), y.data_ptr(), x.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(), M, C,
                                    relu, acc2.data_ptr(), sp))
    nat.check(lib.dmp_bn_bwd_dx(dy.data_ptr(), y.data_ptr(), x.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(),
                                gamma.data_ptr(), acc2.data_ptr(), float(M), dx.data_ptr(), nat.ptr(dres), M, C, relu, sp))
    nat.check(lib.dmp_bn_bwd_params(acc2.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), C, 0, sp))
    torch.cuda.synchronize()
    scale_g = max(1.0, (M ** 0.5) * 0.02)
    _close(dbeta, b_r.grad, 2e-2, 0.05 * scale_g, "bn dbeta")
    _close(dgamma, g_r.grad, 2e-2, 0.05 * scale_g, "bn dgamma")
    _close(dx, xr.grad, 3e-2, 3e-2, "bn dx", 1e-5 if relu else 0.0)
    if res:
        _close(dres, rr.grad, 1e-2, 1e-2, "bn dres", 1e-5 if relu else 0.0)


@pytest.mark.parametrize("N,H,C,stride", [(8, 32, 32, 1), (4, 16, 144, 2), (2, 8, 960, 1), (3, 9, 24, 2), (128, 32, 96, 1), (16, 32, 144, 2), (5, 6, 576, 2)])
@pytest.mark.parametrize("wdt", [torch.float32, torch.bfloat16])
def test_dwconv3x3(N, H, C, stride, wdt):
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(N * H + C)
    W = H
    x = torch.randn(N, H, W, C, device="cuda").bfloat16()
    w = (torch.randn(C, 1, 3, 3, device="cuda") * 0.3).to(wdt)
    Ho = (H + 2 - 3) // stride + 1
    y = torch.empty(N, Ho, Ho, C, device="cuda", dtype=torch.bfloat16)
    sp = _sp()
    wd = 0 if wdt == torch.float32 else 1
    nat.check(lib.dmp_dwconv3x3_fwd(x.data_ptr(), w.data_ptr(), wd, y.data_ptr(), N, H, W, C, stride, sp))
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_()
    wr = w.float().clone().requires_grad_()
    yr = F.conv2d(xr, wr, None, stride, 1, 1, C)
    torch.cuda.synchronize()
    _close(y, yr.permute(0, 2, 3, 1), 2e-2, 2e-2, "dw fwd")
    dy = torch.randn(N, Ho, Ho, C, device="cuda").bfloat16()
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dx = torch.empty_like(x)
    dw = torch.empty(C, 1, 3, 3, device="cuda")
    nat.check(lib.dmp_dwconv3x3_dgrad(dy.data_ptr(), w.data_ptr(), wd, dx.data_ptr(), N, H, W, C, stride, sp))
    nat.check(lib.dmp_dwconv3x3_wgrad(dy.data_ptr(), x.data_ptr(), dw.data_ptr(), N, H, W, C, stride, 0, sp))
    torch.cuda.synchronize()
    _close(dx, xr.grad.permute(0, 2, 3, 1), 2e-2, 2e-2, "dw dgrad")
    _close(dw, wr.grad, 1e-2, 1e-2 * max(1.0, (N * Ho * Ho) ** 0.5 * 0.1), "dw wgrad")


@pytest.mark.parametrize("layout", [0, 1])
@pytest.mark.parametrize("N,H,Cout,k,stride,pad", [(16, 32, 32, 3, 1, 1), (4, 32, 64, 3, 1, 1), (2, 17, 32, 3, 2, 1)])
def test_stem_conv(layout, N, H, Cout, k, stride, pad):
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(N + H + Cout)
    x = torch.randn(N, 3, H, H, device="cuda")
    if layout == 1:
        xin = x.permute(0, 2, 3, 1).contiguous().bfloat16()
        x = xin.float().permute(0, 3, 1, 2).contiguous()
    else:
        xin = x
    w = torch.randn(Cout, 3, k, k, device="cuda") * 0.2
    Ho = (H + 2 * pad - k) // stride + 1
    y = torch.empty(N, Ho, Ho, Cout, device="cuda", dtype=torch.bfloat16)
    sp = _sp()
    nat.check(lib.dmp_stem_conv_fwd(xin.data_ptr(), layout, w.data_ptr(), 0, y.data_ptr(), N, 3, H, H, Cout, k, k, stride,
                                    pad, sp))
    wr = w.clone().requires_grad_()
    yr = F.conv2d(x, wr, None, stride, pad)
    torch.cuda.synchronize()
    _close(y, yr.permute(0, 2, 3, 1), 2e-2, 2e-2, "stem fwd")
    dy = torch.randn(N, Ho, Ho, Cout, device="cuda").bfloat16()
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dw = torch.empty_like(w)
    nat.check(lib.dmp_stem_conv_wgrad(dy.data_ptr(), xin.data_ptr(), layout, dw.data_ptr(), N, 3, H, H, Cout, k, k, stride,
                                      pad, 0, sp))
    torch.cuda.synchronize()
    _close(dw, wr.grad, 1e-2, 1e-2 * max(1.0, (N * Ho * Ho) ** 0.5 * 0.1), "stem wgrad")


def test_pool_linear_elementwise():
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(0)
    sp = _sp()
    N, HW, Cc = 128, 16, 1280
    x = torch.randn(N, HW, Cc, device="cuda").bfloat16()
    for relu in (0, 1):
        y = torch.empty(N, Cc, device="cuda", dtype=torch.bfloat16)
        nat.check(lib.dmp_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, HW, Cc, relu, sp))
        xr = x.float().clone().requires_grad_()
        yr = (F.relu(xr) if relu else xr).mean(1)
        dy = torch.randn(N, Cc, device="cuda").bfloat16()
        yr.backward(dy.float())
        dx = torch.empty_like(x)
        nat.check(lib.dmp_avgpool_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), N, HW, Cc, relu, sp))
        torch.cuda.synchronize()
        _close(y, yr, 1e-2, 1e-2, "avgpool fwd")
        _close(dx, xr.grad, 1e-2, 1e-3, "avgpool bwd")
    # skinny linear
    B, K, Nc = 128, 1280, 10
    xl = torch.randn(B, K, device="cuda").bfloat16()
    w = torch.randn(Nc, K, device="cuda") * 0.05
    bias = torch.randn(Nc, device="cuda")
    out = torch.empty(B, Nc, device="cuda")
    nat.check(lib.dmp_linear_small_fwd(xl.data_ptr(), w.data_ptr(), 0, bias.data_ptr(), out.data_ptr(), B, Nc, K, sp))
    xr = xl.float().clone().requires_grad_(); wr = w.clone().requires_grad_(); br = bias.clone().requires_grad_()
    outr = F.linear(xr, wr, br)
    dyl = torch.randn(B, Nc, device="cuda")
    outr.backward(dyl)
    dxl = torch.empty_like(xl); dwl = torch.empty_like(w); dbl = torch.empty_like(bias)
    nat.check(lib.dmp_linear_small_bwd(dyl.data_ptr(), xl.data_ptr(), w.data_ptr(), 0, dxl.data_ptr(), dwl.data_ptr(),
                                       dbl.data_ptr(), B, Nc, K, 0, sp))
    torch.cuda.synchronize()
    _close(out, outr, 1e-3, 1e-3, "linear fwd")
    _close(dxl, xr.grad, 1e-2, 1e-2, "linear dx")
    _close(dwl, wr.grad, 1e-3, 1e-3, "linear dw")
    _close(dbl, br.grad, 1e-3, 1e-3, "linear db")
    # elementwise
    a = torch.randn(4096 * 8, device="cuda").bfloat16(); b = torch.randn_like(a)
    o = torch.empty_like(a)
    nat.check(lib.dmp_add_act(a.data_ptr(), b.data_ptr(), o.data_ptr(), a.numel(), 1, sp))
    g = torch.empty_like(a)
    nat.check(lib.dmp_relu_bwd(b.data_ptr(), o.data_ptr(), g.data_ptr(), a.numel(), sp))
    torch.cuda.synchronize()
    _close(o, F.relu(a.float() + b.float()), 1e-2, 1e-2, "add_act")
    _close(g, torch.where(o.float() > 0, b.float(), torch.zeros_like(b.float())), 0, 0, "relu_bwd")


@pytest.mark.parametrize("M,C", [(128 * 32 * 32, 96), (128 * 16 * 16, 144), (2048, 1280), (1000, 24)])
@pytest.mark.parametrize("relu,res", [(1, 1), (0, 0), (1, 0)])
def test_batchnorm_fused_single_launch(M, C, relu, res):
    """The cooperative one-launch BN kernels against torch, three calls in a row (exercises the parity workspace)."""
    nat = _lib()
    lib = nat.kernels()
    sp = _sp()
    ws_f = torch.zeros(lib.dmp_bn_fused_ws_bytes(C) // 8, device="cuda", dtype=torch.float64)
    ws_b = torch.zeros_like(ws_f)
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda") * 0.1
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    rm_r, rv_r = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros((), device="cuda", dtype=torch.int64)
    for it in range(3):
        torch.manual_seed(M + C + it)
        x = (torch.randn(M, C, device="cuda") * 2 + 0.5).bfloat16()
        resid = torch.randn(M, C, device="cuda").bfloat16() if res else None
        y = torch.empty_like(x)
        mean = torch.empty(C, device="cuda"); invstd = torch.empty_like(mean)
        rc = lib.dmp_bn_fwd_fused(x.data_ptr(), 1, nat.ptr(resid), y.data_ptr(), M, C, gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                                  rv.data_ptr(), nbt.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), ws_f.data_ptr(), relu, sp)
        assert rc == 0, nat.kernels().dmp_last_error()
        xr = x.float().clone().requires_grad_()
        g_r, b_r = gamma.clone().requires_grad_(), beta.clone().requires_grad_()
        rr = resid.float().clone().requires_grad_() if res else None
        z = F.batch_norm(xr, rm_r, rv_r, g_r, b_r, True, 0.1, 1e-5)
        if res:
            z = z + rr
        yr = F.relu(z) if relu else z
        torch.cuda.synchronize()
        _close(y, yr, 2e-2, 2e-2, "fused bn y (call %d)" % it)
        _close(rm, rm_r, 1e-4, 1e-5, "fused bn running_mean")
        _close(rv, rv_r, 1e-3, 1e-4, "fused bn running_var")
        dy = torch.randn(M, C, device="cuda").bfloat16()
        yr.backward(dy.float())
        dx = torch.empty_like(x); dres = torch.empty_like(x) if res else None
        dgamma = torch.full((C,), 1.0, device="cuda"); dbeta = torch.full((C,), -2.0, device="cuda")
        # with a residual the mask comes from the saved output, without one it is recomputed from x (y = NULL)
        rc = lib.dmp_bn_bwd_fused(dy.data_ptr(), y.data_ptr() if res else None, x.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(),
                                  gamma.data_ptr(), beta.data_ptr(), dx.data_ptr(), nat.ptr(dres), dgamma.data_ptr(), dbeta.data_ptr(), 1,
                                  M, C, relu, ws_b.data_ptr(), sp)
        assert rc == 0
        torch.cuda.synchronize()
        scale_g = max(1.0, (M ** 0.5) * 0.02)
        # per-channel sums over M elements; a handful of ReLU-mask flips move a sum by O(1) each
        _close(dbeta + 2.0, b_r.grad, 2e-2, 0.05 * scale_g + 0.01 * b_r.grad.abs().max().item(), "fused bn dbeta (accumulated)")
        _close(dgamma - 1.0, g_r.grad, 2e-2, 0.05 * scale_g + 0.01 * g_r.grad.abs().max().item(), "fused bn dgamma (accumulated)")
        _close(dx, xr.grad, 3e-2, 3e-2, "fused bn dx", 1e-5 if relu else 0.0)
        if res:
            _close(dres, rr.grad, 1e-2, 1e-2, "fused bn dres", 1e-5 if relu else 0.0)
    assert int(nbt.item()) == 3


@pytest.mark.parametrize("N,H,Cin,Cout,k", [(4, 32, 64, 128, 3), (8, 8, 128, 64, 3), (16, 4, 256, 256, 3), (2, 14, 64, 64, 3), (3, 7, 128, 192, 3),
                                             (1, 56, 64, 64, 3), (2, 16, 64, 32, 5)])
def test_conv_igemm_tcgen05(N, H, Cin, Cout, k):
    """Implicit-GEMM dense convolution (fwd, dgrad via the flipped filter, wgrad) vs F.conv2d in fp32."""
    nat = _lib()
    lib = nat.kernels()
    sp = _sp()
    torch.manual_seed(N * H + Cin + Cout)
    pad = k // 2
    x = torch.randn(N, H, H, Cin, device="cuda").bfloat16()
    w = (torch.randn(Cout, Cin, k, k, device="cuda") * (1.0 / (Cin * k * k) ** 0.5)).bfloat16()
    wp = w.permute(0, 2, 3, 1).reshape(Cout, k * k * Cin).contiguous()
    y = torch.empty(N, H, H, Cout, device="cuda", dtype=torch.bfloat16)
    nat.check(lib.dmp_conv_igemm_fwd(x.data_ptr(), wp.data_ptr(), None, y.data_ptr(), N, H, H, Cin, Cout, k, k, pad, sp), "igemm fwd")
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_()
    wr = w.float().clone().requires_grad_()
    yr = F.conv2d(xr, wr, None, 1, pad)
    torch.cuda.synchronize()
    _close(y, yr.permute(0, 2, 3, 1), 2e-2, 2e-2, "igemm fwd")
    dy = torch.randn(N, H, H, Cout, device="cuda").bfloat16()
    yr.backward(dy.float().permute(0, 3, 1, 2))
    if Cout % 64 == 0:
        wd = w.permute(1, 2, 3, 0).flip(1, 2).reshape(Cin, k * k * Cout).contiguous()
        dx = torch.empty(N, H, H, Cin, device="cuda", dtype=torch.bfloat16)
        nat.check(lib.dmp_conv_igemm_fwd(dy.data_ptr(), wd.data_ptr(), None, dx.data_ptr(), N, H, H, Cout, Cin, k, k, pad, sp), "igemm dgrad")
        torch.cuda.synchronize()
        _close(dx, xr.grad.permute(0, 2, 3, 1), 2e-2, 2e-2 * xr.grad.abs().max().item(), "igemm dgrad")
    dwp = torch.zeros(Cout, k * k * Cin, device="cuda")
    nat.check(lib.dmp_conv_igemm_wgrad(dy.data_ptr(), x.data_ptr(), dwp.data_ptr(), N, H, H, Cin, Cout, k, k, pad, sp), "igemm wgrad")
    torch.cuda.synchronize()
    dw = dwp.view(Cout, k, k, Cin).permute(0, 3, 1, 2)
    _close(dw, wr.grad, 1e-2, 1e-2 * wr.grad.abs().max().item(), "igemm wgrad")


# ------------------------------------------------------------------------------------------------ BN statistics from the producer's epilogue
@pytest.mark.parametrize("M,N,K", [(4096, 144, 24), (131072, 96, 16), (8192, 384, 64), (2048, 960, 160), (1000, 40, 24), (8192, 32, 192)])
def test_gemm_colstats_and_bn_apply_sums(M, N, K):
    """1x1-conv GEMM whose epilogue accumulates the BatchNorm batch sums, followed by the single-pass apply kernel, against
    F.batch_norm (train mode) of the same bf16 conv output in fp32.  Run twice: the sums buffer must clean itself."""
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(M + N + K)
    a = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    b = (torch.randn(N, K, device="cuda") * 0.5 + 0.05).bfloat16()
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    sums = torch.zeros(int(lib.dmp_bn_sums_bytes(N)) // 8, dtype=torch.float64, device="cuda")
    gamma, beta = torch.rand(N, device="cuda") + 0.5, torch.randn(N, device="cuda") * 0.1
    res = torch.randn(M, N, device="cuda").bfloat16()
    sp = _sp()
    for rep in range(2):
        rm, rv = torch.zeros(N, device="cuda"), torch.ones(N, device="cuda")
        nbt = torch.zeros((), dtype=torch.int64, device="cuda")
        mean, invstd = torch.empty(N, device="cuda"), torch.empty(N, device="cuda")
        out = torch.empty_like(y)
        nat.check(lib.dmp_gemm_bf16_colstats(a.data_ptr(), K, b.data_ptr(), K, y.data_ptr(), N, M, N, K, sums.data_ptr(), sp), "gemm+stats")
        torch.cuda.synchronize()
        yref = (a.float() @ b.float().t())
        _close(y, yref, 2e-2, 2e-2, "gemm out")
        yf = y.float()
        s = sums[: 2 * N].clone()
        assert torch.allclose(s[:N], yf.double().sum(0), rtol=1e-4, atol=1e-2), "column sums"
        assert torch.allclose(s[N:], (yf.double() ** 2).sum(0), rtol=1e-4, atol=1e-2), "column sums of squares"
        nat.check(lib.dmp_bn_apply_sums(y.data_ptr(), 1, res.data_ptr(), out.data_ptr(), M, N, gamma.data_ptr(), beta.data_ptr(),
                                        rm.data_ptr(), rv.data_ptr(), nbt.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(),
                                        sums.data_ptr(), 1, sp), "apply sums")
        torch.cuda.synchronize()
        rm2, rv2 = torch.zeros(N, device="cuda"), torch.ones(N, device="cuda")
        want = F.relu(F.batch_norm(yf, rm2, rv2, gamma, beta, True, 0.1, 1e-5) + res.float())
        _close(out, want, 2e-2, 2e-2, "bn out")
        assert torch.allclose(rm, rm2, atol=1e-4) and torch.allclose(rv, rv2, rtol=1e-3, atol=1e-4)
        assert int(nbt) == 1
        assert torch.allclose(mean, yf.mean(0), atol=1e-4)
        assert float(sums.abs().max()) == 0.0 and sums.view(torch.int32)[-2:].abs().max().item() == 0, "sums buffer not cleaned"


@pytest.mark.parametrize("N,H,C,stride", [(8, 32, 32, 1), (4, 16, 144, 2), (2, 8, 960, 1), (3, 9, 24, 2), (128, 32, 96, 1)])
def test_dwconv_colstats(N, H, C, stride):
    nat = _lib()
    lib = nat.kernels()
    torch.manual_seed(N + H + C)
    x = torch.randn(N, H, H, C, device="cuda").bfloat16()
    w = (torch.randn(C, 1, 3, 3, device="cuda") * 0.3).bfloat16()
    Ho = (H + 2 - 3) // stride + 1
    y = torch.empty(N, Ho, Ho, C, device="cuda", dtype=torch.bfloat16)
    y0 = torch.empty_like(y)
    sums = torch.zeros(int(lib.dmp_bn_sums_bytes(C)) // 8, dtype=torch.float64, device="cuda")
    sp = _sp()
    nat.check(lib.dmp_dwconv3x3_fwd(x.data_ptr(), w.data_ptr(), 1, y0.data_ptr(), N, H, H, C, stride, sp))
    nat.check(lib.dmp_dwconv3x3_fwd_colstats(x.data_ptr(), w.data_ptr(), 1, y.data_ptr(), N, H, H, C, stride, sums.data_ptr(), sp))
    torch.cuda.synchronize()
    assert torch.equal(y, y0)
    yf = y.float().reshape(-1, C).double()
    assert torch.allclose(sums[:C], yf.sum(0), rtol=1e-4, atol=1e-2)
    assert torch.allclose(sums[C: 2 * C], (yf ** 2).sum(0), rtol=1e-4, atol=1e-2)
