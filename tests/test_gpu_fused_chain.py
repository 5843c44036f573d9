"""conv -> bn -> relu chain fusions: BatchNorm reductions produced by the neighbouring convolutions' kernels, normalisation
applied while the depthwise stencil loads its input.  Every kernel is called through the C ABI and compared with a plain
PyTorch fp32 restatement of the same maths; the chain as a whole is compared with autograd on the unfused reference ops.
(The chain is the reference's Block: /root/reference/code/distributed_training/model/mobilenetv2.py:31-36.)"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _nat():
    from distributed_model_parallel_b200 import _native

    return _native


def _sp():
    return torch.cuda.current_stream().cuda_stream


def _close(a, b, rtol, atol, what="", max_bad_frac=0.0):
    a, b = a.float(), b.float()
    err = (a - b).abs()
    bad = (err > atol + rtol * b.abs()).sum().item()
    assert bad <= max_bad_frac * a.numel(), "%s: %d/%d mismatches, max abs err %.4g (ref max %.4g)" % (
        what, bad, a.numel(), err.max().item(), b.abs().max().item())


def _sums_of(t2d, lib):
    """{double sums[2C]; ticket} buffer holding the column sums / sums of squares of a [M, C] tensor."""
    C = t2d.shape[1]
    buf = torch.zeros(int(lib.dmp_bn_sums_bytes(C)) // 8, dtype=torch.float64, device=t2d.device)
    buf[:C] = t2d.double().sum(0)
    buf[C:2 * C] = (t2d.double() ** 2).sum(0)
    return buf


DW_CASES = [(8, 32, 32, 1), (4, 16, 144, 2), (2, 8, 960, 1), (3, 10, 48, 1), (128, 32, 96, 1), (16, 16, 192, 2)]


@pytest.mark.parametrize("N,H,C,stride", DW_CASES)
def test_dw_forward_normalises_on_load(N, H, C, stride):
    nat = _nat()
    lib = nat.kernels()
    torch.manual_seed(N + H + C)
    x = (torch.randn(N, H, H, C, device="cuda") * 1.5 + 0.3).bfloat16()               # raw expand output, NHWC
    w = (torch.randn(C, 1, 3, 3, device="cuda") * 0.3).bfloat16()
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda") * 0.2
    Ho = (H + 2 - 3) // stride + 1
    for rep in range(2):                                                              # twice: the sums buffers clean themselves
        in_sums = _sums_of(x.view(-1, C).float(), lib)
        out_stats = torch.zeros(2 * C, dtype=torch.float64, device="cuda")
        y = torch.empty(N, Ho, Ho, C, device="cuda", dtype=torch.bfloat16)
        rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
        nbt = torch.zeros((), dtype=torch.int64, device="cuda")
        mean, invstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
        nat.check(lib.dmp_dwconv3x3_fwd_bnin_colstats(x.data_ptr(), w.data_ptr(), y.data_ptr(), N, H, H, C, stride, in_sums.data_ptr(),
                                                      gamma.data_ptr(), beta.data_ptr(), 1e-5, 0.1, rm.data_ptr(), rv.data_ptr(),
                                                      nbt.data_ptr(), mean.data_ptr(), invstd.data_ptr(), out_stats.data_ptr(), _sp()))
        torch.cuda.synchronize()
        xf = x.float().permute(0, 3, 1, 2)
        rm2, rv2 = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
        a1 = F.relu(F.batch_norm(xf, rm2, rv2, gamma, beta, True, 0.1, 1e-5))
        want = F.conv2d(a1, w.float(), None, stride, 1, 1, C).permute(0, 2, 3, 1)
        _close(y, want, 2e-2, 2e-2, "dw(bn1 on load) output")
        yf = y.float().view(-1, C)
        assert torch.allclose(out_stats[:C], yf.double().sum(0), rtol=1e-4, atol=1e-2)
        assert torch.allclose(out_stats[C:], (yf.double() ** 2).sum(0), rtol=1e-4, atol=1e-2)
        assert torch.allclose(mean, xf.mean((0, 2, 3)), atol=1e-4) and torch.allclose(rm, rm2, atol=1e-4)
        assert torch.allclose(rv, rv2, rtol=1e-3, atol=1e-4) and int(nbt) == 1
        assert float(in_sums.abs().max()) == 0.0 and in_sums.view(torch.int32)[-2:].abs().max().item() == 0, "sums not cleaned"


@pytest.mark.parametrize("N,H,C,stride", DW_CASES)
def test_dw_backward_with_bn1_fused(N, H, C, stride):
    """dgrad (+ ReLU mask + bn1 reductions), wgrad (activation re-created on load) and the elementwise bn1 dx pass, against
    autograd through relu(bn(x)) -> depthwise conv in fp32."""
    nat = _nat()
    lib = nat.kernels()
    torch.manual_seed(3 * N + H + C)
    x = (torch.randn(N, H, H, C, device="cuda") * 1.5 + 0.3).bfloat16()
    w = (torch.randn(C, 1, 3, 3, device="cuda") * 0.3).bfloat16()
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda") * 0.2
    Ho = (H + 2 - 3) // stride + 1
    dy = (torch.randn(N, Ho, Ho, C, device="cuda") * 0.5).bfloat16()
    # fp32 reference
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wr, gr, br = w.float().requires_grad_(True), gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    a1 = F.relu(F.batch_norm(xr, None, None, gr, br, True, 0.1, 1e-5))
    a1.retain_grad()
    F.conv2d(a1, wr, None, stride, 1, 1, C).backward(dy.float().permute(0, 3, 1, 2))
    mean = x.float().view(-1, C).mean(0)
    invstd = 1.0 / torch.sqrt(x.float().view(-1, C).var(0, unbiased=False) + 1e-5)
    # kernels
    dw = torch.empty(C, 1, 3, 3, device="cuda")
    nat.check(lib.dmp_dwconv3x3_wgrad_bnin(dy.data_ptr(), x.data_ptr(), dw.data_ptr(), N, H, H, C, stride, 0, mean.data_ptr(),
                                           invstd.data_ptr(), gamma.data_ptr(), beta.data_ptr(), _sp()))
    bs = torch.zeros(int(lib.dmp_bn_sums_bytes(C)) // 8, dtype=torch.float64, device="cuda")
    g = torch.empty_like(x)
    nat.check(lib.dmp_dwconv3x3_dgrad_bnbwd(dy.data_ptr(), w.data_ptr(), g.data_ptr(), N, H, H, C, stride, x.data_ptr(), mean.data_ptr(),
                                            invstd.data_ptr(), gamma.data_ptr(), beta.data_ptr(), bs.data_ptr(), _sp()))
    torch.cuda.synchronize()
    scale = dw.abs().max().item()
    _close(dw, wr.grad, 2e-2, 2e-2 * max(scale, 1.0), "dw weight gradient")
    _close(g.permute(0, 3, 1, 2), a1.grad * (a1 > 0), 2e-2, 2e-2, "masked data gradient", max_bad_frac=2e-3)
    gf = g.float().view(-1, C)
    xh = (x.float().view(-1, C) - mean) * invstd
    assert torch.allclose(bs[:C], gf.double().sum(0), rtol=1e-3, atol=5e-2), "sum g"
    assert torch.allclose(bs[C:2 * C], (gf * xh).double().sum(0), rtol=1e-3, atol=5e-2), "sum g*xhat"
    dgam, dbet = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    nat.check(lib.dmp_bn_bwd_dx_sums(g.data_ptr(), x.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), g.data_ptr(),
                                     dgam.data_ptr(), dbet.data_ptr(), 0, N * H * H, C, bs.data_ptr(), _sp()))
    torch.cuda.synchronize()
    M = N * H * H
    _close(g.permute(0, 3, 1, 2), xr.grad, 3e-2, 3e-2 * xr.grad.abs().max().item(), "bn1 dx", max_bad_frac=2e-3)
    _close(dgam, gr.grad, 2e-2, 0.05 * M ** 0.5, "dgamma")
    _close(dbet, br.grad, 2e-2, 0.05 * M ** 0.5, "dbeta")
    assert float(bs.abs().max()) == 0.0, "backward sums not cleaned"


@pytest.mark.parametrize("M,N,K,relu", [(4096, 144, 24, 1), (131072, 96, 16, 1), (8192, 384, 64, 1), (2048, 960, 160, 1),
                                         (1000, 40, 24, 0), (8192, 192, 32, 1), (32768, 64, 96, 1)])
def test_gemm_dgrad_with_bn_backward_reductions(M, N, K, relu):
    """G = mask * (dY[M,K] W[K,N]) and {sum G, sum G*xhat}: the NN data-gradient GEMM whose epilogue does the reduction phase
    of the backward of the act(BatchNorm(y_raw)) that produced the convolution's input."""
    nat = _nat()
    lib = nat.kernels()
    torch.manual_seed(M + N + K)
    dy = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(K, N, device="cuda") * 0.3).bfloat16()
    y_raw = (torch.randn(M, N, device="cuda") * 1.2 + 0.2).bfloat16()
    gamma, beta = torch.rand(N, device="cuda") + 0.5, torch.randn(N, device="cuda") * 0.2
    mean = y_raw.float().mean(0)
    invstd = 1.0 / torch.sqrt(y_raw.float().var(0, unbiased=False) + 1e-5)
    stats = torch.zeros(int(lib.dmp_bn_sums_bytes(N)) // 8, dtype=torch.float64, device="cuda")
    g = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    nat.check(lib.dmp_gemm_bf16_dgrad_bnbwd(dy.data_ptr(), K, w.data_ptr(), N, g.data_ptr(), N, M, N, K, y_raw.data_ptr(), N,
                                            mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), beta.data_ptr(), relu,
                                            stats.data_ptr(), _sp()))
    torch.cuda.synchronize()
    full = dy.float() @ w.float()
    a = (y_raw.float() - mean) * invstd * gamma + beta
    want = full * (a > 0) if relu else full
    _close(g, want, 2e-2, 2e-2, "masked gradient", max_bad_frac=2e-3)
    gf = g.float()
    xh = (y_raw.float() - mean) * invstd
    assert torch.allclose(stats[:N], gf.double().sum(0), rtol=1e-3, atol=5e-2), "sum g"
    assert torch.allclose(stats[N:2 * N], (gf * xh).double().sum(0), rtol=1e-3, atol=5e-2), "sum g*xhat"


def _block_pair(cin, cout, expansion, stride, seed):
    from distributed_model_parallel_b200.models.mobilenetv2 import Block

    torch.manual_seed(seed)
    blk = Block(cin, cout, expansion, stride).cuda()
    with torch.no_grad():
        for m in blk.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.weight.uniform_(0.5, 1.5)
                m.bias.normal_(0, 0.2)
    return blk


def _run_block(blk, x, dy=None):
    xk = x.clone().requires_grad_(True)
    out = blk(xk.to(memory_format=torch.channels_last))
    if dy is None:
        dy = torch.randn_like(out.float()) * 0.1
    out.backward(dy.to(out.dtype))
    return out, xk.grad, dy


@pytest.mark.parametrize("cin,cout,expansion,stride,hw", [(16, 24, 6, 1, 16), (24, 24, 6, 1, 16), (32, 64, 6, 2, 16), (32, 16, 1, 1, 8)])
def test_block_fused_chain_matches_unfused_kernels_and_reference(cin, cout, expansion, stride, hw):
    """One inverted-residual block through the fused chain against (a) the same block through the UNFUSED kernels (same bf16
    roundings: tight bound) and (b) the plain PyTorch restatement in fp32 (bf16-level bound): output, input gradient, every
    parameter gradient, BatchNorm running statistics."""
    import copy
    from distributed_model_parallel_b200.ops import functional as OF

    blk = _block_pair(cin, cout, expansion, stride, 7)
    unf, ref = copy.deepcopy(blk), copy.deepcopy(blk)
    torch.manual_seed(11)
    x = torch.randn(32, cin, hw, hw, device="cuda")
    OF.path_counts.clear()
    saved = (OF._FUSE_CHAIN, OF._FUSE_BN_BWD, OF._EPILOGUE_STATS)
    try:
        OF._FUSE_CHAIN = OF._FUSE_BN_BWD = OF._EPILOGUE_STATS = True   # the fusions are opt-in (DMP_FUSE_BN_CHAIN / DMP_FUSE_BN_BWD)
        out, dx, dy = _run_block(blk, x)
    finally:
        OF._FUSE_CHAIN, OF._FUSE_BN_BWD, OF._EPILOGUE_STATS = saved
    assert OF.path_counts["conv_dw:stencil+bn_in+bnstats"] == 1, dict(OF.path_counts)
    assert OF.path_counts["bn:bwd_dx_from_sums"] >= 2 and OF.path_counts["conv_pw:dgrad+bnbwd"] == 1, dict(OF.path_counts)
    try:
        OF._FUSE_CHAIN = OF._FUSE_BN_BWD = OF._EPILOGUE_STATS = False
        out_u, dx_u, _ = _run_block(unf, x, dy)
    finally:
        OF._FUSE_CHAIN, OF._FUSE_BN_BWD, OF._EPILOGUE_STATS = saved
    _close(out, out_u, 2e-2, 2e-2, "block output vs unfused kernels", max_bad_frac=1e-3)
    _close(dx, dx_u, 3e-2, 3e-2 * dx_u.abs().max().item(), "input gradient vs unfused kernels", max_bad_frac=0.015)  # ReLU-mask flips
    for (n, p), (_, q) in zip(blk.named_parameters(), unf.named_parameters()):
        # per-channel sums over thousands of bf16 gradients whose ReLU masks can differ in a few places between the two paths
        _close(p.grad, q.grad, 3e-2, 8e-2 * max(q.grad.abs().max().item(), 1e-3), "grad of %s vs unfused kernels" % n, max_bad_frac=0.05)
    for (n, b), (_, c) in zip(blk.named_buffers(), unf.named_buffers()):
        if b.dtype.is_floating_point:
            _close(b, c, 1e-2, 1e-3, "buffer %s vs unfused kernels" % n)
    with OF.use_reference_ops():
        xr = x.clone().requires_grad_(True)
        outr = ref(xr)
        outr.backward(dy)
    _close(out, outr, 5e-2, 5e-2, "block output", max_bad_frac=1e-3)
    _close(dx, xr.grad, 8e-2, 8e-2 * xr.grad.abs().max().item(), "block input gradient", max_bad_frac=5e-3)
    for (n, p), (_, q) in zip(blk.named_parameters(), ref.named_parameters()):
        tol = 0.12 * max(q.grad.abs().max().item(), 1e-3)
        _close(p.grad, q.grad, 6e-2, tol, "grad of %s" % n, max_bad_frac=0.1)  # per-channel sums: see tests/test_gpu_models.py on their noise
    for (n, b), (_, c) in zip(blk.named_buffers(), ref.named_buffers()):
        if b.dtype.is_floating_point:
            _close(b, c, 2e-2, 2e-2, "buffer %s" % n)
        else:
            assert int(b) == int(c), n
