"""Max-pool and the transformer row-wise kernels (LayerNorm, GELU, softmax), the batched tcgen05 GEMM and the attention op
built from them, each against a plain PyTorch fp32 restatement.  None of these has a counterpart in the reference: max-pool
serves the ResNet / VGG family it lists (/root/reference/code/distributed_training/data_parallel.py:58-73), the rest the
BERT-base pipeline config of BASELINE.json."""
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


@pytest.mark.parametrize("N,H,C,k,s,p", [(4, 16, 64, 2, 2, 0), (2, 56, 64, 3, 2, 1), (3, 15, 24, 3, 2, 1), (2, 8, 512, 2, 2, 0), (1, 9, 8, 3, 1, 1)])
def test_maxpool_forward_backward(N, H, C, k, s, p):
    from distributed_model_parallel_b200.ops import functional as OF

    torch.manual_seed(N + H + C)
    x = torch.randn(N, C, H, H, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    xk = x.clone().requires_grad_(True)
    y = OF.max_pool2d(xk, k, s, p)
    xr = x.float().requires_grad_(True)
    yr = F.max_pool2d(xr, k, s, p)
    assert torch.equal(y.float(), yr)
    dy = torch.randn_like(yr)
    y.backward(dy.bfloat16())
    yr.backward(dy.bfloat16().float())
    # ties between equal bf16 values may pick a different tap than torch: compare where the window has a unique maximum
    _close(xk.grad, xr.grad, 1e-2, 1e-2, "maxpool dx", max_bad_frac=2e-3)


@pytest.mark.parametrize("rows,H,res", [(512, 768, True), (100, 768, False), (64, 1024, True), (33, 64, True), (7, 2048, False)])
def test_layernorm_forward_backward(rows, H, res):
    nat = _nat()
    lib = nat.kernels()
    torch.manual_seed(rows + H)
    x = torch.randn(rows, H, device="cuda").bfloat16()
    r = torch.randn(rows, H, device="cuda").bfloat16() if res else None
    gamma, beta = torch.rand(H, device="cuda") + 0.5, torch.randn(H, device="cuda") * 0.1
    y, z = torch.empty_like(x), torch.empty_like(x)
    mean, rstd = torch.empty(rows, device="cuda"), torch.empty(rows, device="cuda")
    nat.check(lib.dmp_layernorm_fwd(x.data_ptr(), nat.ptr(r), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), z.data_ptr() if res else None,
                                    mean.data_ptr(), rstd.data_ptr(), rows, H, 1e-12, _sp()))
    zin = (x.float() + r.float()).bfloat16().float() if res else x.float()
    zr = zin.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    yr = F.layer_norm(zr, (H,), gr, br, 1e-12)
    torch.cuda.synchronize()
    _close(y, yr, 2e-2, 2e-2, "layernorm y")
    if res:
        assert torch.equal(z.float(), zin)
    dy = (torch.randn(rows, H, device="cuda") * 0.5).bfloat16()
    yr.backward(dy.float())
    dz = torch.empty_like(x)
    dg, db = torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
    zz = z if res else x
    nat.check(lib.dmp_layernorm_bwd(dy.data_ptr(), zz.data_ptr(), gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dz.data_ptr(),
                                    dg.data_ptr(), db.data_ptr(), rows, H, _sp()))
    torch.cuda.synchronize()
    _close(dz, zr.grad, 2e-2, 2e-2 * zr.grad.abs().max().item(), "layernorm dz")
    _close(dg, gr.grad, 1e-2, 1e-2 * max(1.0, gr.grad.abs().max().item()), "layernorm dgamma")
    _close(db, br.grad, 1e-2, 1e-2 * max(1.0, br.grad.abs().max().item()), "layernorm dbeta")


def test_gelu_and_softmax():
    nat = _nat()
    lib = nat.kernels()
    torch.manual_seed(0)
    x = (torch.randn(1000, 3072, device="cuda") * 2).bfloat16()
    y = torch.empty_like(x)
    nat.check(lib.dmp_gelu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _sp()))
    xr = x.float().requires_grad_(True)
    yr = F.gelu(xr)
    _close(y, yr, 1e-2, 1e-2, "gelu")
    dy = torch.randn_like(yr).bfloat16()
    yr.backward(dy.float())
    dx = torch.empty_like(x)
    nat.check(lib.dmp_gelu_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), x.numel(), _sp()))
    _close(dx, xr.grad, 1e-2, 1e-2, "gelu bwd")
    for rows, L in ((96 * 128, 128), (50, 512), (7, 8)):
        s = (torch.randn(rows, L, device="cuda") * 3).bfloat16()
        p = torch.empty_like(s)
        nat.check(lib.dmp_softmax_fwd(s.data_ptr(), p.data_ptr(), rows, L, 0.125, _sp()))
        sr = s.float().requires_grad_(True)
        pr = torch.softmax(sr * 0.125, -1)
        _close(p, pr, 2e-2, 1e-3, "softmax")
        dp = torch.randn(rows, L, device="cuda").bfloat16()
        ds = torch.empty_like(s)
        nat.check(lib.dmp_softmax_bwd(dp.data_ptr(), p.data_ptr(), ds.data_ptr(), rows, L, 0.125, _sp()))
        (p.float().detach().requires_grad_(False))
        pr.backward(dp.float())
        _close(ds, sr.grad, 3e-2, 2e-3, "softmax bwd")


@pytest.mark.parametrize("B,S,heads,d", [(2, 128, 4, 64), (3, 64, 12, 64), (1, 256, 2, 64), (2, 40, 3, 32)])
def test_attention_matches_sdpa(B, S, heads, d):
    """Forward and all three input gradients of the batched-GEMM attention against scaled_dot_product_attention in fp32."""
    from distributed_model_parallel_b200.ops import functional as OF

    torch.manual_seed(B * S + heads)
    Hd = heads * d
    qkv = (torch.randn(B, S, 3 * Hd, device="cuda") * 0.7).bfloat16()
    qk = qkv.clone().requires_grad_(True)
    OF.path_counts.clear()
    out = OF.attention(qk, heads)
    assert OF.path_counts["attention:tcgen05_batched"] == 1
    qr = qkv.float().requires_grad_(True)
    q, k, v = qr.view(B, S, 3, heads, d).permute(2, 0, 3, 1, 4)
    ref = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(B, S, Hd)
    _close(out, ref, 3e-2, 3e-2, "attention output")
    do = (torch.randn(B, S, Hd, device="cuda") * 0.5).bfloat16()
    out.backward(do)
    ref.backward(do.float())
    _close(qk.grad, qr.grad, 5e-2, 5e-2 * qr.grad.abs().max().item(), "attention dqkv", max_bad_frac=1e-3)


@pytest.mark.parametrize("layout", [0, 1, 2])
def test_batched_gemm_strided_operands(layout):
    """nh x nb GEMMs in one launch with operands addressed by strides inside larger tensors."""
    nat = _nat()
    lib = nat.kernels()
    torch.manual_seed(layout)
    nh, nb, M, N, K = 3, 2, 200, 72, 136
    if layout == 0:
        A = torch.randn(nb, nh, M, K, device="cuda").bfloat16(); Bm = torch.randn(nb, nh, N, K, device="cuda").bfloat16()
        ref = A.float() @ Bm.float().transpose(-1, -2)
        lda, ldb = K, K
        a_sh, a_sb, b_sh, b_sb = M * K, nh * M * K, N * K, nh * N * K
    elif layout == 1:
        A = torch.randn(nb, nh, M, K, device="cuda").bfloat16(); Bm = torch.randn(nb, nh, K, N, device="cuda").bfloat16()
        ref = A.float() @ Bm.float()
        lda, ldb = K, N
        a_sh, a_sb, b_sh, b_sb = M * K, nh * M * K, K * N, nh * K * N
    else:
        A = torch.randn(nb, nh, K, M, device="cuda").bfloat16(); Bm = torch.randn(nb, nh, K, N, device="cuda").bfloat16()
        ref = A.float().transpose(-1, -2) @ Bm.float()
        lda, ldb = M, N
        a_sh, a_sb, b_sh, b_sb = K * M, nh * K * M, K * N, nh * K * N
    D = torch.zeros(nb, M, nh, N, device="cuda", dtype=torch.bfloat16)   # heads interleaved inside the rows, like the context tensor
    nat.check(lib.dmp_gemm_bf16_batched(layout, A.data_ptr(), lda, a_sh, a_sb, Bm.data_ptr(), ldb, b_sh, b_sb, D.data_ptr(), nh * N, N,
                                        M * nh * N, M, N, K, nh, nb, _sp()))
    torch.cuda.synchronize()
    _close(D.permute(0, 2, 1, 3), ref, 2e-2, 2e-2 * ref.abs().max().item(), "batched gemm layout %d" % layout)


@pytest.mark.parametrize("N,H,Cin,Cout,k,stride", [(4, 16, 64, 128, 3, 2), (2, 32, 128, 256, 3, 2), (3, 8, 256, 512, 3, 2), (2, 14, 64, 64, 3, 2),
                                                     (2, 16, 64, 64, 3, 1), (2, 12, 128, 64, 5, 1), (8, 64, 3, 64, 7, 2), (4, 32, 3, 64, 3, 1)])
def test_dense_conv_paths_without_library_kernels(N, H, Cin, Cout, k, stride):
    """Dense KxK convolutions of the ResNet / VGG family (stride 1 and 2, 7x7 stem) through the op layer: forward, data gradient
    and weight gradient against F.conv2d in fp32, and no library convolution on the path."""
    from distributed_model_parallel_b200.ops import functional as OF

    torch.manual_seed(N + H + Cin + Cout + k)
    conv = torch.nn.Conv2d(Cin, Cout, k, stride, k // 2, bias=False).cuda()
    x = torch.randn(N, Cin, H, H, device="cuda")
    stem = Cin <= 4
    xk = x.clone().requires_grad_(not stem)
    OF.path_counts.clear()
    y = OF.conv_bn_act(xk if stem else xk.bfloat16().contiguous(memory_format=torch.channels_last), conv, None, relu=False)
    assert "conv_generic:cudnn" not in OF.path_counts, dict(OF.path_counts)
    if stem and k == 7:  # large stems: im2col rows + tensor-core GEMMs, not the per-tap kernels
        assert OF.path_counts.get("conv_stem:im2col+tcgen05"), dict(OF.path_counts)
    xr = x.bfloat16().float().requires_grad_(True) if not stem else x.clone().requires_grad_(True)
    wr = conv.weight.detach().bfloat16().float().requires_grad_(True) if not stem else conv.weight.detach().clone().requires_grad_(True)
    yr = F.conv2d(xr, wr, None, stride, k // 2)
    _close(y, yr, 2e-2, 2e-2 * yr.abs().max().item(), "conv forward")
    dy = torch.randn_like(yr) * 0.5
    y.backward(dy.to(y.dtype))
    yr.backward(dy.bfloat16().float())
    torch.cuda.synchronize()
    _close(conv.weight.grad, wr.grad, 2e-2, 2e-2 * wr.grad.abs().max().item(), "conv weight gradient")
    if not stem:
        _close(xk.grad, xr.grad, 2e-2, 2e-2 * xr.grad.abs().max().item(), "conv data gradient")
