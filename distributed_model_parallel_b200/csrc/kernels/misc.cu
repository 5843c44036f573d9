// Small ops around the GEMM/BN core: stem convolution (Cin <= 4), global average pool, skinny
// classifier linear, elementwise add/ReLU.  Reference call sites:
//   stem conv 3->32    /root/reference/code/distributed_training/model/mobilenetv2.py:52,69
//   avg_pool2d(4)+view mobilenetv2.py:73-74,155-157 (Reshape1)
//   Linear(1280,10)    mobilenetv2.py:57,75
#include "../common.cuh"

namespace {

__device__ __forceinline__ void load8(const bf16* p, float* f) {
    int4 raw = *(const int4*)p;
    bf16x8_to_float(*(bf16x8*)&raw, f);
}
__device__ __forceinline__ void store8(bf16* p, const float* f) {
    bf16x8 b = float_to_bf16x8(f);
    *(int4*)p = *(int4*)&b;
}

// ----------------------------------------------------------------------------- stem convolution
// x: layout 0 = fp32 NCHW (raw images), 1 = bf16 NHWC.
__device__ __forceinline__ float stem_x(const void* x, int layout, int n, int ci, int h, int w, int Cin, int H, int W) {
    if (layout == 0) return ((const float*)x)[(((long long)n * Cin + ci) * H + h) * W + w];
    return __bfloat162float(((const bf16*)x)[(((long long)n * H + h) * W + w) * Cin + ci]);
}

// y[n,ho,wo,co] = sum_{ci,kh,kw} x[n,ci,ho*s+kh-p,wo*s+kw-p] * w[co,ci,kh,kw];  thread = (pixel, 8 couts)
template <typename WT>
__global__ void stem_fwd_kernel(const void* __restrict__ x, int layout, const WT* __restrict__ w, bf16* __restrict__ y,
                                int N, int Cin, int H, int W, int Cout, int KH, int KW, int stride, int pad, int Ho,
                                int Wo) {
    extern __shared__ float sw[];  // [T][Cout]
    const int T = Cin * KH * KW;
    for (int i = threadIdx.x; i < T * Cout; i += blockDim.x) {
        const int co = i / T, tap = i % T;
        sw[tap * Cout + co] = to_f32<WT>(w[i]);
    }
    __syncthreads();
    const int G = Cout >> 3;
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long total = (long long)N * Ho * Wo * G;
    if (idx >= total) return;
    const int cg = (int)(idx % G);
    long long pix = idx / G;
    const int wo = (int)(pix % Wo); pix /= Wo;
    const int ho = (int)(pix % Ho);
    const int n = (int)(pix / Ho);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int ci = 0; ci < Cin; ++ci)
        for (int kh = 0; kh < KH; ++kh) {
            const int hi = ho * stride + kh - pad;
            if (hi < 0 || hi >= H) continue;
            for (int kw = 0; kw < KW; ++kw) {
                const int wi = wo * stride + kw - pad;
                if (wi < 0 || wi >= W) continue;
                const float xv = stem_x(x, layout, n, ci, hi, wi, Cin, H, W);
                const float* wr = sw + ((ci * KH + kh) * KW + kw) * Cout + cg * 8;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += xv * wr[j];
            }
        }
    store8(y + (((long long)n * Ho + ho) * Wo + wo) * Cout + cg * 8, acc);
}

// 3x3 stems with a compile-time tap count: every tap is fetched unconditionally from a clamped address and zeroed by its
// bounds predicate, so the T loads of a pixel are independent and in flight together.  (With the loads under their bounds
// branches -- the generic kernel above -- each one waits for the previous: 27 serial L2 round trips per pixel.)
template <int CIN, int KS, typename WT>
__global__ void __launch_bounds__(256)
stem_fwd_unrolled_kernel(const void* __restrict__ x, int layout, const WT* __restrict__ w, bf16* __restrict__ y, int N, int H,
                         int W, int Cout, int stride, int pad, int Ho, int Wo) {
    extern __shared__ float sw[];  // [T][Cout]
    constexpr int T = CIN * KS * KS;
    for (int i = threadIdx.x; i < T * Cout; i += blockDim.x) {
        const int co = i / T, tap = i % T;
        sw[tap * Cout + co] = to_f32<WT>(w[i]);
    }
    __syncthreads();
    const int G = Cout >> 3;
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long total = (long long)N * Ho * Wo * G;
    if (idx >= total) return;
    const int cg = (int)(idx % G);
    long long pix = idx / G;
    const int wo = (int)(pix % Wo); pix /= Wo;
    const int ho = (int)(pix % Ho);
    const int n = (int)(pix / Ho);
    float xv[T];
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
        for (int kh = 0; kh < KS; ++kh)
#pragma unroll
            for (int kw = 0; kw < KS; ++kw) {
                const int hi = ho * stride + kh - pad, wi = wo * stride + kw - pad;
                const bool ok = (unsigned)hi < (unsigned)H && (unsigned)wi < (unsigned)W;
                const float v = stem_x(x, layout, n, ci, min(max(hi, 0), H - 1), min(max(wi, 0), W - 1), CIN, H, W);
                xv[(ci * KS + kh) * KS + kw] = ok ? v : 0.f;
            }
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const float4 w0 = *(const float4*)(sw + t * Cout + cg * 8), w1 = *(const float4*)(sw + t * Cout + cg * 8 + 4);
        acc[0] += xv[t] * w0.x; acc[1] += xv[t] * w0.y; acc[2] += xv[t] * w0.z; acc[3] += xv[t] * w0.w;
        acc[4] += xv[t] * w1.x; acc[5] += xv[t] * w1.y; acc[6] += xv[t] * w1.z; acc[7] += xv[t] * w1.w;
    }
    store8(y + (((long long)n * Ho + ho) * Wo + wo) * Cout + cg * 8, acc);
}

// dw[co,tap] += sum_pix dy[pix,co] * x[pix (+) tap];  thread = (tap, 8 couts), block strides over pixels
__global__ void stem_wgrad_kernel(const bf16* __restrict__ dy, const void* __restrict__ x, int layout,
                                  float* __restrict__ dw, int N, int Cin, int H, int W, int Cout, int KH, int KW,
                                  int stride, int pad, int Ho, int Wo) {
    const int T = Cin * KH * KW;
    const int G = Cout >> 3;
    const int tc = blockIdx.y * blockDim.x + threadIdx.x;  // (tap, 8 couts) pairs beyond one block go to further blockIdx.y slices
    const int tap = tc / G, cg = tc % G;
    if (tap >= T) return;
    const int ci = tap / (KH * KW), kh = (tap / KW) % KH, kw = tap % KW;
    const long long npix = (long long)N * Ho * Wo;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    constexpr int U = 8;  // pixels in flight per thread: the loop body is two dependent-free global loads per pixel
    for (long long pix0 = blockIdx.x; pix0 < npix; pix0 += (long long)U * gridDim.x) {
        float xv[U];
        int4 gr[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long pix = pix0 + (long long)u * gridDim.x;
            xv[u] = 0.f;
            gr[u] = make_int4(0, 0, 0, 0);
            if (pix < npix) {
                long long t = pix;
                const int wo = (int)(t % Wo); t /= Wo;
                const int ho = (int)(t % Ho);
                const int n = (int)(t / Ho);
                const int hi = ho * stride + kh - pad, wi = wo * stride + kw - pad;
                if (hi >= 0 && hi < H && wi >= 0 && wi < W) {
                    xv[u] = stem_x(x, layout, n, ci, hi, wi, Cin, H, W);
                    gr[u] = *(const int4*)(dy + pix * Cout + cg * 8);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float g[8];
            bf16x8_to_float(*(bf16x8*)&gr[u], g);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += g[j] * xv[u];
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&dw[(long long)(cg * 8 + j) * T + tap], acc[j]);
}


// 3x3 stems (27 or fewer taps): lane = pixel, warp = group of NCO output channels.  A thread keeps the whole [tap][NCO couts]
// slab of the weight gradient in registers and walks pixels with a grid stride, so every load is coalesced along the pixel axis
// (x) or shared by the block's warps through L1 (dy), and each loaded value feeds NCO (x) or up to 27 (dy) FMAs.
// The taps of a pixel are fetched unconditionally from clamped addresses and zeroed by their bounds predicates: under their
// bounds branches the 27 loads ran one after the other (186 us for the MobileNetV2 stem at batch 128 inside the step).
template <int CIN, int KS, int NCO>
__global__ void __launch_bounds__(256)
stem_wgrad_reg_kernel(const bf16* __restrict__ dy, const void* __restrict__ x, int layout, float* __restrict__ dw, int N,
                      int H, int W, int Cout, int stride, int pad, int Ho, int Wo) {
    constexpr int T = CIN * KS * KS;
    static_assert(NCO == 2 || NCO == 4, "NCO");
    const int lane = threadIdx.x & 31, cog = blockIdx.y * 8 + (threadIdx.x >> 5);  // 8 warps = 8*NCO couts per blockIdx.y
    const long long npix = (long long)N * Ho * Wo;
    float acc[T][NCO];
#pragma unroll
    for (int t = 0; t < T; ++t)
#pragma unroll
        for (int j = 0; j < NCO; ++j) acc[t][j] = 0.f;
    for (long long pix = (long long)blockIdx.x * 32 + lane; pix < npix; pix += (long long)gridDim.x * 32) {
        const unsigned up = (unsigned)pix;  // npix < 2^31 is checked by the host
        const int wo = (int)(up % (unsigned)Wo);
        const unsigned r = up / (unsigned)Wo;
        const int ho = (int)(r % (unsigned)Ho);
        const int n = (int)(r / (unsigned)Ho);
        float g[NCO];
        if (NCO == 4) {
            const uint2 graw = *(const uint2*)(dy + pix * Cout + cog * 4);
            g[0] = __uint_as_float(graw.x << 16); g[1] = __uint_as_float(graw.x & 0xffff0000u);
            g[2] = __uint_as_float(graw.y << 16); g[NCO - 1] = __uint_as_float(graw.y & 0xffff0000u);
        } else {
            const unsigned graw = *(const unsigned*)(dy + pix * Cout + cog * 2);
            g[0] = __uint_as_float(graw << 16); g[1] = __uint_as_float(graw & 0xffff0000u);
        }
        float xv[T];
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
            for (int kh = 0; kh < KS; ++kh)
#pragma unroll
                for (int kw = 0; kw < KS; ++kw) {
                    const int hi = ho * stride + kh - pad, wi = wo * stride + kw - pad;
                    const bool ok = (unsigned)hi < (unsigned)H && (unsigned)wi < (unsigned)W;
                    const float v = stem_x(x, layout, n, ci, min(max(hi, 0), H - 1), min(max(wi, 0), W - 1), CIN, H, W);
                    xv[(ci * KS + kh) * KS + kw] = ok ? v : 0.f;
                }
#pragma unroll
        for (int t = 0; t < T; ++t)
#pragma unroll
            for (int j = 0; j < NCO; ++j) acc[t][j] += g[j] * xv[t];
    }
#pragma unroll
    for (int t = 0; t < T; ++t)
#pragma unroll
        for (int j = 0; j < NCO; ++j) {
            float v = acc[t][j];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) atomicAdd(&dw[(long long)(cog * NCO + j) * T + t], v);
        }
}

// ----------------------------------------------------------------------------- stem convolution as im2col + tcgen05 GEMM
// Large stems (7x7: 147 taps) as GEMMs: the input patch of every output pixel is written once as a bf16 row [pixel][Kp]
// (column t = (ci*KH + kh)*KW + kw, columns >= Cin*KH*KW zero, Kp a multiple of 8 so that rows are 16-byte aligned for TMA), then
// forward is cols x W^T and the weight gradient dY^T x cols on the tensor cores (dmp_gemm_bf16).  The per-tap kernels above
// issue 2 loads per 8 FMAs and took 2.6 ms / 15.3 ms (forward / weight gradient) on ResNet-50's stem at batch 64.
__global__ void __launch_bounds__(256)
stem_im2col_kernel(const void* __restrict__ x, int layout, bf16* __restrict__ cols, int N, int Cin, int H, int W, int KH, int KW,
                   int stride, int pad, int Ho, int Wo, int Kp) {
    const int G = Kp >> 3;
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long total = (long long)N * Ho * Wo * G;
    if (idx >= total) return;
    const int g = (int)(idx % G);
    long long pix = idx / G;
    const int wo = (int)(pix % Wo);
    long long r = pix / Wo;
    const int ho = (int)(r % Ho);
    const int n = (int)(r / Ho);
    const int T = Cin * KH * KW, KK = KH * KW;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int t = g * 8 + j;
        const int tc = t < T ? t : 0;
        const int ci = tc / KK, rr = tc - ci * KK;
        const int kh = rr / KW, kw = rr - kh * KW;
        const int hi = ho * stride + kh - pad, wi = wo * stride + kw - pad;
        const bool ok = t < T && (unsigned)hi < (unsigned)H && (unsigned)wi < (unsigned)W;
        const float val = stem_x(x, layout, n, ci, min(max(hi, 0), H - 1), min(max(wi, 0), W - 1), Cin, H, W);
        v[j] = ok ? val : 0.f;
    }
    store8(cols + pix * Kp + g * 8, v);
}

// dst[r][c] (bf16, row pitch dst_pitch elements) = src[r][c] for c < cols, 0 up to dst_pitch.  src: fp32 (0) or bf16 (1), dense.
__global__ void pack_rows_bf16_kernel(const void* __restrict__ src, int src_dtype, bf16* __restrict__ dst, int rows, int cols,
                                      int dst_pitch) {
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (idx >= (long long)rows * dst_pitch) return;
    const int c = (int)(idx % dst_pitch);
    const long long r = idx / dst_pitch;
    float v = 0.f;
    if (c < cols) v = src_dtype == 0 ? ((const float*)src)[r * cols + c] : __bfloat162float(((const bf16*)src)[r * cols + c]);
    dst[idx] = __float2bfloat16_rn(v);
}

// Filter packing for the implicit-GEMM convolutions in ONE launch: w[Cout][Cin][KH][KW] (fp32 master or bf16 shadow) ->
//   mode 0: out[Cout][KH][KW][Cin]                       forward / weight-gradient operand order
//   mode 1: out[Cin][KH][KW][Cout]                       stride-2 data gradient (taps selected per output parity)
//   mode 2: out[Cin][KH-1-kh][KW-1-kw][Cout]             stride-1 data gradient (flipped taps, channels swapped)
// (permute + flip + contiguous through the tensor library were 2-3 generic copy kernels per convolution and direction:
// 54 launches, 14 % of a ResNet-18 step.)
__global__ void pack_filter_kernel(const void* __restrict__ w, int w_dtype, bf16* __restrict__ out, int Cout, int Cin, int KH,
                                   int KW, int mode) {
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long total = (long long)Cout * Cin * KH * KW;
    if (idx >= total) return;
    int co, ci, kh, kw;
    long long t = idx;
    if (mode == 0) {
        ci = (int)(t % Cin); t /= Cin;
        kw = (int)(t % KW); t /= KW;
        kh = (int)(t % KH);
        co = (int)(t / KH);
    } else {
        co = (int)(t % Cout); t /= Cout;
        kw = (int)(t % KW); t /= KW;
        kh = (int)(t % KH);
        ci = (int)(t / KH);
        if (mode == 2) { kh = KH - 1 - kh; kw = KW - 1 - kw; }
    }
    const long long src = (((long long)co * Cin + ci) * KH + kh) * KW + kw;
    const float v = w_dtype == 0 ? ((const float*)w)[src] : __bfloat162float(((const bf16*)w)[src]);
    out[idx] = __float2bfloat16_rn(v);
}

// ----------------------------------------------------------------------------- global average pool
__global__ void avgpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int N, int HW, int C, int relu) {
    const int P = C >> 3;
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (idx >= (long long)N * P) return;
    const int cv = (int)(idx % P), n = (int)(idx / P);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int i = 0; i < HW; ++i) {
        float f[8];
        load8(x + ((long long)n * HW + i) * C + cv * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += relu ? fmaxf(f[j], 0.f) : f[j];
    }
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    store8(y + (long long)n * C + cv * 8, acc);
}
// dx[n,i,c] = dy[n,c] / HW  (masked by x > 0 when the pool had a fused ReLU in front)
__global__ void avgpool_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, bf16* __restrict__ dx, int N,
                                   int HW, int C, int relu) {
    const int P = C >> 3;
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (idx >= (long long)N * HW * P) return;
    const int cv = (int)(idx % P);
    const long long row = idx / P;  // n*HW + i
    const int n = (int)(row / HW);
    float g[8];
    load8(dy + (long long)n * C + cv * 8, g);
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= inv;
    if (relu) {
        float f[8];
        load8(x + row * C + cv * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = f[j] > 0.f ? g[j] : 0.f;
    }
    store8(dx + row * C + cv * 8, g);
}

// ----------------------------------------------------------------------------- skinny linear (N small, e.g. 10 classes)
// y[b,n] = sum_k x[b,k] * w[n,k] + bias[n];  one warp per (b,n)
template <typename WT>
__global__ void linear_small_fwd_kernel(const bf16* __restrict__ x, const WT* __restrict__ w, const float* __restrict__ bias,
                                        float* __restrict__ y, int B, int N, int K) {
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (gw >= B * N) return;
    const int b = gw / N, n = gw % N;
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) acc += __bfloat162float(x[(long long)b * K + k]) * to_f32<WT>(w[(long long)n * K + k]);
    acc = warp_sum(acc);
    if (lane == 0) y[(long long)b * N + n] = acc + (bias ? bias[n] : 0.f);
}
// dx[b,k] = sum_n dy[b,n] * w[n,k]
template <typename WT>
__global__ void linear_small_dx_kernel(const float* __restrict__ dy, const WT* __restrict__ w, bf16* __restrict__ dx, int B,
                                       int N, int K) {
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (idx >= (long long)B * K) return;
    const int b = (int)(idx / K), k = (int)(idx % K);
    float acc = 0.f;
    for (int n = 0; n < N; ++n) acc += dy[(long long)b * N + n] * to_f32<WT>(w[(long long)n * K + k]);
    dx[idx] = __float2bfloat16_rn(acc);
}
// dw[n,k] = sum_b dy[b,n] * x[b,k];  db[n] = sum_b dy[b,n]
__global__ void linear_small_dw_kernel(const float* __restrict__ dy, const bf16* __restrict__ x, float* __restrict__ dw,
                                       float* __restrict__ db, int B, int N, int K, int accumulate) {
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (idx < (long long)N * K) {
        const int n = (int)(idx / K), k = (int)(idx % K);
        float acc = 0.f;
        for (int b = 0; b < B; ++b) acc += dy[(long long)b * N + n] * __bfloat162float(x[(long long)b * K + k]);
        dw[idx] = accumulate ? dw[idx] + acc : acc;
    } else if (db != nullptr && idx < (long long)N * K + N) {
        const int n = (int)(idx - (long long)N * K);
        float acc = 0.f;
        for (int b = 0; b < B; ++b) acc += dy[(long long)b * N + n];
        db[n] = accumulate ? db[n] + acc : acc;
    }
}

// ----------------------------------------------------------------------------- elementwise
// y = act(a + b) (b nullable), bf16; relu_bwd: dx = (y > 0) ? dy : 0
__global__ void add_act_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, bf16* __restrict__ y, long long nvec,
                               int relu) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += stride) {
        float f[8];
        load8(a + v * 8, f);
        if (b != nullptr) {
            float g[8];
            load8(b + v * 8, g);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] += g[j];
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        store8(y + v * 8, f);
    }
}
__global__ void relu_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ y, bf16* __restrict__ dx, long long nvec) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += stride) {
        float g[8], f[8];
        load8(dy + v * 8, g);
        load8(y + v * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = f[j] > 0.f ? g[j] : 0.f;
        store8(dx + v * 8, g);
    }
}

}  // namespace

DMP_API int dmp_stem_conv_fwd(const void* x, int x_layout, const void* w, int w_dtype, void* y, int N, int Cin, int H, int W,
                              int Cout, int KH, int KW, int stride, int pad, cudaStream_t stream) {
    DMP_REQUIRE(Cin >= 1 && Cin <= 4 && Cout % 8 == 0, "dmp_stem_conv_fwd: needs Cin<=4 and Cout%%8==0 (Cin=%d Cout=%d)", Cin,
                Cout);
    const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
    const int T = Cin * KH * KW;
    const size_t smem = (size_t)T * Cout * sizeof(float);
    DMP_REQUIRE(smem <= 48 * 1024, "dmp_stem_conv_fwd: filter too large for shared memory (%zu bytes)", smem);
    const long long total = (long long)N * Ho * Wo * (Cout / 8);
    const int grid = dmp_ceil_div(total, 256);
    if (KH == 3 && KW == 3 && Cin == 3) {
        if (w_dtype == 0)
            stem_fwd_unrolled_kernel<3, 3, float><<<grid, 256, smem, stream>>>(x, x_layout, (const float*)w, (bf16*)y, N, H, W, Cout,
                                                                              stride, pad, Ho, Wo);
        else
            stem_fwd_unrolled_kernel<3, 3, bf16><<<grid, 256, smem, stream>>>(x, x_layout, (const bf16*)w, (bf16*)y, N, H, W, Cout,
                                                                             stride, pad, Ho, Wo);
        DMP_LAUNCH_CHECK();
        DMP_COUNT_LAUNCH(1);
        return 0;
    }
    if (w_dtype == 0)
        stem_fwd_kernel<float><<<grid, 256, smem, stream>>>(x, x_layout, (const float*)w, (bf16*)y, N, Cin, H, W, Cout, KH, KW,
                                                            stride, pad, Ho, Wo);
    else
        stem_fwd_kernel<bf16><<<grid, 256, smem, stream>>>(x, x_layout, (const bf16*)w, (bf16*)y, N, Cin, H, W, Cout, KH, KW,
                                                           stride, pad, Ho, Wo);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_stem_conv_wgrad(const void* dy, const void* x, int x_layout, float* dw, int N, int Cin, int H, int W,
                                int Cout, int KH, int KW, int stride, int pad, int accumulate, cudaStream_t stream) {
    DMP_REQUIRE(Cin >= 1 && Cin <= 4 && Cout % 8 == 0, "dmp_stem_conv_wgrad: needs Cin<=4 and Cout%%8==0");
    const int T = Cin * KH * KW;
    const int threads = T * (Cout / 8);
    const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
    if (!accumulate) DMP_CHECK_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)T * Cout, stream));
    if (KH == 3 && KW == 3 && (Cin == 3 || Cin == 1) && Cout % 16 == 0 && (long long)N * Ho * Wo < (1ll << 31)) {
        // 2 couts per warp: 54 accumulators, two 256-thread blocks per SM; one block column per 16 couts, a whole number of
        // waves of (blocks per SM) x SMs
        const long long npix = (long long)N * Ho * Wo;
        const int ny = Cout / 16;
        long long gx = (npix + 32 * 4 - 1) / (32 * 4);  // >= 4 pixels per lane, else the final reduction dominates
        const long long cap = (long long)dmp_num_sms() * 2 / ny;
        if (gx > cap) gx = cap;
        if (gx < 1) gx = 1;
        const dim3 grid((unsigned)gx, (unsigned)ny);
        if (Cin == 3)
            stem_wgrad_reg_kernel<3, 3, 2><<<grid, 256, 0, stream>>>((const bf16*)dy, x, x_layout, dw, N, H, W, Cout, stride, pad, Ho, Wo);
        else
            stem_wgrad_reg_kernel<1, 3, 2><<<grid, 256, 0, stream>>>((const bf16*)dy, x, x_layout, dw, N, H, W, Cout, stride, pad, Ho, Wo);
        DMP_LAUNCH_CHECK();
        DMP_COUNT_LAUNCH(1);
        return 0;
    }
    const int slices = (threads + 511) / 512;  // <= 512 threads per block: the kernel needs ~100 registers per thread
    const int block = (((threads + slices - 1) / slices + 31) / 32) * 32;
    const dim3 grid(dmp_num_sms() * 4 / slices + 1, slices);
    stem_wgrad_kernel<<<grid, block, 0, stream>>>((const bf16*)dy, x, x_layout, dw, N, Cin, H, W, Cout, KH, KW, stride, pad,
                                                  Ho, Wo);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_stem_im2col(const void* x, int x_layout, void* cols, int N, int Cin, int H, int W, int KH, int KW, int stride, int pad,
                            int Kp, cudaStream_t stream) {
    DMP_REQUIRE(Cin >= 1 && Cin <= 4 && Kp % 8 == 0 && Kp >= Cin * KH * KW, "dmp_stem_im2col: needs Cin<=4 and Kp a multiple of 8 >= Cin*KH*KW");
    const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
    const long long total = (long long)N * Ho * Wo * (Kp / 8);
    stem_im2col_kernel<<<dmp_ceil_div(total, 256), 256, 0, stream>>>(x, x_layout, (bf16*)cols, N, Cin, H, W, KH, KW, stride, pad, Ho, Wo, Kp);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_pack_rows_bf16(const void* src, int src_dtype, void* dst, int rows, int cols, int dst_pitch, cudaStream_t stream) {
    DMP_REQUIRE(rows > 0 && cols > 0 && dst_pitch >= cols, "dmp_pack_rows_bf16: bad shape");
    const long long total = (long long)rows * dst_pitch;
    pack_rows_bf16_kernel<<<dmp_ceil_div(total, 256), 256, 0, stream>>>(src, src_dtype, (bf16*)dst, rows, cols, dst_pitch);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_pack_filter(const void* w, int w_dtype, void* out, int Cout, int Cin, int KH, int KW, int mode, cudaStream_t stream) {
    DMP_REQUIRE(Cout > 0 && Cin > 0 && KH > 0 && KW > 0 && mode >= 0 && mode <= 2, "dmp_pack_filter: bad arguments");
    const long long total = (long long)Cout * Cin * KH * KW;
    pack_filter_kernel<<<dmp_ceil_div(total, 256), 256, 0, stream>>>(w, w_dtype, (bf16*)out, Cout, Cin, KH, KW, mode);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_avgpool_fwd(const void* x, void* y, int N, int HW, int C, int relu, cudaStream_t stream) {
    DMP_REQUIRE(C % 8 == 0, "dmp_avgpool_fwd: C %% 8 != 0");
    const long long total = (long long)N * (C / 8);
    avgpool_fwd_kernel<<<dmp_ceil_div(total, 128), 128, 0, stream>>>((const bf16*)x, (bf16*)y, N, HW, C, relu);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
DMP_API int dmp_avgpool_bwd(const void* dy, const void* x, void* dx, int N, int HW, int C, int relu, cudaStream_t stream) {
    DMP_REQUIRE(C % 8 == 0, "dmp_avgpool_bwd: C %% 8 != 0");
    const long long total = (long long)N * HW * (C / 8);
    avgpool_bwd_kernel<<<dmp_ceil_div(total, 256), 256, 0, stream>>>((const bf16*)dy, (const bf16*)x, (bf16*)dx, N, HW, C,
                                                                     relu);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_linear_small_fwd(const void* x, const void* w, int w_dtype, const float* bias, float* y, int B, int N, int K,
                                 cudaStream_t stream) {
    const long long warps = (long long)B * N;
    const int grid = dmp_ceil_div(warps * 32, 256);
    if (w_dtype == 0) linear_small_fwd_kernel<float><<<grid, 256, 0, stream>>>((const bf16*)x, (const float*)w, bias, y, B, N, K);
    else linear_small_fwd_kernel<bf16><<<grid, 256, 0, stream>>>((const bf16*)x, (const bf16*)w, bias, y, B, N, K);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
DMP_API int dmp_linear_small_bwd(const float* dy, const void* x, const void* w, int w_dtype, void* dx, float* dw, float* db,
                                 int B, int N, int K, int accumulate, cudaStream_t stream) {
    if (dx != nullptr) {
        const int grid = dmp_ceil_div((long long)B * K, 256);
        if (w_dtype == 0) linear_small_dx_kernel<float><<<grid, 256, 0, stream>>>(dy, (const float*)w, (bf16*)dx, B, N, K);
        else linear_small_dx_kernel<bf16><<<grid, 256, 0, stream>>>(dy, (const bf16*)w, (bf16*)dx, B, N, K);
        DMP_LAUNCH_CHECK();
        DMP_COUNT_LAUNCH(1);
    }
    if (dw != nullptr) {
        const int grid = dmp_ceil_div((long long)N * K + N, 256);
        linear_small_dw_kernel<<<grid, 256, 0, stream>>>(dy, (const bf16*)x, dw, db, B, N, K, accumulate);
        DMP_LAUNCH_CHECK();
        DMP_COUNT_LAUNCH(1);
    }
    return 0;
}

DMP_API int dmp_add_act(const void* a, const void* b, void* y, long long n, int relu, cudaStream_t stream) {
    DMP_REQUIRE(n % 8 == 0, "dmp_add_act: n %% 8 != 0");
    const long long nvec = n / 8;
    int grid = dmp_ceil_div(nvec, 256 * 2);
    const int cap = dmp_num_sms() * 8;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    add_act_kernel<<<grid, 256, 0, stream>>>((const bf16*)a, (const bf16*)b, (bf16*)y, nvec, relu);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
DMP_API int dmp_relu_bwd(const void* dy, const void* y, void* dx, long long n, cudaStream_t stream) {
    DMP_REQUIRE(n % 8 == 0, "dmp_relu_bwd: n %% 8 != 0");
    const long long nvec = n / 8;
    int grid = dmp_ceil_div(nvec, 256 * 2);
    const int cap = dmp_num_sms() * 8;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    relu_bwd_kernel<<<grid, 256, 0, stream>>>((const bf16*)dy, (const bf16*)y, (bf16*)dx, nvec);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
