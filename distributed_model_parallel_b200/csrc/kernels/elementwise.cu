// Pooling and transformer row-wise kernels (sm_100a): max-pool forward/backward for the ResNet / VGG configs the reference
// lists but never runs (/root/reference/code/distributed_training/data_parallel.py:58-73), and LayerNorm, GELU and masked
// softmax forward/backward for the BERT-base pipeline config (BASELINE.json config 4; no counterpart in the reference).
// All activations are bf16; statistics and accumulations are fp32.  Every kernel moves 16-byte vectors (8 bf16).
#include "../common.cuh"

namespace {

__device__ __forceinline__ void ld8(const bf16* p, float* f) {
    int4 raw = *(const int4*)p;
    bf16x8_to_float(*(bf16x8*)&raw, f);
}
__device__ __forceinline__ void st8(bf16* p, const float* f) {
    bf16x8 b = float_to_bf16x8(f);
    *(int4*)p = *(int4*)&b;
}

// ---------------------------------------------------------------------------------------------------------- max pool (NHWC)
// y[n,ho,wo,c] = max over the k x k window (stride s, padding p; padded positions never win).  idx keeps the winning tap
// (kh*k + kw) per output element so that the backward needs neither x nor y.
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx, int N, int H, int W, int C,
                   int Ho, int Wo, int k, int s, int p) {
    const int P = C >> 3;
    const long long i = blockIdx.x * 256ll + threadIdx.x;
    if (i >= (long long)N * Ho * Wo * P) return;
    const int cv = (int)(i % P);
    long long t = i / P;
    const int wo = (int)(t % Wo); t /= Wo;
    const int ho = (int)(t % Ho);
    const int n = (int)(t / Ho);
    float best[8];
    int arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
    for (int kh = 0; kh < k; ++kh) {
        const int hi = ho * s + kh - p;
        if (hi < 0 || hi >= H) continue;
        for (int kw = 0; kw < k; ++kw) {
            const int wi = wo * s + kw - p;
            if (wi < 0 || wi >= W) continue;
            float f[8];
            ld8(x + (((long long)n * H + hi) * W + wi) * C + cv * 8, f);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (f[j] > best[j]) { best[j] = f[j]; arg[j] = kh * k + kw; }
        }
    }
    const long long o = (((long long)n * Ho + ho) * Wo + wo) * C + cv * 8;
    st8(y + o, best);
    uint2 packed;
    packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    *(uint2*)(idx + o) = packed;
}

// dx[n,h,w,c] = sum of dy over the windows that contain (h,w) and whose winner is (h,w): a gather, no atomics.
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ idx, bf16* __restrict__ dx, int N, int H, int W,
                   int C, int Ho, int Wo, int k, int s, int p) {
    const int P = C >> 3;
    const long long i = blockIdx.x * 256ll + threadIdx.x;
    if (i >= (long long)N * H * W * P) return;
    const int cv = (int)(i % P);
    long long t = i / P;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    // windows ho with ho*s - p <= h <= ho*s - p + k - 1
    int ho_lo = h + p - k + 1;
    ho_lo = ho_lo <= 0 ? 0 : (ho_lo + s - 1) / s;
    int ho_hi = (h + p) / s;
    if (ho_hi >= Ho) ho_hi = Ho - 1;
    int wo_lo = w + p - k + 1;
    wo_lo = wo_lo <= 0 ? 0 : (wo_lo + s - 1) / s;
    int wo_hi = (w + p) / s;
    if (wo_hi >= Wo) wo_hi = Wo - 1;
    for (int ho = ho_lo; ho <= ho_hi; ++ho)
        for (int wo = wo_lo; wo <= wo_hi; ++wo) {
            const int tap = (h - (ho * s - p)) * k + (w - (wo * s - p));
            const long long o = (((long long)n * Ho + ho) * Wo + wo) * C + cv * 8;
            const uint2 packed = *(const uint2*)(idx + o);
            float g[8];
            ld8(dy + o, g);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int a = ((j < 4 ? packed.x : packed.y) >> (8 * (j & 3))) & 0xff;
                if (a == tap) acc[j] += g[j];
            }
        }
    st8(dx + (((long long)n * H + h) * W + w) * C + cv * 8, acc);
}

// ---------------------------------------------------------------------------------------------------------- LayerNorm
// One warp per row of H elements (H % 8 == 0, H <= 2048): y = (x (+ res) - mean) * rstd * gamma + beta.  The row lives in
// registers between the three passes (8 vectors of 8 per lane).  The pre-normalisation sum z = x + res is what backward
// needs, so it is written out (bf16) when a residual is given.  mean / rstd are saved in fp32.
__global__ void __launch_bounds__(256)
layernorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res, const float* __restrict__ gamma,
                     const float* __restrict__ beta, bf16* __restrict__ y, bf16* __restrict__ z, float* __restrict__ mean,
                     float* __restrict__ rstd, long long rows, int H, float eps) {
    const long long row = (blockIdx.x * 256ll + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    constexpr int KV = 8;
    const int nvec = H >> 3;
    float v[KV][8];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < KV; ++k) {
        const int i = lane + 32 * k;
        if (i < nvec) {
            ld8(x + row * H + i * 8, v[k]);
            if (res != nullptr) {
                float g[8];
                ld8(res + row * H + i * 8, g);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[k][j] = __bfloat162float(__float2bfloat16_rn(v[k][j] + g[j]));
                if (z != nullptr) st8(z + row * H + i * 8, v[k]);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) s += v[k][j];
        }
    }
    const float mu = warp_sum(s) / (float)H;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < KV; ++k)
        if (lane + 32 * k < nvec) {
#pragma unroll
            for (int j = 0; j < 8; ++j) q += (v[k][j] - mu) * (v[k][j] - mu);
        }
    const float rs = rsqrtf(warp_sum(q) / (float)H + eps);
    if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
#pragma unroll
    for (int k = 0; k < KV; ++k) {
        const int i = lane + 32 * k;
        if (i < nvec) {
            float g8[8], b8[8], o[8];
            *(float4*)g8 = *(const float4*)(gamma + i * 8); *(float4*)(g8 + 4) = *(const float4*)(gamma + i * 8 + 4);
            *(float4*)b8 = *(const float4*)(beta + i * 8); *(float4*)(b8 + 4) = *(const float4*)(beta + i * 8 + 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[k][j] - mu) * rs * g8[j] + b8[j];
            st8(y + row * H + i * 8, o);
        }
    }
}

// dz = rstd * (g - mean_H(g) - xhat * mean_H(g * xhat)) with g = dy * gamma; per-block partial dgamma / dbeta are accumulated
// in shared memory over the block's rows and flushed with one fp32 atomic per column and block.
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z, const float* __restrict__ gamma,
                     const float* __restrict__ mean, const float* __restrict__ rstd, bf16* __restrict__ dz,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, long long rows, int H, int rows_per_warp) {
    extern __shared__ float sm[];  // [2][H]
    for (int i = threadIdx.x; i < 2 * H; i += 256) sm[i] = 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int nvec = H >> 3;
    const long long row0 = ((long long)blockIdx.x * 8 + wib) * rows_per_warp;
    // a lane owns columns (lane + 32k) * 8 .. +7: the parameter-gradient partials of up to 4 such vectors stay in registers
    // across all of the warp's rows; wider rows (H > 1024) spill over to shared-memory atomics per element
    constexpr int KV = 4;
    float ag[KV][8], ab[KV][8];
#pragma unroll
    for (int k = 0; k < KV; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) ag[k][j] = ab[k][j] = 0.f;
    for (int r = 0; r < rows_per_warp; ++r) {
        const long long row = row0 + r;
        if (row >= rows) break;
        const float mu = mean[row], rs = rstd[row];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < KV + 2; ++k) {
            const int i = lane + 32 * k;
            if (i >= nvec) break;
            float g[8], zz[8], gm[8];
            ld8(dy + row * H + i * 8, g);
            ld8(z + row * H + i * 8, zz);
            *(float4*)gm = *(const float4*)(gamma + i * 8); *(float4*)(gm + 4) = *(const float4*)(gamma + i * 8 + 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (zz[j] - mu) * rs, gg = g[j] * gm[j];
                s1 += gg;
                s2 += gg * xh;
                if (k < KV) { ag[k][j] += g[j] * xh; ab[k][j] += g[j]; }
                else { atomicAdd(&sm[i * 8 + j], g[j] * xh); atomicAdd(&sm[H + i * 8 + j], g[j]); }
            }
        }
        for (int i = lane + 32 * (KV + 2); i < nvec; i += 32) {  // very wide rows
            float g[8], zz[8], gm[8];
            ld8(dy + row * H + i * 8, g);
            ld8(z + row * H + i * 8, zz);
            *(float4*)gm = *(const float4*)(gamma + i * 8); *(float4*)(gm + 4) = *(const float4*)(gamma + i * 8 + 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (zz[j] - mu) * rs, gg = g[j] * gm[j];
                s1 += gg;
                s2 += gg * xh;
                atomicAdd(&sm[i * 8 + j], g[j] * xh);
                atomicAdd(&sm[H + i * 8 + j], g[j]);
            }
        }
        s1 = warp_sum(s1) / (float)H;
        s2 = warp_sum(s2) / (float)H;
        for (int i = lane; i < nvec; i += 32) {
            float g[8], zz[8], gm[8];
            ld8(dy + row * H + i * 8, g);
            ld8(z + row * H + i * 8, zz);
            *(float4*)gm = *(const float4*)(gamma + i * 8); *(float4*)(gm + 4) = *(const float4*)(gamma + i * 8 + 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (zz[j] - mu) * rs;
                g[j] = rs * (g[j] * gm[j] - s1 - xh * s2);
            }
            st8(dz + row * H + i * 8, g);
        }
    }
#pragma unroll
    for (int k = 0; k < KV; ++k) {
        const int i = lane + 32 * k;
        if (i < nvec) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(&sm[i * 8 + j], ag[k][j]);
                atomicAdd(&sm[H + i * 8 + j], ab[k][j]);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < H; i += 256) {
        if (sm[i] != 0.f) atomicAdd(&dgamma[i], sm[i]);
        if (sm[H + i] != 0.f) atomicAdd(&dbeta[i], sm[H + i]);
    }
}

// ---------------------------------------------------------------------------------------------------------- GELU (erf form)
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad_f(float x) {
    const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752f));
    return cdf + x * 0.3989422804014327f * __expf(-0.5f * x * x);
}
__global__ void __launch_bounds__(256) gelu_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long nvec) {
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += (long long)gridDim.x * 256) {
        float f[8];
        ld8(x + i * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = gelu_f(f[j]);
        st8(y + i * 8, f);
    }
}
__global__ void __launch_bounds__(256)
gelu_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, bf16* __restrict__ dx, long long nvec) {
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += (long long)gridDim.x * 256) {
        float g[8], f[8];
        ld8(dy + i * 8, g);
        ld8(x + i * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] *= gelu_grad_f(f[j]);
        st8(dx + i * 8, g);
    }
}

// ---------------------------------------------------------------------------------------------------------- row softmax
// p = softmax(scale * s) over rows of L (L % 8 == 0, L <= 4096), one warp per row; in place allowed.  Backward:
// ds = scale * p * (dp - sum(dp * p)).
__global__ void __launch_bounds__(256)
softmax_fwd_kernel(const bf16* __restrict__ s, bf16* __restrict__ p, long long rows, int L, float scale) {
    const long long row = (blockIdx.x * 256ll + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const int nvec = L >> 3;
    float m = -INFINITY;
    for (int i = lane; i < nvec; i += 32) {
        float f[8];
        ld8(s + row * L + i * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) m = fmaxf(m, f[j] * scale);
    }
    m = warp_max(m);
    float sum = 0.f;
    for (int i = lane; i < nvec; i += 32) {
        float f[8];
        ld8(s + row * L + i * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += __expf(f[j] * scale - m);
    }
    const float inv = 1.f / warp_sum(sum);
    for (int i = lane; i < nvec; i += 32) {
        float f[8];
        ld8(s + row * L + i * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = __expf(f[j] * scale - m) * inv;
        st8(p + row * L + i * 8, f);
    }
}
__global__ void __launch_bounds__(256)
softmax_bwd_kernel(const bf16* __restrict__ dp, const bf16* __restrict__ p, bf16* __restrict__ ds, long long rows, int L,
                   float scale) {
    const long long row = (blockIdx.x * 256ll + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const int nvec = L >> 3;
    float dot = 0.f;
    for (int i = lane; i < nvec; i += 32) {
        float a[8], b[8];
        ld8(dp + row * L + i * 8, a);
        ld8(p + row * L + i * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) dot += a[j] * b[j];
    }
    dot = warp_sum(dot);
    for (int i = lane; i < nvec; i += 32) {
        float a[8], b[8];
        ld8(dp + row * L + i * 8, a);
        ld8(p + row * L + i * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = scale * b[j] * (a[j] - dot);
        st8(ds + row * L + i * 8, a);
    }
}

}  // namespace

// x, y: bf16 NHWC; idx: uint8 [N,Ho,Wo,C] (winning tap).  C % 8 == 0, k <= 15.
DMP_API int dmp_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                            cudaStream_t stream) {
    DMP_REQUIRE(C % 8 == 0 && k >= 1 && k <= 15 && stride >= 1 && pad >= 0 && pad < k, "dmp_maxpool_fwd: bad arguments (C=%d k=%d)", C, k);
    const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    const long long total = (long long)N * Ho * Wo * (C / 8);
    maxpool_fwd_kernel<<<dmp_ceil_div(total, 256), 256, 0, stream>>>((const bf16*)x, (bf16*)y, (uint8_t*)idx, N, H, W, C, Ho, Wo, k,
                                                                      stride, pad);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
DMP_API int dmp_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride, int pad,
                            cudaStream_t stream) {
    DMP_REQUIRE(C % 8 == 0 && k >= 1 && k <= 15 && stride >= 1, "dmp_maxpool_bwd: bad arguments");
    const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    const long long total = (long long)N * H * W * (C / 8);
    maxpool_bwd_kernel<<<dmp_ceil_div(total, 256), 256, 0, stream>>>((const bf16*)dy, (const uint8_t*)idx, (bf16*)dx, N, H, W, C, Ho,
                                                                      Wo, k, stride, pad);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// y = LayerNorm(x (+ res)); z (nullable unless res != null and backward is needed) receives x + res.
DMP_API int dmp_layernorm_fwd(const void* x, const void* res, const float* gamma, const float* beta, void* y, void* z, float* mean,
                              float* rstd, long long rows, int H, float eps, cudaStream_t stream) {
    DMP_REQUIRE(rows > 0 && H > 0 && H % 8 == 0 && H <= 2048, "dmp_layernorm_fwd: H=%d must be a multiple of 8, at most 2048", H);
    layernorm_fwd_kernel<<<dmp_ceil_div(rows * 32, 256), 256, 0, stream>>>((const bf16*)x, (const bf16*)res, gamma, beta, (bf16*)y,
                                                                            (bf16*)z, mean, rstd, rows, H, eps);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
// dz: gradient w.r.t. the normalised input z (= dx = dres); dgamma / dbeta: fp32 [H], accumulated (caller zeroes or chains).
DMP_API int dmp_layernorm_bwd(const void* dy, const void* z, const float* gamma, const float* mean, const float* rstd, void* dz,
                              float* dgamma, float* dbeta, long long rows, int H, cudaStream_t stream) {
    DMP_REQUIRE(rows > 0 && H > 0 && H % 8 == 0 && H <= 6144, "dmp_layernorm_bwd: H=%d unsupported", H);
    long long warps = rows;
    int rpw = 1;
    const long long target = (long long)dmp_num_sms() * 8 * 8;  // ~8 blocks of 8 warps per SM
    if (warps > target) { rpw = (int)((rows + target - 1) / target); warps = (rows + rpw - 1) / rpw; }
    const int grid = dmp_ceil_div(warps, 8);
    layernorm_bwd_kernel<<<grid, 256, 2 * H * sizeof(float), stream>>>((const bf16*)dy, (const bf16*)z, gamma, mean, rstd, (bf16*)dz,
                                                                       dgamma, dbeta, rows, H, rpw);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_gelu_fwd(const void* x, void* y, long long n, cudaStream_t stream) {
    DMP_REQUIRE(n > 0 && n % 8 == 0, "dmp_gelu_fwd: n must be a multiple of 8");
    long long blocks = (n / 8 + 255) / 256;
    if (blocks > (long long)dmp_num_sms() * 16) blocks = (long long)dmp_num_sms() * 16;
    gelu_fwd_kernel<<<(int)blocks, 256, 0, stream>>>((const bf16*)x, (bf16*)y, n / 8);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
DMP_API int dmp_gelu_bwd(const void* dy, const void* x, void* dx, long long n, cudaStream_t stream) {
    DMP_REQUIRE(n > 0 && n % 8 == 0, "dmp_gelu_bwd: n must be a multiple of 8");
    long long blocks = (n / 8 + 255) / 256;
    if (blocks > (long long)dmp_num_sms() * 16) blocks = (long long)dmp_num_sms() * 16;
    gelu_bwd_kernel<<<(int)blocks, 256, 0, stream>>>((const bf16*)dy, (const bf16*)x, (bf16*)dx, n / 8);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_softmax_fwd(const void* s, void* p, long long rows, int L, float scale, cudaStream_t stream) {
    DMP_REQUIRE(rows > 0 && L > 0 && L % 8 == 0, "dmp_softmax_fwd: L=%d must be a positive multiple of 8", L);
    softmax_fwd_kernel<<<dmp_ceil_div(rows * 32, 256), 256, 0, stream>>>((const bf16*)s, (bf16*)p, rows, L, scale);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
DMP_API int dmp_softmax_bwd(const void* dp, const void* p, void* ds, long long rows, int L, float scale, cudaStream_t stream) {
    DMP_REQUIRE(rows > 0 && L > 0 && L % 8 == 0, "dmp_softmax_bwd: L=%d must be a positive multiple of 8", L);
    softmax_bwd_kernel<<<dmp_ceil_div(rows * 32, 256), 256, 0, stream>>>((const bf16*)dp, (const bf16*)p, (bf16*)ds, rows, L, scale);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
