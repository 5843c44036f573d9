// Depthwise 3x3 convolution (groups == channels), channels-last, pad 1, stride 1 or 2: forward, data
// gradient and weight gradient.  This is the op behind the 17 `groups=planes` convolutions of the
// reference MobileNetV2 (/root/reference/code/distributed_training/model/mobilenetv2.py:19,33).  It has
// no reduction over channels, so it is a CUDA-core stencil (not a tensor-core GEMM): each thread owns
// 8 contiguous channels (one 16-byte vector) of one output pixel, so every global access is a 128-bit
// coalesced load/store and neighbouring pixels are served from L1/L2.
#include "../common.cuh"
#include <stdlib.h>

namespace {

constexpr int DW_THREADS = 256;

// 8 channels x 9 taps = 72 contiguous weights of the [C,1,3,3] tensor, fetched with 16-byte loads
// (ch0 is a multiple of 8, so the 144-byte / 288-byte block is 16-byte aligned).
template <typename WT>
__device__ __forceinline__ void load_w72(const WT* __restrict__ w, int ch0, float* out /*[8][9]*/);
template <>
__device__ __forceinline__ void load_w72<bf16>(const bf16* __restrict__ w, int ch0, float* out) {
    const int4* p = (const int4*)(w + ch0 * 9);
#pragma unroll
    for (int i = 0; i < 9; ++i) {
        int4 raw = __ldg(p + i);
        bf16x8_to_float(*(bf16x8*)&raw, out + i * 8);
    }
}
template <>
__device__ __forceinline__ void load_w72<float>(const float* __restrict__ w, int ch0, float* out) {
    const float4* p = (const float4*)(w + ch0 * 9);
#pragma unroll
    for (int i = 0; i < 18; ++i) {
        float4 v = __ldg(p + i);
        out[i * 4] = v.x; out[i * 4 + 1] = v.y; out[i * 4 + 2] = v.z; out[i * 4 + 3] = v.w;
    }
}

__device__ __forceinline__ void load8(const bf16* p, float* f) {
    int4 raw = *(const int4*)p;
    bf16x8_to_float(*(bf16x8*)&raw, f);
}
__device__ __forceinline__ void store8(bf16* p, const float* f) {
    bf16x8 b = float_to_bf16x8(f);
    *(int4*)p = *(int4*)&b;
}

// Fusions with the BatchNorm layers around a depthwise convolution (the expand -> bn1 -> relu -> depthwise -> bn2 chain of
// /root/reference/code/distributed_training/model/mobilenetv2.py:32-33):
//   BNIN == 1  the input is the RAW output y1 of the previous convolution; relu(bn1(y1)) is applied while loading, with the
//              coefficients derived from the batch sums the producer's epilogue accumulated (`sums`, double[2C] + int ticket).
//              Block 0 publishes mean / invstd / running statistics, the last block to have read the sums clears them.
//              The normalised activation is never written to memory.
//   BNIN == 2  same transform from saved mean / invstd (backward passes re-create the activation on the fly).
//   BNBWD      (data gradient) the result is d(relu(bn1(y1))); the ReLU mask and xhat are recomputed from y1 at the output
//              pixel, the masked gradient g is stored and sum(g), sum(g * xhat) are accumulated into `bwd_sums` (double[2C]).
struct DwBn {
    const double* sums;          // BNIN 1
    double inv_count;            // BNIN 1: 1 / (N*H*W of the input)
    double count;
    const float* gamma;
    const float* beta;
    float eps, momentum;
    float* running_mean;
    float* running_var;
    long long* num_batches_tracked;
    float* mean_out;             // BNIN 1: published by block 0
    float* invstd_out;
    const float* mean;           // BNIN 2 / BNBWD
    const float* invstd;
    const bf16* y_raw;           // BNBWD: the BatchNorm input at the OUTPUT resolution of this kernel
    double* bwd_sums;            // BNBWD
};

// scale / shift (and mean / invstd) of channels ch0..ch0+7
template <int BNIN>
__device__ __forceinline__ void dw_bn_coeffs(const DwBn& a, int ch0, int C, float* sc, float* sh, float* mu, float* is) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = ch0 + j;
        float m, r;
        if (BNIN == 1) {
            const double md = a.sums[c] * a.inv_count;
            double var = a.sums[C + c] * a.inv_count - md * md;
            if (var < 0.0) var = 0.0;
            m = (float)md;
            r = 1.0f / sqrtf((float)var + a.eps);
        } else {
            m = a.mean[c];
            r = a.invstd[c];
        }
        const float g = a.gamma ? a.gamma[c] : 1.f, b = a.beta ? a.beta[c] : 0.f;
        sc[j] = g * r;
        sh[j] = b - m * g * r;
        mu[j] = m;
        is[j] = r;
    }
}

// BNIN == 1 bookkeeping, reached by every thread of the block: block 0 publishes the statistics, the last block clears the sums.
__device__ __forceinline__ void dw_bn_publish_and_release(const DwBn& a, int C) {
    __shared__ int s_last;
    if (blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += DW_THREADS) {
            const double m = a.sums[c] / a.count;
            double var = a.sums[C + c] / a.count - m * m;
            if (var < 0.0) var = 0.0;
            a.mean_out[c] = (float)m;
            a.invstd_out[c] = (float)(1.0 / sqrt(var + (double)a.eps));
            if (a.running_mean != nullptr) {
                const double unbiased = a.count > 1.0 ? var * a.count / (a.count - 1.0) : var;
                a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * (float)m;
                a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * (float)unbiased;
            }
        }
        if (threadIdx.x == 0 && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        int* ticket = (int*)(a.sums + 2 * (long long)C);
        s_last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        double* sm = const_cast<double*>(a.sums);
        for (int i = threadIdx.x; i < 2 * C; i += DW_THREADS) sm[i] = 0.0;
        if (threadIdx.x == 0) *(int*)(sm + 2 * (long long)C) = 0;
    }
}

// Block-level combine of per-thread {sum[8], sumsq[8]} partials of channels ch0..ch0+7 into global double accumulators.
// Channel vectors repeat over the block's threads with period P = C/8, so thread g < min(P, 256) folds threads g, g+P, ...
// out of shared memory and issues the 16 atomics of its channel vector.  Must be reached by every thread of the block.
__device__ __forceinline__ void dw_block_channel_reduce(const float* vals, int ch0, int C, double* __restrict__ accum) {
    __shared__ float s_part[DW_THREADS * 17];  // padded rows: 17 floats per thread
    __shared__ int s_ch0[DW_THREADS];
#pragma unroll
    for (int i = 0; i < 16; ++i) s_part[threadIdx.x * 17 + i] = vals[i];
    s_ch0[threadIdx.x] = ch0;
    __syncthreads();
    const int P = C >> 3;
    const int ngroups = P < DW_THREADS ? P : DW_THREADS;
    for (int item = threadIdx.x; item < ngroups * 16; item += DW_THREADS) {
        const int g = item >> 4, i = item & 15;
        float sacc = 0.f;
        for (int t = g; t < DW_THREADS; t += P) sacc += s_part[t * 17 + i];
        atomicAdd(&accum[(long long)(i >> 3) * C + s_ch0[g] + (i & 7)], (double)sacc);
    }
}

// y[n,ho,wo,c] = sum_{kh,kw} x[n, ho*s+kh-1, wo*s+kw-1, c] * w[c,kh,kw]
template <typename WT>
__global__ void __launch_bounds__(DW_THREADS)
dw_fwd_kernel(const bf16* __restrict__ x, const WT* __restrict__ w, bf16* __restrict__ y, int N, int H, int W, int C,
              int Ho, int Wo, int stride) {
    const int P = C >> 3;
    const long long idx = blockIdx.x * (long long)DW_THREADS + threadIdx.x;
    const long long total = (long long)N * Ho * Wo * P;
    if (idx >= total) return;
    const int cv = (int)(idx % P);
    long long pix = idx / P;
    const int wo = (int)(pix % Wo); pix /= Wo;
    const int ho = (int)(pix % Ho);
    const int n = (int)(pix / Ho);
    const int ch0 = cv * 8;
    float wt[72];
    load_w72<WT>(w, ch0, wt);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
        const int hi = ho * stride + kh - 1;
        if (hi < 0 || hi >= H) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
            const int wi = wo * stride + kw - 1;
            if (wi < 0 || wi >= W) continue;
            float f[8];
            load8(x + (((long long)n * H + hi) * W + wi) * C + ch0, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j] * wt[j * 9 + kh * 3 + kw];
        }
    }
    store8(y + (((long long)n * Ho + ho) * Wo + wo) * C + ch0, acc);
}

// dx[n,h,w,c] = sum_{kh,kw} dy[n,(h+1-kh)/s,(w+1-kw)/s,c] * w[c,kh,kw]   (only exact divisions)
template <typename WT>
__global__ void __launch_bounds__(DW_THREADS)
dw_dgrad_kernel(const bf16* __restrict__ dy, const WT* __restrict__ w, bf16* __restrict__ dx, int N, int H, int W, int C,
                int Ho, int Wo, int stride) {
    const int P = C >> 3;
    const long long idx = blockIdx.x * (long long)DW_THREADS + threadIdx.x;
    const long long total = (long long)N * H * W * P;
    if (idx >= total) return;
    const int cv = (int)(idx % P);
    long long pix = idx / P;
    const int wi = (int)(pix % W); pix /= W;
    const int hi = (int)(pix % H);
    const int n = (int)(pix / H);
    const int ch0 = cv * 8;
    float wt[72];
    load_w72<WT>(w, ch0, wt);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
        const int th = hi + 1 - kh;
        if (th < 0 || (th % stride) != 0) continue;
        const int ho = th / stride;
        if (ho >= Ho) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
            const int tw = wi + 1 - kw;
            if (tw < 0 || (tw % stride) != 0) continue;
            const int wo = tw / stride;
            if (wo >= Wo) continue;
            float f[8];
            load8(dy + (((long long)n * Ho + ho) * Wo + wo) * C + ch0, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j] * wt[j * 9 + kh * 3 + kw];
        }
    }
    store8(dx + (((long long)n * H + hi) * W + wi) * C + ch0, acc);
}

// Stride-2 data gradient by output parity.  With pad 1 and even H, W the 2x2 block of dx at rows 2a..2a+1, columns 2b..2b+1
// reads exactly dy[a..a+1][b..b+1] and uses each of the nine taps once:
//   dx[2a  ][2b  ] = dy[a][b] w11                       dx[2a  ][2b+1] = dy[a][b+1] w10 + dy[a][b] w12
//   dx[2a+1][2b  ] = dy[a+1][b] w01 + dy[a][b] w21      dx[2a+1][2b+1] = dy[a+1][b+1] w00 + dy[a+1][b] w02 + dy[a][b+1] w20 + dy[a][b] w22
// One thread writes two such blocks (2 x 4 pixels of its 8 channels) from six 16-byte loads -- no divisibility tests, 0.75
// loads per output instead of 2.25, and the 72 weights are fetched once per eight outputs instead of once per output.
template <typename WT, bool BNBWD = false>
__global__ void __launch_bounds__(DW_THREADS, 2)
dw_dgrad_s2_kernel(const bf16* __restrict__ dy, const WT* __restrict__ w, bf16* __restrict__ dx, int N, int H, int W, int C,
                   int Ho, int Wo, const DwBn aux) {
    constexpr int TQ = 2;
    const int P = C >> 3;
    const int tiles_w = (Wo + TQ - 1) / TQ;
    const long long idx = blockIdx.x * (long long)DW_THREADS + threadIdx.x;
    const long long total = (long long)N * Ho * tiles_w * P;
    const bool active = idx < total;
    if (!BNBWD && !active) return;
    const int ch0 = (int)(idx % P) * 8;
    float st[BNBWD ? 16 : 1];
    if (BNBWD) {
#pragma unroll
        for (int i = 0; i < 16; ++i) st[i] = 0.f;
    }
    if (active) {
        long long t = idx / P;
        const int tw = (int)(t % tiles_w); t /= tiles_w;
        const int a = (int)(t % Ho);
        const int n = (int)(t / Ho);
        const int b0 = tw * TQ;
        float wt[72];
        load_w72<WT>(w, ch0, wt);
        float sc[8], sh[8], mu[8], is[8];
        if (BNBWD) dw_bn_coeffs<2>(aux, ch0, C, sc, sh, mu, is);
        float d0[TQ + 1][8], d1[TQ + 1][8];
#pragma unroll
        for (int c = 0; c <= TQ; ++c) {
            const int col = b0 + c;
            const bool okc = col < Wo;
            if (okc) load8(dy + (((long long)n * Ho + a) * Wo + col) * C + ch0, d0[c]);
            else {
#pragma unroll
                for (int j = 0; j < 8; ++j) d0[c][j] = 0.f;
            }
            if (okc && a + 1 < Ho) load8(dy + (((long long)n * Ho + a + 1) * Wo + col) * C + ch0, d1[c]);
            else {
#pragma unroll
                for (int j = 0; j < 8; ++j) d1[c][j] = 0.f;
            }
        }
#pragma unroll
        for (int q = 0; q < TQ; ++q) {
            const int b = b0 + q;
            if (b >= Wo) continue;
            float o[4][8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float* k = wt + j * 9;
                o[0][j] = d0[q][j] * k[4];
                o[1][j] = d0[q + 1][j] * k[3] + d0[q][j] * k[5];
                o[2][j] = d1[q][j] * k[1] + d0[q][j] * k[7];
                o[3][j] = d1[q + 1][j] * k[0] + d1[q][j] * k[2] + d0[q + 1][j] * k[6] + d0[q][j] * k[8];
            }
            const long long off = (((long long)n * H + 2 * a) * W + 2 * b) * C + ch0;
            const long long offs[4] = {off, off + C, off + (long long)W * C, off + (long long)W * C + C};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (BNBWD) {  // the results are d(relu(bn(y1))): mask from y1, store g, accumulate the BatchNorm-backward sums
                    float yv[8];
                    load8(aux.y_raw + offs[e], yv);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float g = (yv[j] * sc[j] + sh[j] > 0.f) ? o[e][j] : 0.f;
                        g = __bfloat162float(__float2bfloat16_rn(g));
                        o[e][j] = g;
                        st[j] += g;
                        st[8 + j] += g * ((yv[j] - mu[j]) * is[j]);
                    }
                }
                store8(dx + offs[e], o[e]);
            }
        }
    }
    if (BNBWD) dw_block_channel_reduce(st, ch0, C, aux.bwd_sums);
}

// Strip version of the forward stencil: one thread produces TW horizontally adjacent output pixels of its 8 channels,
// so every input vector of a row is loaded once and reused by up to 3 outputs (4.5 instead of 9 loads per output at
// stride 1).  FLIP=true mirrors the taps: with stride 1 the data gradient is the same stencil applied to dy.
// STATS=true additionally accumulates, per channel, the sum and sum of squares of the bf16-rounded outputs into
// stats[0..C) / stats[C..2C) (double): the BatchNorm batch statistics come out of the convolution itself.
template <typename WT, int STRIDE, int TW, bool FLIP, bool STATS, int BNIN = 0, bool BNBWD = false>
__global__ void __launch_bounds__(DW_THREADS, 2)  // 2 blocks/SM: the few spilled registers cost less than half occupancy
dw_strip_kernel(const bf16* __restrict__ x, const WT* __restrict__ w, bf16* __restrict__ y, int N, int H, int W, int C,
                int Ho, int Wo, double* __restrict__ stats, const DwBn aux) {
    // One work item (a strip of TW output pixels of one channel vector) per thread, many more blocks than SMs.  A persistent
    // variant (each thread keeps its channel vector and weights across items) measured 83 us against 53 us on 128x32x32x144:
    // block turnover hides the load latency of this register-bound kernel better than a loop inside 16 resident warps does.
    constexpr int NC = (TW - 1) * STRIDE + 3;  // input columns feeding the strip
    constexpr bool ALL_THREADS = STATS || BNBWD || BNIN == 1;  // block-wide steps follow: no early exit
    const int P = C >> 3;
    const int tiles_w = (Wo + TW - 1) / TW;
    const long long idx = blockIdx.x * (long long)DW_THREADS + threadIdx.x;
    const long long total = (long long)N * Ho * tiles_w * P;
    const bool active = idx < total;
    if (!ALL_THREADS && !active) return;
    const int cv = (int)(idx % P);
    long long t = idx / P;
    const int wt_ = (int)(t % tiles_w); t /= tiles_w;
    const int ho = (int)(t % Ho);
    const int n = (int)(t / Ho);
    const int ch0 = cv * 8;
    const int wo0 = wt_ * TW;
    float st[(STATS || BNBWD) ? 16 : 1];
    if (STATS || BNBWD) {
#pragma unroll
        for (int i = 0; i < 16; ++i) st[i] = 0.f;
    }
    float sc[8], sh[8], mu[8], is[8];
    if (BNIN != 0 && active) dw_bn_coeffs<BNIN>(aux, ch0, C, sc, sh, mu, is);
    if (BNBWD && active) dw_bn_coeffs<2>(aux, ch0, C, sc, sh, mu, is);
    if (active) {
        float wt[72];
        load_w72<WT>(w, ch0, wt);
        float acc[TW][8];
#pragma unroll
        for (int i = 0; i < TW; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
        const int wi0 = wo0 * STRIDE - 1;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * STRIDE + kh - 1;
            if (hi < 0 || hi >= H) continue;
            const bf16* row = x + (((long long)n * H + hi) * W) * C + ch0;
            float col[NC][8];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int wi = wi0 + c;
                if (wi >= 0 && wi < W) {
                    load8(row + (long long)wi * C, col[c]);
                    if (BNIN != 0) {  // the convolution's real input is relu(bn(y1)); the zero padding stays zero
#pragma unroll
                        for (int j = 0; j < 8; ++j) col[c][j] = fmaxf(col[c][j] * sc[j] + sh[j], 0.f);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) col[c][j] = 0.f;
                }
            }
#pragma unroll
            for (int i = 0; i < TW; ++i)
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) {
                    const int tap = FLIP ? ((2 - kh) * 3 + (2 - kw)) : (kh * 3 + kw);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] += col[i * STRIDE + kw][j] * wt[j * 9 + tap];
                }
        }
#pragma unroll
        for (int i = 0; i < TW; ++i)
            if (wo0 + i < Wo) {
                if (BNBWD) {  // acc = d(relu(bn(y1))) at this pixel: mask, store g, accumulate sum g and sum g * xhat
                    float yv[8];
                    load8(aux.y_raw + (((long long)n * Ho + ho) * Wo + wo0 + i) * C + ch0, yv);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float g = (yv[j] * sc[j] + sh[j] > 0.f) ? acc[i][j] : 0.f;
                        g = __bfloat162float(__float2bfloat16_rn(g));
                        acc[i][j] = g;
                        st[j] += g;
                        st[8 + j] += g * ((yv[j] - mu[j]) * is[j]);
                    }
                }
                store8(y + (((long long)n * Ho + ho) * Wo + wo0 + i) * C + ch0, acc[i]);
                if (STATS) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float q = __bfloat162float(__float2bfloat16_rn(acc[i][j]));
                        st[j] += q;
                        st[8 + j] += q * q;
                    }
                }
            }
    }
    if (STATS) dw_block_channel_reduce(st, ch0, C, stats);  // reached by every thread of the block (idle tail threads add zeros)
    if (BNBWD) dw_block_channel_reduce(st, ch0, C, aux.bwd_sums);
    if (BNIN == 1) dw_bn_publish_and_release(aux, C);  // this block read its coefficients long ago: the ticket goes last
}

// Branch-free strip stencil (forward, and the stride-1 data gradient with FLIP): the default path of the training step.
// What the kernel above does under branches -- skip rows and columns that fall into the padding -- this one does with clamped
// addresses and a select: all 3 x NC input vectors of a strip are fetched by independent loads issued back to back (the
// compiler cannot hoist a load out of the branch that guards it, so the branchy version paid up to 18 L2 round trips one after
// the other), the taps of the block's own channels are staged once per block in shared memory as fp32 (72 registers per thread
// in the version above, which spilled), and all index arithmetic is 32-bit (the 64-bit divisions were function calls).
// Shared memory: the taps of the block's channels, overlaid by the statistics scratch once they are consumed.
constexpr int DW_STAT_SCRATCH_BYTES = DW_THREADS * 17 * 4;

template <typename WT, int STRIDE, int TW, bool FLIP, bool STATS>
__global__ void __launch_bounds__(DW_THREADS, 2)
dw_strip_bf_kernel(const bf16* __restrict__ x, const WT* __restrict__ w, bf16* __restrict__ y, int N, int H, int W, int C,
                   int Ho, int Wo, double* __restrict__ stats, int cvb) {
    constexpr int NC = (TW - 1) * STRIDE + 3;
    extern __shared__ __align__(16) float dw_smem[];
    float* s_w = dw_smem;  // [9][CB]: the taps of this block's CB = 8 * cvb channels
    const int CB = cvb * 8;
    {
        const int cbase = (int)(blockIdx.x % (((unsigned)C >> 3) / (unsigned)cvb)) * CB;
        for (int i = threadIdx.x; i < 9 * CB; i += DW_THREADS) {
            const int c = i / 9, tap = i - c * 9;
            s_w[(FLIP ? 8 - tap : tap) * CB + c] = to_f32<WT>(w[(size_t)cbase * 9 + i]);
        }
    }
    // A block owns `cvb` adjacent channel vectors (cvb divides C/8, at most 8: 128 contiguous bytes per pixel) and
    // DW_THREADS / cvb strips.  With the channel vector as the fastest index over ALL of C (the general kernel's mapping) every
    // block touches every channel and ends with 2*C atomics on the same 2*C addresses as every other block; here a block adds
    // 16 * cvb values and a channel's accumulator is shared by 1 / (C/8/cvb) of the blocks.
    const unsigned P = (unsigned)C >> 3;
    const unsigned tiles_w = (unsigned)(Wo + TW - 1) / TW;
    const unsigned nstrips = (unsigned)N * Ho * tiles_w;  // < 2^31, checked by the host
    const unsigned ucvb = (unsigned)cvb, spb = (unsigned)DW_THREADS / ucvb, cvgroups = P / ucvb;
    const unsigned cvl = threadIdx.x % ucvb, sl = threadIdx.x / ucvb;
    const unsigned strip = (blockIdx.x / cvgroups) * spb + sl;
    const bool active = sl < spb && strip < nstrips;
    const unsigned cv = (blockIdx.x % cvgroups) * ucvb + cvl;
    unsigned t = active ? strip : 0u;
    const int wt_ = (int)(t % tiles_w); t /= tiles_w;
    const int ho = (int)(t % (unsigned)Ho);
    const int n = (int)(t / (unsigned)Ho);
    const int ch0 = (int)cv * 8;
    const int wo0 = wt_ * TW;
    const int wi0 = wo0 * STRIDE - 1;
    int4 raw[3][NC];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
        const int hi = ho * STRIDE + kh - 1;
        const bool okh = (unsigned)hi < (unsigned)H;
        const bf16* row = x + ((size_t)((size_t)n * H + min(max(hi, 0), H - 1)) * W) * C + ch0;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int wi = wi0 + c;
            const bool ok = okh && (unsigned)wi < (unsigned)W;
            int4 v = *(const int4*)(row + (size_t)min(max(wi, 0), W - 1) * C);
            if (!ok) v = make_int4(0, 0, 0, 0);
            raw[kh][c] = v;
        }
    }
    __syncthreads();  // weights staged
    float acc[TW][8];
#pragma unroll
    for (int i = 0; i < TW; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
        float wk[3][8];
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
            const float* wp = s_w + (kh * 3 + kw) * CB + (int)cvl * 8;
            const float4 a = *(const float4*)wp, b = *(const float4*)(wp + 4);
            wk[kw][0] = a.x; wk[kw][1] = a.y; wk[kw][2] = a.z; wk[kw][3] = a.w;
            wk[kw][4] = b.x; wk[kw][5] = b.y; wk[kw][6] = b.z; wk[kw][7] = b.w;
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            float v[8];
            bf16x8_to_float(*(const bf16x8*)&raw[kh][c], v);
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                // input column c feeds output i with i * STRIDE + kw == c
                if ((c - kw) >= 0 && (c - kw) % STRIDE == 0 && (c - kw) / STRIDE < TW) {
                    const int i = (c - kw) / STRIDE;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] += v[j] * wk[kw][j];
                }
            }
        }
    }
    float st[STATS ? 16 : 1];
    if (STATS) {
#pragma unroll
        for (int i = 0; i < 16; ++i) st[i] = 0.f;
    }
    if (active) {
#pragma unroll
        for (int i = 0; i < TW; ++i)
            if (wo0 + i < Wo) {
                store8(y + (((size_t)n * Ho + ho) * Wo + wo0 + i) * C + ch0, acc[i]);
                if (STATS) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float q = __bfloat162float(__float2bfloat16_rn(acc[i][j]));
                        st[j] += q;
                        st[8 + j] += q * q;
                    }
                }
            }
    }
    if (STATS) {
        __syncthreads();  // every thread is done with the staged weights: the scratch below overlays them
        float* s_part = dw_smem;                              // [DW_THREADS][17]
#pragma unroll
        for (int i = 0; i < 16; ++i) s_part[threadIdx.x * 17 + i] = st[i];  // idle threads hold zeros
        __syncthreads();
        const int cvg0 = (int)((blockIdx.x % cvgroups) * ucvb) * 8;   // first channel of this block
        for (int item = threadIdx.x; item < cvb * 16; item += DW_THREADS) {
            const int g = item >> 4, i = item & 15;
            float sacc = 0.f;
            for (unsigned q = 0; q < spb; ++q) sacc += s_part[(q * ucvb + g) * 17 + i];
            atomicAdd(&stats[(long long)(i >> 3) * C + cvg0 + g * 8 + (i & 7)], (double)sacc);
        }
    }
}

// Launches dw_strip_bf_kernel when the problem fits its limits (weights <= 64 KB of shared memory, < 2^31 threads); returns
// false when the caller should use the general kernel.
template <bool FLIP, bool STATS>
static bool launch_dw_strip_bf(const void* x, const void* w, int w_dtype, void* y, int N, int H, int W, int C, int Ho, int Wo,
                               int stride, double* stats, cudaStream_t stream, int* rc_out) {
    static const bool enabled = [] { const char* e = getenv("DMP_DW_BRANCHFREE"); return e == nullptr || e[0] != '0'; }();
    *rc_out = 0;
    if (!enabled) return false;
    size_t smem = (size_t)9 * 64 * sizeof(float);  // the taps of at most 8 channel vectors
    if (STATS && smem < (size_t)DW_STAT_SCRATCH_BYTES) smem = DW_STAT_SCRATCH_BYTES;
    const int TW = stride == 1 ? 4 : 2;
    const long long nstrips = (long long)N * Ho * ((Wo + TW - 1) / TW);
    if (smem > 64 * 1024 || nstrips >= (1ll << 31) || (stride != 1 && stride != 2)) return false;
    const int P = C / 8;
    int cvb = 8;
    while (cvb > 1 && P % cvb != 0) --cvb;  // largest divisor of C/8 up to 8
    const int spb = DW_THREADS / cvb;
    const long long grid_ll = ((nstrips + spb - 1) / spb) * (P / cvb);
    if (grid_ll >= (1ll << 31)) return false;
    const int grid = (int)grid_ll;
#define DW_BF_LAUNCH(WT, S, T)                                                                                                   \
    do {                                                                                                                         \
        auto kern = dw_strip_bf_kernel<WT, S, T, FLIP, STATS>;                                                                   \
        if (smem > 48 * 1024) {                                                                                                  \
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);                  \
            if (e != cudaSuccess) { dmp_set_error("dw_strip_bf: cudaFuncSetAttribute -> %s", cudaGetErrorString(e)); *rc_out = (int)e; return true; } \
        }                                                                                                                        \
        kern<<<grid, DW_THREADS, smem, stream>>>((const bf16*)x, (const WT*)w, (bf16*)y, N, H, W, C, Ho, Wo, stats, cvb);        \
    } while (0)
    if (stride == 1) {
        if (w_dtype == 0) DW_BF_LAUNCH(float, 1, 4); else DW_BF_LAUNCH(bf16, 1, 4);
    } else {
        if (w_dtype == 0) DW_BF_LAUNCH(float, 2, 2); else DW_BF_LAUNCH(bf16, 2, 2);
    }
#undef DW_BF_LAUNCH
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { dmp_set_error("dw_strip_bf launch -> %s", cudaGetErrorString(e)); *rc_out = (int)e; }
    return true;
}

// Strip version of the weight gradient.  A warp owns TWO adjacent channel vectors (lane & 1) and its 16 lane pairs walk
// 16 different (n, ho, strip) work items at a time, so a lane pair reads 32 contiguous bytes (one full sector) and all
// partial sums that must be combined live in the same warp: 4 xor-shuffle rounds fold the 16 items, then lanes 0/1 issue
// the 72 fp32 atomics of their channel vector.  (An earlier version combined threads through shared-memory atomics and
// spent most of its time serialised on bank conflicts.)
template <int STRIDE, int TW, bool BNIN = false>
__global__ void __launch_bounds__(DW_THREADS)
dw_wgrad_strip_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, float* __restrict__ dw, int N, int H, int W,
                      int C, int Ho, int Wo, const DwBn aux) {
    constexpr int NC = (TW - 1) * STRIDE + 3;
    const int P2 = C >> 4;                       // pairs of channel vectors (host guarantees C % 16 == 0)
    const int lane = threadIdx.x & 31, sub = lane & 1, li = lane >> 1;
    const long long gw = (blockIdx.x * (long long)DW_THREADS + threadIdx.x) >> 5;
    const long long nw = ((long long)gridDim.x * DW_THREADS) >> 5;   // multiple of P2 (host guarantees)
    const int pw = (int)(gw % P2);
    const int ch0 = (pw * 2 + sub) * 8;
    const int tiles_w = (Wo + TW - 1) / TW;
    const long long nitems = (long long)N * Ho * tiles_w;
    const long long istride = (nw / P2) * 16;
    float acc[72];
#pragma unroll
    for (int i = 0; i < 72; ++i) acc[i] = 0.f;
    float sc[8], sh[8], mu[8], is[8];
    if (BNIN) dw_bn_coeffs<2>(aux, ch0, C, sc, sh, mu, is);  // x holds the raw y1: the convolution saw relu(bn(y1))
    (void)mu; (void)is;
    for (long long item = (gw / P2) * 16 + li; item < nitems; item += istride) {
        long long t = item;
        const int wt_ = (int)(t % tiles_w); t /= tiles_w;
        const int ho = (int)(t % Ho);
        const int n = (int)(t / Ho);
        const int wo0 = wt_ * TW;
        // operands stay packed (8 bf16 = one int4) until the FMA that consumes them
        int4 g[TW];
#pragma unroll
        for (int i = 0; i < TW; ++i)
            g[i] = (wo0 + i < Wo) ? *(const int4*)(dy + (((long long)n * Ho + ho) * Wo + wo0 + i) * C + ch0) : make_int4(0, 0, 0, 0);
        const int wi0 = wo0 * STRIDE - 1;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * STRIDE + kh - 1;
            if (hi < 0 || hi >= H) continue;
            const bf16* row = x + (((long long)n * H + hi) * W) * C + ch0;
            int4 col[NC];
            bool cok[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int wi = wi0 + c;
                cok[c] = (wi >= 0 && wi < W);
                col[c] = cok[c] ? *(const int4*)(row + (long long)wi * C) : make_int4(0, 0, 0, 0);
            }
#pragma unroll
            for (int i = 0; i < TW; ++i) {
                float gf[8];
                bf16x8_to_float(*(bf16x8*)&g[i], gf);
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) {
                    float cf[8];
                    bf16x8_to_float(*(bf16x8*)&col[i * STRIDE + kw], cf);
                    if (BNIN) {
                        if (!cok[i * STRIDE + kw]) continue;  // zero padding of the normalised activation
#pragma unroll
                        for (int j = 0; j < 8; ++j) cf[j] = fmaxf(cf[j] * sc[j] + sh[j], 0.f);
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j * 9 + kh * 3 + kw] += gf[j] * cf[j];
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 72; ++i) {
        float v = acc[i];
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        acc[i] = v;
    }
    if (li == 0) {
        float* d = dw + (long long)ch0 * 9;
#pragma unroll
        for (int i = 0; i < 72; ++i) atomicAdd(d + i, acc[i]);
    }
}

// dw[c,kh,kw] = sum_{n,ho,wo} dy[n,ho,wo,c] * x[n,ho*s+kh-1,wo*s+kw-1,c]
// Each thread keeps 8x9 partial sums for its channel vector over a strided subset of output pixels,
// partials are combined per block in shared memory and flushed with one fp32 atomic per (block, weight).
__global__ void __launch_bounds__(DW_THREADS)
dw_wgrad_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, float* __restrict__ dw, int N, int H, int W,
                int C, int Ho, int Wo, int stride) {
    extern __shared__ float sm[];  // [groups_in_block][72]
    const int P = C >> 3;
    const int groups = P < DW_THREADS ? P : DW_THREADS;
    for (int i = threadIdx.x; i < groups * 72; i += DW_THREADS) sm[i] = 0.f;
    __syncthreads();
    const long long tid = blockIdx.x * (long long)DW_THREADS + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * DW_THREADS;  // multiple of P (host guarantees)
    const int cv = (int)(tid % P);
    const long long pstride = nthreads / P;
    const long long npix = (long long)N * Ho * Wo;
    const int ch0 = cv * 8;
    float acc[72];
#pragma unroll
    for (int i = 0; i < 72; ++i) acc[i] = 0.f;
    for (long long pix = tid / P; pix < npix; pix += pstride) {
        long long t = pix;
        const int wo = (int)(t % Wo); t /= Wo;
        const int ho = (int)(t % Ho);
        const int n = (int)(t / Ho);
        float g[8];
        load8(dy + pix * C + ch0, g);
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int hi = ho * stride + kh - 1;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int wi = wo * stride + kw - 1;
                if (wi < 0 || wi >= W) continue;
                float f[8];
                load8(x + (((long long)n * H + hi) * W + wi) * C + ch0, f);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j * 9 + kh * 3 + kw] += g[j] * f[j];
            }
        }
    }
    // group slot inside this block: threads with equal cv share a slot
    const int slot = (P < DW_THREADS) ? cv : threadIdx.x;
#pragma unroll
    for (int i = 0; i < 72; ++i) atomicAdd(&sm[slot * 72 + i], acc[i]);
    __syncthreads();
    __shared__ int s_cv[DW_THREADS];
    s_cv[(P < DW_THREADS) ? cv : threadIdx.x] = cv;  // identical value from every writer of a slot
    __syncthreads();
    for (int i = threadIdx.x; i < groups * 72; i += DW_THREADS) {
        const int sl = i / 72, k = i % 72;
        const float v = sm[i];
        if (v != 0.f) atomicAdd(&dw[(long long)s_cv[sl] * 72 + k], v);
    }
}

}  // namespace

// grid of the strip weight-gradient kernel: the total warp count must be a multiple of P2 = C/16 (pairs of channel vectors)
static int dw_wgrad_grid(int N, int Ho, int Wo, int C, int stride) {
    const int P2 = C / 16;
    const int warps_per_block = DW_THREADS / 32;
    int a = P2, b = warps_per_block;
    while (b) { int t = a % b; a = b; b = t; }
    const int q = P2 / a;
    const int TWs = stride == 1 ? 4 : 2;
    const long long nitems = (long long)N * Ho * ((Wo + TWs - 1) / TWs);
    long long want = (nitems * P2 / 16 + (long long)warps_per_block * 4 - 1) / ((long long)warps_per_block * 4);  // ~4 items per lane pair
    long long cap = (long long)dmp_num_sms();  // 160-190 registers per thread: one 256-thread block per SM
    // DMP_DW_WGRAD_MAX_BLOCKS (0 = none; default a third of the SMs): the kernel runs on a side stream next to the data-gradient
    // chain, and a grid that takes every SM stalls the chain until it drains (see wgrad_max_ctas in gemm_tcgen05.cu)
    static const int side_cap = [] { const char* e = getenv("DMP_DW_WGRAD_MAX_BLOCKS"); return e ? atoi(e) : dmp_num_sms() / 3; }();
    if (side_cap > 0 && cap > side_cap) cap = side_cap;
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    long long g = (want / q) * q;  // largest multiple of q under the cap (at least q)
    if (g < q) g = q;
    return (int)g;
}

static int dw_check(const char* name, int N, int H, int W, int C, int stride) {
    DMP_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0, "%s: empty tensor", name);
    DMP_REQUIRE(C % 8 == 0, "%s: C=%d must be a multiple of 8", name, C);
    DMP_REQUIRE(stride == 1 || stride == 2, "%s: stride %d unsupported", name, stride);
    return 0;
}

// w_dtype: 0 = fp32 weights [C,1,3,3], 1 = bf16.  x,y bf16 NHWC.
DMP_API int dmp_dwconv3x3_fwd(const void* x, const void* w, int w_dtype, void* y, int N, int H, int W, int C, int stride,
                              cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_fwd", N, H, W, C, stride)) return rc;
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    {
        int rc = 0;
        if (launch_dw_strip_bf<false, false>(x, w, w_dtype, y, N, H, W, C, Ho, Wo, stride, nullptr, stream, &rc)) {
            if (rc == 0) DMP_COUNT_LAUNCH(1);
            return rc;
        }
    }
    if (stride == 1) {
        const long long total = (long long)N * Ho * ((Wo + 3) / 4) * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0) dw_strip_kernel<float, 1, 4, false, false><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const float*)w, (bf16*)y, N, H, W, C, Ho, Wo, nullptr, DwBn{});
        else dw_strip_kernel<bf16, 1, 4, false, false><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const bf16*)w, (bf16*)y, N, H, W, C, Ho, Wo, nullptr, DwBn{});
    } else {
        const long long total = (long long)N * Ho * ((Wo + 1) / 2) * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0) dw_strip_kernel<float, 2, 2, false, false><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const float*)w, (bf16*)y, N, H, W, C, Ho, Wo, nullptr, DwBn{});
        else dw_strip_kernel<bf16, 2, 2, false, false><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const bf16*)w, (bf16*)y, N, H, W, C, Ho, Wo, nullptr, DwBn{});
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// Depthwise forward that also accumulates the BatchNorm batch sums of its (bf16-rounded) output: stats = double[2C]
// {sum, sum of squares}, accumulated (the consumer dmp_bn_apply_sums re-zeroes them).
DMP_API int dmp_dwconv3x3_fwd_colstats(const void* x, const void* w, int w_dtype, void* y, int N, int H, int W, int C, int stride,
                                       double* stats, cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_fwd_colstats", N, H, W, C, stride)) return rc;
    DMP_REQUIRE(stats != nullptr, "dmp_dwconv3x3_fwd_colstats: null stats");
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    {
        int rc = 0;
        if (launch_dw_strip_bf<false, true>(x, w, w_dtype, y, N, H, W, C, Ho, Wo, stride, stats, stream, &rc)) {
            if (rc == 0) DMP_COUNT_LAUNCH(1);
            return rc;
        }
    }
    if (stride == 1) {
        const long long total = (long long)N * Ho * ((Wo + 3) / 4) * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0) dw_strip_kernel<float, 1, 4, false, true><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const float*)w, (bf16*)y, N, H, W, C, Ho, Wo, stats, DwBn{});
        else dw_strip_kernel<bf16, 1, 4, false, true><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const bf16*)w, (bf16*)y, N, H, W, C, Ho, Wo, stats, DwBn{});
    } else {
        const long long total = (long long)N * Ho * ((Wo + 1) / 2) * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0) dw_strip_kernel<float, 2, 2, false, true><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const float*)w, (bf16*)y, N, H, W, C, Ho, Wo, stats, DwBn{});
        else dw_strip_kernel<bf16, 2, 2, false, true><<<grid, DW_THREADS, 0, stream>>>((const bf16*)x, (const bf16*)w, (bf16*)y, N, H, W, C, Ho, Wo, stats, DwBn{});
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_dwconv3x3_dgrad(const void* dy, const void* w, int w_dtype, void* dx, int N, int H, int W, int C,
                                int stride, cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_dgrad", N, H, W, C, stride)) return rc;
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    if (stride == 1) {  // dx = dy (*) flipped taps: the forward strip stencil applied to dy (Ho == H, Wo == W)
        int rc = 0;
        if (launch_dw_strip_bf<true, false>(dy, w, w_dtype, dx, N, H, W, C, H, W, 1, nullptr, stream, &rc)) {
            if (rc == 0) DMP_COUNT_LAUNCH(1);
            return rc;
        }
        const long long total = (long long)N * H * ((W + 3) / 4) * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0) dw_strip_kernel<float, 1, 4, true, false><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const float*)w, (bf16*)dx, N, H, W, C, H, W, nullptr, DwBn{});
        else dw_strip_kernel<bf16, 1, 4, true, false><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)w, (bf16*)dx, N, H, W, C, H, W, nullptr, DwBn{});
    } else if (stride == 2 && H % 2 == 0 && W % 2 == 0) {
        const long long total = (long long)N * Ho * ((Wo + 1) / 2) * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0) dw_dgrad_s2_kernel<float><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const float*)w, (bf16*)dx, N, H, W, C, Ho, Wo, DwBn{});
        else dw_dgrad_s2_kernel<bf16><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)w, (bf16*)dx, N, H, W, C, Ho, Wo, DwBn{});
    } else {
        const long long total = (long long)N * H * W * (C / 8);
        const int grid = dmp_ceil_div(total, DW_THREADS);
        if (w_dtype == 0)
            dw_dgrad_kernel<float><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const float*)w, (bf16*)dx, N, H, W, C, Ho, Wo, stride);
        else
            dw_dgrad_kernel<bf16><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)w, (bf16*)dx, N, H, W, C, Ho, Wo, stride);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// dw: fp32 [C,1,3,3]; overwritten unless accumulate != 0.
DMP_API int dmp_dwconv3x3_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int stride,
                                int accumulate, cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_wgrad", N, H, W, C, stride)) return rc;
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    if (!accumulate) DMP_CHECK_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * 9 * (size_t)C, stream));
    if (C % 16 == 0) {
        const int grid = dw_wgrad_grid(N, Ho, Wo, C, stride);
        if (stride == 1) dw_wgrad_strip_kernel<1, 4><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x, dw, N, H, W, C, Ho, Wo, DwBn{});
        else dw_wgrad_strip_kernel<2, 2><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x, dw, N, H, W, C, Ho, Wo, DwBn{});
    } else {
        const int P = C / 8;
        int a = P, b = DW_THREADS;
        while (b) { int t = a % b; a = b; b = t; }
        const int q = P / a;
        const long long npix = (long long)N * Ho * Wo;
        long long want = (npix * P + (long long)DW_THREADS * 16 - 1) / ((long long)DW_THREADS * 16);
        const long long cap = (long long)dmp_num_sms() * 2;
        if (want > cap) want = cap;
        if (want < 1) want = 1;
        const int grid = (int)(((want + q - 1) / q) * q);
        const int groups = P < DW_THREADS ? P : DW_THREADS;
        const size_t smem = (size_t)groups * 72 * sizeof(float);
        DMP_REQUIRE(smem <= 48 * 1024, "dmp_dwconv3x3_wgrad: C=%d (not a multiple of 16) is too wide for the fallback kernel", C);
        dw_wgrad_kernel<<<grid, DW_THREADS, smem, stream>>>((const bf16*)dy, (const bf16*)x, dw, N, H, W, C, Ho, Wo, stride);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}


// ---------------------------------------------------------------------------------------------------------------------
// Fused entry points for the expand -> bn1 -> relu -> depthwise -> bn2 chain (bf16 weights, bf16 NHWC activations).
// ---------------------------------------------------------------------------------------------------------------------
static DwBn dw_bn_args(const double* in_sums, double count, const float* gamma, const float* beta, float eps, float momentum,
                       float* running_mean, float* running_var, long long* nbt, float* mean_out, float* invstd_out,
                       const float* mean, const float* invstd, const void* y_raw, double* bwd_sums) {
    DwBn a;
    a.sums = in_sums; a.count = count; a.inv_count = count > 0 ? 1.0 / count : 0.0;
    a.gamma = gamma; a.beta = beta; a.eps = eps; a.momentum = momentum;
    a.running_mean = running_mean; a.running_var = running_var; a.num_batches_tracked = nbt;
    a.mean_out = mean_out; a.invstd_out = invstd_out; a.mean = mean; a.invstd = invstd;
    a.y_raw = (const bf16*)y_raw; a.bwd_sums = bwd_sums;
    return a;
}

// y = depthwise3x3(relu(bn1(x_raw))) where bn1's batch sums `in_sums` (double[2C] + ticket, see dmp_bn_sums_bytes) were
// accumulated by the producer of x_raw; also publishes bn1's mean / invstd / running statistics, clears `in_sums`, and
// accumulates the batch sums of y into `out_stats` (double[2C], for the BatchNorm that follows the depthwise convolution).
DMP_API int dmp_dwconv3x3_fwd_bnin_colstats(const void* x_raw, const void* w_bf16, void* y, int N, int H, int W, int C, int stride,
                                            double* in_sums, const float* gamma, const float* beta, float eps, float momentum,
                                            float* running_mean, float* running_var, long long* num_batches_tracked,
                                            float* mean_out, float* invstd_out, double* out_stats, cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_fwd_bnin_colstats", N, H, W, C, stride)) return rc;
    DMP_REQUIRE(in_sums && mean_out && invstd_out && out_stats, "dmp_dwconv3x3_fwd_bnin_colstats: null buffers");
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    const DwBn a = dw_bn_args(in_sums, (double)N * H * W, gamma, beta, eps, momentum, running_mean, running_var, num_batches_tracked,
                              mean_out, invstd_out, nullptr, nullptr, nullptr, nullptr);
    if (stride == 1) {
        const long long total = (long long)N * Ho * ((Wo + 3) / 4) * (C / 8);
        dw_strip_kernel<bf16, 1, 4, false, true, 1><<<dmp_ceil_div(total, DW_THREADS), DW_THREADS, 0, stream>>>(
            (const bf16*)x_raw, (const bf16*)w_bf16, (bf16*)y, N, H, W, C, Ho, Wo, out_stats, a);
    } else {
        const long long total = (long long)N * Ho * ((Wo + 1) / 2) * (C / 8);
        dw_strip_kernel<bf16, 2, 2, false, true, 1><<<dmp_ceil_div(total, DW_THREADS), DW_THREADS, 0, stream>>>(
            (const bf16*)x_raw, (const bf16*)w_bf16, (bf16*)y, N, H, W, C, Ho, Wo, out_stats, a);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// g = mask * depthwise_dgrad(dy) with mask = (bn1(y_raw) > 0), stored bf16, and bwd_sums[0..C) += sum g,
// bwd_sums[C..2C) += sum g * xhat: the data gradient of the depthwise convolution fused with the reduction phase of the
// backward of the relu(bn1(.)) that produced its input.  y_raw: [N,H,W,C] (the raw input of bn1).  H, W even for stride 2.
DMP_API int dmp_dwconv3x3_dgrad_bnbwd(const void* dy, const void* w_bf16, void* g, int N, int H, int W, int C, int stride,
                                      const void* y_raw, const float* mean, const float* invstd, const float* gamma,
                                      const float* beta, double* bwd_sums, cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_dgrad_bnbwd", N, H, W, C, stride)) return rc;
    DMP_REQUIRE(y_raw && mean && invstd && bwd_sums, "dmp_dwconv3x3_dgrad_bnbwd: null buffers");
    DMP_REQUIRE(stride == 1 || (H % 2 == 0 && W % 2 == 0), "dmp_dwconv3x3_dgrad_bnbwd: stride 2 needs even H, W");
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    const DwBn a = dw_bn_args(nullptr, (double)N * H * W, gamma, beta, 0.f, 0.f, nullptr, nullptr, nullptr, nullptr, nullptr, mean,
                              invstd, y_raw, bwd_sums);
    if (stride == 1) {
        const long long total = (long long)N * H * ((W + 3) / 4) * (C / 8);
        dw_strip_kernel<bf16, 1, 4, true, false, 0, true><<<dmp_ceil_div(total, DW_THREADS), DW_THREADS, 0, stream>>>(
            (const bf16*)dy, (const bf16*)w_bf16, (bf16*)g, N, H, W, C, H, W, nullptr, a);
    } else {
        const long long total = (long long)N * Ho * ((Wo + 1) / 2) * (C / 8);
        dw_dgrad_s2_kernel<bf16, true><<<dmp_ceil_div(total, DW_THREADS), DW_THREADS, 0, stream>>>(
            (const bf16*)dy, (const bf16*)w_bf16, (bf16*)g, N, H, W, C, Ho, Wo, a);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// dw += / = sum dy * relu(bn1(x_raw)) taps: the depthwise weight gradient with the normalised input re-created on the fly.
DMP_API int dmp_dwconv3x3_wgrad_bnin(const void* dy, const void* x_raw, float* dw, int N, int H, int W, int C, int stride,
                                     int accumulate, const float* mean, const float* invstd, const float* gamma, const float* beta,
                                     cudaStream_t stream) {
    if (int rc = dw_check("dmp_dwconv3x3_wgrad_bnin", N, H, W, C, stride)) return rc;
    DMP_REQUIRE(C % 16 == 0, "dmp_dwconv3x3_wgrad_bnin: C=%d must be a multiple of 16", C);
    DMP_REQUIRE(mean && invstd, "dmp_dwconv3x3_wgrad_bnin: null statistics");
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    if (!accumulate) DMP_CHECK_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * 9 * (size_t)C, stream));
    const DwBn a = dw_bn_args(nullptr, (double)N * H * W, gamma, beta, 0.f, 0.f, nullptr, nullptr, nullptr, nullptr, nullptr, mean,
                              invstd, nullptr, nullptr);
    const int grid = dw_wgrad_grid(N, Ho, Wo, C, stride);
    if (stride == 1) dw_wgrad_strip_kernel<1, 4, true><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x_raw, dw, N, H, W, C, Ho, Wo, a);
    else dw_wgrad_strip_kernel<2, 2, true><<<grid, DW_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x_raw, dw, N, H, W, C, Ho, Wo, a);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
