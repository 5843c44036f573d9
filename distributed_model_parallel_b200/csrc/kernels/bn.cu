// BatchNorm (train mode) for channels-last activations x[M, C] (M = N*H*W rows, C contiguous).
// Replaces the 57 nn.BatchNorm2d layers of the reference model
// (/root/reference/code/distributed_training/model/mobilenetv2.py:18,20,22,28,53,56) and is split so that
// SyncBatchNorm can insert its cross-GPU combine between the statistics and the finalize step
// (torch/nn/modules/_functions.py:39-122 is the behaviour being matched):
//   bn_stats      : per-channel sum / sum-of-squares (double accumulators in global scratch)
//   bn_finalize   : mean, invstd, running-stat update, fused scale/shift
//   bn_apply      : y = act(x*scale + shift (+ residual))
//   bn_bwd_reduce : per-channel sum(dy_m), sum(dy_m * xhat)       (dy_m = dy masked by the fused ReLU)
//   bn_bwd_dx     : dx = gamma*invstd*(dy_m - s1/n - xhat*s2/n), optional d(residual) = dy_m
//   bn_bwd_params : dgamma, dbeta
// The activation tensor is treated as a flat array of 16-byte vectors.  The grid is sized so that
// (threads * VEC) is a multiple of C, therefore every thread sees the same VEC channels on every
// iteration and keeps its partial sums / affine coefficients in registers.
#include "../common.cuh"
#include <stdlib.h>
#include <atomic>

namespace {

constexpr int BN_THREADS = 256;

template <typename T>
struct Vec;
template <>
struct Vec<bf16> {
    static constexpr int N = 8;
    __device__ static int4 load_raw(const bf16* p, long long v) { return *(const int4*)(p + v * 8); }
    __device__ static void unpack(const int4& raw, float* out) { bf16x8_to_float(*(const bf16x8*)&raw, out); }
    __device__ static void load(const bf16* p, long long v, float* out) {
        int4 raw = *(const int4*)(p + v * 8);
        bf16x8 b = *(bf16x8*)&raw;
        bf16x8_to_float(b, out);
    }
    __device__ static void store(bf16* p, long long v, const float* in) {
        bf16x8 b = float_to_bf16x8(in);
        *(int4*)(p + v * 8) = *(int4*)&b;
    }
};
template <>
struct Vec<float> {
    static constexpr int N = 4;
    __device__ static int4 load_raw(const float* p, long long v) { return *(const int4*)(p + v * 4); }
    __device__ static void unpack(const int4& raw, float* out) {
        out[0] = __int_as_float(raw.x); out[1] = __int_as_float(raw.y); out[2] = __int_as_float(raw.z); out[3] = __int_as_float(raw.w);
    }
    __device__ static void load(const float* p, long long v, float* out) {
        float4 r = *(const float4*)(p + v * 4);
        out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
    }
    __device__ static void store(float* p, long long v, const float* in) {
        *(float4*)(p + v * 4) = make_float4(in[0], in[1], in[2], in[3]);
    }
};

// Block-level reduction of per-thread channel partials into global double accumulators.
// vals[k*VEC + j] is the k-th statistic of channel (ch0 + j).  Threads whose ch0 is equal are combined
// through shared memory; one double atomic per (block, channel, statistic) reaches global memory.
template <int VEC, int NSTAT>
__device__ void block_channel_reduce(const float* vals, int ch0, int C, double* __restrict__ accum, float* sm) {
    // sm: [BN_THREADS][NSTAT*VEC] floats
    float* mine = sm + threadIdx.x * (NSTAT * VEC);
#pragma unroll
    for (int i = 0; i < NSTAT * VEC; ++i) mine[i] = vals[i];
    __shared__ int s_ch0[BN_THREADS];
    s_ch0[threadIdx.x] = ch0;
    __syncthreads();
    // groups repeat with period P = C / VEC threads (or all threads distinct if P >= blockDim)
    const int P = C / VEC;
    const int ngroups = P < BN_THREADS ? P : BN_THREADS;
    for (int item = threadIdx.x; item < ngroups * NSTAT * VEC; item += BN_THREADS) {
        const int g = item / (NSTAT * VEC);  // first thread of this group
        const int i = item % (NSTAT * VEC);
        float s = 0.f;
        for (int t = g; t < BN_THREADS; t += P) s += sm[t * (NSTAT * VEC) + i];
        const int k = i / VEC, j = i % VEC;
        atomicAdd(&accum[(long long)k * C + s_ch0[g] + j], (double)s);
    }
}


// Slab mapping of the cooperative BatchNorm kernels: a block owns `cvb` adjacent channel vectors (cvb divides C / VEC, at most 8:
// 128 contiguous bytes of a row) and BN_THREADS / cvb rows per sweep; thread t has channel vector t % cvb and row t / cvb.  Every
// block then contributes NSTAT * VEC * cvb values, to accumulators shared only with the blocks of the same slab -- with the
// flat mapping (consecutive threads over the whole tensor) every block ended with NSTAT * C atomics on the same NSTAT * C
// addresses as every other block.  vals[k*VEC + j]: k-th statistic of the thread's channel j.  sm: [BN_THREADS][NSTAT*VEC + 1].
template <int VEC, int NSTAT>
__device__ void block_slab_reduce(const float* vals, int cvb, int C, int slab_ch0, double* __restrict__ accum, float* sm) {
    constexpr int W = NSTAT * VEC;
    float* mine = sm + threadIdx.x * (W + 1);
#pragma unroll
    for (int i = 0; i < W; ++i) mine[i] = vals[i];  // threads without a row hold zeros
    __syncthreads();
    const int items = cvb * W;                      // <= 128
    const int rpb = BN_THREADS / cvb;
    int T = 1;                                      // threads per item: a power of two, item-major so partners are adjacent lanes
    while (T * 2 * items <= BN_THREADS && T < 32) T *= 2;
    const int item = threadIdx.x / T, k = threadIdx.x % T;
    float s = 0.f;
    if (item < items) {
        const int g = item / W, i = item % W;
        for (int q = k; q < rpb; q += T) s += sm[(q * cvb + g) * (W + 1) + i];
    }
    for (int o = T >> 1; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (item < items && k == 0) {
        const int g = item / W, i = item % W;
        atomicAdd(&accum[(long long)(i / VEC) * C + slab_ch0 + g * VEC + (i % VEC)], (double)s);
    }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const T* __restrict__ x, long long nvec, int C, double* __restrict__ accum) {
    constexpr int VEC = Vec<T>::N;
    extern __shared__ float sm[];
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)((tid * VEC) % C);
    float acc[2 * VEC];
#pragma unroll
    for (int i = 0; i < 2 * VEC; ++i) acc[i] = 0.f;
    for (long long v = tid; v < nvec; v += stride) {
        float f[VEC];
        Vec<T>::load(x, v, f);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            acc[j] += f[j];
            acc[VEC + j] += f[j] * f[j];
        }
    }
    block_channel_reduce<VEC, 2>(acc, ch0, C, accum, sm);
}

__global__ void bn_finalize_kernel(const double* __restrict__ accum, double count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, float momentum, float eps, float* __restrict__ mean,
                                   float* __restrict__ invstd, float* __restrict__ scale, float* __restrict__ shift,
                                   int C) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double m = accum[c] / count;
    double var = accum[C + c] / count - m * m;
    if (var < 0.0) var = 0.0;
    float is = (float)(1.0 / sqrt(var + (double)eps));
    float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    mean[c] = (float)m;
    invstd[c] = is;
    scale[c] = g * is;
    shift[c] = b - (float)m * g * is;
    if (running_mean != nullptr) {
        double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const T* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                const T* __restrict__ residual, T* __restrict__ y, long long nvec, int C, int relu) {
    constexpr int VEC = Vec<T>::N;
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)((tid * VEC) % C);
    float sc[VEC], sh[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        sc[j] = scale[ch0 + j];
        sh[j] = shift[ch0 + j];
    }
    for (long long v = tid; v < nvec; v += stride) {
        float f[VEC];
        Vec<T>::load(x, v, f);
#pragma unroll
        for (int j = 0; j < VEC; ++j) f[j] = f[j] * sc[j] + sh[j];
        if (residual != nullptr) {
            float r[VEC];
            Vec<T>::load(residual, v, r);
#pragma unroll
            for (int j = 0; j < VEC; ++j) f[j] += r[j];
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        Vec<T>::store(y, v, f);
    }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x,
                     const float* __restrict__ mean, const float* __restrict__ invstd, long long nvec, int C, int relu,
                     double* __restrict__ accum) {
    constexpr int VEC = Vec<T>::N;
    extern __shared__ float sm[];
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)((tid * VEC) % C);
    float mu[VEC], is[VEC], acc[2 * VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        mu[j] = mean[ch0 + j];
        is[j] = invstd[ch0 + j];
        acc[j] = 0.f;
        acc[VEC + j] = 0.f;
    }
    for (long long v = tid; v < nvec; v += stride) {
        float g[VEC], xv[VEC];
        Vec<T>::load(dy, v, g);
        Vec<T>::load(x, v, xv);
        if (relu) {
            float yv[VEC];
            Vec<T>::load(y, v, yv);
#pragma unroll
            for (int j = 0; j < VEC; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            acc[j] += g[j];
            acc[VEC + j] += g[j] * (xv[j] - mu[j]) * is[j];
        }
    }
    block_channel_reduce<VEC, 2>(acc, ch0, C, accum, sm);
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_dx_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x,
                 const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                 const double* __restrict__ accum, double count, T* __restrict__ dx, T* __restrict__ dres,
                 long long nvec, int C, int relu) {
    constexpr int VEC = Vec<T>::N;
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)((tid * VEC) % C);
    float mu[VEC], is[VEC], a[VEC], b[VEC], c[VEC];
    const double inv_count = 1.0 / count;  // one fp64 division per thread; per-channel divisions cost more than the data pass
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        mu[j] = mean[ch0 + j];
        is[j] = invstd[ch0 + j];
        a[j] = (gamma ? gamma[ch0 + j] : 1.f) * is[j];
        b[j] = (float)(accum[ch0 + j] * inv_count);
        c[j] = (float)(accum[C + ch0 + j] * inv_count);
    }
    for (long long v = tid; v < nvec; v += stride) {
        float g[VEC], xv[VEC];
        Vec<T>::load(dy, v, g);
        Vec<T>::load(x, v, xv);
        if (relu) {
            float yv[VEC];
            Vec<T>::load(y, v, yv);
#pragma unroll
            for (int j = 0; j < VEC; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        }
        if (dres != nullptr) Vec<T>::store(dres, v, g);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            float xh = (xv[j] - mu[j]) * is[j];
            g[j] = a[j] * (g[j] - b[j] - xh * c[j]);
        }
        Vec<T>::store(dx, v, g);
    }
}

__global__ void bn_bwd_params_kernel(const double* __restrict__ accum_local, float* __restrict__ dgamma,
                                     float* __restrict__ dbeta, int C, int accumulate) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float db = (float)accum_local[c], dg = (float)accum_local[C + c];
    if (dgamma) dgamma[c] = accumulate ? dgamma[c] + dg : dg;
    if (dbeta) dbeta[c] = accumulate ? dbeta[c] + db : db;
}

// ---------------------------------------------------------------------------------------------------------
// Fused single-launch BatchNorm (cooperative grid): statistics -> grid barrier -> normalise, and for backward
// reductions -> grid barrier -> dx.  The second pass re-reads the activation from L2 (the largest MobileNetV2 tensor
// is 38 MB, L2 is 126 MB).  `ws` is a per-layer workspace {double accum[2][2C]; int parity}: call k accumulates into
// accum[parity] and clears accum[parity^1] for call k+1, then flips the parity -- all on the device, so a captured
// CUDA graph can be replayed without host-side bookkeeping.  Replaces 5 launches (memset, stats, finalize, counter
// increment, apply) by one, and 4 (memset, reduce, dx, params) by one.
// ---------------------------------------------------------------------------------------------------------
}  // namespace
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
namespace {

// The per-block double atomics can be spread over NSETS accumulator copies (block b uses copy b % NSETS, every thread adds
// the copies of its channels after the grid barrier).  Measured on B200: NSETS = 8 made the MobileNetV2 step 15% SLOWER
// (6.37 vs 5.52 ms) -- the extra dependent L2 loads after the barrier and the 8x larger clear cost more than the atomic
// serialisation they remove -- so a single copy is used.
constexpr int NSETS = 1;

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_fwd_fused_kernel(const T* __restrict__ x, const T* __restrict__ residual, T* __restrict__ y, long long nvec, int C,
                    double count, const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ running_mean,
                    float* __restrict__ running_var, long long* __restrict__ num_batches_tracked, float momentum, float eps,
                    float* __restrict__ mean_out, float* __restrict__ invstd_out, double* __restrict__ ws, int relu) {
    constexpr int VEC = Vec<T>::N;
    extern __shared__ float sm[];
    cg::grid_group grid = cg::this_grid();
    int* parity_p = (int*)(ws + 4 * (long long)NSETS * C);
    const int parity = *parity_p;
    double* accum = ws + (long long)parity * 2 * NSETS * C;
    double* other = ws + (long long)(parity ^ 1) * 2 * NSETS * C;
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)((tid * VEC) % C);
    float acc[2 * VEC];
#pragma unroll
    for (int i = 0; i < 2 * VEC; ++i) acc[i] = 0.f;
    constexpr int U = 4;  // independent 16-byte loads in flight per thread (HBM latency x bandwidth needs MBs in flight)
    for (long long v0 = tid; v0 < nvec; v0 += U * stride) {
        int4 raw[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) raw[u] = Vec<T>::load_raw(x, v0 + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                float f[VEC];
                Vec<T>::unpack(raw[u], f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    acc[j] += f[j];
                    acc[VEC + j] += f[j] * f[j];
                }
            }
    }
    block_channel_reduce<VEC, 2>(acc, ch0, C, accum + (long long)(blockIdx.x % NSETS) * 2 * C, sm);
    grid.sync();
    const double inv_count = 1.0 / count;
    float sc[VEC], sh[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const int c = ch0 + j;
        double s1 = 0.0, s2 = 0.0;
#pragma unroll
        for (int k = 0; k < NSETS; ++k) {
            s1 += accum[(long long)k * 2 * C + c];
            s2 += accum[(long long)k * 2 * C + C + c];
        }
        // per-thread path: fp64 only for the cancellation-prone s2/n - m^2 (multiplies), the root and reciprocal in fp32 --
        // fp64 divisions and square roots are long instruction sequences and, at 8 channels x 3 per thread, used to cost
        // more than the data pass of a small layer.  Block 0 below still publishes mean / invstd with full fp64 math.
        const double m = s1 * inv_count;
        double var = s2 * inv_count - m * m;
        if (var < 0.0) var = 0.0;
        const float is = 1.0f / sqrtf((float)var + eps);
        const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
        sc[j] = g * is;
        sh[j] = b - (float)m * g * is;
    }
    if (blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += BN_THREADS) {
            double s1 = 0.0, s2 = 0.0;
            for (int k = 0; k < NSETS; ++k) {
                s1 += accum[(long long)k * 2 * C + c];
                s2 += accum[(long long)k * 2 * C + C + c];
            }
            const double m = s1 / count;
            double var = s2 / count - m * m;
            if (var < 0.0) var = 0.0;
            mean_out[c] = (float)m;
            invstd_out[c] = (float)(1.0 / sqrt(var + (double)eps));
            if (running_mean != nullptr) {
                const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
            }
        }
        if (threadIdx.x == 0) {
            if (num_batches_tracked != nullptr) *num_batches_tracked += 1;
            *parity_p = parity ^ 1;
        }
    }
    for (long long i = tid; i < 2 * (long long)NSETS * C; i += stride) other[i] = 0.0;  // ready for the next call
    for (long long v0 = tid; v0 < nvec; v0 += U * stride) {
        int4 raw[U], rres[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                raw[u] = Vec<T>::load_raw(x, v0 + u * stride);
                if (residual != nullptr) rres[u] = Vec<T>::load_raw(residual, v0 + u * stride);
            }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                float f[VEC];
                Vec<T>::unpack(raw[u], f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) f[j] = f[j] * sc[j] + sh[j];
                if (residual != nullptr) {
                    float r[VEC];
                    Vec<T>::unpack(rres[u], r);
#pragma unroll
                    for (int j = 0; j < VEC; ++j) f[j] += r[j];
                }
                if (relu) {
#pragma unroll
                    for (int j = 0; j < VEC; ++j) f[j] = fmaxf(f[j], 0.f);
                }
                Vec<T>::store(y, v0 + u * stride, f);
            }
    }
}

// BatchNorm forward when the batch sums already exist (written by the producing convolution's epilogue, see
// dmp_gemm_bf16_colstats / dmp_dwconv3x3_fwd_colstats): one plain elementwise pass, no second read of x for statistics and no
// grid-wide barrier.  sums = double[2C] {sum, sum of squares} followed by one int ticket.  Every block derives scale/shift for
// its own channels; block 0 also publishes mean / invstd / running statistics; the last block to have read the sums (ticket)
// zeroes them again, so the buffer is ready for the producer's next accumulation (CUDA-graph replay included).
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_sums_kernel(const T* __restrict__ x, const T* __restrict__ residual, T* __restrict__ y, long long nvec, int C,
                     double count, const float* __restrict__ gamma, const float* __restrict__ beta,
                     float* __restrict__ running_mean, float* __restrict__ running_var,
                     long long* __restrict__ num_batches_tracked, float momentum, float eps, float* __restrict__ mean_out,
                     float* __restrict__ invstd_out, double* __restrict__ sums, int relu) {
    constexpr int VEC = Vec<T>::N;
    constexpr int U = 4;
    __shared__ int s_last;
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)(((unsigned)tid * (unsigned)VEC) % (unsigned)C);  // grid * BN_THREADS * VEC stays far below 2^31
    const double inv_count = 1.0 / count;
    float sc[VEC], sh[VEC];
    bool odd = false;  // folds every loaded sum into the barrier predicate below
    extern __shared__ float s_coef[];  // [2][C] scale / shift when the launch provides it (C <= BN_APPLY_SMEM_MAX_C)
    const bool coop = (relu & 8) != 0;
    if (coop) {
        // The block derives each channel's coefficients ONCE (C / 256 channels per thread) and its threads pick theirs out of
        // shared memory; every thread doing its own 8 channels repeated the same 16 dependent fp64 operations 256 times per
        // block and was the longest part of a launch on the small layers.
        for (int c = threadIdx.x; c < C; c += BN_THREADS) {
            const double m = sums[c] * inv_count;
            double var = sums[C + c] * inv_count - m * m;
            if (var < 0.0) var = 0.0;
            const float is = 1.0f / sqrtf((float)var + eps);
            const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
            const float scv = g * is, shv = b - (float)m * g * is;
            s_coef[c] = scv;
            s_coef[C + c] = shv;
            odd |= !(scv == scv) && !(shv == shv) && (scv == shv);  // never true; depends on both loads
        }
    } else {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const int c = ch0 + j;
            const double m = sums[c] * inv_count;
            double var = sums[C + c] * inv_count - m * m;
            if (var < 0.0) var = 0.0;
            const float is = 1.0f / sqrtf((float)var + eps);
            const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
            sc[j] = g * is;
            sh[j] = b - (float)m * g * is;
            odd |= !(sc[j] == sc[j]) && !(sh[j] == sh[j]) && (sc[j] == sh[j]);  // never true; depends on both loads
        }
    }
    // One channel per thread over the first ceil(C / BN_THREADS) blocks publishes mean / invstd / running statistics (a single
    // block looping over C channels of double divisions and square roots was the slowest block of every launch).
    if (tid < C) {
        const int c = (int)tid;
        const double m = sums[c] * inv_count;
        double var = sums[C + c] * inv_count - m * m;
        if (var < 0.0) var = 0.0;
        mean_out[c] = (float)m;
        const double isd = 1.0 / sqrt(var + (double)eps);
        invstd_out[c] = (float)isd;
        odd |= !(isd == isd);
        if (running_mean != nullptr) {
            const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
        }
        if (c == 0 && num_batches_tracked != nullptr) *num_batches_tracked += 1;
    }
    // Everything this block needs from `sums` has ARRIVED in registers once every thread has evaluated `odd` (the barrier
    // predicate forces that use), so the ticket needs no memory fence: the last block to take one clears the buffer for the
    // producer's next accumulation (CUDA-graph replay included), and a cleared value can no longer reach any thread.
    const int any_odd = __syncthreads_or(odd ? 1 : 0);
    if (coop) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            sc[j] = s_coef[ch0 + j];
            sh[j] = s_coef[C + ch0 + j];
        }
    }
    if (threadIdx.x == 0) {
        int* ticket = (int*)(sums + 2 * (long long)C);
        s_last = (atomicAdd(ticket, 1 + any_odd) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        for (int i = threadIdx.x; i < 2 * C; i += BN_THREADS) sums[i] = 0.0;
        if (threadIdx.x == 0) *(int*)(sums + 2 * (long long)C) = 0;
    }
    for (long long v0 = tid; v0 < nvec; v0 += U * stride) {
        int4 raw[U], rres[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                raw[u] = Vec<T>::load_raw(x, v0 + u * stride);
                if (residual != nullptr) rres[u] = Vec<T>::load_raw(residual, v0 + u * stride);
            }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                float f[VEC];
                Vec<T>::unpack(raw[u], f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) f[j] = f[j] * sc[j] + sh[j];
                if (residual != nullptr) {
                    float r[VEC];
                    Vec<T>::unpack(rres[u], r);
#pragma unroll
                    for (int j = 0; j < VEC; ++j) f[j] += r[j];
                }
                if (relu & 1) {
#pragma unroll
                    for (int j = 0; j < VEC; ++j) f[j] = fmaxf(f[j], 0.f);
                }
                Vec<T>::store(y, v0 + u * stride, f);
            }
    }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS, 2)  // two blocks per SM: at 148 registers (one block) the whole step lost 8 %
bn_bwd_fused_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, const float* __restrict__ mean,
                    const float* __restrict__ invstd, const float* __restrict__ gamma, const float* __restrict__ beta, double count,
                    T* __restrict__ dx, T* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    int accumulate_params, long long nvec, int C, int relu, double* __restrict__ ws, int cvb) {
    constexpr int VEC = Vec<T>::N;
    constexpr int U = 4;
    extern __shared__ float sm[];
    cg::grid_group grid = cg::this_grid();
    int* parity_p = (int*)(ws + 4 * (long long)NSETS * C);
    const int parity = *parity_p;
    double* accum = ws + (long long)parity * 2 * NSETS * C;
    double* other = ws + (long long)(parity ^ 1) * 2 * NSETS * C;
    // slab mapping (see block_slab_reduce): this thread's vectors are vfirst, vfirst + stride, ... (all of one channel vector)
    const long long gtid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long gstride = (long long)gridDim.x * BN_THREADS;
    const int P = C / VEC, cvgroups = P / cvb, rpb = BN_THREADS / cvb;
    const int cvl = threadIdx.x % cvb, rl = threadIdx.x / cvb;
    const int cv = (blockIdx.x % cvgroups) * cvb + cvl;
    const int row_blocks = gridDim.x / cvgroups;
    const long long stride = (long long)row_blocks * rpb * P;
    const long long tid = rl < rpb ? ((long long)(blockIdx.x / cvgroups) * rpb + rl) * P + cv : nvec;  // no row: nothing to do
    const int ch0 = cv * VEC;
    // ReLU mask: from the saved output when given, otherwise recomputed from x (y = relu(xhat*gamma + beta) without a
    // residual), which saves reading a third tensor in both passes
    const bool mask_from_x = relu && (y == nullptr);
    float mu[VEC], is[VEC], gs[VEC], bs[VEC], acc[2 * VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        mu[j] = mean[ch0 + j];
        is[j] = invstd[ch0 + j];
        gs[j] = (gamma ? gamma[ch0 + j] : 1.f);
        bs[j] = (beta ? beta[ch0 + j] : 0.f);
        acc[j] = 0.f;
        acc[VEC + j] = 0.f;
    }
    for (long long v0 = tid; v0 < nvec; v0 += U * stride) {
        int4 rg[U], rx[U], ry[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                rg[u] = Vec<T>::load_raw(dy, v0 + u * stride);
                rx[u] = Vec<T>::load_raw(x, v0 + u * stride);
                if (relu && !mask_from_x) ry[u] = Vec<T>::load_raw(y, v0 + u * stride);
            }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                float g[VEC], xv[VEC];
                Vec<T>::unpack(rg[u], g);
                Vec<T>::unpack(rx[u], xv);
                if (relu) {
                    if (mask_from_x) {
#pragma unroll
                        for (int j = 0; j < VEC; ++j) g[j] = ((xv[j] - mu[j]) * is[j] * gs[j] + bs[j]) > 0.f ? g[j] : 0.f;
                    } else {
                        float yv[VEC];
                        Vec<T>::unpack(ry[u], yv);
#pragma unroll
                        for (int j = 0; j < VEC; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
                    }
                }
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    acc[j] += g[j];
                    acc[VEC + j] += g[j] * (xv[j] - mu[j]) * is[j];
                }
            }
    }
    block_slab_reduce<VEC, 2>(acc, cvb, C, (int)(blockIdx.x % cvgroups) * cvb * VEC, accum, sm);
    grid.sync();
    const double inv_count = 1.0 / count;
    float a[VEC], b[VEC], c[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        double s1 = 0.0, s2 = 0.0;
#pragma unroll
        for (int k = 0; k < NSETS; ++k) {
            s1 += accum[(long long)k * 2 * C + ch0 + j];
            s2 += accum[(long long)k * 2 * C + C + ch0 + j];
        }
        a[j] = gs[j] * is[j];
        b[j] = (float)(s1 * inv_count);
        c[j] = (float)(s2 * inv_count);
    }
    if (blockIdx.x == 0) {
        for (int ch = threadIdx.x; ch < C; ch += BN_THREADS) {
            double s1 = 0.0, s2 = 0.0;
            for (int k = 0; k < NSETS; ++k) {
                s1 += accum[(long long)k * 2 * C + ch];
                s2 += accum[(long long)k * 2 * C + C + ch];
            }
            const float db = (float)s1, dg = (float)s2;
            if (dgamma) dgamma[ch] = (accumulate_params & 1) ? dgamma[ch] + dg : dg;
            if (dbeta) dbeta[ch] = (accumulate_params & 1) ? dbeta[ch] + db : db;
        }
        if (threadIdx.x == 0) *parity_p = parity ^ 1;
    }
    for (long long i = gtid; i < 2 * (long long)NSETS * C; i += gstride) other[i] = 0.0;
    // second pass in REVERSE order: the tail of the first pass is what the L2 still holds (tensors larger than the L2 would
    // otherwise be evicted just ahead of their re-use, front to back)
    const long long n_iter = (nvec - tid + U * stride - 1) / (U * stride);
    const bool reverse = (accumulate_params & 2) != 0;
    for (long long k = 0; k < n_iter; ++k) {
        const long long it = reverse ? n_iter - 1 - k : k;
        const long long v0 = tid + it * (U * stride);
        int4 rg[U], rx[U], ry[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                rg[u] = Vec<T>::load_raw(dy, v0 + u * stride);
                rx[u] = Vec<T>::load_raw(x, v0 + u * stride);
                if (relu && !mask_from_x) ry[u] = Vec<T>::load_raw(y, v0 + u * stride);
            }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                float g[VEC], xv[VEC];
                Vec<T>::unpack(rg[u], g);
                Vec<T>::unpack(rx[u], xv);
                if (relu) {
                    if (mask_from_x) {
#pragma unroll
                        for (int j = 0; j < VEC; ++j) g[j] = ((xv[j] - mu[j]) * is[j] * gs[j] + bs[j]) > 0.f ? g[j] : 0.f;
                    } else {
                        float yv[VEC];
                        Vec<T>::unpack(ry[u], yv);
#pragma unroll
                        for (int j = 0; j < VEC; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
                    }
                }
                if (dres != nullptr) Vec<T>::store(dres, v0 + u * stride, g);
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    const float xh = (xv[j] - mu[j]) * is[j];
                    g[j] = a[j] * (g[j] - b[j] - xh * c[j]);
                }
                Vec<T>::store(dx, v0 + u * stride, g);
            }
    }
}

// Shared-memory-resident variant of bn_bwd_fused_kernel (bf16 tensors that fit on chip: 32 or 48 bytes per 8 elements, up to
// ~190 KB per SM): dy, x (and y when the ReLU mask needs it) are fetched ONCE, with cp.async straight into shared memory -- every
// load of a thread is in flight at the same time and costs no registers -- and both passes (reductions, then dx after the grid
// barrier) read them back from there.  The two-pass kernel re-reads both tensors after the barrier and keeps only 8 loads per
// thread in flight; for the layers of the 8x8 / 4x4 stages (1-10 MB) that second round trip was a third of the launch.
// vpt = vectors per thread (runtime); shared memory: [vpt][NT][BN_THREADS] int4 followed by the reduction scratch.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}

__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_resident_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ y, const bf16* __restrict__ x,
                       const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                       const float* __restrict__ beta, double count, bf16* __restrict__ dx, bf16* __restrict__ dres,
                       float* __restrict__ dgamma, float* __restrict__ dbeta, int accumulate_params, long long nvec, int C, int relu,
                       double* __restrict__ ws, int vpt, int cvb) {
    constexpr int VEC = 8;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cg::grid_group grid = cg::this_grid();
    const bool mask_from_x = relu && (y == nullptr);
    const bool need_y = relu && !mask_from_x;
    const int NT = need_y ? 3 : 2;
    int4* stash = (int4*)smem_raw;                                                      // [vpt][NT][BN_THREADS]
    float* sm = (float*)(smem_raw + (size_t)vpt * NT * BN_THREADS * sizeof(int4));      // block_channel_reduce scratch
    const long long gtid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long gstride = (long long)gridDim.x * BN_THREADS;
    const int P = C / VEC, cvgroups = P / cvb, rpb = BN_THREADS / cvb;
    const int cvl = threadIdx.x % cvb, rl = threadIdx.x / cvb;
    const int cv = (blockIdx.x % cvgroups) * cvb + cvl;
    const long long stride = (long long)(gridDim.x / cvgroups) * rpb * P;
    const long long tid = rl < rpb ? ((long long)(blockIdx.x / cvgroups) * rpb + rl) * P + cv : nvec;
    for (int u = 0; u < vpt; ++u) {
        const long long v = tid + u * stride;
        if (v < nvec) {
            int4* slot = stash + ((size_t)u * NT) * BN_THREADS + threadIdx.x;
            cp_async16(slot, dy + v * 8);
            cp_async16(slot + BN_THREADS, x + v * 8);
            if (need_y) cp_async16(slot + 2 * BN_THREADS, y + v * 8);
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    int* parity_p = (int*)(ws + 4 * (long long)NSETS * C);
    const int parity = *parity_p;
    double* accum = ws + (long long)parity * 2 * NSETS * C;
    double* other = ws + (long long)(parity ^ 1) * 2 * NSETS * C;
    const int ch0 = cv * VEC;
    float mu[VEC], is[VEC], gs[VEC], bs[VEC], acc[2 * VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        mu[j] = mean[ch0 + j];
        is[j] = invstd[ch0 + j];
        gs[j] = (gamma ? gamma[ch0 + j] : 1.f);
        bs[j] = (beta ? beta[ch0 + j] : 0.f);
        acc[j] = 0.f;
        acc[VEC + j] = 0.f;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");  // each thread reads back only what it fetched itself
    for (int u = 0; u < vpt; ++u) {
        if (tid + u * stride < nvec) {
            const int4* slot = stash + ((size_t)u * NT) * BN_THREADS + threadIdx.x;
            float g[VEC], xv[VEC];
            Vec<bf16>::unpack(slot[0], g);
            Vec<bf16>::unpack(slot[BN_THREADS], xv);
            if (relu) {
                if (mask_from_x) {
#pragma unroll
                    for (int j = 0; j < VEC; ++j) g[j] = ((xv[j] - mu[j]) * is[j] * gs[j] + bs[j]) > 0.f ? g[j] : 0.f;
                } else {
                    float yv[VEC];
                    Vec<bf16>::unpack(slot[2 * BN_THREADS], yv);
#pragma unroll
                    for (int j = 0; j < VEC; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
                }
            }
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                acc[j] += g[j];
                acc[VEC + j] += g[j] * (xv[j] - mu[j]) * is[j];
            }
        }
    }
    block_slab_reduce<VEC, 2>(acc, cvb, C, (int)(blockIdx.x % cvgroups) * cvb * VEC, accum, sm);
    grid.sync();
    const double inv_count = 1.0 / count;
    float a[VEC], b[VEC], c[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        double s1 = 0.0, s2 = 0.0;
#pragma unroll
        for (int k = 0; k < NSETS; ++k) {
            s1 += accum[(long long)k * 2 * C + ch0 + j];
            s2 += accum[(long long)k * 2 * C + C + ch0 + j];
        }
        a[j] = gs[j] * is[j];
        b[j] = (float)(s1 * inv_count);
        c[j] = (float)(s2 * inv_count);
    }
    if (gtid < C) {  // one channel per thread over the first blocks
        const int ch = (int)gtid;
        double s1 = 0.0, s2 = 0.0;
        for (int k = 0; k < NSETS; ++k) {
            s1 += accum[(long long)k * 2 * C + ch];
            s2 += accum[(long long)k * 2 * C + C + ch];
        }
        const float db = (float)s1, dg = (float)s2;
        if (dgamma) dgamma[ch] = accumulate_params ? dgamma[ch] + dg : dg;
        if (dbeta) dbeta[ch] = accumulate_params ? dbeta[ch] + db : db;
        if (ch == 0) *parity_p = parity ^ 1;
    }
    for (long long i = gtid; i < 2 * (long long)NSETS * C; i += gstride) other[i] = 0.0;
    for (int u = 0; u < vpt; ++u) {
        const long long v = tid + u * stride;
        if (v < nvec) {
            const int4* slot = stash + ((size_t)u * NT) * BN_THREADS + threadIdx.x;
            float g[VEC], xv[VEC];
            Vec<bf16>::unpack(slot[0], g);
            Vec<bf16>::unpack(slot[BN_THREADS], xv);
            if (relu) {
                if (mask_from_x) {
#pragma unroll
                    for (int j = 0; j < VEC; ++j) g[j] = ((xv[j] - mu[j]) * is[j] * gs[j] + bs[j]) > 0.f ? g[j] : 0.f;
                } else {
                    float yv[VEC];
                    Vec<bf16>::unpack(slot[2 * BN_THREADS], yv);
#pragma unroll
                    for (int j = 0; j < VEC; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
                }
            }
            if (dres != nullptr) Vec<bf16>::store(dres, v, g);
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const float xh = (xv[j] - mu[j]) * is[j];
                g[j] = a[j] * (g[j] - b[j] - xh * c[j]);
            }
            Vec<bf16>::store(dx, v, g);
        }
    }
}

// BatchNorm backward when the two reductions already exist: g is the incoming gradient with the ReLU mask applied, and
// sums = {sum g, sum g * xhat} (double[2C] + int ticket) were accumulated by the kernel that produced g (the data-gradient
// kernel of the consuming convolution: dmp_gemm_bf16_dgrad_bnbwd / dmp_dwconv3x3_dgrad_bnbwd).  One elementwise pass:
//   dx = gamma * invstd * (g - s1/M - xhat * s2/M)        (dx may alias g)
// Block 0 writes dgamma = s2, dbeta = s1 (or accumulates); the last block to have read the sums clears them (ticket).
template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_dx_sums_kernel(const T* g, const T* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ invstd,
                      const float* __restrict__ gamma, double count, T* dx, float* __restrict__ dgamma, float* __restrict__ dbeta,
                      int accumulate_params, long long nvec, int C, double* __restrict__ sums) {
    constexpr int VEC = Vec<T>::N;
    constexpr int U = 4;
    __shared__ int s_last;
    const long long tid = blockIdx.x * (long long)BN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * BN_THREADS;
    const int ch0 = (int)((tid * VEC) % C);
    const double inv_count = 1.0 / count;
    float mu[VEC], is[VEC], a[VEC], b[VEC], c[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        mu[j] = mean[ch0 + j];
        is[j] = invstd[ch0 + j];
        a[j] = (gamma ? gamma[ch0 + j] : 1.f) * is[j];
        b[j] = (float)(sums[ch0 + j] * inv_count);
        c[j] = (float)(sums[C + ch0 + j] * inv_count);
    }
    if (blockIdx.x == 0) {
        for (int ch = threadIdx.x; ch < C; ch += BN_THREADS) {
            const float db = (float)sums[ch], dg = (float)sums[C + ch];
            if (dgamma) dgamma[ch] = accumulate_params ? dgamma[ch] + dg : dg;
            if (dbeta) dbeta[ch] = accumulate_params ? dbeta[ch] + db : db;
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        int* ticket = (int*)(sums + 2 * (long long)C);
        s_last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (s_last) {
        for (int i = threadIdx.x; i < 2 * C; i += BN_THREADS) sums[i] = 0.0;
        if (threadIdx.x == 0) *(int*)(sums + 2 * (long long)C) = 0;
    }
    for (long long v0 = tid; v0 < nvec; v0 += U * stride) {
        int4 rg[U], rx[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                rg[u] = Vec<T>::load_raw(g, v0 + u * stride);
                rx[u] = Vec<T>::load_raw(x, v0 + u * stride);
            }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v0 + u * stride < nvec) {
                float gv[VEC], xv[VEC];
                Vec<T>::unpack(rg[u], gv);
                Vec<T>::unpack(rx[u], xv);
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    const float xh = (xv[j] - mu[j]) * is[j];
                    gv[j] = a[j] * (gv[j] - b[j] - xh * c[j]);
                }
                Vec<T>::store(dx, v0 + u * stride, gv);
            }
    }
}

// grid for a cooperative launch: a multiple of q (channel alignment) that fits co-resident on the device
template <typename K>
int bn_coop_grid(K kernel, long long nvec, int C, int vec, size_t smem) {
    long long per_block = (long long)BN_THREADS * vec;
    long long a = C, b = per_block;
    while (b) { long long t = a % b; a = b; b = t; }
    const int q = (int)(C / a);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, BN_THREADS, smem) != cudaSuccess || per_sm < 1) return -1;
    if (per_sm > 2) per_sm = 2;
    long long cap = (long long)dmp_num_sms() * per_sm;
    // DMP_BN_COOP_MAX_BLOCKS: a cooperative grid needs all its blocks resident at once; next to a side-stream kernel that
    // holds part of the SMs, a grid sized for the whole GPU waits at its barrier until that kernel has drained
    // (0 = none; default two blocks on each of the two thirds of the SMs the side-stream weight-gradient kernels leave free)
    static const int coop_cap = [] { const char* e = getenv("DMP_BN_COOP_MAX_BLOCKS"); return e ? atoi(e) : 2 * (dmp_num_sms() - dmp_num_sms() / 3); }();
    if (coop_cap > 0 && cap > coop_cap) cap = coop_cap;
    // 8 16-byte vectors per thread.  Measured (bench/bn_phase_bench.py): fewer, fatter blocks are slower across the board
    // (16 per thread: +2..5 us per launch); thinner ones (4) gain ~1 us on narrow layers in isolation but cost wide layers
    // 2 us (every block ends with 2*C double atomics) and the whole step came out slower (5.05 vs 4.96 ms).
    const int vecs_per_thread = 8;
    long long want = (nvec + BN_THREADS * vecs_per_thread - 1) / (BN_THREADS * vecs_per_thread);
    if (want > cap) want = cap;
    long long g = (want / q) * q;
    if (g < q) g = q;
    if (g > cap) return -1;  // cannot satisfy alignment and co-residency together: caller falls back to the split path
    return (int)g;
}

// Largest cooperative grid of `kernel` that is co-resident next to the side-stream kernels: occupancy (at most two blocks per
// SM) times the SMs left to the chain (DMP_BN_COOP_MAX_BLOCKS, see bn_coop_grid).  -1 if the kernel cannot run at all.
template <typename K>
long long bn_coop_cap(K kernel, size_t smem) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, BN_THREADS, smem) != cudaSuccess || per_sm < 1) return -1;
    if (per_sm > 2) per_sm = 2;
    long long cap = (long long)dmp_num_sms() * per_sm;
    static const int coop_cap = [] { const char* e = getenv("DMP_BN_COOP_MAX_BLOCKS"); return e ? atoi(e) : 2 * (dmp_num_sms() - dmp_num_sms() / 3); }();
    if (coop_cap > 0 && cap > coop_cap) cap = coop_cap;
    return cap;
}

int bn_grid(long long nvec, int C, int vec, int vecs_per_thread = 8) {
    // smallest multiple q of blocks such that (blocks * BN_THREADS * vec) % C == 0
    long long per_block = (long long)BN_THREADS * vec;
    long long a = C, b = per_block;
    while (b) { long long t = a % b; a = b; b = t; }
    int q = (int)(C / a);
    long long want = (nvec + BN_THREADS * vecs_per_thread - 1) / (BN_THREADS * vecs_per_thread);
    long long cap = (long long)dmp_num_sms() * 2;  // fewer, longer-lived blocks: each pays 2*C double atomics at its end
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    long long g = ((want + q - 1) / q) * q;
    return (int)g;
}

// Shared-memory-resident BatchNorm backward (bn_bwd_resident_kernel).  Returns 0 when launched, 1 when the tensor does not fit
// on chip (the caller uses the two-pass kernel), negative on a CUDA error.  DMP_BN_BWD_RESIDENT=0 disables.
int bn_coop_sms() {  // SMs a cooperative BatchNorm grid is sized for (see DMP_BN_COOP_MAX_BLOCKS in bn_coop_grid)
    static const int cap = [] { const char* e = getenv("DMP_BN_COOP_MAX_BLOCKS"); return e ? atoi(e) : 2 * (dmp_num_sms() - dmp_num_sms() / 3); }();
    const int sms = dmp_num_sms();
    return (cap > 0 && cap / 2 < sms) ? (cap / 2 > 0 ? cap / 2 : 1) : sms;
}

// Grid of a slab-mapped cooperative kernel (block_slab_reduce): cvb = largest divisor of C / vec up to 8, C / vec / cvb slabs,
// `rows_per_thread` rows per thread where the co-residency cap allows.  Returns the grid (a multiple of the slab count) or -1
// when even one block per slab does not fit the cap.
int bn_slab_grid(long long M, int C, int vec, long long cap_blocks, int rows_per_thread, int* cvb_out) {
    const int P = C / vec;
    int cvb = 8;
    while (cvb > 1 && P % cvb != 0) --cvb;
    const int cvgroups = P / cvb, rpb = BN_THREADS / cvb;
    if (cvgroups > cap_blocks) return -1;
    long long row_blocks = (M + (long long)rpb * rows_per_thread - 1) / ((long long)rpb * rows_per_thread);
    if (row_blocks * cvgroups > cap_blocks) row_blocks = cap_blocks / cvgroups;
    if (row_blocks < 1) row_blocks = 1;
    *cvb_out = cvb;
    return (int)(row_blocks * cvgroups);
}

int bn_bwd_try_resident(const void* dy, const void* y, const void* x, const float* mean, const float* invstd, const float* gamma,
                        const float* beta, double count, void* dx, void* dres, float* dgamma, float* dbeta, int accumulate_params,
                        long long nvec, int C, int relu, double* ws, cudaStream_t stream) {
    static const bool enabled = [] { const char* e = getenv("DMP_BN_BWD_RESIDENT"); return e == nullptr || e[0] != '0'; }();
    if (!enabled) return 1;
    // Measured inside a graph (bench/chain_bench.py, B200): 9.0 -> 7.0 us up to ~2 MB per tensor (8192 x 64 ... 2048 x 320),
    // no change at 6 MB (8192 x 384: 15 us either way, bound by the arithmetic of 16 resident warps per SM), slightly slower
    // for few rows of many channels (2048 x 960).  Larger tensors keep the two-pass kernel.
    if (nvec * 16 > (3ll << 20)) return 1;
    const int NT = (relu && y != nullptr) ? 3 : 2;
    const size_t scratch = (size_t)BN_THREADS * (2 * 8 + 1) * sizeof(float);
    const size_t per_vec = (size_t)NT * BN_THREADS * sizeof(int4);
    const long long M = nvec / (C / 8);
    const int sms = bn_coop_sms();
    static std::atomic<unsigned long long> configured{0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!(configured.load(std::memory_order_acquire) & (1ull << (dev & 63)))) {
        DMP_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_resident_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured.fetch_or(1ull << (dev & 63), std::memory_order_release);
    }
    // two blocks per SM with up to ~96 KB each, else one block with up to 200 KB
    for (int per_sm = 2; per_sm >= 1; --per_sm) {
        const size_t budget = per_sm == 2 ? 100 * 1024 : 200 * 1024;
        const int vpt_max = (int)((budget - scratch) / per_vec);
        int cvb = 1;
        const int g = bn_slab_grid(M, C, 8, (long long)sms * per_sm, 4, &cvb);  // 4 rows per thread where the cap allows
        if (g < 0) continue;
        const int rpb = BN_THREADS / cvb, row_blocks = g / ((C / 8) / cvb);
        const long long vpt = (M + (long long)row_blocks * rpb - 1) / ((long long)row_blocks * rpb);
        if (vpt > vpt_max) continue;
        const size_t smem = (size_t)vpt * per_vec + scratch;
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, bn_bwd_resident_kernel, BN_THREADS, smem) != cudaSuccess || occ < 1) continue;
        if ((long long)occ * dmp_num_sms() < g) continue;
        int vpt_i = (int)vpt;
        void* args[] = {(void*)&dy, (void*)&y, (void*)&x, (void*)&mean, (void*)&invstd, (void*)&gamma, (void*)&beta, (void*)&count,
                        (void*)&dx, (void*)&dres, (void*)&dgamma, (void*)&dbeta, (void*)&accumulate_params, (void*)&nvec, (void*)&C,
                        (void*)&relu, (void*)&ws, (void*)&vpt_i, (void*)&cvb};
        DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)bn_bwd_resident_kernel, dim3((unsigned)g), dim3(BN_THREADS), args, smem, stream));
        return 0;
    }
    return 1;
}

int bn_check(const char* name, long long M, int C, int dtype) {
    const int vec = dtype == 1 ? 8 : 4;
    DMP_REQUIRE(M > 0 && C > 0, "%s: empty tensor", name);
    DMP_REQUIRE(C % vec == 0, "%s: C=%d must be a multiple of %d", name, C, vec);
    DMP_REQUIRE(C <= 8192, "%s: C=%d too large", name, C);
    return 0;
}

}  // namespace

// dtype: 0 = fp32 activations, 1 = bf16 activations.  accum: double[2*C] scratch, zeroed here.
DMP_API int dmp_bn_stats(const void* x, int dtype, long long M, int C, double* accum, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_stats", M, C, dtype)) return rc;
    DMP_CHECK_CUDA(cudaMemsetAsync(accum, 0, sizeof(double) * 2 * C, stream));
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    const int grid = bn_grid(nvec, C, vec);
    const size_t smem = (size_t)BN_THREADS * 2 * vec * sizeof(float);
    if (dtype == 1) bn_stats_kernel<bf16><<<grid, BN_THREADS, smem, stream>>>((const bf16*)x, nvec, C, accum);
    else bn_stats_kernel<float><<<grid, BN_THREADS, smem, stream>>>((const float*)x, nvec, C, accum);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_bn_finalize(const double* accum, double count, const float* gamma, const float* beta, float* running_mean,
                            float* running_var, float momentum, float eps, float* mean, float* invstd, float* scale,
                            float* shift, int C, cudaStream_t stream) {
    bn_finalize_kernel<<<dmp_ceil_div(C, 128), 128, 0, stream>>>(accum, count, gamma, beta, running_mean, running_var,
                                                                 momentum, eps, mean, invstd, scale, shift, C);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_bn_apply(const void* x, int dtype, const float* scale, const float* shift, const void* residual, void* y,
                         long long M, int C, int relu, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_apply", M, C, dtype)) return rc;
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    const int grid = bn_grid(nvec, C, vec);
    if (dtype == 1)
        bn_apply_kernel<bf16><<<grid, BN_THREADS, 0, stream>>>((const bf16*)x, scale, shift, (const bf16*)residual,
                                                               (bf16*)y, nvec, C, relu);
    else
        bn_apply_kernel<float><<<grid, BN_THREADS, 0, stream>>>((const float*)x, scale, shift, (const float*)residual,
                                                                (float*)y, nvec, C, relu);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// accum: double[2*C] scratch, zeroed here; receives {sum dy_m, sum dy_m*xhat}. y may be null when relu == 0.
DMP_API int dmp_bn_bwd_reduce(const void* dy, const void* y, const void* x, int dtype, const float* mean,
                              const float* invstd, long long M, int C, int relu, double* accum, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_bwd_reduce", M, C, dtype)) return rc;
    DMP_REQUIRE(!relu || y != nullptr, "dmp_bn_bwd_reduce: relu mask needs the saved output");
    DMP_CHECK_CUDA(cudaMemsetAsync(accum, 0, sizeof(double) * 2 * C, stream));
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    const int grid = bn_grid(nvec, C, vec);
    const size_t smem = (size_t)BN_THREADS * 2 * vec * sizeof(float);
    if (dtype == 1)
        bn_bwd_reduce_kernel<bf16><<<grid, BN_THREADS, smem, stream>>>((const bf16*)dy, (const bf16*)y, (const bf16*)x,
                                                                       mean, invstd, nvec, C, relu, accum);
    else
        bn_bwd_reduce_kernel<float><<<grid, BN_THREADS, smem, stream>>>((const float*)dy, (const float*)y,
                                                                        (const float*)x, mean, invstd, nvec, C, relu,
                                                                        accum);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// accum_global: the (possibly cross-GPU reduced) sums with their total element count.
DMP_API int dmp_bn_bwd_dx(const void* dy, const void* y, const void* x, int dtype, const float* mean, const float* invstd,
                          const float* gamma, const double* accum_global, double count, void* dx, void* dres, long long M,
                          int C, int relu, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_bwd_dx", M, C, dtype)) return rc;
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    const int grid = bn_grid(nvec, C, vec);
    if (dtype == 1)
        bn_bwd_dx_kernel<bf16><<<grid, BN_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)y, (const bf16*)x, mean,
                                                                invstd, gamma, accum_global, count, (bf16*)dx, (bf16*)dres,
                                                                nvec, C, relu);
    else
        bn_bwd_dx_kernel<float><<<grid, BN_THREADS, 0, stream>>>((const float*)dy, (const float*)y, (const float*)x, mean,
                                                                 invstd, gamma, accum_global, count, (float*)dx,
                                                                 (float*)dres, nvec, C, relu);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_bn_bwd_params(const double* accum_local, float* dgamma, float* dbeta, int C, int accumulate,
                              cudaStream_t stream) {
    bn_bwd_params_kernel<<<dmp_ceil_div(C, 128), 128, 0, stream>>>(accum_local, dgamma, dbeta, C, accumulate);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}


// Workspace size (bytes) of the fused BatchNorm kernels for C channels; must be zero-initialised once.
DMP_API long long dmp_bn_fused_ws_bytes(int C) { return (long long)(4 * (long long)NSETS * C + 2) * 8; }

// One-launch train-mode BatchNorm forward: y = act((x-mean)*invstd*gamma + beta (+ residual)); also writes mean/invstd
// (saved for backward), updates running statistics and the batch counter.  Returns 1 if the shape cannot be launched
// cooperatively (caller uses dmp_bn_stats/finalize/apply), 0 on success.
DMP_API int dmp_bn_fwd_fused(const void* x, int dtype, const void* residual, void* y, long long M, int C, const float* gamma,
                             const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                             float momentum, float eps, float* mean, float* invstd, double* ws, int relu, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_fwd_fused", M, C, dtype)) return rc;
    const int vec = dtype == 1 ? 8 : 4;
    long long nvec = M * C / vec;
    size_t smem = (size_t)BN_THREADS * 2 * vec * sizeof(float);
    double count = (double)M;
    void* args[] = {(void*)&x, (void*)&residual, (void*)&y, (void*)&nvec, (void*)&C, (void*)&count, (void*)&gamma, (void*)&beta,
                    (void*)&running_mean, (void*)&running_var, (void*)&num_batches_tracked, (void*)&momentum, (void*)&eps,
                    (void*)&mean, (void*)&invstd, (void*)&ws, (void*)&relu};
    if (dtype == 1) {
        const int grid = bn_coop_grid(bn_fwd_fused_kernel<bf16>, nvec, C, vec, smem);
        if (grid < 0) return 1;
        DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)bn_fwd_fused_kernel<bf16>, dim3(grid), dim3(BN_THREADS), args, smem, stream));
    } else {
        const int grid = bn_coop_grid(bn_fwd_fused_kernel<float>, nvec, C, vec, smem);
        if (grid < 0) return 1;
        DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)bn_fwd_fused_kernel<float>, dim3(grid), dim3(BN_THREADS), args, smem, stream));
    }
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// Bytes of the {double sums[2C]; int ticket} buffer dmp_bn_apply_sums and the *_colstats producers share (zero it once).
DMP_API long long dmp_bn_sums_bytes(int C) { return (long long)(2 * (long long)C + 1) * 8; }

// Train-mode BatchNorm forward from sums produced by the convolution epilogue; see bn_apply_sums_kernel.
DMP_API int dmp_bn_apply_sums(const void* x, int dtype, const void* residual, void* y, long long M, int C, const float* gamma,
                              const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                              float momentum, float eps, float* mean, float* invstd, double* sums, int relu,
                              cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_apply_sums", M, C, dtype)) return rc;
    DMP_REQUIRE(sums && mean && invstd, "dmp_bn_apply_sums: null sums/mean/invstd");
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    // no reduction at the end of a block here, so thin blocks: 4 vectors per thread are one batch of loads in flight
    const int grid = bn_grid(nvec, C, vec, 4);
    // bit 0: ReLU; bit 3: coefficients derived once per block into shared memory (2*C floats, C <= 4096)
    const int flags = (relu ? 1 : 0) | (C <= 4096 ? 8 : 0);
    const size_t smem = C <= 4096 ? (size_t)2 * C * sizeof(float) : 0;
    if (dtype == 1)
        bn_apply_sums_kernel<bf16><<<grid, BN_THREADS, smem, stream>>>((const bf16*)x, (const bf16*)residual, (bf16*)y, nvec, C, (double)M,
                                                                       gamma, beta, running_mean, running_var, num_batches_tracked,
                                                                       momentum, eps, mean, invstd, sums, flags);
    else
        bn_apply_sums_kernel<float><<<grid, BN_THREADS, smem, stream>>>((const float*)x, (const float*)residual, (float*)y, nvec, C, (double)M,
                                                                        gamma, beta, running_mean, running_var, num_batches_tracked,
                                                                        momentum, eps, mean, invstd, sums, flags);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// BatchNorm backward from reductions produced by the kernel that wrote g (see bn_bwd_dx_sums_kernel).  dx may alias g.
DMP_API int dmp_bn_bwd_dx_sums(const void* g, const void* x, int dtype, const float* mean, const float* invstd, const float* gamma,
                               void* dx, float* dgamma, float* dbeta, int accumulate_params, long long M, int C, double* sums,
                               cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_bwd_dx_sums", M, C, dtype)) return rc;
    DMP_REQUIRE(sums && mean && invstd, "dmp_bn_bwd_dx_sums: null sums/mean/invstd");
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    const int grid = bn_grid(nvec, C, vec);
    if (dtype == 1)
        bn_bwd_dx_sums_kernel<bf16><<<grid, BN_THREADS, 0, stream>>>((const bf16*)g, (const bf16*)x, mean, invstd, gamma, (double)M,
                                                                     (bf16*)dx, dgamma, dbeta, accumulate_params, nvec, C, sums);
    else
        bn_bwd_dx_sums_kernel<float><<<grid, BN_THREADS, 0, stream>>>((const float*)g, (const float*)x, mean, invstd, gamma, (double)M,
                                                                      (float*)dx, dgamma, dbeta, accumulate_params, nvec, C, sums);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// One-launch BatchNorm backward: dx (and optional d(residual)), dgamma/dbeta (written or accumulated).
// y == null with relu != 0: the mask is recomputed from x, gamma, beta (only valid when the forward had no residual).
DMP_API int dmp_bn_bwd_fused(const void* dy, const void* y, const void* x, int dtype, const float* mean, const float* invstd,
                             const float* gamma, const float* beta, void* dx, void* dres, float* dgamma, float* dbeta,
                             int accumulate_params, long long M, int C, int relu, double* ws, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_bwd_fused", M, C, dtype)) return rc;
    const int vec = dtype == 1 ? 8 : 4;
    long long nvec = M * C / vec;
    size_t smem = (size_t)BN_THREADS * (2 * vec + 1) * sizeof(float);
    double count = (double)M;
    int cvb = 1;
    void* args[] = {(void*)&dy, (void*)&y, (void*)&x, (void*)&mean, (void*)&invstd, (void*)&gamma, (void*)&beta, (void*)&count, (void*)&dx,
                    (void*)&dres, (void*)&dgamma, (void*)&dbeta, (void*)&accumulate_params, (void*)&nvec, (void*)&C, (void*)&relu,
                    (void*)&ws, (void*)&cvb};
    if (dtype == 1) {
        if (int rc = bn_bwd_try_resident(dy, y, x, mean, invstd, gamma, beta, count, dx, dres, dgamma, dbeta, accumulate_params, nvec, C,
                                         relu, ws, stream)) {
            if (rc < 0) return rc;
        } else {
            DMP_COUNT_LAUNCH(1);
            return 0;
        }
        const int grid = bn_slab_grid(M, C, vec, bn_coop_cap(bn_bwd_fused_kernel<bf16>, smem), 8, &cvb);
        if (grid < 0) return 1;
        // the second pass walks backwards when dy + x exceed what the 126 MB L2 keeps of the first pass (DMP_BN_BWD_REVERSE_MB,
        // default 64; 0 = never)
        static const long long rev_mb = [] { const char* e = getenv("DMP_BN_BWD_REVERSE_MB"); return e ? atoll(e) : 64ll; }();
        if (rev_mb > 0 && nvec * 32 > (rev_mb << 20)) accumulate_params |= 2;
        DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)bn_bwd_fused_kernel<bf16>, dim3(grid), dim3(BN_THREADS), args, smem, stream));
    } else {
        const int grid = bn_slab_grid(M, C, vec, bn_coop_cap(bn_bwd_fused_kernel<float>, smem), 8, &cvb);
        if (grid < 0) return 1;
        DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)bn_bwd_fused_kernel<float>, dim3(grid), dim3(BN_THREADS), args, smem, stream));
    }
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// =====================================================================================================
// Allocation-free variant: batch statistics + finalize with the accumulator held inside the library.
//
// Who needs this: hosts that must not (or cannot) hand the library any extra device memory and in which
// nothing may call cudaMalloc once the process is set up -- for example a serving or training runtime that
// reserved the whole HBM for its own arena at start-up (a later cudaMalloc would simply fail), or a
// latency-critical loop run under a "no allocator calls in steady state" rule because cudaMalloc/cudaFree
// synchronise the device.  dmp_bn_stats needs a double[2*C] accumulator from its caller; such callers have
// nowhere to take one from.  Here the accumulator is a fixed array that is part of the loaded module.
//
// Cost and limits:
//   * the array is DMP_BN_NOALLOC_MAX_C * 2 doubles = 64 KiB and every GPU that loads libdmp_b200.so carries
//     it until the process exits, whether or not this entry point is ever used;
//   * C > DMP_BN_NOALLOC_MAX_C is refused with an error (use dmp_bn_stats + dmp_bn_finalize instead);
//   * there is ONE array per GPU, shared by every call: it still holds the sums of the previous call, so each
//     call clears the 2*C doubles it is about to use (on `stream`) before accumulating; two calls whose
//     kernels can overlap on the same GPU (different streams without an ordering between them, or several
//     host threads) race on it and produce garbage -- serialise them;
//   * the first call made on each device looks up the array's device address (cudaGetSymbolAddress with that
//     device current) and caches it.  Make that first call during start-up, before any stream capture begins.
// Apart from where the accumulator lives the arithmetic is that of dmp_bn_stats followed by dmp_bn_finalize.
// =====================================================================================================
#include <mutex>

#define DMP_BN_NOALLOC_MAX_C 4096

static __device__ __align__(256) double g_bn_noalloc_accum[2 * DMP_BN_NOALLOC_MAX_C];

namespace {
std::mutex g_bn_noalloc_mu;
double* g_bn_noalloc_addr[64] = {nullptr};  // per device: device address of g_bn_noalloc_accum

// In host code `g_bn_noalloc_accum` is only a handle for the symbol, never a usable pointer: the address a
// kernel may dereference has to be asked for, once per device, while that device is current.
int bn_noalloc_accum(double** out) {
    int dev = -1;
    DMP_CHECK_CUDA(cudaGetDevice(&dev));
    DMP_REQUIRE(dev >= 0 && dev < 64, "dmp_bn_stats_finalize_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_bn_noalloc_mu);
    if (g_bn_noalloc_addr[dev] == nullptr) {
        void* p = nullptr;
        DMP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bn_noalloc_accum));
        DMP_REQUIRE(p != nullptr, "dmp_bn_stats_finalize_noalloc: cudaGetSymbolAddress returned a null address");
        g_bn_noalloc_addr[dev] = (double*)p;
    }
    *out = g_bn_noalloc_addr[dev];
    return 0;
}
}  // namespace

DMP_API int dmp_bn_noalloc_max_channels() { return DMP_BN_NOALLOC_MAX_C; }

// Train-mode BatchNorm statistics of x[M, C] (dtype 0 = fp32, 1 = bf16) and everything derived from them:
// mean, invstd, fused scale/shift, and the running-stat update when running_mean/var are given.  Inputs and
// outputs are the caller's; no device memory is allocated or freed.  See the block comment above for the
// sharing / ordering rules.
DMP_API int dmp_bn_stats_finalize_noalloc(const void* x, int dtype, long long M, int C, const float* gamma, const float* beta,
                                          float* running_mean, float* running_var, float momentum, float eps, float* mean,
                                          float* invstd, float* scale, float* shift, cudaStream_t stream) {
    if (int rc = bn_check("dmp_bn_stats_finalize_noalloc", M, C, dtype)) return rc;
    DMP_REQUIRE(C <= DMP_BN_NOALLOC_MAX_C,
                "dmp_bn_stats_finalize_noalloc: C=%d exceeds the built-in accumulator (max %d channels); use dmp_bn_stats "
                "with a caller-provided accumulator instead",
                C, DMP_BN_NOALLOC_MAX_C);
    DMP_REQUIRE(x && mean && invstd && scale && shift, "dmp_bn_stats_finalize_noalloc: null input/output pointer");
    double* accum = nullptr;
    if (int rc = bn_noalloc_accum(&accum)) return rc;
    // the array still holds the previous call's sums: clear exactly the part this call accumulates into
    DMP_CHECK_CUDA(cudaMemsetAsync(accum, 0, sizeof(double) * 2 * (size_t)C, stream));
    const int vec = dtype == 1 ? 8 : 4;
    const long long nvec = M * C / vec;
    const int grid = bn_grid(nvec, C, vec);
    const size_t smem = (size_t)BN_THREADS * 2 * vec * sizeof(float);
    if (dtype == 1) bn_stats_kernel<bf16><<<grid, BN_THREADS, smem, stream>>>((const bf16*)x, nvec, C, accum);
    else bn_stats_kernel<float><<<grid, BN_THREADS, smem, stream>>>((const float*)x, nvec, C, accum);
    DMP_LAUNCH_CHECK();
    bn_finalize_kernel<<<dmp_ceil_div(C, 128), 128, 0, stream>>>(accum, (double)M, gamma, beta, running_mean, running_var,
                                                                 momentum, eps, mean, invstd, scale, shift, C);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(2);
    return 0;
}
