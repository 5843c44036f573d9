// SyncBatchNorm with the cross-GPU combine INSIDE the BatchNorm kernel: one cooperative launch per direction.
//
//   forward : local per-channel sums -> grid barrier -> block 0: flag barrier with the peers, peer-loads of every rank's partial
//             sums, total published -> grid barrier -> normalise (+residual, +ReLU), running statistics
//   backward: local {sum g, sum g*xhat} -> same combine -> dx (+dres); dgamma / dbeta from the LOCAL sums (they are averaged
//             with every other gradient by the DDP bucket all-reduce, torch/nn/modules/_functions.py:139-179)
//
// torch does this with an all_gather between three kernels forward and an all_reduce between two kernels backward
// (torch/nn/modules/_functions.py:39-122, 139-179; the reference only mentions "prepare for sync batch norm layers",
// /root/reference/Readme.md:151).  Here the ~2C doubles per rank are read straight out of the peers' memory by the same
// kernel that produced the local part, between its two passes over the activation.
//
// The partial sums live in a peer-visible workspace ws = { double acc[2][2C]; int parity } per (layer, direction); call k uses
// acc[k & 1] and, AFTER its cross-GPU barrier (every peer has therefore finished call k-1 and its reads of acc[(k-1) & 1]),
// zeroes acc[(k-1) & 1] for call k+1.  Grids are capped at one block per SM so that a spinning bucket all-reduce kernel and
// this kernel can always be co-resident (two kernels that wait for peers must never wait for each other's SM resources).
#include "comm_common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace {

constexpr int SBN_THREADS = 256;

__device__ __forceinline__ void ld8(const bf16* p, long long v, float* f) {
    int4 raw = *(const int4*)(p + v * 8);
    bf16x8_to_float(*(bf16x8*)&raw, f);
}
__device__ __forceinline__ void st8(bf16* p, long long v, const float* f) {
    bf16x8 b = float_to_bf16x8(f);
    *(int4*)(p + v * 8) = *(int4*)&b;
}

// per-thread {stat0[8], stat1[8]} of channels ch0..ch0+7 -> double atomics into accum[2C]
__device__ void sbn_block_reduce(const float* vals, int ch0, int C, double* __restrict__ accum, float* sm) {
    float* mine = sm + threadIdx.x * 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) mine[i] = vals[i];
    __shared__ int s_ch0[SBN_THREADS];
    s_ch0[threadIdx.x] = ch0;
    __syncthreads();
    const int P = C / 8;
    const int ngroups = P < SBN_THREADS ? P : SBN_THREADS;
    for (int item = threadIdx.x; item < ngroups * 16; item += SBN_THREADS) {
        const int g = item / 16, i = item % 16;
        float s = 0.f;
        for (int t = g; t < SBN_THREADS; t += P) s += sm[t * 16 + i];
        atomicAdd(&accum[(long long)(i / 8) * C + s_ch0[g] + (i % 8)], (double)s);
    }
}

// Block 0, after the grid barrier: wait for every rank's partial sums, add them up in rank order, publish the totals.
// Returns false if a peer did not show up in time (status word set).
__device__ bool sbn_combine(const PeerPtrs& ws_peers, const PeerPtrs& flags, uint32_t* epoch_mem, uint32_t* status, int rank, int W,
                            int parity, int C, double* __restrict__ total) {
    const uint32_t epoch = epoch_mem[0];
    __syncthreads();
    if (!barrier_ranks(flags, rank, W, epoch + 1, status, true)) return false;
    for (int c = threadIdx.x; c < 2 * C; c += SBN_THREADS) {
        double s = 0.0;
        for (int r = 0; r < W; ++r) s += ((const double*)ws_peers.p[r])[(long long)parity * 2 * C + c];
        total[c] = s;
    }
    if (threadIdx.x == 0) epoch_mem[0] = epoch + 1;
    __threadfence();
    return true;
}

__global__ void __launch_bounds__(SBN_THREADS)
syncbn_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ residual, bf16* __restrict__ y, long long nvec, int C,
                  double count_local, const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ running_mean,
                  float* __restrict__ running_var, long long* __restrict__ num_batches_tracked, float momentum, float eps,
                  float* __restrict__ mean_out, float* __restrict__ invstd_out, PeerPtrs ws_peers, PeerPtrs flags, uint32_t* epoch_mem,
                  uint32_t* status, int rank, int W, double* __restrict__ total, int relu) {
    extern __shared__ float sm[];
    cg::grid_group grid = cg::this_grid();
    double* ws = (double*)ws_peers.p[rank];
    int* parity_p = (int*)(ws + 4 * (long long)C);
    const int parity = *parity_p;
    double* accum = ws + (long long)parity * 2 * C;
    const long long tid = blockIdx.x * (long long)SBN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * SBN_THREADS;
    const int ch0 = (int)((tid * 8) % C);
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    for (long long v = tid; v < nvec; v += stride) {
        float f[8];
        ld8(x, v, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc[j] += f[j];
            acc[8 + j] += f[j] * f[j];
        }
    }
    sbn_block_reduce(acc, ch0, C, accum, sm);
    grid.sync();
    if (blockIdx.x == 0) sbn_combine(ws_peers, flags, epoch_mem, status, rank, W, parity, C, total);
    grid.sync();
    const double inv_count = 1.0 / (count_local * W);  // every rank holds the same number of rows (equal batches)
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = ch0 + j;
        const double m = total[c] * inv_count;
        double var = total[C + c] * inv_count - m * m;
        if (var < 0.0) var = 0.0;
        const float is = 1.0f / sqrtf((float)var + eps);
        const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
        sc[j] = g * is;
        sh[j] = b - (float)m * g * is;
    }
    if (blockIdx.x == 0) {
        const double n = count_local * W;
        for (int c = threadIdx.x; c < C; c += SBN_THREADS) {
            const double m = total[c] / n;
            double var = total[C + c] / n - m * m;
            if (var < 0.0) var = 0.0;
            mean_out[c] = (float)m;
            invstd_out[c] = (float)(1.0 / sqrt(var + (double)eps));
            if (running_mean != nullptr) {
                const double unbiased = n > 1.0 ? var * n / (n - 1.0) : var;
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
            }
        }
        if (threadIdx.x == 0) {
            if (num_batches_tracked != nullptr) *num_batches_tracked += 1;
            *parity_p = parity ^ 1;
        }
    }
    // the previous call's accumulator: every peer is past this call's barrier, so nobody reads it any more
    double* other = ws + (long long)(parity ^ 1) * 2 * C;
    for (long long i = tid; i < 2 * (long long)C; i += stride) other[i] = 0.0;
    for (long long v = tid; v < nvec; v += stride) {
        float f[8];
        ld8(x, v, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = f[j] * sc[j] + sh[j];
        if (residual != nullptr) {
            float r[8];
            ld8(residual, v, r);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] += r[j];
        }
        if (relu) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        st8(y, v, f);
    }
}

__global__ void __launch_bounds__(SBN_THREADS)
syncbn_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ y, const bf16* __restrict__ x, const float* __restrict__ mean,
                  const float* __restrict__ invstd, const float* __restrict__ gamma, double count_local, bf16* __restrict__ dx,
                  bf16* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta, int accumulate_params, long long nvec,
                  int C, int relu, PeerPtrs ws_peers, PeerPtrs flags, uint32_t* epoch_mem, uint32_t* status, int rank, int W,
                  double* __restrict__ total) {
    extern __shared__ float sm[];
    cg::grid_group grid = cg::this_grid();
    double* ws = (double*)ws_peers.p[rank];
    int* parity_p = (int*)(ws + 4 * (long long)C);
    const int parity = *parity_p;
    double* accum = ws + (long long)parity * 2 * C;
    const long long tid = blockIdx.x * (long long)SBN_THREADS + threadIdx.x;
    const long long stride = (long long)gridDim.x * SBN_THREADS;
    const int ch0 = (int)((tid * 8) % C);
    float mu[8], is[8], gs[8], acc[16];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mu[j] = mean[ch0 + j];
        is[j] = invstd[ch0 + j];
        gs[j] = gamma ? gamma[ch0 + j] : 1.f;
        acc[j] = acc[8 + j] = 0.f;
    }
    for (long long v = tid; v < nvec; v += stride) {
        float g[8], xv[8];
        ld8(dy, v, g);
        ld8(x, v, xv);
        if (relu) {
            float yv[8];
            ld8(y, v, yv);
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc[j] += g[j];
            acc[8 + j] += g[j] * (xv[j] - mu[j]) * is[j];
        }
    }
    sbn_block_reduce(acc, ch0, C, accum, sm);
    grid.sync();
    if (blockIdx.x == 0) {
        // parameter gradients come from the LOCAL sums
        for (int ch = threadIdx.x; ch < C; ch += SBN_THREADS) {
            const float db = (float)accum[ch], dg = (float)accum[C + ch];
            if (dgamma) dgamma[ch] = accumulate_params ? dgamma[ch] + dg : dg;
            if (dbeta) dbeta[ch] = accumulate_params ? dbeta[ch] + db : db;
        }
        sbn_combine(ws_peers, flags, epoch_mem, status, rank, W, parity, C, total);
        if (threadIdx.x == 0) *parity_p = parity ^ 1;
    }
    grid.sync();
    const double inv_count = 1.0 / (count_local * W);
    float a[8], b[8], c[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        a[j] = gs[j] * is[j];
        b[j] = (float)(total[ch0 + j] * inv_count);
        c[j] = (float)(total[C + ch0 + j] * inv_count);
    }
    double* other = ws + (long long)(parity ^ 1) * 2 * C;
    for (long long i = tid; i < 2 * (long long)C; i += stride) other[i] = 0.0;
    for (long long v = tid; v < nvec; v += stride) {
        float g[8], xv[8];
        ld8(dy, v, g);
        ld8(x, v, xv);
        if (relu) {
            float yv[8];
            ld8(y, v, yv);
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        }
        if (dres != nullptr) st8(dres, v, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float xh = (xv[j] - mu[j]) * is[j];
            g[j] = a[j] * (g[j] - b[j] - xh * c[j]);
        }
        st8(dx, v, g);
    }
}

// cooperative grid: a multiple of q blocks (so that every thread keeps its channels), at most ONE block per SM
template <typename K>
int sbn_grid(K kernel, long long nvec, int C, size_t smem) {
    long long a = C, b = (long long)SBN_THREADS * 8;
    while (b) { long long t = a % b; a = b; b = t; }
    const int q = (int)(C / a);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, SBN_THREADS, smem) != cudaSuccess || per_sm < 1) return -1;
    const long long cap = dmp_num_sms();
    long long want = (nvec + SBN_THREADS * 8 - 1) / (SBN_THREADS * 8);
    if (want > cap) want = cap;
    long long g = (want / q) * q;
    if (g < q) g = q;
    if (g > cap) return -1;
    return (int)g;
}

PeerPtrs make_peers(void* const* ptrs, int W) {
    PeerPtrs p;
    for (int i = 0; i < DMP_MAX_RANKS; ++i) p.p[i] = i < W ? ptrs[i] : nullptr;
    return p;
}

}  // namespace

// Bytes of the peer-visible workspace of one (layer, direction): double acc[2][2C] + parity word.  Zero it once.
DMP_API long long dmp_syncbn_ws_bytes(int C) { return (long long)(4 * (long long)C + 2) * 8; }

// ws_ptrs[W]: every rank's workspace (peer-mapped); state_ptrs[W]: every rank's communicator state (flag pads first, see
// dmp_comm_state_bytes); total: local double[2C] scratch that receives the cross-rank sums.  count_local = rows on this rank
// (equal on every rank).  Returns 1 if the shape cannot be launched cooperatively (caller uses the split path).
DMP_API int dmp_syncbn_fwd(const void* x, const void* residual, void* y, long long M, int C, const float* gamma, const float* beta,
                           float* running_mean, float* running_var, long long* num_batches_tracked, float momentum, float eps,
                           float* mean, float* invstd, void* const* ws_ptrs, void* const* state_ptrs, int rank, int W, double* total,
                           int relu, cudaStream_t stream) {
    DMP_REQUIRE(M > 0 && C > 0 && C % 8 == 0 && W >= 1 && W <= DMP_MAX_RANKS, "dmp_syncbn_fwd: bad arguments (C=%d W=%d)", C, W);
    long long nvec = M * C / 8;
    const size_t smem = (size_t)SBN_THREADS * 16 * sizeof(float);
    const int grid = sbn_grid(syncbn_fwd_kernel, nvec, C, smem);
    if (grid < 0) return 1;
    PeerPtrs wsp = make_peers(ws_ptrs, W), flags = make_peers(state_ptrs, W);
    uint32_t* epoch_mem = (uint32_t*)state_ptrs[rank] + FLAGS_WORDS;
    uint32_t* status = epoch_mem + DMP_COMM_MAX_BLOCKS;
    double count = (double)M;
    void* args[] = {(void*)&x, (void*)&residual, (void*)&y, (void*)&nvec, (void*)&C, (void*)&count, (void*)&gamma, (void*)&beta,
                    (void*)&running_mean, (void*)&running_var, (void*)&num_batches_tracked, (void*)&momentum, (void*)&eps, (void*)&mean,
                    (void*)&invstd, (void*)&wsp, (void*)&flags, (void*)&epoch_mem, (void*)&status, (void*)&rank, (void*)&W,
                    (void*)&total, (void*)&relu};
    DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)syncbn_fwd_kernel, dim3(grid), dim3(SBN_THREADS), args, smem, stream));
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// y (the forward output) is only read when relu != 0.  dgamma / dbeta receive the LOCAL sums (written or accumulated).
DMP_API int dmp_syncbn_bwd(const void* dy, const void* y, const void* x, const float* mean, const float* invstd, const float* gamma,
                           void* dx, void* dres, float* dgamma, float* dbeta, int accumulate_params, long long M, int C, int relu,
                           void* const* ws_ptrs, void* const* state_ptrs, int rank, int W, double* total, cudaStream_t stream) {
    DMP_REQUIRE(M > 0 && C > 0 && C % 8 == 0 && W >= 1 && W <= DMP_MAX_RANKS, "dmp_syncbn_bwd: bad arguments (C=%d W=%d)", C, W);
    DMP_REQUIRE(!relu || y != nullptr, "dmp_syncbn_bwd: the ReLU mask needs the forward output");
    long long nvec = M * C / 8;
    const size_t smem = (size_t)SBN_THREADS * 16 * sizeof(float);
    const int grid = sbn_grid(syncbn_bwd_kernel, nvec, C, smem);
    if (grid < 0) return 1;
    PeerPtrs wsp = make_peers(ws_ptrs, W), flags = make_peers(state_ptrs, W);
    uint32_t* epoch_mem = (uint32_t*)state_ptrs[rank] + FLAGS_WORDS;
    uint32_t* status = epoch_mem + DMP_COMM_MAX_BLOCKS;
    double count = (double)M;
    void* args[] = {(void*)&dy, (void*)&y, (void*)&x, (void*)&mean, (void*)&invstd, (void*)&gamma, (void*)&count, (void*)&dx, (void*)&dres,
                    (void*)&dgamma, (void*)&dbeta, (void*)&accumulate_params, (void*)&nvec, (void*)&C, (void*)&relu, (void*)&wsp,
                    (void*)&flags, (void*)&epoch_mem, (void*)&status, (void*)&rank, (void*)&W, (void*)&total};
    DMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)syncbn_bwd_kernel, dim3(grid), dim3(SBN_THREADS), args, smem, stream));
    DMP_COUNT_LAUNCH(1);
    return 0;
}
