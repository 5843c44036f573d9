// Communication kernels over NVLink / NVSwitch peer memory (sm_100a).
//
// Every kernel here moves data with ordinary ld/st on peer-mapped pointers (cudaIpc or
// cudaDeviceEnablePeerAccess mappings) or multimem.* on an NVLS multicast mapping -- there is no NCCL
// call on these paths.  They implement the three communication-bound paths of the north star
// (SURVEY.md section 2.6):
//   E3  DDP bucket all-reduce, one-shot or two-shot, fused with 1/W scale, bf16->fp32 cast and the
//       SGD-momentum update                                   (torch Reducer + default_hooks.py:26-32)
//   D2  DataParallel broadcast_coalesced of params/buffers    (torch/nn/parallel/comm.py:50-67)
//   D4/D6 DataParallel gather and ReduceAddCoalesced          (torch/nn/parallel/_functions.py:47-123)
// plus the small fp64 all-reduce SyncBatchNorm needs between its statistics and apply kernels (E5).
//
// Cross-GPU synchronisation: each rank owns a flag pad `flags[block][src_rank]` in peer-visible memory.
// A block-level barrier writes a monotonically increasing epoch into every peer's pad with
// st.release.sys and spins with ld.acquire.sys on its own pad.  The epoch lives in device memory and is
// advanced by the kernel itself, so a captured CUDA graph can be replayed.  Spins are bounded: on timeout
// the kernel records an error in `status` and returns instead of hanging the GPU.
#include "../common.cuh"
#include <type_traits>
#include <stdlib.h>

#include "comm_common.cuh"

namespace {

template <typename T>
struct VecIO;  // 16-byte vector load -> floats (or doubles)
template <>
struct VecIO<float> {
    static constexpr int N = 4;
    typedef float acc_t;
    __device__ static void load(const void* base, long long v, acc_t* o) {
        int4 r = *(const int4*)((const char*)base + v * 16);  // plain ld.global: coherent for peer memory
        o[0] = __int_as_float(r.x); o[1] = __int_as_float(r.y); o[2] = __int_as_float(r.z); o[3] = __int_as_float(r.w);
    }
    __device__ static void store(void* base, long long v, const acc_t* o) {
        *(float4*)((char*)base + v * 16) = make_float4(o[0], o[1], o[2], o[3]);
    }
};
template <>
struct VecIO<bf16> {
    static constexpr int N = 8;
    typedef float acc_t;
    __device__ static void load(const void* base, long long v, acc_t* o) {
        int4 r = *(const int4*)((const char*)base + v * 16);
        bf16x8_to_float(*(bf16x8*)&r, o);
    }
    __device__ static void store(void* base, long long v, const acc_t* o) {
        bf16x8 b = float_to_bf16x8(o);
        *(int4*)((char*)base + v * 16) = *(int4*)&b;
    }
};
template <>
struct VecIO<double> {
    static constexpr int N = 2;
    typedef double acc_t;
    __device__ static void load(const void* base, long long v, acc_t* o) {
        double2 r = *(const double2*)((const char*)base + v * 16);
        o[0] = r.x; o[1] = r.y;
    }
    __device__ static void store(void* base, long long v, const acc_t* o) {
        *(double2*)((char*)base + v * 16) = make_double2(o[0], o[1]);
    }
};

// Sum of vector v over all ranks in rank order (identical order everywhere -> bitwise identical results).
template <typename T>
__device__ __forceinline__ void reduce_vec(const PeerPtrs& src, int W, long long v, typename VecIO<T>::acc_t* acc) {
    constexpr int N = VecIO<T>::N;
    typename VecIO<T>::acc_t tmp[DMP_MAX_RANKS > 8 ? 8 : DMP_MAX_RANKS][N];
    // issue the loads of up to 8 ranks before consuming them (NVLink latency ~2 us: keep requests in flight)
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = 0;
    for (int r0 = 0; r0 < W; r0 += 8) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (r0 + r < W) VecIO<T>::load(src.p[r0 + r], v, tmp[r]);
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (r0 + r < W) {
#pragma unroll
                for (int j = 0; j < N; ++j) acc[j] += tmp[r][j];
            }
    }
}

struct SgdArgs {
    float* master;       // fp32 params (local)            -- null => no fused optimizer
    float* momentum;     // fp32 momentum (local)
    bf16* shadow;        // bf16 copy of params (local), nullable
    const float* hp;     // device {lr, momentum, weight_decay, grad_scale(unused here)}
    int nesterov;
};

template <int N>
__device__ __forceinline__ void sgd_update_vec(const SgdArgs& s, long long elem0, const float* g, float* p_out) {
    const float lr = s.hp[0], mu = s.hp[1], wd = s.hp[2];
    float p[N], m[N];
#pragma unroll
    for (int j = 0; j < N; j += 4) {
        float4 pv = *(const float4*)(s.master + elem0 + j);
        float4 mv = *(const float4*)(s.momentum + elem0 + j);
        p[j] = pv.x; p[j + 1] = pv.y; p[j + 2] = pv.z; p[j + 3] = pv.w;
        m[j] = mv.x; m[j + 1] = mv.y; m[j + 2] = mv.z; m[j + 3] = mv.w;
    }
#pragma unroll
    for (int j = 0; j < N; ++j) {
        const float d = g[j] + wd * p[j];
        m[j] = mu * m[j] + d;
        p[j] -= lr * (s.nesterov ? (d + mu * m[j]) : m[j]);
        p_out[j] = p[j];
    }
#pragma unroll
    for (int j = 0; j < N; j += 4) {
        *(float4*)(s.master + elem0 + j) = make_float4(p[j], p[j + 1], p[j + 2], p[j + 3]);
        *(float4*)(s.momentum + elem0 + j) = make_float4(m[j], m[j + 1], m[j + 2], m[j + 3]);
    }
    if (s.shadow != nullptr) {
#pragma unroll
        for (int j = 0; j < N; j += 2) *(bf162*)(s.shadow + elem0 + j) = __floats2bfloat162_rn(p[j], p[j + 1]);
    }
}

// ------------------------------------------------------------------------------------------------
// one-shot all-reduce: every rank reads all W copies of the whole buffer.
//   mode 0: averaged gradient -> `out` (local), then copied back over the local input after the 2nd barrier
//   mode 1: fused SGD on local master/momentum/shadow; the local gradient buffer is zeroed afterwards
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(DMP_COMM_THREADS)
allreduce_oneshot_kernel(PeerPtrs bufs, PeerPtrs flags, uint32_t* epoch_mem, uint32_t* status, int rank, int W,
                         long long nvec, float scale, void* out, SgdArgs sgd, int zero_after) {
    constexpr int N = VecIO<T>::N;
    typedef typename VecIO<T>::acc_t acc_t;
    uint32_t epoch = epoch_mem[blockIdx.x];
    __syncthreads();
    // inputs were written by earlier kernels of this stream: the kernel boundary already published them
    if (!barrier_ranks(flags, rank, W, epoch + 1, status, false)) return;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long first = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    for (long long v = first; v < nvec; v += stride) {
        acc_t acc[N];
        reduce_vec<T>(bufs, W, v, acc);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] *= (acc_t)scale;
        if constexpr (!std::is_same<T, double>::value) {
            if (sgd.master != nullptr) {
                float pnew[N];
                sgd_update_vec<N>(sgd, v * N, acc, pnew);
                continue;
            }
        }
        VecIO<T>::store(out, v, acc);
    }
    if (!barrier_ranks(flags, rank, W, epoch + 2, status, false)) return;  // "done reading": nothing to publish
    // every peer finished reading my buffer: now it may be overwritten
    void* mine = bufs.p[rank];
    if (sgd.master == nullptr) {
        if (out != mine)
            for (long long v = first; v < nvec; v += stride) ((int4*)mine)[v] = ((const int4*)out)[v];
    } else if (zero_after) {
        for (long long v = first; v < nvec; v += stride) ((int4*)mine)[v] = make_int4(0, 0, 0, 0);
    }
    if (threadIdx.x == 0) epoch_mem[blockIdx.x] = epoch + 2;
}

// ------------------------------------------------------------------------------------------------
// two-shot all-reduce: rank r reduces slice r (reduce-scatter by peer loads), applies the update, and writes
// the result of slice r into every peer (all-gather by peer stores).  In place and race free because slice r
// of every buffer is only ever read and written by rank r.
//   mode 0: averaged gradient written to all peers' gradient buffers
//   mode 1: fused SGD on slice r of master/momentum; new fp32 params (and bf16 shadow) written to all peers'
//           param buffers (`params` / `shadows` are peer pointer tables); gradient slices zeroed on all peers
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(DMP_COMM_THREADS)
allreduce_twoshot_kernel(PeerPtrs bufs, PeerPtrs flags, uint32_t* epoch_mem, uint32_t* status, int rank, int W,
                         long long nvec, float scale, SgdArgs sgd, PeerPtrs params, PeerPtrs shadows, int zero_after) {
    constexpr int N = VecIO<T>::N;
    uint32_t epoch = epoch_mem[blockIdx.x];
    __syncthreads();
    if (!barrier_ranks(flags, rank, W, epoch + 1, status, false)) return;
    const long long per = (nvec + W - 1) / W;
    const long long lo = per * rank;
    long long hi = lo + per;
    if (hi > nvec) hi = nvec;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; v < hi; v += stride) {
        float acc[N];
        reduce_vec<T>(bufs, W, v, acc);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] *= scale;
        if (sgd.master != nullptr) {
            float pnew[N];
            sgd_update_vec<N>(sgd, v * N, acc, pnew);
            for (int r = 0; r < W; ++r) {
                if (r != rank) {
#pragma unroll
                    for (int j = 0; j < N; j += 4)
                        *(float4*)((float*)params.p[r] + v * N + j) = make_float4(pnew[j], pnew[j + 1], pnew[j + 2], pnew[j + 3]);
                    if (sgd.shadow != nullptr) {
#pragma unroll
                        for (int j = 0; j < N; j += 2)
                            *(bf162*)((bf16*)shadows.p[r] + v * N + j) = __floats2bfloat162_rn(pnew[j], pnew[j + 1]);
                    }
                }
                if (zero_after) ((int4*)bufs.p[r])[v] = make_int4(0, 0, 0, 0);
            }
        } else {
            for (int r = 0; r < W; ++r) VecIO<T>::store(bufs.p[r], v, acc);
        }
    }
    if (!barrier_ranks(flags, rank, W, epoch + 2, status)) return;
    if (threadIdx.x == 0) epoch_mem[blockIdx.x] = epoch + 2;
}

// ------------------------------------------------------------------------------------------------
// NVLS all-reduce (two-shot through the switch): `mc` is the multicast address of a buffer every rank has bound its copy to
// (csrc/comm/multicast.cu).  Rank r pulls slice r with multimem.ld_reduce -- the NVSwitch reads all W copies and returns
// their sum, so each element crosses this GPU's link once instead of W-1 times -- scales it, and pushes it back with
// multimem.st, which the switch replicates into every copy.  In place; slices are disjoint, so there is no hazard between
// the phases of different ranks, and all ranks end up with bitwise identical values (one sum, broadcast).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DMP_COMM_THREADS)
allreduce_nvls_kernel(float* mc, PeerPtrs flags, uint32_t* epoch_mem, uint32_t* status, int rank, int W, long long nvec,
                      float scale) {
    uint32_t epoch = epoch_mem[blockIdx.x];
    __syncthreads();
    if (!barrier_ranks(flags, rank, W, epoch + 1, status, false)) return;  // every rank's input is in place
    const long long per = (nvec + W - 1) / W;
    const long long lo = per * rank;
    long long hi = lo + per;
    if (hi > nvec) hi = nvec;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; v < hi; v += stride) {
        float a, b, c, d;
        float4* p = (float4*)mc + v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(a), "=f"(b), "=f"(c), "=f"(d)
                     : "l"(p)
                     : "memory");
        a *= scale; b *= scale; c *= scale; d *= scale;
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                     : "memory");
    }
    if (!barrier_ranks(flags, rank, W, epoch + 2, status)) return;  // system-scope fence inside: the multicast stores are visible
    if (threadIdx.x == 0) epoch_mem[blockIdx.x] = epoch + 2;
}

// ------------------------------------------------------------------------------------------------
// NVLS all-reduce FUSED with SGD: gradients, fp32 masters and bf16 shadows of a bucket all live in multicast-bound buffers.
// Rank r owns slice r: multimem.ld_reduce pulls the switch-side SUM of the W gradient copies (S/W bytes cross this GPU's link
// instead of S(W-1)/W), the momentum update runs on the local master / momentum, and three multimem.st push the new fp32
// parameters, their bf16 shadow and zeros for the gradient slice into EVERY copy (the switch replicates: 10 bytes per element
// leave this GPU once instead of 6 bytes to each of W-1 peers).  Per GPU and element that is (4 + 10)/W bytes on the wire
// against 10 (W-1)/W for the peer-load/peer-store two-shot kernel above.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DMP_COMM_THREADS)
allreduce_nvls_sgd_kernel(float* mc_grad, float* mc_master, bf16* mc_shadow, const float* __restrict__ master,
                          float* __restrict__ momentum, const float* __restrict__ hp, int nesterov, PeerPtrs flags,
                          uint32_t* epoch_mem, uint32_t* status, int rank, int W, long long n8, float scale) {
    uint32_t epoch = epoch_mem[blockIdx.x];
    __syncthreads();
    if (!barrier_ranks(flags, rank, W, epoch + 1, status, false)) return;  // every rank's gradients are in place
    const long long per = (n8 + W - 1) / W;
    const long long lo = per * rank;
    long long hi = lo + per;
    if (hi > n8) hi = n8;
    const float lr = hp[0], mu = hp[1], wd = hp[2];
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; v < hi; v += stride) {
        float g[8], p[8], m[8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(g[4 * h]), "=f"(g[4 * h + 1]), "=f"(g[4 * h + 2]), "=f"(g[4 * h + 3])
                         : "l"((float4*)mc_grad + 2 * v + h)
                         : "memory");
            const float4 pv = *((const float4*)master + 2 * v + h);
            const float4 mv = *((const float4*)momentum + 2 * v + h);
            p[4 * h] = pv.x; p[4 * h + 1] = pv.y; p[4 * h + 2] = pv.z; p[4 * h + 3] = pv.w;
            m[4 * h] = mv.x; m[4 * h + 1] = mv.y; m[4 * h + 2] = mv.z; m[4 * h + 3] = mv.w;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float d = g[j] * scale + wd * p[j];
            m[j] = mu * m[j] + d;
            p[j] -= lr * (nesterov ? (d + mu * m[j]) : m[j]);
        }
        uint32_t sh[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            bf162 b = __floats2bfloat162_rn(p[2 * j], p[2 * j + 1]);
            sh[j] = *(uint32_t*)&b;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            *((float4*)momentum + 2 * v + h) = make_float4(m[4 * h], m[4 * h + 1], m[4 * h + 2], m[4 * h + 3]);
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"((float4*)mc_master + 2 * v + h),
                         "f"(p[4 * h]), "f"(p[4 * h + 1]), "f"(p[4 * h + 2]), "f"(p[4 * h + 3])
                         : "memory");
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%1,%1,%1};" ::"l"((float4*)mc_grad + 2 * v + h), "f"(0.f)
                         : "memory");
        }
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"((float4*)mc_shadow + v),
                     "f"(__uint_as_float(sh[0])), "f"(__uint_as_float(sh[1])), "f"(__uint_as_float(sh[2])),
                     "f"(__uint_as_float(sh[3]))
                     : "memory");
    }
    if (!barrier_ranks(flags, rank, W, epoch + 2, status)) return;  // system-scope fence inside: the multicast stores are visible
    if (threadIdx.x == 0) epoch_mem[blockIdx.x] = epoch + 2;
}

// ------------------------------------------------------------------------------------------------
// broadcast: the source rank stores its buffer into every destination (peer stores, or one multimem.st
// through an NVLS multicast mapping).  Optional fp32 -> bf16 cast into a second set of destinations.
// With flags == null pointers (single-process DataParallel) ordering comes from stream events instead.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DMP_COMM_THREADS)
broadcast_kernel(const void* src, PeerPtrs dst, int n_dst, void* mc_dst, PeerPtrs dst_bf16, long long n16,
                 PeerPtrs flags, uint32_t* epoch_mem, uint32_t* status, int rank, int W, int is_src) {
    uint32_t epoch = 0;
    const bool synced = epoch_mem != nullptr;
    if (synced) {
        epoch = epoch_mem[blockIdx.x];
        __syncthreads();
        if (!barrier_ranks(flags, rank, W, epoch + 1, status)) return;  // destinations are ready to be overwritten
    }
    if (is_src) {
        const long long stride = (long long)gridDim.x * blockDim.x;
        for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < n16; v += stride) {
            const int4 val = ((const int4*)src)[v];
            if (mc_dst != nullptr) {
                asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"((int4*)mc_dst + v),
                             "f"(__int_as_float(val.x)), "f"(__int_as_float(val.y)), "f"(__int_as_float(val.z)),
                             "f"(__int_as_float(val.w))
                             : "memory");
            } else {
                for (int d = 0; d < n_dst; ++d) ((int4*)dst.p[d])[v] = val;
            }
            if (dst_bf16.p[0] != nullptr) {
                bf162 lo = __floats2bfloat162_rn(__int_as_float(val.x), __int_as_float(val.y));
                bf162 hi = __floats2bfloat162_rn(__int_as_float(val.z), __int_as_float(val.w));
                uint2 o;
                o.x = *(uint32_t*)&lo;
                o.y = *(uint32_t*)&hi;
                for (int d = 0; d < n_dst; ++d)
                    if (dst_bf16.p[d] != nullptr) ((uint2*)dst_bf16.p[d])[v] = o;
            }
        }
    }
    if (synced) {
        if (!barrier_ranks(flags, rank, W, epoch + 2, status)) return;  // data landed everywhere
        if (threadIdx.x == 0) epoch_mem[blockIdx.x] = epoch + 2;
    }
}

// ------------------------------------------------------------------------------------------------
// gather: the output device pulls each replica's chunk with peer loads into one concatenated buffer.
// ------------------------------------------------------------------------------------------------
struct GatherArgs {
    const void* src[DMP_MAX_RANKS];
    long long off16[DMP_MAX_RANKS + 1];  // prefix offsets in 16-byte units
};
__global__ void __launch_bounds__(DMP_COMM_THREADS) gather_peer_kernel(GatherArgs a, int n_src, void* dst) {
    const long long total = a.off16[n_src];
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total; v += stride) {
        int s = 0;
        while (s + 1 < n_src && v >= a.off16[s + 1]) ++s;
        ((int4*)dst)[v] = ((const int4*)a.src[s])[v - a.off16[s]];
    }
}
// scatter (inverse): each destination device pulls its slice; run on the destination with a plain copy kernel
__global__ void __launch_bounds__(DMP_COMM_THREADS) copy16_kernel(const void* src, void* dst, long long n16) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < n16; v += stride)
        ((int4*)dst)[v] = ((const int4*)src)[v];
}

// ------------------------------------------------------------------------------------------------
// reduce-add: the destination device sums W flat buffers (its own + peers, peer loads), scales, and either
// writes the fp32 result or applies the fused SGD update (DataParallel's ReduceAddCoalesced + optimizer).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(DMP_COMM_THREADS)
reduce_add_peer_kernel(PeerPtrs src, int W, long long nvec, float scale, float* out_f32, SgdArgs sgd, int zero_src) {
    constexpr int N = VecIO<T>::N;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += stride) {
        float acc[N];
        reduce_vec<T>(src, W, v, acc);
        if (zero_src) {  // the replicas' gradient arenas start the next iteration from zero (peer stores)
            for (int r = 0; r < W; ++r) ((int4*)src.p[r])[v] = make_int4(0, 0, 0, 0);
        }
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] *= scale;
        if (sgd.master != nullptr) {
            float pnew[N];
            sgd_update_vec<N>(sgd, v * N, acc, pnew);
        } else {
#pragma unroll
            for (int j = 0; j < N; j += 4) *(float4*)(out_f32 + v * N + j) = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
        }
    }
}

int comm_grid(long long nvec, int max_blocks) {
    long long want = (nvec + DMP_COMM_THREADS * 4 - 1) / (DMP_COMM_THREADS * 4);
    if (max_blocks <= 0 || max_blocks > DMP_COMM_MAX_BLOCKS) max_blocks = DMP_COMM_MAX_BLOCKS;
    if (want > max_blocks) want = max_blocks;
    if (want < 1) want = 1;
    return (int)want;
}

PeerPtrs make_ptrs(void* const* p, int n) {
    PeerPtrs r;
    memset(&r, 0, sizeof(r));
    for (int i = 0; i < n && i < DMP_MAX_RANKS; ++i) r.p[i] = p ? p[i] : nullptr;
    return r;
}

}  // namespace

// Size of the per-rank communicator state; layout: flags | epoch | status.  Must be zero-initialised once.
DMP_API long long dmp_comm_state_bytes() { return (long long)(FLAGS_WORDS + DMP_COMM_MAX_BLOCKS + 4) * 4; }
DMP_API int dmp_comm_max_ranks() { return DMP_MAX_RANKS; }
DMP_API int dmp_comm_max_blocks() { return DMP_COMM_MAX_BLOCKS; }

// dtype: 0 fp32, 1 bf16, 2 fp64.  algo: 0 one-shot, 1 two-shot.
// bufs[W]: peer pointers of the symmetric gradient buffer (bufs[rank] is local).  state[W]: peer pointers of
// the communicator state (flags first).  n: element count; must be padded so that n*sizeof(T) % (16*W) == 0.
// Fused SGD (master != null): master/momentum/shadow are local arrays laid out like the bucket; for the
// two-shot algorithm param_peers/shadow_peers[W] are the peer pointers of master/shadow.
DMP_API int dmp_allreduce_fused(int algo, int dtype, void* const* bufs, void* const* state, int rank, int W, long long n,
                                float scale, void* out, float* master, float* momentum, void* shadow, const float* hp_dev,
                                int nesterov, void* const* param_peers, void* const* shadow_peers, int zero_after,
                                int max_blocks, cudaStream_t stream) {
    DMP_REQUIRE(W >= 1 && W <= DMP_MAX_RANKS, "dmp_allreduce_fused: world size %d unsupported", W);
    DMP_REQUIRE(rank >= 0 && rank < W, "dmp_allreduce_fused: bad rank");
    DMP_REQUIRE(dtype >= 0 && dtype <= 2, "dmp_allreduce_fused: bad dtype");
    const int esz = dtype == 0 ? 4 : (dtype == 1 ? 2 : 8);
    DMP_REQUIRE((n * esz) % 16 == 0, "dmp_allreduce_fused: buffer must be a multiple of 16 bytes");
    DMP_REQUIRE(!(master != nullptr && dtype == 2), "dmp_allreduce_fused: fused SGD needs fp32/bf16 gradients");
    const long long nvec = n * esz / 16;
    if (nvec == 0) return 0;
    PeerPtrs b = make_ptrs(bufs, W), f = make_ptrs(state, W);
    uint32_t* st = (uint32_t*)state[rank];
    uint32_t* epoch_mem = st + FLAGS_WORDS;
    uint32_t* status = epoch_mem + DMP_COMM_MAX_BLOCKS;
    SgdArgs s;
    s.master = master; s.momentum = momentum; s.shadow = (bf16*)shadow; s.hp = hp_dev; s.nesterov = nesterov;
    DMP_REQUIRE(master == nullptr || (momentum != nullptr && hp_dev != nullptr), "dmp_allreduce_fused: fused SGD needs momentum and hp");
    if (algo == 0) {
        const int grid = comm_grid(nvec, max_blocks);
        if (master == nullptr) DMP_REQUIRE(out != nullptr, "dmp_allreduce_fused: one-shot without fused SGD needs an out buffer");
        if (dtype == 0) allreduce_oneshot_kernel<float><<<grid, DMP_COMM_THREADS, 0, stream>>>(b, f, epoch_mem, status, rank, W, nvec, scale, out, s, zero_after);
        else if (dtype == 1) allreduce_oneshot_kernel<bf16><<<grid, DMP_COMM_THREADS, 0, stream>>>(b, f, epoch_mem, status, rank, W, nvec, scale, out, s, zero_after);
        else allreduce_oneshot_kernel<double><<<grid, DMP_COMM_THREADS, 0, stream>>>(b, f, epoch_mem, status, rank, W, nvec, scale, out, s, zero_after);
    } else {
        DMP_REQUIRE(dtype != 2, "dmp_allreduce_fused: two-shot supports fp32/bf16");
        DMP_REQUIRE(nvec % W == 0, "dmp_allreduce_fused: two-shot needs the buffer padded to 16*W bytes");
        const int grid = comm_grid(nvec / W, max_blocks);
        PeerPtrs pp = make_ptrs(param_peers, W), sp = make_ptrs(shadow_peers, W);
        if (master != nullptr) DMP_REQUIRE(param_peers != nullptr, "dmp_allreduce_fused: two-shot fused SGD needs param peer pointers");
        if (dtype == 0) allreduce_twoshot_kernel<float><<<grid, DMP_COMM_THREADS, 0, stream>>>(b, f, epoch_mem, status, rank, W, nvec, scale, s, pp, sp, zero_after);
        else allreduce_twoshot_kernel<bf16><<<grid, DMP_COMM_THREADS, 0, stream>>>(b, f, epoch_mem, status, rank, W, nvec, scale, s, pp, sp, zero_after);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// In-place fp32 all-reduce (sum * scale) of the n floats behind the multicast address `mc` (see allreduce_nvls_kernel).
// state[W]: peer pointers of a communicator state, as for dmp_allreduce_fused.  n*4 must be a multiple of 16*W bytes.
DMP_API int dmp_allreduce_nvls(void* mc, void* const* state, int rank, int W, long long n, float scale, int max_blocks,
                               cudaStream_t stream) {
    DMP_REQUIRE(mc != nullptr && W >= 2 && W <= DMP_MAX_RANKS && rank >= 0 && rank < W, "dmp_allreduce_nvls: bad arguments");
    DMP_REQUIRE((n * 4) % (16 * (long long)W) == 0, "dmp_allreduce_nvls: buffer must be padded to 16*W bytes");
    const long long nvec = n / 4;
    if (nvec == 0) return 0;
    PeerPtrs f = make_ptrs(state, W);
    uint32_t* st = (uint32_t*)state[rank];
    uint32_t* epoch_mem = st + FLAGS_WORDS;
    uint32_t* status = epoch_mem + DMP_COMM_MAX_BLOCKS;
    const int grid = comm_grid(nvec / W, max_blocks);
    allreduce_nvls_kernel<<<grid, DMP_COMM_THREADS, 0, stream>>>((float*)mc, f, epoch_mem, status, rank, W, nvec, scale);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// All-reduce (mean) + SGD through the switch: mc_grad / mc_master / mc_shadow are the multicast addresses of the bucket's gradient,
// fp32 master and bf16 shadow buffers (every rank bound its copy), master / momentum the LOCAL unicast views of this rank's
// master copy and its private momentum.  n: element count, a multiple of 8*W.  After the call every rank holds identical masters
// and shadows and a zeroed gradient bucket.
DMP_API int dmp_allreduce_nvls_sgd(void* mc_grad, void* mc_master, void* mc_shadow, const float* master, float* momentum,
                                   const float* hp_dev, int nesterov, void* const* state, int rank, int W, long long n, float scale,
                                   int max_blocks, cudaStream_t stream) {
    DMP_REQUIRE(mc_grad && mc_master && mc_shadow && master && momentum && hp_dev, "dmp_allreduce_nvls_sgd: null buffer");
    DMP_REQUIRE(W >= 2 && W <= DMP_MAX_RANKS && rank >= 0 && rank < W, "dmp_allreduce_nvls_sgd: bad rank / world size");
    DMP_REQUIRE(n % (8 * (long long)W) == 0, "dmp_allreduce_nvls_sgd: element count must be a multiple of 8*W");
    const long long n8 = n / 8;
    if (n8 == 0) return 0;
    PeerPtrs f = make_ptrs(state, W);
    uint32_t* st = (uint32_t*)state[rank];
    uint32_t* epoch_mem = st + FLAGS_WORDS;
    uint32_t* status = epoch_mem + DMP_COMM_MAX_BLOCKS;
    const int grid = comm_grid(n8 / W, max_blocks);
    allreduce_nvls_sgd_kernel<<<grid, DMP_COMM_THREADS, 0, stream>>>((float*)mc_grad, (float*)mc_master, (bf16*)mc_shadow, master, momentum,
                                                                      hp_dev, nesterov, f, epoch_mem, status, rank, W, n8, scale);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// Broadcast `bytes` (multiple of 16) from `src` (on the source rank) into dst[0..n_dst).
// state == null: unsynchronised (single process; caller orders streams).  Otherwise all W ranks call this with
// the same arguments shape; only `is_src` differs.  mc_dst: optional NVLS multicast address covering all dsts.
// dst_bf16: optional second destination set receiving an fp32->bf16 cast of the data (bytes counts fp32 input).
DMP_API int dmp_broadcast(const void* src, void* const* dst, int n_dst, void* mc_dst, void* const* dst_bf16, long long bytes,
                          void* const* state, int rank, int W, int is_src, int max_blocks, cudaStream_t stream) {
    DMP_REQUIRE(bytes % 16 == 0, "dmp_broadcast: size must be a multiple of 16 bytes");
    DMP_REQUIRE(n_dst >= 0 && n_dst <= DMP_MAX_RANKS, "dmp_broadcast: too many destinations");
    if (bytes == 0) return 0;
    const long long n16 = bytes / 16;
    PeerPtrs d = make_ptrs(dst, n_dst), db = make_ptrs(dst_bf16, dst_bf16 ? n_dst : 0), f = make_ptrs(state, state ? W : 0);
    uint32_t* epoch_mem = nullptr;
    uint32_t* status = nullptr;
    if (state != nullptr) {
        uint32_t* st = (uint32_t*)state[rank];
        epoch_mem = st + FLAGS_WORDS;
        status = epoch_mem + DMP_COMM_MAX_BLOCKS;
    }
    const int grid = comm_grid(n16, max_blocks);
    broadcast_kernel<<<grid, DMP_COMM_THREADS, 0, stream>>>(src, d, n_dst, mc_dst, db, n16, f, epoch_mem, status, rank, W, is_src);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// Concatenate n_src flat chunks (bytes[i] each, multiples of 16) into dst on the current device via peer loads.
DMP_API int dmp_gather_peer(void* const* src, const long long* bytes, int n_src, void* dst, int max_blocks, cudaStream_t stream) {
    DMP_REQUIRE(n_src >= 1 && n_src <= DMP_MAX_RANKS, "dmp_gather_peer: bad source count");
    GatherArgs a;
    a.off16[0] = 0;
    for (int i = 0; i < n_src; ++i) {
        DMP_REQUIRE(bytes[i] % 16 == 0, "dmp_gather_peer: chunk %d is not a multiple of 16 bytes", i);
        a.src[i] = src[i];
        a.off16[i + 1] = a.off16[i] + bytes[i] / 16;
    }
    if (a.off16[n_src] == 0) return 0;
    gather_peer_kernel<<<comm_grid(a.off16[n_src], max_blocks), DMP_COMM_THREADS, 0, stream>>>(a, n_src, dst);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

DMP_API int dmp_copy_peer(const void* src, void* dst, long long bytes, int max_blocks, cudaStream_t stream) {
    DMP_REQUIRE(bytes % 16 == 0, "dmp_copy_peer: size must be a multiple of 16 bytes");
    if (bytes == 0) return 0;
    static const bool debug_peer = getenv("DMP_DEBUG_PEER") != nullptr;  // diagnostic print of where both ends live
    if (debug_peer) {
        int dev = -1, can = -1;
        cudaGetDevice(&dev);
        cudaPointerAttributes a, b;
        cudaPointerGetAttributes(&a, src);
        cudaPointerGetAttributes(&b, dst);
        if (a.device != dev) cudaDeviceCanAccessPeer(&can, dev, a.device);
        fprintf(stderr, "[dmp_copy_peer] current device %d, src on %d (type %d), dst on %d (type %d), canAccess(src)=%d stream=%p\n", dev,
                a.device, (int)a.type, b.device, (int)b.type, can, (void*)stream);
    }
    copy16_kernel<<<comm_grid(bytes / 16, max_blocks), DMP_COMM_THREADS, 0, stream>>>(src, dst, bytes / 16);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// out = scale * sum_i src[i]  (fp32 out), or fused SGD on master/momentum/shadow when master != null.  zero_src != 0 also zeroes
// every source buffer after it has been read (persistent DataParallel replicas: their gradient arenas are accumulated into).
DMP_API int dmp_reduce_add_peer_z(int dtype, void* const* src, int W, long long n, float scale, float* out_f32, float* master,
                                  float* momentum, void* shadow, const float* hp_dev, int nesterov, int zero_src, int max_blocks,
                                  cudaStream_t stream);
DMP_API int dmp_reduce_add_peer(int dtype, void* const* src, int W, long long n, float scale, float* out_f32, float* master,
                                float* momentum, void* shadow, const float* hp_dev, int nesterov, int max_blocks,
                                cudaStream_t stream) {
    return dmp_reduce_add_peer_z(dtype, src, W, n, scale, out_f32, master, momentum, shadow, hp_dev, nesterov, 0, max_blocks, stream);
}
DMP_API int dmp_reduce_add_peer_z(int dtype, void* const* src, int W, long long n, float scale, float* out_f32, float* master,
                                  float* momentum, void* shadow, const float* hp_dev, int nesterov, int zero_src, int max_blocks,
                                  cudaStream_t stream) {
    DMP_REQUIRE(W >= 1 && W <= DMP_MAX_RANKS, "dmp_reduce_add_peer: bad source count");
    DMP_REQUIRE(dtype == 0 || dtype == 1, "dmp_reduce_add_peer: dtype must be fp32 or bf16");
    const int esz = dtype == 0 ? 4 : 2;
    DMP_REQUIRE((n * esz) % 16 == 0, "dmp_reduce_add_peer: buffer must be a multiple of 16 bytes");
    const long long nvec = n * esz / 16;
    if (nvec == 0) return 0;
    PeerPtrs s = make_ptrs(src, W);
    SgdArgs a;
    a.master = master; a.momentum = momentum; a.shadow = (bf16*)shadow; a.hp = hp_dev; a.nesterov = nesterov;
    const int grid = comm_grid(nvec, max_blocks);
    if (dtype == 0) reduce_add_peer_kernel<float><<<grid, DMP_COMM_THREADS, 0, stream>>>(s, W, nvec, scale, out_f32, a, zero_src);
    else reduce_add_peer_kernel<bf16><<<grid, DMP_COMM_THREADS, 0, stream>>>(s, W, nvec, scale, out_f32, a, zero_src);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// peer-memory plumbing (host): allocation, CUDA IPC export/import, peer access
// ---------------------------------------------------------------------------------------------------
DMP_API int dmp_symm_malloc(long long bytes, void** out) {
    DMP_REQUIRE(bytes > 0 && out != nullptr, "dmp_symm_malloc: bad arguments");
    DMP_CHECK_CUDA(cudaMalloc(out, (size_t)bytes));
    DMP_CHECK_CUDA(cudaMemset(*out, 0, (size_t)bytes));
    return 0;
}
DMP_API int dmp_symm_free(void* p) {
    if (p) DMP_CHECK_CUDA(cudaFree(p));
    return 0;
}
DMP_API int dmp_ipc_export(void* p, void* handle64) {
    cudaIpcMemHandle_t h;
    DMP_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle64, &h, sizeof(h));
    return 0;
}
DMP_API int dmp_ipc_import(const void* handle64, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    DMP_CHECK_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
DMP_API int dmp_ipc_close(void* p) {
    if (p) DMP_CHECK_CUDA(cudaIpcCloseMemHandle(p));
    return 0;
}
DMP_API int dmp_ipc_handle_bytes() { return (int)sizeof(cudaIpcMemHandle_t); }

// Enables access from the current device to `peer` (idempotent).  Returns 0 if access is possible.
DMP_API int dmp_enable_peer_access(int peer) {
    int dev = 0, can = 0;
    DMP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev == peer) return 0;
    DMP_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer));
    DMP_REQUIRE(can, "device %d cannot access peer %d", dev, peer);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return 0;
    }
    DMP_CHECK_CUDA(e);
    return 0;
}

// Reads the communicator status word (0 = healthy; 1+peer = barrier with that peer timed out).
DMP_API int dmp_comm_status(void* state_local, cudaStream_t stream) {
    uint32_t v = 0;
    uint32_t* status = (uint32_t*)state_local + FLAGS_WORDS + DMP_COMM_MAX_BLOCKS;
    if (cudaMemcpyAsync(&v, status, 4, cudaMemcpyDeviceToHost, stream) != cudaSuccess) return -1;
    if (cudaStreamSynchronize(stream) != cudaSuccess) return -1;
    return (int)v;
}
