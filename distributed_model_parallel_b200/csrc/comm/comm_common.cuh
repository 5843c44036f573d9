// Cross-GPU primitives shared by the communication kernels: peer pointer tables, release/acquire flag accesses and the
// block-level barrier across ranks with a bounded spin.  See collectives.cu for the protocol description.
#pragma once
#include "../common.cuh"

#define DMP_MAX_RANKS 16
#define DMP_COMM_THREADS 512
#define DMP_COMM_MAX_BLOCKS 128

struct PeerPtrs {
    void* p[DMP_MAX_RANKS];
};

// Per-rank communicator state living in device memory (allocated by dmp_comm_state_bytes / Python).
//   flags : uint32 [DMP_COMM_MAX_BLOCKS][DMP_MAX_RANKS]  (peer visible)
//   epoch : uint32 [DMP_COMM_MAX_BLOCKS]                 (local)
//   status: uint32 [4]                                   (local; status[0] != 0 => a barrier timed out)
#define FLAGS_WORDS (DMP_COMM_MAX_BLOCKS * DMP_MAX_RANKS)

namespace {

constexpr unsigned long long SPIN_TIMEOUT_DEFAULT_MS = 30000ull;  // status[1] (milliseconds) overrides when non-zero

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Block-level barrier across all ranks.  `flags.p[r]` is rank r's flag pad, `epoch` the value to publish.
// Every thread's prior global writes are made visible system-wide first.
// fence_sys: the caller's threads wrote data that PEERS will read after this barrier (stores into peer memory or into local
// memory that peers load next); a barrier that only announces "I have finished reading" needs no system-wide fence.
__device__ __forceinline__ bool barrier_ranks(const PeerPtrs& flags, int rank, int W, uint32_t epoch, uint32_t* status,
                                              bool fence_sys = true) {
    if (fence_sys) __threadfence_system();
    __syncthreads();
    int ok = 1;
    if (threadIdx.x < W) {
        const unsigned long long timeout_ns = (status[1] ? (unsigned long long)status[1] : SPIN_TIMEOUT_DEFAULT_MS) * 1000000ull;
        const int peer = threadIdx.x;
        uint32_t* remote = (uint32_t*)flags.p[peer] + blockIdx.x * DMP_MAX_RANKS + rank;
        st_release_sys(remote, epoch);
        const uint32_t* mine = (const uint32_t*)flags.p[rank] + blockIdx.x * DMP_MAX_RANKS + peer;
        unsigned long long t0 = 0;
        unsigned spins = 0;
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if ((++spins & 0x3FF) == 0) {
                const unsigned long long now = globaltimer_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > timeout_ns) {
                    atomicExch(status, 1u + (uint32_t)peer);
                    ok = 0;
                    break;
                }
            }
        }
    }
    return __syncthreads_and(ok) != 0;  // block-uniform verdict
}

}  // namespace
