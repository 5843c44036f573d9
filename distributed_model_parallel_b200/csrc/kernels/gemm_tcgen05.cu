// tcgen05 / TMEM / TMA GEMM for sm_100a.
//
//   D[M,N] (+)= A * B   with bf16 operands, fp32 accumulation in tensor memory.
//
// Three operand-layout variants cover every GEMM-shaped op of the training step
// (1x1 convolutions in NHWC are plain GEMMs; BERT linears):
//   NT  : A[M,K] K-major,            B[N,K] K-major     -> forward   Y  = X * W^T
//   NN  : A[M,K] K-major,            B[K,N] N-contiguous -> data grad dX = dY * W
//   TN  : A[K,M] M-contiguous,       B[K,N] N-contiguous -> weight grad dW = dY^T * X (split-K, fp32 atomics)
//
// Structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0   : TMA producer   (cp.async.bulk.tensor.2d, SWIZZLE_128B, mbarrier complete_tx)
//   warp 1   : TMEM allocator + single elected thread issuing tcgen05.mma.cta_group::1.kind::f16
//   warps 2-5: epilogue       (tcgen05.ld 32x32b -> registers -> bias/convert -> global / red.add)
// A 3-stage smem ring (full/empty mbarriers) decouples TMA from the tensor core; tcgen05.commit releases
// ring slots and finally signals the epilogue.  K tails rely on TMA out-of-bounds zero fill; M/N tails are
// masked in the epilogue.
#include "tcgen05_common.cuh"
#include <atomic>
#include <stdlib.h>


namespace {

struct GemmParams {
    int M, N, K;
    int k_blocks_per_split;  // split-K: each split handles this many BK blocks
    int splits;
    int m_tiles, n_tiles;
    void* D;
    long long ldd;
    const float* bias;  // per-column (N) fp32 bias, nullable
    int epi;            // 0: store bf16, 1: store fp32, 2: atomic add fp32
    int tma_store;      // epi 0 only: stage the bf16 tile in shared memory and store it with TMA (coalesced full lines)
    double* stats;      // ECOL kernels: [2N] per-column sums of the bf16-rounded outputs, accumulated (see the column pass)
    // ECOL == 2 (data-gradient GEMM feeding a BatchNorm backward): the BatchNorm's input y_raw [M, N] and its coefficients
    const bf16* aux;
    long long ld_aux;
    const float *bn_mean, *bn_invstd, *bn_gamma, *bn_beta;
    int relu;
    // batched GEMM (attention): item z = zb * nh + zh selects the 3rd / 4th tensor-map coordinate of A, B and D
    int nh, nbatch;
    long long ldd_h, ldd_b;  // element offsets of D per zh / zb for the epilogues that do not store through TMA
};

// Persistent kernel: CTA b processes work items b, b+gridDim.x, ...  A work item is one 128 x BN output tile of one
// K split.  Two TMEM accumulator buffers let the epilogue of item i overlap the TMA/MMA of item i+1.
// Depth of the TMA -> MMA shared-memory ring.  These GEMMs are memory-bound (K is a handful of 64-wide blocks forward, or a
// very long split pixel axis for the weight gradient), so what matters is bytes in flight per SM: HBM needs roughly
// 7.7 TB/s x ~1 us / 148 SMs = 50 KB outstanding per SM to stay busy, and a 3-stage ring holds at most two stages of loads
// (32-64 KB) in the air.  One CTA per SM owns the whole 227 KB, so the ring takes ~200 KB of it: 6 stages at BN=128, 8 below.
// Narrow tiles (BN <= 64, 20-24 KB per stage) keep two CTAs per SM instead and take 3 stages.
constexpr int STAT_SMEM_BYTES = 8 * 2 * 128 * 4;  // ECOL: per-CTA combine of the epilogue warps' column partials
constexpr int stg_warp_bytes(int bn) { return (32 * (bn * 2 + 16) + 1023) / 1024 * 1024; }  // output staging per epilogue warp
constexpr int gemm_stages(int b_rows, int ecol) {
    const int per = BM * BK * 2 + b_rows * BK * 2;
    if (per <= 24 * 1024 && ecol != 2) return 3;  // + 8-16 KB output staging (+ statistics scratch) stays under half an SM's shared memory
    const int s = (200 * 1024 - (ecol == 2 ? 8 : 4) * stg_warp_bytes(b_rows) - (ecol != 0 ? STAT_SMEM_BYTES : 0)) / per;
    return s > 8 ? 8 : s;
}

// BRES ("resident B"): NT forward GEMMs whose whole B panel [BN x K] fits a fixed shared-memory region load it ONCE per CTA
// (every tile of a CTA shares one n0: the host makes the grid a multiple of n_tiles) and the ring carries A only.  The kernel is
// bound by TMA load issue (profiles/ncu_gemm_fwd_tma_store_r1.txt), and re-fetching the 16 KB panel for every 128-row tile
// was half of that traffic.
constexpr int bres_bytes(int bn) { return bn >= 128 ? 64 * 1024 : 32 * 1024; }   // region for ceil(K/64) blocks of BN x 64 bf16
constexpr int bres_stages(int bn) { return bn >= 128 ? 6 : 3; }                  // A-only ring (16 KB per stage)

template <int BN, bool A_MN, bool B_MN, int ECOL, bool BRES = false>
__global__ void __launch_bounds__(NUM_THREADS, BN <= 64 ? 2 : 1)  // narrow tiles keep two CTAs per SM (<= 168 registers)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmY, const GemmParams p) {
    constexpr int A_BYTES = BM * BK * 2;  // 16 KB
    constexpr int B_ROWS = (B_MN && BN < 64) ? 64 : BN;
    constexpr int B_BYTES = B_ROWS * BK * 2;
    static_assert(!BRES || (!A_MN && !B_MN), "resident B is for the NT forward GEMM");
    constexpr int STAGE_BYTES = BRES ? A_BYTES : A_BYTES + B_BYTES;
    constexpr int STAGES = BRES ? bres_stages(BN) : gemm_stages(B_ROWS, ECOL);  // shadows the common 3-stage default
    constexpr int RES_BYTES = BRES ? bres_bytes(BN) : 0;
    constexpr uint32_t ACC_COLS = BN < 32 ? 32 : BN;
    constexpr uint32_t TMEM_COLS = 2 * ACC_COLS;  // double-buffered accumulator (power of two: 64/128/256)

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    // output staging for the TMA-store epilogue: per epilogue warp 32 rows x BN bf16, in panels of PW columns (128-byte rows,
    // SWIZZLE_128B; 64-byte rows and SWIZZLE_64B when BN == 32)
    constexpr int PW = BN >= 64 ? 64 : BN;
    constexpr int STG_PITCH = BN * 2 + 16;                 // mode 2: padded rows, conflict-free both ways
    constexpr int STG_WARP_BYTES = (32 * STG_PITCH + 1023) / 1024 * 1024;  // >= the 32 * BN * 2 the TMA layout needs, 1 KB aligned
    uint8_t* bres = smem + STAGES * STAGE_BYTES;             // BRES: the CTA's B panel, K block kb at bres + kb * B_BYTES
    uint8_t* stg = bres + RES_BYTES;
    // ECOL == 2: a second staging area of the same shape, filled by TMA with the matching tile of the BatchNorm input y_raw
    constexpr int YSTG_BYTES = ECOL == 2 ? 4 * STG_WARP_BYTES : 0;
    uint8_t* ystg = stg + 4 * STG_WARP_BYTES;
    uint64_t* full_bar = (uint64_t*)(stg + 4 * STG_WARP_BYTES + YSTG_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;   // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;   // [2]
    uint64_t* bres_bar = tmem_empty_bar + 2;        // BRES: the panel has landed
    uint32_t* tmem_ptr_smem = (uint32_t*)(bres_bar + 1);
    uint64_t* ybar = (uint64_t*)((uint8_t*)full_bar + 192);  // [4], ECOL == 2: one per epilogue warp (the barrier block is 256 bytes)
    float* stat_smem = (float*)(stg + 4 * STG_WARP_BYTES + YSTG_BYTES + 256);  // ECOL only (the launch adds STAT_SMEM_BYTES)

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int total_kb = (p.K + BK - 1) / BK;
    const int tiles_mn = p.m_tiles * p.n_tiles;
    const int num_items = tiles_mn * p.splits * p.nh * p.nbatch;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if (p.tma_store == 1) tma_prefetch_desc(&tmD);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tmem_full_bar[b], 1);
            mbar_init(&tmem_empty_bar[b], 4);  // one arrive per epilogue warp
        }
        mbar_init(bres_bar, 1);
        if (ECOL == 2) {
            tma_prefetch_desc(&tmY);
            for (int b = 0; b < 4; ++b) mbar_init(&ybar[b], 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int it = 0;  // global k-block counter -> smem ring position
            if (BRES && blockIdx.x < num_items) {
                const int n0c = ((blockIdx.x % tiles_mn) % p.n_tiles) * BN;  // the same for every tile of this CTA
                mbar_expect_tx(bres_bar, (uint32_t)(total_kb * B_BYTES));
                for (int kb = 0; kb < total_kb; ++kb) tma_load_4d(bres + kb * B_BYTES, &tmB, bres_bar, kb * BK, n0c, 0, 0);
            }
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                const int z = item / (tiles_mn * p.splits), rem = item % (tiles_mn * p.splits);
                const int zh = z % p.nh, zb = z / p.nh;
                const int split = rem / tiles_mn, tile = rem % tiles_mn;
                const int m0 = (tile / p.n_tiles) * BM, n0 = (tile % p.n_tiles) * BN;
                const int kb_begin = split * p.k_blocks_per_split;
                int kb_end = kb_begin + p.k_blocks_per_split;
                if (kb_end > total_kb) kb_end = total_kb;
                for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sa = smem + s * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    const int k0 = kb * BK;
                    // MN-major operands arrive as 64-wide boxes.  A box that lies entirely past M (or N) is not fetched at all:
                    // the TMA unit would still spend a full box of issue slots on it, and these GEMMs are bound by exactly that
                    // (Cout <= 64 halves the weight-gradient's A traffic).  The MMA then reads stale shared memory for those
                    // rows/columns; accumulator rows and columns are independent and the epilogue masks everything >= M, N.
                    if (BRES) {  // the ring carries A only
                        mbar_expect_tx(&full_bar[s], A_BYTES);
                        tma_load_4d(sa, &tmA, &full_bar[s], k0, m0, zh, zb);
                        continue;
                    }
                    constexpr int BOX_BYTES = 64 * BK * 2;
                    uint32_t tx = STAGE_BYTES;
                    const bool a_hi = !A_MN || (m0 + 64 < p.M);
                    if (A_MN && !a_hi) tx -= BOX_BYTES;
                    if (B_MN) {
#pragma unroll
                        for (int c = 1; c < B_ROWS / 64; ++c)
                            if (n0 + c * 64 >= p.N) tx -= BOX_BYTES;
                    }
                    mbar_expect_tx(&full_bar[s], tx);
                    if (A_MN) {
                        tma_load_4d(sa, &tmA, &full_bar[s], m0, k0, zh, zb);  // global A^T [K, M]: two boxes of {64 m, 64 k}
                        if (a_hi) tma_load_4d(sa + A_BYTES / 2, &tmA, &full_bar[s], m0 + 64, k0, zh, zb);
                    } else {
                        tma_load_4d(sa, &tmA, &full_bar[s], k0, m0, zh, zb);  // box {64 k, 128 m}
                    }
                    if (B_MN) {
#pragma unroll
                        for (int c = 0; c < B_ROWS / 64; ++c)
                            if (c == 0 || n0 + c * 64 < p.N) tma_load_4d(sb + c * (64 * BK * 2), &tmB, &full_bar[s], n0 + c * 64, k0, zh, zb);
                    } else {
                        tma_load_4d(sb, &tmB, &full_bar[s], k0, n0, zh, zb);  // box {64 k, BN n}
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            int it = 0, local = 0;
            if (BRES && blockIdx.x < num_items) {
                mbar_wait(bres_bar, 0);  // the resident B panel is in shared memory
                tcgen05_fence_after();
            }
            for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++local) {
                const int split = (item % (tiles_mn * p.splits)) / tiles_mn;
                const int kb_begin = split * p.k_blocks_per_split;
                int kb_end = kb_begin + p.k_blocks_per_split;
                if (kb_end > total_kb) kb_end = total_kb;
                const int buf = local & 1;
                const uint32_t acc_ph = (local >> 1) & 1;
                mbar_wait(&tmem_empty_bar[buf], acc_ph ^ 1);  // epilogue has drained this accumulator
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + buf * ACC_COLS;
                for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
                    const uint32_t sb = BRES ? smem_u32(bres + kb * B_BYTES) : sa + A_BYTES;
                    const int k0 = kb * BK;
                    int ksteps = (p.K - k0 + 15) / 16;  // skip the zero-filled K tail
                    if (ksteps > BK / 16) ksteps = BK / 16;
                    for (int k = 0; k < ksteps; ++k) {
                        // K-major: 16 bf16 = 32 B inside the 128 B swizzle row.  MN-major: 16 k-rows = 2 KB.
                        const uint64_t adesc = A_MN ? make_smem_desc(sa + k * 2048, A_BYTES / 2, 1024)
                                                    : make_smem_desc(sa + k * 32, 16, 1024);
                        const uint64_t bdesc = B_MN ? make_smem_desc(sb + k * 2048, 64 * BK * 2, 1024)
                                                    : make_smem_desc(sb + k * 32, 16, 1024);
                        umma_bf16(tmem_d, adesc, bdesc, idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[s]);  // frees the smem slot once the MMAs above have read it
                }
                umma_commit(&tmem_full_bar[buf]);  // accumulator complete -> epilogue
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue (warps 2..5)
        const int lane_grp = warp & 3;  // TMEM lane quarter this warp may access
        // ECOL (column pass): this CTA's tiles all share one n0 (the host makes gridDim.x a multiple of n_tiles), so a lane keeps
        // the running sums of ITS two columns per panel in registers across all of its tiles and issues its atomics once, at the end.
        constexpr int NCHUNK = BN >= 32 ? BN / 32 : 1;
        constexpr int NPAN = BN / PW;          // output panels per tile (PW columns each)
        constexpr int WPR = PW / 2;            // 32-bit words (column pairs) per staged panel row: 32 or 16
        constexpr int RPI = 32 / WPR;          // staged rows covered by one warp instruction: 1 or 2
        const int cw = lane % WPR, crs = lane / WPR;
        uint32_t cxo[8];                       // byte offset of the lane's word inside a staged row, per swizzle phase t
#pragma unroll
        for (int t = 0; t < 8; ++t) cxo[t] = ((((uint32_t)cw >> 2) ^ (uint32_t)t) << 4) + ((uint32_t)cw & 3) * 4;
        float cs1[NPAN][2], cs2[NPAN][2];      // [panel][column of the pair]: sum / second statistic
        float cmu[NPAN][2], cis[NPAN][2], csc[NPAN][2], csh[NPAN][2];  // ECOL == 2: BatchNorm coefficients of the lane's columns
        uint32_t ycount = 0;                   // ECOL == 2: y_raw slabs consumed so far (parity of ybar)
#pragma unroll
        for (int pn = 0; pn < NPAN; ++pn)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                cs1[pn][e] = cs2[pn][e] = 0.f;
                cmu[pn][e] = cis[pn][e] = csc[pn][e] = csh[pn][e] = 0.f;
            }
        if (ECOL == 2 && blockIdx.x < num_items) {
            const int n0c = ((blockIdx.x % tiles_mn) % p.n_tiles) * BN;
#pragma unroll
            for (int pn = 0; pn < NPAN; ++pn)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int col = n0c + pn * PW + 2 * cw + e;
                    if (col < p.N) {
                        const float g = p.bn_gamma ? p.bn_gamma[col] : 1.f, b = p.bn_beta ? p.bn_beta[col] : 0.f;
                        cmu[pn][e] = p.bn_mean[col];
                        cis[pn][e] = p.bn_invstd[col];
                        csc[pn][e] = g * cis[pn][e];                       // same expressions as the forward apply kernels
                        csh[pn][e] = b - cmu[pn][e] * g * cis[pn][e];
                    }
                }
        }
        int local = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++local) {
            const int tile = item % tiles_mn;
            const int z = item / (tiles_mn * p.splits);
            const int zh = z % p.nh, zb = z / p.nh;
            const long long dz = zh * p.ldd_h + zb * p.ldd_b;  // batch offset of D (elements) for the non-TMA stores
            const int m0 = (tile / p.n_tiles) * BM, n0 = (tile % p.n_tiles) * BN;
            const int buf = local & 1;
            const uint32_t acc_ph = (local >> 1) & 1;
            const int row = m0 + lane_grp * 32 + lane;
            if (ECOL == 2) {
                // fetch this warp's 32 x BN slab of y_raw while the tile's MMAs are still running (the previous tile's column
                // pass, the last reader of ystg, is behind us in program order)
                __syncwarp();
                if (lane == 0 && m0 + lane_grp * 32 < p.M) {
                    uint32_t bytes = 0;
#pragma unroll
                    for (int pn = 0; pn < BN / PW; ++pn)
                        if (n0 + pn * PW < p.N) bytes += 32 * PW * 2;
                    mbar_expect_tx(&ybar[lane_grp], bytes);
#pragma unroll
                    for (int pn = 0; pn < BN / PW; ++pn)
                        if (n0 + pn * PW < p.N)
                            tma_load_4d(ystg + lane_grp * STG_WARP_BYTES + pn * (32 * PW * 2), &tmY, &ybar[lane_grp], n0 + pn * PW,
                                        m0 + lane_grp * 32, 0, 0);
                }
            }
            mbar_wait(&tmem_full_bar[buf], acc_ph);
            tcgen05_fence_after();
            const uint32_t taddr_row = tmem_base + buf * ACC_COLS + ((uint32_t)(lane_grp * 32) << 16);
            uint8_t* my_stg = stg + lane_grp * STG_WARP_BYTES;
            if (p.tma_store == 1) {  // the previous tile's bulk stores must have finished reading this warp's staging area
                if (lane == 0) tma_store_wait_read();
                __syncwarp();
            }
            // The whole accumulator row comes out of TMEM with ONE wait: there is a single epilogue warp per scheduler, so every
            // tcgen05.wait::ld is fully exposed latency (4 serialised waits per tile used to dominate this kernel).  The
            // accumulator buffer goes back to the MMA warp right after the loads, before any conversion or store.
            uint32_t rall[NCHUNK][32];
#pragma unroll
            for (int ci = 0; ci < NCHUNK; ++ci) {
                if (n0 + ci * 32 < p.N) {  // warp-uniform
                    if (BN >= 32) tmem_ld_32(taddr_row + ci * 32, rall[ci]);
                    else tmem_ld_16(taddr_row + ci * 32, rall[ci]);
                }
            }
            tmem_ld_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);  // this warp is done reading the accumulator
#pragma unroll
            for (int c0 = 0; c0 < BN; c0 += 32) {
                if (n0 + c0 >= p.N) break;  // warp-uniform
                uint32_t* r = rall[c0 / 32];
                if (row < p.M) {
                    const int n = n0 + c0;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (p.bias != nullptr) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n + j < p.N) v[j] += p.bias[n + j];
                    }
                    if (p.epi == 0 && p.tma_store == 2) {
                        // padded row-major staging; written back below as whole 2*BN-byte rows
                        uint8_t* rowp = my_stg + lane * STG_PITCH + c0 * 2;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            bf16x8 o = float_to_bf16x8(v + j);
                            *(int4*)(rowp + j * 2) = *(int4*)&o;
                        }
                    } else if (p.epi == 0 && p.tma_store == 1) {
                        // lane = row of the warp's 32 x BN slab; 16-byte chunks are XOR-swizzled exactly as the tensor map expects
                        const int panel = c0 / PW;
                        const int chunk0 = (c0 % PW) >> 3;
                        const int swz = PW == 64 ? (lane & 7) : ((lane >> 1) & 3);
                        uint8_t* rowp = my_stg + panel * (32 * PW * 2) + lane * (PW * 2);
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            bf16x8 o = float_to_bf16x8(v + j);
                            *(int4*)(rowp + (((chunk0 + (j >> 3)) ^ swz) << 4)) = *(int4*)&o;
                        }
                    } else if (p.epi == 0) {
                        bf16* d = (bf16*)p.D + dz + (long long)row * p.ldd + n;
                        if (n + 32 <= p.N && ((p.ldd & 7) == 0)) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8) {
                                bf16x8 o = float_to_bf16x8(v + j);
                                *(int4*)(d + j) = *(int4*)&o;  // one 16-byte store (a struct copy would be split into 4)
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; j += 8) {
                                if (n + j + 8 <= p.N && ((p.ldd & 7) == 0)) {
                                    bf16x8 o = float_to_bf16x8(v + j);
                                    *(int4*)(d + j) = *(int4*)&o;
                                } else
                                    for (int q = 0; q < 8; ++q)
                                        if (n + j + q < p.N) d[j + q] = __float2bfloat16_rn(v[j + q]);
                            }
                        }
                    } else if (p.epi == 1) {
                        float* d = (float*)p.D + dz + (long long)row * p.ldd + n;
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            if (n + j + 4 <= p.N && ((p.ldd & 3) == 0)) *(float4*)(d + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                            else
                                for (int q = 0; q < 4; ++q)
                                    if (n + j + q < p.N) d[j + q] = v[j + q];
                        }
                    } else {
                        float* d = (float*)p.D + dz + (long long)row * p.ldd + n;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n + j < p.N) atomicAdd(d + j, v[j]);
                    }
                }
            }
            if (ECOL != 0) {
                // Column pass over the staged bf16 tile (this warp's 32 rows x BN columns, SWIZZLE_128B / 64B panels): lane cw
                // owns the column pair (2cw, 2cw+1) of every panel, i.e. one 32-bit word per staged row.  Across a warp
                // instruction the words of one row (two rows when PW == 32) cover 128 contiguous, swizzle-permuted bytes: every
                // bank once.  No shuffles, no transpose.  The row loop is fully unrolled over compile-time row numbers, so the
                // XOR swizzle folds into eight (four) per-lane base addresses computed once per kernel and every row is
                // LDS [base_t + imm] + 6 arithmetic instructions (an earlier rolled version recomputed the swizzle per row:
                // ~20 instructions and two branches per row, and the epilogue warps -- one per scheduler -- became the kernel's bound).
                //   ECOL == 1: cs1 += y, cs2 += y^2                      (BatchNorm batch sums of the bf16-rounded output)
                //   ECOL == 2: the tile is dA of a convolution whose input was act(BatchNorm(y_raw)); the ReLU mask is
                //              recomputed from y_raw (coalesced 128-byte row reads), applied IN the staged tile, and
                //              cs1 += g, cs2 += g * xhat -- the two reductions of the BatchNorm backward.
                __syncwarp();
                int rows_valid = p.M - (m0 + lane_grp * 32);
                if (rows_valid > 32) rows_valid = 32;
                const uint32_t stg_s = smem_u32(my_stg);
                const uint32_t ydelta = smem_u32(ystg + lane_grp * STG_WARP_BYTES) - stg_s;  // same swizzled layout, other buffer
                if (ECOL == 2 && rows_valid > 0) {
                    mbar_wait(&ybar[lane_grp], ycount & 1);  // the y_raw slab has landed
                    ++ycount;
                }
#pragma unroll
                for (int pn = 0; pn < NPAN; ++pn) {
                    if (n0 + pn * PW >= p.N) break;  // warp-uniform
                    const bool colok = n0 + pn * PW + 2 * cw < p.N;  // N % 8 == 0: a pair is valid or invalid as a whole
                    const uint32_t pbase = stg_s + pn * (32 * PW * 2) + crs * (PW * 2);
                    float a1 = 0.f, a2 = 0.f, b1 = 0.f, b2 = 0.f;
#pragma unroll
                    for (int r0 = 0; r0 < 32; r0 += RPI) {
                        const int t = PW == 64 ? (r0 & 7) : ((r0 >> 1) & 3);
                        const uint32_t addr = pbase + cxo[t] + r0 * (PW * 2);
                        const bool rowok = r0 + crs < rows_valid;  // all true except in the last row tile
                        uint32_t w;
                        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(addr));
                        float lo = __uint_as_float(w << 16), hi = __uint_as_float(w & 0xffff0000u);
                        if (!rowok) lo = hi = 0.f;  // rows past M were never staged
                        if (ECOL == 1) {
                            a1 += lo; a2 = fmaf(lo, lo, a2);
                            b1 += hi; b2 = fmaf(hi, hi, b2);
                        } else {
                            uint32_t yw;
                            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(yw) : "r"(addr + ydelta));
                            if (!rowok) yw = 0;
                            const float ylo = __uint_as_float(yw << 16), yhi = __uint_as_float(yw & 0xffff0000u);
                            if (p.relu) {
                                if (!(fmaf(ylo, csc[pn][0], csh[pn][0]) > 0.f)) lo = 0.f;
                                if (!(fmaf(yhi, csc[pn][1], csh[pn][1]) > 0.f)) hi = 0.f;
                                const uint32_t wm = (__float_as_uint(lo) >> 16) | (__float_as_uint(hi) & 0xffff0000u);
                                asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(wm) : "memory");
                            }
                            a1 += lo; a2 = fmaf(lo, (ylo - cmu[pn][0]) * cis[pn][0], a2);
                            b1 += hi; b2 = fmaf(hi, (yhi - cmu[pn][1]) * cis[pn][1], b2);
                        }
                    }
                    if (colok) {  // lanes past N read stale staging columns: drop them
                        cs1[pn][0] += a1; cs2[pn][0] += a2;
                        cs1[pn][1] += b1; cs2[pn][1] += b2;
                    }
                }
            }
            if (p.tma_store == 1) fence_proxy_async();  // generic-proxy smem writes -> visible to the TMA (async proxy)
            __syncwarp();
            if (p.tma_store == 2) {
                // coalesced write-back: one warp instruction stores LPR-lane-wide whole rows (16 bytes per lane), so every
                // request is a run of full 32-byte sectors instead of 32 scattered 16-byte pieces
                constexpr int LPR = BN / 8;             // lanes per row (16-byte pieces per row): 16 / 8 / 4
                constexpr int RPI = 32 / LPR;           // rows per instruction
                const int piece = lane % LPR, rsub = lane / LPR;
                const int ncol = n0 + piece * 8;
                bf16* dbase = (bf16*)p.D + dz + (long long)(m0 + lane_grp * 32) * p.ldd + ncol;
#pragma unroll 4
                for (int r0 = 0; r0 < 32; r0 += RPI) {
                    const int rr = r0 + rsub;
                    if (m0 + lane_grp * 32 + rr < p.M && ncol < p.N)
                        *(int4*)(dbase + (long long)rr * p.ldd) = *(const int4*)(my_stg + rr * STG_PITCH + piece * 16);
                }
                __syncwarp();  // the staging rows are rewritten by the next tile
            }
            if (lane == 0) {
                if (p.tma_store == 1 && m0 + lane_grp * 32 < p.M) {
#pragma unroll
                    for (int pn = 0; pn < BN / PW; ++pn)
                        if (n0 + pn * PW < p.N) tma_store_4d(&tmD, my_stg + pn * (32 * PW * 2), n0 + pn * PW, m0 + lane_grp * 32, zh, zb);
                    tma_store_commit();
                }
            }
        }
        if (p.tma_store == 1 && lane == 0) tma_store_wait_all();  // shared memory must outlive the last bulk store
        if (ECOL != 0) {
            // Combine the four epilogue warps' partials (and the two half-warps when PW == 32) in shared memory and issue ONE
            // double atomic per column and statistic per CTA: with every lane adding on its own, ~2400 atomics landed on each of
            // the 48 addresses of a 24-channel layer and the serialisation at L2 cost 35 us.
            float* part = stat_smem + (lane_grp * RPI + crs) * (2 * BN);
#pragma unroll
            for (int pn = 0; pn < NPAN; ++pn)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    part[pn * PW + 2 * cw + e] = cs1[pn][e];
                    part[BN + pn * PW + 2 * cw + e] = cs2[pn][e];
                }
            asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
            const int e = (warp - 2) * 32 + lane;
            if (blockIdx.x < num_items && e < BN) {
                const int col = ((blockIdx.x % tiles_mn) % p.n_tiles) * BN + e;
                float t1 = 0.f, t2 = 0.f;
#pragma unroll
                for (int g = 0; g < 4 * RPI; ++g) {
                    t1 += stat_smem[g * (2 * BN) + e];
                    t2 += stat_smem[g * (2 * BN) + BN + e];
                }
                if (col < p.N) {
                    atomicAdd(p.stats + col, (double)t1);
                    atomicAdd(p.stats + p.N + col, (double)t2);
                }
            }
        }
    }
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------ host side
// 2D bf16 tensor map: dim0 (contiguous) = inner, dim1 = outer with row pitch ld (elements).
// Batch geometry of one operand: nh x nb matrices, element strides between them (0, 0 for a plain 2-D operand).
struct Batch {
    int nh = 1, nb = 1;
    long long stride_h = 0, stride_b = 0;
};

// Every operand is encoded as a 4-D tensor {inner, outer, nh, nb}; plain GEMMs use nh = nb = 1 and coordinates (.., 0, 0).
int make_tmap_2d(CUtensorMap* tm, const void* base, uint64_t inner, uint64_t outer, uint64_t ld_elems, uint32_t box_inner,
                 uint32_t box_outer, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B, const Batch& bt = Batch()) {
    PFN_tmapEncodeTiled enc = get_encode_fn();
    DMP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled driver entry point unavailable");
    DMP_REQUIRE(((uintptr_t)base & 15) == 0, "TMA operand base must be 16-byte aligned");
    DMP_REQUIRE((ld_elems * 2) % 16 == 0, "TMA operand row pitch must be a multiple of 16 bytes (ld=%llu elements)",
                (unsigned long long)ld_elems);
    DMP_REQUIRE((bt.stride_h * 2) % 16 == 0 && (bt.stride_b * 2) % 16 == 0, "TMA batch strides must be multiples of 16 bytes");
    const uint64_t plane = ld_elems * 2 * outer;  // any valid non-zero stride does for an extent-1 dimension
    cuuint64_t gdim[4] = {inner, outer, (cuuint64_t)bt.nh, (cuuint64_t)bt.nb};
    cuuint64_t gstride[3] = {ld_elems * 2, bt.nh > 1 ? (cuuint64_t)bt.stride_h * 2 : plane, bt.nb > 1 ? (cuuint64_t)bt.stride_b * 2 : plane};
    cuuint32_t box[4] = {box_inner, box_outer, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DMP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu ld=%llu)", (int)r,
                (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld_elems);
    return 0;
}

// bf16 epilogue: 1 (default) = swizzled shared-memory staging + TMA store, 2 = padded staging + whole-row 16-byte global
// stores, 0 = registers straight to global (32 scattered 16-byte pieces per instruction).  DMP_GEMM_STORE_MODE overrides.
// Measured on the 131072 x 144 x 24 forward GEMM: 37.7 / 39.2 / 46.5 us.  (Removing the stores altogether only gets to
// ~25 us: with one 64-wide K block per tile the kernel is bound by TMA *load* issue -- every tile re-fetches its B panel
// and a 24-channel row still costs a full 128-byte box row -- see profiles/README.md.)
const int g_gemm_store_mode = [] {
    const char* e = getenv("DMP_GEMM_STORE_MODE");
    return (e != nullptr && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1;
}();

// DMP_WGRAD_MAX_CTAS: grid cap of the split-K weight-gradient GEMM (0 = none; default a third of the SMs).  Measured on the
// captured MobileNetV2 step at batch 128 (profiles/sm_partition_sweep.txt): uncapped 4.28 ms, 96 CTAs 4.16, 64 4.10, 48 4.07;
// with the depthwise weight gradient and the cooperative BatchNorm grids sized to match, 3.93 ms.  Neutral at batch 512.
int wgrad_max_ctas() {
    static const int v = [] { const char* e = getenv("DMP_WGRAD_MAX_CTAS"); return e ? atoi(e) : dmp_num_sms() / 3; }();
    return v;
}

// DMP_GEMM_SMALLK_BN64: largest K for which forward / data-gradient GEMMs with N > 64 use 64-wide tiles (0 = never).
// Measured (bench/chain_bench.py, MobileNetV2 shapes, B200): 131072 x 144 x 24 forward + statistics 40.0 -> 24.3 us,
// 131072 x 96 x 16 21.0 -> 15.1, 32768 x 192 x 32 13.6 -> 8.4; whole step 3.67 -> 3.53 ms with K <= 160 (no further gain beyond).
int small_k_bn64_max_k() {
    static const int v = [] { const char* e = getenv("DMP_GEMM_SMALLK_BN64"); return e ? atoi(e) : 160; }();
    return v;
}

int small_k_bn64_min_m() {
    static const int v = [] { const char* e = getenv("DMP_GEMM_SMALLK_MIN_M"); return e ? atoi(e) : 0; }();
    return v;
}

// DMP_GEMM_CHAIN_MAX_SMS: SMs the forward / data-gradient GEMMs size their persistent grid for (0 = all).
int chain_max_sms() {
    static const int v = [] { const char* e = getenv("DMP_GEMM_CHAIN_MAX_SMS"); return e ? atoi(e) : 0; }();
    return v;
}

template <int BN, bool A_MN, bool B_MN, int ECOL = 0, bool BRES = false>
int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int splits, cudaStream_t stream,
                const Batch& dbatch = Batch()) {
    constexpr int B_ROWS = (B_MN && BN < 64) ? 64 : BN;
    constexpr int ring_bytes = BRES ? bres_stages(BN) * (BM * BK * 2) + bres_bytes(BN)
                                    : gemm_stages(B_ROWS, ECOL) * (BM * BK * 2 + B_ROWS * BK * 2);
    constexpr int smem_bytes = ring_bytes + 1024 /*align slack*/ + (ECOL == 2 ? 8 : 4) * stg_warp_bytes(BN) /*output (+ y_raw) staging*/ +
                               256 /*barriers*/ + (ECOL != 0 ? STAT_SMEM_BYTES : 0);
    auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, ECOL, BRES>;
    // per template instantiation and per device; DataParallel replicas call this from one thread per device, so the "already
    // configured" mask is atomic (setting the attribute twice is harmless, a torn read-modify-write is not)
    static std::atomic<unsigned long long> configured_dev_mask{0};
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    if (!(configured_dev_mask.load(std::memory_order_acquire) & bit)) {
        DMP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        configured_dev_mask.fetch_or(bit, std::memory_order_release);
    }
    GemmParams q = p;
    // bf16 outputs leave through TMA when the rows are 16-byte aligned: box = {one panel of PW columns, one epilogue warp's 32 rows}
    CUtensorMap tmD = tmA;
    q.tma_store = 0;
    if (p.epi == 0 && (p.ldd & 7) == 0 && (p.N & 7) == 0 && ((uintptr_t)p.D & 15) == 0 && (g_gemm_store_mode != 0 || ECOL != 0)) {
        q.tma_store = ECOL != 0 ? 1 : g_gemm_store_mode;  // the column pass reads the swizzled staging of mode 1
        if (q.tma_store == 1) {
            constexpr int PW = BN >= 64 ? 64 : BN;
            if (int rc = make_tmap_2d(&tmD, p.D, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ldd, PW, 32,
                                      PW == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, dbatch))
                return rc;
        }
    }
    q.m_tiles = dmp_ceil_div(p.M, BM);
    q.n_tiles = dmp_ceil_div(p.N, BN);
    q.splits = splits;
    const long long items = (long long)q.m_tiles * q.n_tiles * splits * q.nh * q.nbatch;
    const int ctas_per_sm = smem_bytes <= 110 * 1024 ? 2 : 1;
    long long grid = (long long)dmp_num_sms() * ctas_per_sm;
    if (grid > items) grid = items;
    // weight-gradient GEMMs run on a side stream next to the data-gradient chain: a grid that fills every SM (one CTA with
    // ~200 KB of shared memory each) leaves the chain's kernels nowhere to run until it drains, so it may be capped
    if (A_MN && p.epi == 2 && wgrad_max_ctas() > 0 && grid > wgrad_max_ctas()) grid = wgrad_max_ctas();
    // ... and the chain's own GEMMs then take the SMs that are left (a statically scheduled CTA that has to wait for an SM
    // delays all of its tiles)
    if (!(A_MN && p.epi == 2) && chain_max_sms() > 0 && grid > (long long)chain_max_sms() * ctas_per_sm) grid = (long long)chain_max_sms() * ctas_per_sm;
    if (ECOL != 0 || BRES) {  // every tile of a CTA must have the same n0: the item stride has to be a multiple of n_tiles
        DMP_REQUIRE(splits == 1 && q.n_tiles <= grid, "gemm statistics epilogue: %d column tiles do not fit the grid", q.n_tiles);
        grid = grid / q.n_tiles * q.n_tiles;
    }
    if (ECOL != 0) DMP_REQUIRE(q.tma_store == 1, "gemm column-pass epilogue needs the TMA-store staging (N and ldd multiples of 8, 16-byte aligned D)");
    CUtensorMap tmY = tmA;
    if (ECOL == 2) {
        constexpr int PW = BN >= 64 ? 64 : BN;
        if (int rc = make_tmap_2d(&tmY, p.aux, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ld_aux, PW, 32,
                                  PW == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B))
            return rc;
    }
    kern<<<(int)grid, NUM_THREADS, smem_bytes, stream>>>(tmA, tmB, tmD, tmY, q);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

template <bool A_MN, bool B_MN>
int dispatch_bn(int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int splits, cudaStream_t s,
                const Batch& dbatch = Batch()) {
    switch (bn) {
        case 32: return launch_gemm<32, A_MN, B_MN>(tmA, tmB, p, splits, s, dbatch);
        case 64: return launch_gemm<64, A_MN, B_MN>(tmA, tmB, p, splits, s, dbatch);
        default: return launch_gemm<128, A_MN, B_MN>(tmA, tmB, p, splits, s, dbatch);
    }
}

}  // namespace

// layout: 0 = NT (A[M,K], B[N,K]); 1 = NN (A[M,K], B[K,N]); 2 = TN (A[K,M], B[K,N])
// out_mode: 0 = bf16 store, 1 = fp32 store, 2 = fp32 atomic accumulate (D must be pre-zeroed / hold the addend)
// splits <= 0 lets the library choose (only out_mode 2 may use more than one split).
DMP_API int dmp_gemm_bf16(int layout, const void* A, long long lda, const void* B, long long ldb, void* D, long long ldd,
                          int M, int N, int K, const float* bias, int out_mode, int splits, cudaStream_t stream) {
    DMP_REQUIRE(layout >= 0 && layout <= 2, "dmp_gemm_bf16: bad layout %d", layout);
    DMP_REQUIRE(M > 0 && N > 0 && K > 0, "dmp_gemm_bf16: empty problem %dx%dx%d", M, N, K);
    DMP_REQUIRE(out_mode >= 0 && out_mode <= 2, "dmp_gemm_bf16: bad out_mode %d", out_mode);
    const bool a_mn = (layout == 2);
    const bool b_mn = (layout != 0);
    int bn = N > 64 ? 128 : (N > 32 ? 64 : 32);
    if (b_mn && bn < 64) bn = 64;
    // forward / data-gradient GEMMs with a short reduction: 64-wide tiles, two CTAs per SM (see dmp_gemm_bf16_colstats)
    if (layout != 2 && out_mode == 0 && bn == 128 && K <= small_k_bn64_max_k() && M >= small_k_bn64_min_m()) bn = 64;

    CUtensorMap tmA, tmB;
    int rc;
    if (a_mn) rc = make_tmap_2d(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK);
    else rc = make_tmap_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM);
    if (rc) return rc;
    if (b_mn) rc = make_tmap_2d(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK);
    else rc = make_tmap_2d(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, (uint32_t)bn);
    if (rc) return rc;

    const int total_kb = dmp_ceil_div(K, BK);
    if (out_mode != 2) splits = 1;
    if (splits <= 0) {
        const int tiles = dmp_ceil_div(M, BM) * dmp_ceil_div(N, bn);
        int want = dmp_ceil_div(2 * dmp_num_sms(), tiles);
        int max_by_k = total_kb / 4 > 0 ? total_kb / 4 : 1;  // at least 4 k-blocks per split
        splits = want < max_by_k ? want : max_by_k;
        if (splits < 1) splits = 1;
    }
    if (splits > total_kb) splits = total_kb;
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.k_blocks_per_split = dmp_ceil_div(total_kb, splits);
    splits = dmp_ceil_div(total_kb, p.k_blocks_per_split);  // no empty splits
    p.D = D; p.ldd = ldd; p.bias = bias; p.epi = out_mode; p.stats = nullptr;
    p.aux = nullptr; p.ld_aux = 0; p.bn_mean = p.bn_invstd = p.bn_gamma = p.bn_beta = nullptr; p.relu = 0;
    p.nh = p.nbatch = 1; p.ldd_h = p.ldd_b = 0;

    if (!a_mn && !b_mn) {
        // resident-B variant: whole [bn x K] panel in shared memory, loaded once per CTA (DMP_GEMM_BRES=0 disables).  Measured:
        // 131072x144x24 36.7 -> 24.1 us, 131072x128x64 17.1 -> 13.2 us, 32768x192x32 11.2 -> 9.3 us (bench/gemm_sweep.py).
        static const bool use_bres = [] { const char* e = getenv("DMP_GEMM_BRES"); return e == nullptr || e[0] != '0'; }();
        if (use_bres && splits == 1 && (long long)total_kb * bn * BK * 2 <= bres_bytes(bn) && dmp_ceil_div(N, bn) <= dmp_num_sms()) {
            switch (bn) {
                case 32: return launch_gemm<32, false, false, 0, true>(tmA, tmB, p, 1, stream);
                case 64: return launch_gemm<64, false, false, 0, true>(tmA, tmB, p, 1, stream);
                default: return launch_gemm<128, false, false, 0, true>(tmA, tmB, p, 1, stream);
            }
        }
        return dispatch_bn<false, false>(bn, tmA, tmB, p, splits, stream);
    }
    if (!a_mn && b_mn) return dispatch_bn<false, true>(bn, tmA, tmB, p, splits, stream);
    return dispatch_bn<true, true>(bn, tmA, tmB, p, splits, stream);
}

// D[M,N] (bf16) = A[M,K] * B[N,K]^T, and stats[0..N) += column sums, stats[N..2N) += column sums of squares of the bf16-rounded
// outputs (double, accumulated: the caller zeroes or chains them).  This is the 1x1-convolution forward with the BatchNorm
// batch statistics produced by the GEMM epilogue, so the normalisation that follows is a single elementwise pass.
DMP_API int dmp_gemm_bf16_colstats(const void* A, long long lda, const void* B, long long ldb, void* D, long long ldd, int M,
                                   int N, int K, double* stats, cudaStream_t stream) {
    DMP_REQUIRE(M > 0 && N > 0 && K > 0 && stats != nullptr, "dmp_gemm_bf16_colstats: bad arguments");
    int bn = N > 64 ? 128 : (N > 32 ? 64 : 32);
    // One K block (K <= 64) and many rows: the tile is all epilogue (TMEM -> registers -> staging -> TMA store, plus the column
    // statistics), and a 128-wide tile keeps one CTA -- four epilogue warps -- per SM.  64-wide tiles run two CTAs per SM.
    if (bn == 128 && K <= small_k_bn64_max_k() && M >= small_k_bn64_min_m()) bn = 64;
    CUtensorMap tmA, tmB;
    if (int rc = make_tmap_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM)) return rc;
    if (int rc = make_tmap_2d(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, (uint32_t)bn)) return rc;
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.k_blocks_per_split = dmp_ceil_div(K, BK);
    p.D = D; p.ldd = ldd; p.bias = nullptr; p.epi = 0; p.stats = stats;
    p.aux = nullptr; p.ld_aux = 0; p.bn_mean = p.bn_invstd = p.bn_gamma = p.bn_beta = nullptr; p.relu = 0;
    p.nh = p.nbatch = 1; p.ldd_h = p.ldd_b = 0;
    static const bool use_bres = [] { const char* e = getenv("DMP_GEMM_BRES"); return e == nullptr || e[0] != '0'; }();
    if (use_bres && (long long)p.k_blocks_per_split * bn * BK * 2 <= bres_bytes(bn) && dmp_ceil_div(N, bn) <= dmp_num_sms()) {
        switch (bn) {
            case 32: return launch_gemm<32, false, false, 1, true>(tmA, tmB, p, 1, stream);
            case 64: return launch_gemm<64, false, false, 1, true>(tmA, tmB, p, 1, stream);
            default: return launch_gemm<128, false, false, 1, true>(tmA, tmB, p, 1, stream);
        }
    }
    switch (bn) {
        case 32: return launch_gemm<32, false, false, 1>(tmA, tmB, p, 1, stream);
        case 64: return launch_gemm<64, false, false, 1>(tmA, tmB, p, 1, stream);
        default: return launch_gemm<128, false, false, 1>(tmA, tmB, p, 1, stream);
    }
}

// Batched GEMM: nh x nb independent problems D_z = A_z * B_z of one shape, operands addressed as base + zh*stride_h + zb*stride_b
// (elements), bf16 output.  One persistent launch walks all tiles of all problems.  This is what attention is made of
// (BASELINE.json config 4, no counterpart in the reference): scores = Q K^T (layout 0), context = P V (layout 1),
// dV = P^T dO and dK = dS^T Q (layout 2), dP = dO V^T (0), dQ = dS K (1), with z = (head, batch) walking the fused
// [B, S, 3, heads, d] projection in place -- no per-head copies.
DMP_API int dmp_gemm_bf16_batched(int layout, const void* A, long long lda, long long a_sh, long long a_sb, const void* B, long long ldb,
                                  long long b_sh, long long b_sb, void* D, long long ldd, long long d_sh, long long d_sb, int M, int N,
                                  int K, int nh, int nb, cudaStream_t stream) {
    DMP_REQUIRE(layout >= 0 && layout <= 2, "dmp_gemm_bf16_batched: bad layout %d", layout);
    DMP_REQUIRE(M > 0 && N > 0 && K > 0 && nh > 0 && nb > 0, "dmp_gemm_bf16_batched: empty problem");
    DMP_REQUIRE((ldd & 7) == 0 && (N & 7) == 0 && ((uintptr_t)D & 15) == 0 && (d_sh & 7) == 0 && (d_sb & 7) == 0,
                "dmp_gemm_bf16_batched: D must be 16-byte aligned with N, ldd and batch strides multiples of 8");
    const bool a_mn = (layout == 2), b_mn = (layout != 0);
    int bn = N > 64 ? 128 : (N > 32 ? 64 : 32);
    if (b_mn && bn < 64) bn = 64;
    Batch ba, bb, bd;
    ba.nh = bb.nh = bd.nh = nh; ba.nb = bb.nb = bd.nb = nb;
    ba.stride_h = a_sh; ba.stride_b = a_sb; bb.stride_h = b_sh; bb.stride_b = b_sb; bd.stride_h = d_sh; bd.stride_b = d_sb;
    CUtensorMap tmA, tmB;
    int rc;
    if (a_mn) rc = make_tmap_2d(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK, CU_TENSOR_MAP_SWIZZLE_128B, ba);
    else rc = make_tmap_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM, CU_TENSOR_MAP_SWIZZLE_128B, ba);
    if (rc) return rc;
    if (b_mn) rc = make_tmap_2d(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK, CU_TENSOR_MAP_SWIZZLE_128B, bb);
    else rc = make_tmap_2d(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, (uint32_t)bn, CU_TENSOR_MAP_SWIZZLE_128B, bb);
    if (rc) return rc;
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.k_blocks_per_split = dmp_ceil_div(K, BK);
    p.D = D; p.ldd = ldd; p.bias = nullptr; p.epi = 0; p.stats = nullptr;
    p.aux = nullptr; p.ld_aux = 0; p.bn_mean = p.bn_invstd = p.bn_gamma = p.bn_beta = nullptr; p.relu = 0;
    p.nh = nh; p.nbatch = nb; p.ldd_h = d_sh; p.ldd_b = d_sb;
    if (!a_mn && !b_mn) return dispatch_bn<false, false>(bn, tmA, tmB, p, 1, stream, bd);
    if (!a_mn && b_mn) return dispatch_bn<false, true>(bn, tmA, tmB, p, 1, stream, bd);
    return dispatch_bn<true, true>(bn, tmA, tmB, p, 1, stream, bd);
}

// Data-gradient GEMM of a 1x1 convolution whose INPUT was act(BatchNorm(y_raw)):  G[M,N] (bf16) = mask * (A[M,K] * B[K,N])
// with mask = (y_raw * scale + shift > 0) when relu != 0 (else 1), and stats[0..N) += sum_m G, stats[N..2N) += sum_m G * xhat
// (xhat = (y_raw - mean) * invstd): the two reductions the BatchNorm backward needs come out of the producer of its incoming
// gradient, so that backward is one elementwise pass (dmp_bn_bwd_dx_sums) instead of reduce -> grid barrier -> dx.
// A = dY [M,K] K-major, B = W [K,N] N-contiguous (layout NN of dmp_gemm_bf16).  N, ldd, ld_y multiples of 8.
DMP_API int dmp_gemm_bf16_dgrad_bnbwd(const void* A, long long lda, const void* B, long long ldb, void* G, long long ldd, int M, int N,
                                      int K, const void* y_raw, long long ld_y, const float* mean, const float* invstd,
                                      const float* gamma, const float* beta, int relu, double* stats, cudaStream_t stream) {
    DMP_REQUIRE(M > 0 && N > 0 && K > 0 && stats != nullptr && y_raw != nullptr && mean != nullptr && invstd != nullptr,
                "dmp_gemm_bf16_dgrad_bnbwd: bad arguments");
    DMP_REQUIRE((ld_y & 7) == 0 && ((uintptr_t)y_raw & 15) == 0, "dmp_gemm_bf16_dgrad_bnbwd: y_raw rows must be 16-byte aligned");
    int bn = N > 64 ? 128 : 64;
    CUtensorMap tmA, tmB;
    if (int rc = make_tmap_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM)) return rc;
    if (int rc = make_tmap_2d(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK)) return rc;
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.k_blocks_per_split = dmp_ceil_div(K, BK);
    p.D = G; p.ldd = ldd; p.bias = nullptr; p.epi = 0; p.stats = stats;
    p.aux = (const bf16*)y_raw; p.ld_aux = ld_y; p.bn_mean = mean; p.bn_invstd = invstd; p.bn_gamma = gamma; p.bn_beta = beta;
    p.relu = relu;
    p.nh = p.nbatch = 1; p.ldd_h = p.ldd_b = 0;
    if (bn == 64) return launch_gemm<64, false, true, 2>(tmA, tmB, p, 1, stream);
    return launch_gemm<128, false, true, 2>(tmA, tmB, p, 1, stream);
}

// ------------------------------------------------------------------ CUDA-core oracle (tests / debugging only)
__global__ void gemm_ref_kernel(int layout, const bf16* A, long long lda, const bf16* B, long long ldb, float* D,
                                long long ldd, int M, int N, int K, const float* bias, int accumulate) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    int m = blockIdx.y * blockDim.y + threadIdx.y;
    if (m >= M || n >= N) return;
    float acc = 0.f;
    for (int k = 0; k < K; ++k) {
        float a = __bfloat162float(layout == 2 ? A[(long long)k * lda + m] : A[(long long)m * lda + k]);
        float b = __bfloat162float(layout == 0 ? B[(long long)n * ldb + k] : B[(long long)k * ldb + n]);
        acc += a * b;
    }
    if (bias) acc += bias[n];
    if (accumulate) D[(long long)m * ldd + n] += acc;
    else D[(long long)m * ldd + n] = acc;
}
DMP_API int dmp_gemm_ref_f32out(int layout, const void* A, long long lda, const void* B, long long ldb, float* D,
                                long long ldd, int M, int N, int K, const float* bias, int accumulate,
                                cudaStream_t stream) {
    dim3 block(32, 8), grid(dmp_ceil_div(N, 32), dmp_ceil_div(M, 8));
    gemm_ref_kernel<<<grid, block, 0, stream>>>(layout, (const bf16*)A, lda, (const bf16*)B, ldb, D, ldd, M, N, K, bias,
                                                accumulate);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
