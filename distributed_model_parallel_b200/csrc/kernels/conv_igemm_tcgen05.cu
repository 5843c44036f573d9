// Implicit-GEMM dense convolution (3x3 / KxK, stride 1, zero padding) on tcgen05 / TMEM, fed by 4-D TMA tiles.
//
// NOT IN REFERENCE: the reference model has a single dense 3x3 convolution (its stem, Cin = 3, handled by the direct
// stem kernel); this kernel serves the ResNet-18/50 and VGG-16 configurations of BASELINE.json (SURVEY.md work-list W18).
//
// The im2col matrix is never materialised.  With channels-last activations x[N,H,W,C], one (kh,kw) tap of a 128-pixel
// output patch (BNI images x BH rows x BW columns) and 64 input channels is exactly one 4-D TMA box
// {64, BW, BH, BNI} at coordinates {c0, w0+kw-pad, h0+kh-pad, n0}: out-of-image elements are zero-filled by the TMA unit
// (that is the padding) and the box lands in shared memory as 128 rows of 128 bytes in SWIZZLE_128B order -- the K-major
// A-operand layout tcgen05.mma expects.  The K loop runs over taps x (Cin/64) such boxes, the B operand is the weight
// pre-permuted to [Cout][kh][kw][Cin].
//   fwd   : y  = conv(x, W)                      A = x boxes,  B = Wp[Cout, KH*KW*Cin]
//   dgrad : dx = conv(dy, flipped W^T)  (stride 1: the same kernel with A = dy boxes, B = Wd[Cin, KH*KW*Cout])
//   wgrad : dWp[co, tap, ci] = sum_pixels dy[p,co] * x[p (+) tap, ci]   -- both operands MN-major 4-D boxes of 64 pixels,
//           split over pixel patches, fp32 atomics into dWp
// Same warp roles / pipelines as gemm_tcgen05.cu (TMA producer, single-thread MMA issuer, 4 epilogue warps, persistent
// CTAs, double-buffered TMEM accumulators).
#include "tcgen05_common.cuh"

namespace {

struct ConvParams {
    int Nimg, Ho, Wo;            // output geometry (== input geometry of the A operand: stride 1, "same" padding)
    int Cin, Cout;               // channels of the A operand / output columns
    int KH, KW, pad;
    int BW, BH, BNI;             // pixel patch (BW*BH*BNI = 128 for fwd/dgrad, 64 for wgrad)
    int tiles_w, tiles_h, tiles_n;
    int n_tiles;                 // output-column tiles
    int m_tiles;                 // fwd: pixel patches; wgrad: Cout tiles
    int splits, patches_per_split;  // wgrad only
    void* D;
    long long ldd;
    const float* bias;
    // generalisation to strided convolutions and their data gradient: an item belongs to a class c (forward: one class; stride-2
    // data gradient: the four output-pixel parities), a class has its own tap list -- (weight tap index, patch offset dh, dw) --
    // and writes output pixel (a * oscale + ph[c], b * oscale + pw[c]) of an OH x OW image for patch pixel (a, b)
    int astride;                 // element stride of the A-operand boxes (forward of a strided convolution: its stride; else 1)
    int ncls, oscale, OH, OW;
    int cls_ntaps[4], cls_ph[4], cls_pw[4];
    signed char cls_tap[4][49], cls_dh[4][49], cls_dw[4][49];
    int kcrs;                    // wgrad: write dW as [Cout][Cin][KH][KW] (the parameter's own layout) instead of [Cout][tap][Cin]
    int wstride;                 // wgrad: stride of the convolution (x boxes are read with this element stride)
};

// ------------------------------------------------------------------ forward / stride-1 data gradient
template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ConvParams p) {
    constexpr int A_BYTES = BM * BK * 2;
    constexpr int B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr uint32_t ACC_COLS = BN < 32 ? 32 : BN;
    constexpr uint32_t TMEM_COLS = 2 * ACC_COLS;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + STAGES * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_ptr_smem = (uint32_t*)(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cchunks = p.Cin / BK;
    const int tiles_mn = p.m_tiles * p.n_tiles;
    const int num_items = tiles_mn * p.ncls;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tmem_full_bar[b], 1);
            mbar_init(&tmem_empty_bar[b], 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                const int cls = item / tiles_mn, rem = item % tiles_mn;
                const int mt = rem / p.n_tiles, n0 = (rem % p.n_tiles) * BN;
                const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h, tn = mt / (p.tiles_w * p.tiles_h);
                const int total_kb = p.cls_ntaps[cls] * cchunks;
                for (int kb = 0; kb < total_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sa = smem + s * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    const int ti = kb / cchunks, cc = kb % cchunks;
                    const int tap = p.cls_tap[cls][ti];
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                    // one shifted pixel patch x 64 channels (every astride-th pixel); negative / past-the-edge coordinates are
                    // zero-filled by the TMA unit = the convolution's padding
                    tma_load_4d(sa, &tmA, &full_bar[s], cc * BK, tw * p.BW * p.astride + p.cls_dw[cls][ti],
                                th * p.BH * p.astride + p.cls_dh[cls][ti], tn * p.BNI);
                    tma_load_2d(sb, &tmB, &full_bar[s], tap * p.Cin + cc * BK, n0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BN, 0, 0);
            int it = 0, local = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++local) {
                const int total_kb = p.cls_ntaps[item / tiles_mn] * cchunks;
                const int buf = local & 1;
                const uint32_t acc_ph = (local >> 1) & 1;
                mbar_wait(&tmem_empty_bar[buf], acc_ph ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + buf * ACC_COLS;
                for (int kb = 0; kb < total_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
                    const uint32_t sb = sa + A_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        umma_bf16(tmem_d, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                                  (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&tmem_full_bar[buf]);
            }
        }
    } else {
        const int lane_grp = warp & 3;
        int local = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++local) {
            const int cls = item / tiles_mn, rem = item % tiles_mn;
            const int mt = rem / p.n_tiles, n0 = (rem % p.n_tiles) * BN;
            const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h, tn = mt / (p.tiles_w * p.tiles_h);
            const int buf = local & 1;
            const uint32_t acc_ph = (local >> 1) & 1;
            // accumulator row r of the tile = pixel (ni, hi, wi) of the patch, in TMA box order
            const int r = lane_grp * 32 + lane;
            const int wi = r % p.BW, hi = (r / p.BW) % p.BH, ni = r / (p.BW * p.BH);
            const int n = tn * p.BNI + ni, ho = th * p.BH + hi, wo = tw * p.BW + wi;
            const bool valid = n < p.Nimg && ho < p.Ho && wo < p.Wo;
            const long long row = ((long long)n * p.OH + ho * p.oscale + p.cls_ph[cls]) * p.OW + wo * p.oscale + p.cls_pw[cls];
            mbar_wait(&tmem_full_bar[buf], acc_ph);
            tcgen05_fence_after();
            const uint32_t taddr_row = tmem_base + buf * ACC_COLS + ((uint32_t)(lane_grp * 32) << 16);
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                if (n0 + c0 >= p.Cout) break;
                uint32_t rr[32];
                tmem_ld_32(taddr_row + c0, rr);
                tmem_ld_wait();
                if (valid) {
                    const int nn = n0 + c0;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(rr[j]);
                    if (p.bias != nullptr) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (nn + j < p.Cout) v[j] += p.bias[nn + j];
                    }
                    bf16* d = (bf16*)p.D + row * p.ldd + nn;
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        if (nn + j + 8 <= p.Cout) {
                            bf16x8 o = float_to_bf16x8(v + j);
                            *(int4*)(d + j) = *(int4*)&o;
                        } else {
                            for (int q = 0; q < 8; ++q)
                                if (nn + j + q < p.Cout) d[j + q] = __float2bfloat16_rn(v[j + q]);
                        }
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
        }
    }
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------ weight gradient
// item = (co tile, ci tile, tap, split); K loop over 64-pixel patches.  A = dy boxes {64 co, BW, BH, BNI} (two per 128 co),
// B = x boxes {64 ci, BW, BH, BNI} shifted by the tap (BN/64 per tile): both MN-major.
template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ConvParams p) {
    constexpr int A_BYTES = BM * BK * 2;
    constexpr int B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr uint32_t ACC_COLS = BN;
    constexpr uint32_t TMEM_COLS = 2 * ACC_COLS;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = (uint64_t*)(smem + STAGES * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_ptr_smem = (uint32_t*)(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int taps = p.KH * p.KW;
    const int tiles_mn = p.m_tiles * p.n_tiles;
    const int num_items = tiles_mn * taps * p.splits;
    const int total_patches = p.tiles_n * p.tiles_h * p.tiles_w;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tmem_full_bar[b], 1);
            mbar_init(&tmem_empty_bar[b], 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                const int tile = item % tiles_mn, rest = item / tiles_mn;
                const int tap = rest % taps, split = rest / taps;
                const int m0 = (tile / p.n_tiles) * BM, n0 = (tile % p.n_tiles) * BN;
                const int kh = tap / p.KW, kw = tap % p.KW;
                const int pb0 = split * p.patches_per_split;
                int pb1 = pb0 + p.patches_per_split;
                if (pb1 > total_patches) pb1 = total_patches;
                for (int pb = pb0; pb < pb1; ++pb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sa = smem + s * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    const int tw = pb % p.tiles_w, th = (pb / p.tiles_w) % p.tiles_h, tn = pb / (p.tiles_w * p.tiles_h);
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                    tma_load_4d(sa, &tmA, &full_bar[s], m0, tw * p.BW, th * p.BH, tn * p.BNI);
                    tma_load_4d(sa + A_BYTES / 2, &tmA, &full_bar[s], m0 + 64, tw * p.BW, th * p.BH, tn * p.BNI);
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c)
                        tma_load_4d(sb + c * (64 * BK * 2), &tmB, &full_bar[s], n0 + c * 64, tw * p.BW * p.wstride + kw - p.pad,
                                    th * p.BH * p.wstride + kh - p.pad, tn * p.BNI);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BN, 1, 1);
            int it = 0, local = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++local) {
                const int split = (item / tiles_mn) / taps;
                const int pb0 = split * p.patches_per_split;
                int pb1 = pb0 + p.patches_per_split;
                if (pb1 > total_patches) pb1 = total_patches;
                const int buf = local & 1;
                const uint32_t acc_ph = (local >> 1) & 1;
                mbar_wait(&tmem_empty_bar[buf], acc_ph ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + buf * ACC_COLS;
                for (int pb = pb0; pb < pb1; ++pb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
                    const uint32_t sb = sa + A_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        umma_bf16(tmem_d, make_smem_desc(sa + k * 2048, A_BYTES / 2, 1024),
                                  make_smem_desc(sb + k * 2048, 64 * BK * 2, 1024), idesc, (pb > pb0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&tmem_full_bar[buf]);
            }
        }
    } else {
        const int lane_grp = warp & 3;
        int local = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++local) {
            const int tile = item % tiles_mn, rest = item / tiles_mn;
            const int tap = rest % taps;
            const int m0 = (tile / p.n_tiles) * BM, n0 = (tile % p.n_tiles) * BN;
            const int buf = local & 1;
            const uint32_t acc_ph = (local >> 1) & 1;
            const int row = m0 + lane_grp * 32 + lane;  // output channel
            mbar_wait(&tmem_full_bar[buf], acc_ph);
            tcgen05_fence_after();
            const uint32_t taddr_row = tmem_base + buf * ACC_COLS + ((uint32_t)(lane_grp * 32) << 16);
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                if (n0 + c0 >= p.Cin) break;
                uint32_t rr[32];
                tmem_ld_32(taddr_row + c0, rr);
                tmem_ld_wait();
                if (row < p.Cout) {
                    if (p.kcrs) {  // straight into the parameter's [Cout][Cin][KH][KW] gradient: no repacking pass afterwards
                        float* d = (float*)p.D + (long long)row * p.ldd + (long long)(n0 + c0) * taps + tap;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n0 + c0 + j < p.Cin) atomicAdd(d + (long long)j * taps, __uint_as_float(rr[j]));
                    } else {
                        float* d = (float*)p.D + (long long)row * p.ldd + (long long)tap * p.Cin + n0 + c0;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n0 + c0 + j < p.Cin) atomicAdd(d + j, __uint_as_float(rr[j]));
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
        }
    }
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------ host helpers
int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// 4-D bf16 map over a channels-last activation [N, H, W, C] with box {64, bw, bh, bni}
// estride > 1: the box takes every estride-th pixel along W and H (bw x bh pixels land in shared memory either way)
int make_tmap_act(CUtensorMap* tm, const void* base, int N, int H, int W, int C, int bw, int bh, int bni, int estride = 1) {
    PFN_tmapEncodeTiled enc = get_encode_fn();
    DMP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled driver entry point unavailable");
    DMP_REQUIRE(((uintptr_t)base & 15) == 0 && (C % 8) == 0, "conv operand must be 16-byte aligned with C %% 8 == 0");
    DMP_REQUIRE(bw * estride <= 256 && bh * estride <= 256, "conv patch %dx%d with stride %d exceeds the TMA box limit", bw, bh, estride);
    cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t gstride[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)(bw * estride), (cuuint32_t)(bh * estride), (cuuint32_t)bni};
    cuuint32_t estr[4] = {1, (cuuint32_t)estride, (cuuint32_t)estride, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DMP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (4-D activation map) failed with CUresult %d", (int)r);
    return 0;
}

int make_tmap_w(CUtensorMap* tm, const void* base, int rows, long long kdim, int box_rows) {
    PFN_tmapEncodeTiled enc = get_encode_fn();
    DMP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled driver entry point unavailable");
    cuuint64_t gdim[2] = {(cuuint64_t)kdim, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)kdim * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DMP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (weight map) failed with CUresult %d", (int)r);
    return 0;
}

void patch_shape(int Ho, int Wo, int pixels, int* bw, int* bh, int* bni) {
    int w = next_pow2(Wo);
    if (w > pixels) w = pixels;
    int h = next_pow2(Ho);
    if (h > pixels / w) h = pixels / w;
    *bw = w; *bh = h; *bni = pixels / (w * h);
}

template <typename K>
int set_smem(K kern, int bytes) {
    // (all instantiations share one function-pointer type, so a static "already configured" flag here would be shared between
    //  different kernels; the attribute call is cheap and idempotent)
    DMP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    return 0;
}

}  // namespace

namespace {

template <int BN_unused = 0>
int launch_conv_igemm(const CUtensorMap& tmA, const CUtensorMap& tmB, ConvParams& p, int bn, cudaStream_t stream) {
    const long long items = (long long)p.m_tiles * p.n_tiles * p.ncls;
    const int smem_bytes = STAGES * (BM * BK * 2 + bn * BK * 2) + 1024 + 256;
    long long grid = (long long)dmp_num_sms() * (smem_bytes <= 110 * 1024 ? 2 : 1);
    if (grid > items) grid = items;
    if (bn == 128) {
        if (int rc = set_smem(conv_igemm_kernel<128>, smem_bytes)) return rc;
        conv_igemm_kernel<128><<<(int)grid, NUM_THREADS, smem_bytes, stream>>>(tmA, tmB, p);
    } else if (bn == 64) {
        if (int rc = set_smem(conv_igemm_kernel<64>, smem_bytes)) return rc;
        conv_igemm_kernel<64><<<(int)grid, NUM_THREADS, smem_bytes, stream>>>(tmA, tmB, p);
    } else {
        if (int rc = set_smem(conv_igemm_kernel<32>, smem_bytes)) return rc;
        conv_igemm_kernel<32><<<(int)grid, NUM_THREADS, smem_bytes, stream>>>(tmA, tmB, p);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

}  // namespace

// y[N,Ho,Wo,Cout] = conv(x[N,H,W,Cin], wp), KH x KW taps, padding `pad`, stride `stride` (1 or 2): Ho = (H + 2 pad - KH) / stride + 1.
// wp: bf16 [Cout, KH*KW*Cin] (weight permuted to [Cout][kh][kw][Cin]).  Cin % 64 == 0, Cout % 8 == 0, KH*KW <= 49.
// With stride 2 the A-operand boxes take every second pixel (TMA element strides): still no im2col buffer.
DMP_API int dmp_conv_igemm_fwd_strided(const void* x, const void* wp, const float* bias, void* y, int N, int H, int W, int Cin, int Cout,
                                       int KH, int KW, int pad, int stride, cudaStream_t stream) {
    DMP_REQUIRE(Cin % 64 == 0 && Cout % 8 == 0, "dmp_conv_igemm_fwd: needs Cin %% 64 == 0 and Cout %% 8 == 0 (Cin=%d Cout=%d)", Cin, Cout);
    DMP_REQUIRE(stride >= 1 && stride <= 4 && KH * KW <= 49 && pad >= 0, "dmp_conv_igemm_fwd: unsupported geometry (k=%dx%d stride %d)", KH, KW, stride);
    const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
    DMP_REQUIRE(Ho > 0 && Wo > 0, "dmp_conv_igemm_fwd: empty output");
    ConvParams p;
    memset(&p, 0, sizeof(p));
    p.Nimg = N; p.Ho = Ho; p.Wo = Wo; p.Cin = Cin; p.Cout = Cout; p.KH = KH; p.KW = KW; p.pad = pad;
    p.astride = stride; p.ncls = 1; p.oscale = 1; p.OH = Ho; p.OW = Wo;
    p.cls_ntaps[0] = KH * KW;
    for (int kh = 0; kh < KH; ++kh)
        for (int kw = 0; kw < KW; ++kw) {
            const int t = kh * KW + kw;
            p.cls_tap[0][t] = (signed char)t; p.cls_dh[0][t] = (signed char)(kh - pad); p.cls_dw[0][t] = (signed char)(kw - pad);
        }
    patch_shape(Ho, Wo, 128, &p.BW, &p.BH, &p.BNI);
    p.tiles_w = dmp_ceil_div(Wo, p.BW); p.tiles_h = dmp_ceil_div(Ho, p.BH); p.tiles_n = dmp_ceil_div(N, p.BNI);
    p.m_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
    const int bn = Cout > 64 ? 128 : (Cout > 32 ? 64 : 32);
    p.n_tiles = dmp_ceil_div(Cout, bn);
    p.D = y; p.ldd = Cout; p.bias = bias;
    CUtensorMap tmA, tmB;
    if (int rc = make_tmap_act(&tmA, x, N, H, W, Cin, p.BW, p.BH, p.BNI, stride)) return rc;
    if (int rc = make_tmap_w(&tmB, wp, Cout, (long long)KH * KW * Cin, bn)) return rc;
    return launch_conv_igemm(tmA, tmB, p, bn, stream);
}

DMP_API int dmp_conv_igemm_fwd(const void* x, const void* wp, const float* bias, void* y, int N, int H, int W, int Cin, int Cout,
                               int KH, int KW, int pad, cudaStream_t stream) {
    DMP_REQUIRE(KH == 2 * pad + 1 && KW == 2 * pad + 1, "dmp_conv_igemm_fwd: only 'same' stride-1 convolutions (K = 2*pad+1)");
    return dmp_conv_igemm_fwd_strided(x, wp, bias, y, N, H, W, Cin, Cout, KH, KW, pad, 1, stream);
}

// Data gradient of a stride-2 convolution: dx[N,H,W,Cin] from dy[N,H/2,W/2,Cout] (H, W even; Ho = H/2 requires KH = 2 pad + 1 or
// 2 pad + 2).  The four output-pixel parities (h % 2, w % 2) are four stride-1 implicit GEMMs over dy, each with the subset of
// taps whose (h + pad - kh) is even -- 1, 2, 2 and 4 taps for a 3x3 filter -- so no multiplication by an inserted zero is done.
// wd: bf16 [Cin, KH*KW*Cout] = the weight permuted to [Cin][kh][kw][Cout].  Cout % 64 == 0, Cin % 8 == 0.
DMP_API int dmp_conv_igemm_dgrad_s2(const void* dy, const void* wd, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                    int pad, cudaStream_t stream) {
    DMP_REQUIRE(Cout % 64 == 0 && Cin % 8 == 0, "dmp_conv_igemm_dgrad_s2: needs Cout %% 64 == 0 and Cin %% 8 == 0");
    DMP_REQUIRE(H % 2 == 0 && W % 2 == 0 && KH * KW <= 49, "dmp_conv_igemm_dgrad_s2: needs even H, W");
    const int Ho = (H + 2 * pad - KH) / 2 + 1, Wo = (W + 2 * pad - KW) / 2 + 1;
    DMP_REQUIRE(Ho == H / 2 && Wo == W / 2, "dmp_conv_igemm_dgrad_s2: output of the convolution must be H/2 x W/2");
    ConvParams p;
    memset(&p, 0, sizeof(p));
    // "A operand" = dy with Cout channels; "output columns" = Cin
    p.Nimg = N; p.Ho = Ho; p.Wo = Wo; p.Cin = Cout; p.Cout = Cin; p.KH = KH; p.KW = KW; p.pad = pad;
    p.astride = 1; p.ncls = 4; p.oscale = 2; p.OH = H; p.OW = W;
    for (int c = 0; c < 4; ++c) {
        const int ph = c >> 1, pw = c & 1;
        p.cls_ph[c] = ph; p.cls_pw[c] = pw;
        int nt = 0;
        for (int kh = 0; kh < KH; ++kh) {
            if (((ph + pad - kh) % 2 + 2) % 2 != 0) continue;
            for (int kw = 0; kw < KW; ++kw) {
                if (((pw + pad - kw) % 2 + 2) % 2 != 0) continue;
                p.cls_tap[c][nt] = (signed char)(kh * KW + kw);
                // dx[2a+ph] takes dy[ho] with 2 ho + kh - pad = 2a + ph  ->  ho = a + (ph + pad - kh) / 2 (exact; may be negative)
                const int nh = ph + pad - kh, nw = pw + pad - kw;
                p.cls_dh[c][nt] = (signed char)(nh >= 0 ? nh / 2 : -((-nh) / 2));
                p.cls_dw[c][nt] = (signed char)(nw >= 0 ? nw / 2 : -((-nw) / 2));
                ++nt;
            }
        }
        p.cls_ntaps[c] = nt;
    }
    patch_shape(Ho, Wo, 128, &p.BW, &p.BH, &p.BNI);
    p.tiles_w = dmp_ceil_div(Wo, p.BW); p.tiles_h = dmp_ceil_div(Ho, p.BH); p.tiles_n = dmp_ceil_div(N, p.BNI);
    p.m_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
    const int bn = Cin > 64 ? 128 : (Cin > 32 ? 64 : 32);
    p.n_tiles = dmp_ceil_div(Cin, bn);
    p.D = dx; p.ldd = Cin; p.bias = nullptr;
    CUtensorMap tmA, tmB;
    if (int rc = make_tmap_act(&tmA, dy, N, Ho, Wo, Cout, p.BW, p.BH, p.BNI, 1)) return rc;
    if (int rc = make_tmap_w(&tmB, wd, Cin, (long long)KH * KW * Cout, bn)) return rc;
    return launch_conv_igemm(tmA, tmB, p, bn, stream);
}

// dW += sum over output pixels of dy[p, co] * x[p * stride (+) tap, ci].  dy: [N,Ho,Wo,Cout], x: [N,H,W,Cin] bf16 channels-last.
// kcrs = 0: dw is fp32 [Cout][KH*KW][Cin]; kcrs = 1: dw is fp32 [Cout][Cin][KH][KW], the layout of the nn.Conv2d parameter, so the
// kernel accumulates straight into param.grad.  dw must hold the addend (zeros to overwrite).
DMP_API int dmp_conv_igemm_wgrad_strided(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                         int pad, int stride, int kcrs, cudaStream_t stream) {
    DMP_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0, "dmp_conv_igemm_wgrad: channel counts must be multiples of 8");
    DMP_REQUIRE(stride >= 1 && stride <= 4 && KH * KW <= 49, "dmp_conv_igemm_wgrad: unsupported geometry");
    const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
    ConvParams p;
    memset(&p, 0, sizeof(p));
    p.Nimg = N; p.Ho = Ho; p.Wo = Wo; p.Cin = Cin; p.Cout = Cout; p.KH = KH; p.KW = KW; p.pad = pad;
    p.kcrs = kcrs; p.wstride = stride;
    patch_shape(Ho, Wo, 64, &p.BW, &p.BH, &p.BNI);
    p.tiles_w = dmp_ceil_div(Wo, p.BW); p.tiles_h = dmp_ceil_div(Ho, p.BH); p.tiles_n = dmp_ceil_div(N, p.BNI);
    const int bn = Cin > 64 ? 128 : 64;
    p.m_tiles = dmp_ceil_div(Cout, BM);
    p.n_tiles = dmp_ceil_div(Cin, bn);
    p.D = dw; p.ldd = (long long)KH * KW * Cin;
    const int total_patches = p.tiles_w * p.tiles_h * p.tiles_n;
    const int base_items = p.m_tiles * p.n_tiles * KH * KW;
    int splits = dmp_ceil_div(2 * dmp_num_sms(), base_items);
    int max_by_k = total_patches / 4 > 0 ? total_patches / 4 : 1;
    if (splits > max_by_k) splits = max_by_k;
    if (splits < 1) splits = 1;
    p.patches_per_split = dmp_ceil_div(total_patches, splits);
    p.splits = dmp_ceil_div(total_patches, p.patches_per_split);
    CUtensorMap tmA, tmB;
    if (int rc = make_tmap_act(&tmA, dy, N, Ho, Wo, Cout, p.BW, p.BH, p.BNI, 1)) return rc;
    if (int rc = make_tmap_act(&tmB, x, N, H, W, Cin, p.BW, p.BH, p.BNI, stride)) return rc;
    const long long items = (long long)base_items * p.splits;
    const int smem_bytes = STAGES * (BM * BK * 2 + bn * BK * 2) + 1024 + 256;
    long long grid = (long long)dmp_num_sms() * (smem_bytes <= 110 * 1024 ? 2 : 1);
    if (grid > items) grid = items;
    if (bn == 128) {
        if (int rc = set_smem(conv_wgrad_kernel<128>, smem_bytes)) return rc;
        conv_wgrad_kernel<128><<<(int)grid, NUM_THREADS, smem_bytes, stream>>>(tmA, tmB, p);
    } else {
        if (int rc = set_smem(conv_wgrad_kernel<64>, smem_bytes)) return rc;
        conv_wgrad_kernel<64><<<(int)grid, NUM_THREADS, smem_bytes, stream>>>(tmA, tmB, p);
    }
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// stride-1 'same' convolution, dwp in [Cout][kh][kw][Cin] order (kept for callers of the round-1 ABI)
DMP_API int dmp_conv_igemm_wgrad(const void* dy, const void* x, float* dwp, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                 int pad, cudaStream_t stream) {
    DMP_REQUIRE(KH == 2 * pad + 1 && KW == 2 * pad + 1, "dmp_conv_igemm_wgrad: only 'same' stride-1 convolutions");
    return dmp_conv_igemm_wgrad_strided(dy, x, dwp, N, H, W, Cin, Cout, KH, KW, pad, 1, 0, stream);
}
