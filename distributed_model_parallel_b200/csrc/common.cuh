// Shared device/host helpers for the dmp_b200 kernel library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define DMP_API extern "C" __attribute__((visibility("default")))

// ---------------------------------------------------------------------------------------------
// host-side error plumbing: every entry point returns 0 on success or a negative/CUDA error code
// and leaves a human readable message retrievable with dmp_last_error().
// ---------------------------------------------------------------------------------------------
void dmp_set_error(const char* fmt, ...);

#define DMP_CHECK_CUDA(expr)                                                               \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            dmp_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                                \
        }                                                                                  \
    } while (0)

#define DMP_REQUIRE(cond, ...)                                                             \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            dmp_set_error(__VA_ARGS__);                                                    \
            return -1;                                                                     \
        }                                                                                  \
    } while (0)

#define DMP_LAUNCH_CHECK() DMP_CHECK_CUDA(cudaGetLastError())

// launch counter (bench.py reports it as gpu_launches)
extern unsigned long long g_dmp_launch_count;
#define DMP_COUNT_LAUNCH(n) (g_dmp_launch_count += (n))

static inline int dmp_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline int dmp_num_sms() {
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return 148;
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev] = v;
    }
    return sms[dev];
}

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// 16-byte vector of 8 bf16
struct __align__(16) bf16x8 {
    bf162 v[4];
};

__device__ __forceinline__ void bf16x8_to_float(const bf16x8& in, float* out) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 f = __bfloat1622float2(in.v[i]);
        out[2 * i] = f.x;
        out[2 * i + 1] = f.y;
    }
}
__device__ __forceinline__ bf16x8 float_to_bf16x8(const float* in) {
    bf16x8 o;
#pragma unroll
    for (int i = 0; i < 4; ++i) o.v[i] = __floats2bfloat162_rn(in[2 * i], in[2 * i + 1]);
    return o;
}

// streaming (read-once) 16B load / store
__device__ __forceinline__ int4 ld_stream_16(const void* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_16(void* p, const int4& v) {
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

// generic typed load/store to float
template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

#endif  // __CUDACC__
