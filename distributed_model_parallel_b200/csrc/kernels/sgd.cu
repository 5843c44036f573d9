// Fused multi-tensor SGD (momentum + weight decay) over a flat parameter arena, plus flat casts.
// Replaces the per-tensor foreach SGD the reference runs after backward
// (/root/reference/code/distributed_training/data_parallel.py:90-91,111; model_parallel.py:105,131,146).
// One launch updates fp32 master weights, the momentum buffer and (optionally) the bf16 shadow copy
// that the bf16 compute kernels read; the gradient may be fp32 or bf16 and is scaled on the fly.
#include "../common.cuh"

namespace {

// hp = {lr, momentum, weight_decay, grad_scale}; read from device memory so a captured CUDA graph
// sees learning-rate schedule updates without re-capture.
template <typename G>
__global__ void sgd_flat_kernel(float* __restrict__ p, bf16* __restrict__ shadow, G* __restrict__ g,
                                float* __restrict__ mom, long long n, const float* __restrict__ hp, int nesterov,
                                int zero_grad) {
    const float lr = hp[0], mu = hp[1], wd = hp[2], gs = hp[3];
    const long long n4 = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 pv = ((float4*)p)[i];
        float4 mv = ((float4*)mom)[i];
        float gv[4];
        if (sizeof(G) == 4) {
            float4 t = ((const float4*)g)[i];
            gv[0] = t.x; gv[1] = t.y; gv[2] = t.z; gv[3] = t.w;
        } else {
            uint2 t = ((const uint2*)g)[i];
            float2 a = __bfloat1622float2(*(bf162*)&t.x), b = __bfloat1622float2(*(bf162*)&t.y);
            gv[0] = a.x; gv[1] = a.y; gv[2] = b.x; gv[3] = b.y;
        }
        float pa[4] = {pv.x, pv.y, pv.z, pv.w};
        float ma[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float d = gv[j] * gs + wd * pa[j];
            ma[j] = mu * ma[j] + d;
            float upd = nesterov ? (d + mu * ma[j]) : ma[j];
            pa[j] -= lr * upd;
        }
        ((float4*)p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
        ((float4*)mom)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
        if (shadow != nullptr) {
            bf162 lo = __floats2bfloat162_rn(pa[0], pa[1]), hi = __floats2bfloat162_rn(pa[2], pa[3]);
            uint2 o;
            o.x = *(uint32_t*)&lo;
            o.y = *(uint32_t*)&hi;
            ((uint2*)shadow)[i] = o;
        }
        if (zero_grad) {
            if (sizeof(G) == 4) ((float4*)g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            else ((uint2*)g)[i] = make_uint2(0u, 0u);
        }
    }
    // scalar tail
    for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
        float gvv = to_f32<G>(g[i]);
        float d = gvv * gs + wd * p[i];
        float m = mu * mom[i] + d;
        mom[i] = m;
        float upd = nesterov ? (d + mu * m) : m;
        float pn = p[i] - lr * upd;
        p[i] = pn;
        if (shadow != nullptr) shadow[i] = __float2bfloat16_rn(pn);
        if (zero_grad) g[i] = from_f32<G>(0.f);
    }
}

template <typename S, typename D>
__global__ void cast_flat_kernel(const S* __restrict__ s, D* __restrict__ d, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) d[i] = from_f32<D>(to_f32<S>(s[i]));
}

}  // namespace

// grad_dtype: 0 = fp32, 1 = bf16.  hp_dev: device pointer to {lr, momentum, weight_decay, grad_scale}.
DMP_API int dmp_sgd_flat(float* master, void* shadow_bf16, void* grad, int grad_dtype, float* momentum, long long n,
                         const float* hp_dev, int nesterov, int zero_grad, cudaStream_t stream) {
    if (n <= 0) return 0;
    DMP_REQUIRE(master && grad && momentum && hp_dev, "dmp_sgd_flat: null pointer");
    DMP_REQUIRE((((uintptr_t)master | (uintptr_t)momentum | (uintptr_t)grad | (uintptr_t)shadow_bf16) & 15) == 0,
                "dmp_sgd_flat: buffers must be 16-byte aligned");
    int blocks = dmp_ceil_div(n / 4 + 1, 256);
    int cap = dmp_num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (grad_dtype == 0)
        sgd_flat_kernel<float><<<blocks, 256, 0, stream>>>(master, (bf16*)shadow_bf16, (float*)grad, momentum, n,
                                                           hp_dev, nesterov, zero_grad);
    else
        sgd_flat_kernel<bf16><<<blocks, 256, 0, stream>>>(master, (bf16*)shadow_bf16, (bf16*)grad, momentum, n,
                                                          hp_dev, nesterov, zero_grad);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}

// direction: 0 = fp32 -> bf16, 1 = bf16 -> fp32
DMP_API int dmp_cast_flat(const void* src, void* dst, long long n, int direction, cudaStream_t stream) {
    if (n <= 0) return 0;
    int blocks = dmp_ceil_div(n, 256 * 4);
    int cap = dmp_num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (direction == 0) cast_flat_kernel<float, bf16><<<blocks, 256, 0, stream>>>((const float*)src, (bf16*)dst, n);
    else cast_flat_kernel<bf16, float><<<blocks, 256, 0, stream>>>((const bf16*)src, (float*)dst, n);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(1);
    return 0;
}
