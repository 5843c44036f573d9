// Fused softmax cross-entropy forward + backward + top-1/top-k accuracy.
// One pass produces what the reference computes with nn.CrossEntropyLoss()(logits, targets),
// loss.backward() and its accuracy() helper
// (/root/reference/code/distributed_training/utils.py:58-65,215-229; data_parallel.py:109-115):
//   row_loss[b]  = logsumexp(x_b) - x_b[t_b]
//   dlogits[b,:] = (softmax(x_b) - onehot(t_b)) * grad_scale          (grad_scale = 1/B for 'mean')
//   out[0] = mean loss, out[1] = #rows whose argmax == target, out[2] = #rows with target in top-5
// One warp per row; the final reduction over rows is a second single-block kernel so the result is
// deterministic (no float atomics).
#include "../common.cuh"

namespace {

template <typename T>
__global__ void xent_rows_kernel(const T* __restrict__ logits, const long long* __restrict__ targets,
                                 T* __restrict__ dlogits, float* __restrict__ row_loss, float* __restrict__ row_top1,
                                 float* __restrict__ row_top5, int B, int C, float grad_scale) {
    const int warps_per_block = blockDim.x >> 5;
    const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= B) return;
    const T* x = logits + (long long)row * C;
    const int t = (int)targets[row];
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, to_f32<T>(x[c]));
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(to_f32<T>(x[c]) - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    const float xt = (t >= 0 && t < C) ? to_f32<T>(x[t]) : 0.f;
    // rank of the target logit: number of classes strictly greater (ties broken towards lower index, like topk)
    float greater = 0.f;
    for (int c = lane; c < C; c += 32) {
        float v = to_f32<T>(x[c]);
        greater += (v > xt || (v == xt && c < t)) ? 1.f : 0.f;
    }
    greater = warp_sum(greater);
    if (dlogits != nullptr) {
        T* dx = dlogits + (long long)row * C;
        const float inv = 1.f / se;
        for (int c = lane; c < C; c += 32) {
            float pr = __expf(to_f32<T>(x[c]) - mx) * inv;
            dx[c] = from_f32<T>((pr - (c == t ? 1.f : 0.f)) * grad_scale);
        }
    }
    if (lane == 0) {
        row_loss[row] = lse - xt;
        row_top1[row] = greater < 0.5f ? 1.f : 0.f;
        row_top5[row] = greater < 4.5f ? 1.f : 0.f;
    }
}

__global__ void xent_finalize_kernel(const float* __restrict__ row_loss, const float* __restrict__ row_top1,
                                     const float* __restrict__ row_top5, float* __restrict__ out, int B) {
    __shared__ float sh[3][32];
    float a = 0.f, b = 0.f, c = 0.f;
    for (int i = threadIdx.x; i < B; i += blockDim.x) {
        a += row_loss[i];
        b += row_top1[i];
        c += row_top5[i];
    }
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sh[0][w] = a; sh[1][w] = b; sh[2][w] = c; }
    __syncthreads();
    if (w == 0) {
        const int nw = blockDim.x >> 5;
        a = l < nw ? sh[0][l] : 0.f;
        b = l < nw ? sh[1][l] : 0.f;
        c = l < nw ? sh[2][l] : 0.f;
        a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
        if (l == 0) { out[0] = a / (float)B; out[1] = b; out[2] = c; }
    }
}

}  // namespace

// dtype: 0 = fp32 logits, 1 = bf16 logits. `rows` is scratch/aux of 3*B floats (per-row loss, top1, top5 flags).
// `out` receives {mean loss, top-1 hits, top-5 hits}. dlogits may be null (evaluation).
DMP_API int dmp_xent_fwd_bwd(const void* logits, const long long* targets, void* dlogits, float* rows, float* out, int B,
                             int C, int dtype, float grad_scale, cudaStream_t stream) {
    DMP_REQUIRE(B > 0 && C > 0, "dmp_xent_fwd_bwd: empty input");
    const int wpb = 8;
    dim3 grid(dmp_ceil_div(B, wpb));
    if (dtype == 0)
        xent_rows_kernel<float><<<grid, wpb * 32, 0, stream>>>((const float*)logits, targets, (float*)dlogits, rows,
                                                               rows + B, rows + 2 * B, B, C, grad_scale);
    else
        xent_rows_kernel<bf16><<<grid, wpb * 32, 0, stream>>>((const bf16*)logits, targets, (bf16*)dlogits, rows, rows + B,
                                                              rows + 2 * B, B, C, grad_scale);
    DMP_LAUNCH_CHECK();
    xent_finalize_kernel<<<1, 256, 0, stream>>>(rows, rows + B, rows + 2 * B, out, B);
    DMP_LAUNCH_CHECK();
    DMP_COUNT_LAUNCH(2);
    return 0;
}
