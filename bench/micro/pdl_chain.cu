// Micro-benchmark: a chain of small dependent kernels replayed from a CUDA graph, with ordinary stream-order edges versus
// programmatic dependent launch (griddepcontrol.launch_dependents at entry, griddepcontrol.wait before the first read).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pdl_chain pdl_chain.cu && ./pdl_chain
#include <cuda_runtime.h>
#include <stdio.h>
#include <vector>

template <bool PDL>
__global__ void __launch_bounds__(256) step_kernel(const float4* __restrict__ in, float4* __restrict__ out, long long n4, float s) {
    if (PDL) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __shared__ float sh[256];
    sh[threadIdx.x] = s * threadIdx.x;   // stand-in for a prologue that does not depend on the previous kernel
    __syncthreads();
    float add = sh[(threadIdx.x + 1) & 255] * 0.f + s;
    if (PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 v = in[i];
        v.x += add; v.y += add; v.z += add; v.w += add;
        out[i] = v;
    }
}

template <bool PDL>
static void launch(const float4* in, float4* out, long long n4, int grid, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = PDL ? 1 : 0;
    cudaLaunchKernelEx(&cfg, step_kernel<PDL>, in, out, n4, 1.0f);
}

template <bool PDL>
static float run(float4* a, float4* b, long long n4, int grid, int chain, cudaStream_t st) {
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < chain; ++i) launch<PDL>((i & 1) ? b : a, (i & 1) ? a : b, n4, grid, st);
    cudaStreamEndCapture(st, &g);
    if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { printf("instantiate failed: %s\n", cudaGetErrorString(cudaGetLastError())); return -1; }
    for (int i = 0; i < 3; ++i) cudaGraphLaunch(ge, st);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
    for (int i = 0; i < 10; ++i) cudaGraphLaunch(ge, st);
    cudaEventRecord(e1, st);
    cudaStreamSynchronize(st);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
    return ms / 10 / chain * 1000.f;  // us per kernel
}

int main() {
    cudaStream_t st; cudaStreamCreate(&st);
    const int chain = 200;
    for (long long bytes : {65536LL, 1LL << 20, 4LL << 20, 16LL << 20, 64LL << 20}) {
        long long n4 = bytes / 16;
        float4 *a, *b; cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMemset(a, 0, bytes);
        for (int grid : {148, 296, 1184}) {
            float t0 = run<false>(a, b, n4, grid, chain, st);
            float t1 = run<true>(a, b, n4, grid, chain, st);
            printf("bytes %9lld grid %5d: plain %.2f us/kernel, pdl %.2f us/kernel (saves %.2f)\n", bytes, grid, t0, t1, t0 - t1);
        }
        // correctness of the PDL chain: every element must equal `chain` * 10 replays ... (value check on one element)
        cudaFree(a); cudaFree(b);
    }
    // value check: 1 MB, one graph launch of the PDL chain from zeros
    long long bytes = 1 << 20, n4 = bytes / 16;
    float4 *a, *b; cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMemset(a, 0, bytes); cudaMemset(b, 0, bytes);
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < chain; ++i) launch<true>((i & 1) ? b : a, (i & 1) ? a : b, n4, 296, st);
    cudaGraph_t g; cudaGraphExec_t ge; cudaStreamEndCapture(st, &g); cudaGraphInstantiate(&ge, g, 0);
    cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
    std::vector<float> h(n4 * 4); cudaMemcpy(h.data(), a, bytes, cudaMemcpyDeviceToHost);
    int bad = 0; for (float v : h) bad += (v != (float)chain);
    printf("value check: %d wrong of %lld (expect %d everywhere), last error %s\n", bad, n4 * 4, chain, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
