// Library-wide state: last-error string, launch counter, version / capability probes.
#include "common.cuh"
#include <stdarg.h>

static thread_local char g_err[1024] = "";
unsigned long long g_dmp_launch_count = 0;

void dmp_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

DMP_API const char* dmp_last_error() { return g_err; }
DMP_API unsigned long long dmp_launch_count() { return g_dmp_launch_count; }
DMP_API void dmp_reset_launch_count() { g_dmp_launch_count = 0; }
DMP_API int dmp_version() { return 100; }

// Returns compute capability major*10+minor of the current device, or a negative CUDA error.
DMP_API int dmp_device_cc() {
    int dev = 0, maj = 0, min = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) return -1;
    return maj * 10 + min;
}

// L2 flush helper for benchmarks: writes `bytes` of a caller-provided buffer.
__global__ void l2_flush_kernel(int4* p, long long n16) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    long long stride = (long long)gridDim.x * blockDim.x;
    int4 z = make_int4(0, 0, 0, 0);
    for (; i < n16; i += stride) p[i] = z;
}
DMP_API int dmp_l2_flush(void* buf, long long bytes, cudaStream_t stream) {
    long long n16 = bytes / 16;
    if (n16 <= 0) return 0;
    l2_flush_kernel<<<dmp_num_sms() * 4, 256, 0, stream>>>((int4*)buf, n16);
    DMP_LAUNCH_CHECK();
    return 0;
}
