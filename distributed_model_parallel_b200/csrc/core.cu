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

// ---------------------------------------------------------------------------------------------------------
// Stream ordering helpers for the native Reducer launch path (csrc/runtime/reducer.cpp keeps itself free of CUDA and calls
// these through function pointers).  Events come from a small per-device ring: a wait captures the event's state at the time
// of the cudaStreamWaitEvent call, so re-recording an event later is harmless, eagerly and inside a stream capture alike.
// ---------------------------------------------------------------------------------------------------------
#include <mutex>
namespace {
constexpr int EV_RING = 32;
std::mutex g_ev_mu;
cudaEvent_t g_ev[64][EV_RING];
int g_ev_pos[64] = {0};
bool g_ev_init[64] = {false};

int next_event(cudaEvent_t* out) {
    int dev = 0;
    DMP_CHECK_CUDA(cudaGetDevice(&dev));
    DMP_REQUIRE(dev >= 0 && dev < 64, "stream helpers: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_ev_mu);
    if (!g_ev_init[dev]) {
        for (int i = 0; i < EV_RING; ++i) DMP_CHECK_CUDA(cudaEventCreateWithFlags(&g_ev[dev][i], cudaEventDisableTiming));
        g_ev_init[dev] = true;
    }
    *out = g_ev[dev][g_ev_pos[dev]];
    g_ev_pos[dev] = (g_ev_pos[dev] + 1) % EV_RING;
    return 0;
}
}  // namespace

// `comm` waits for everything issued so far on `cur` and on each of the side streams.  While `cur` is being captured, side
// streams that are not part of the capture are skipped (they hold no work of the captured region, and a capturing stream may
// not wait on an event recorded outside its capture).
DMP_API int dmp_stream_fork(cudaStream_t cur, cudaStream_t comm, const cudaStream_t* side, int n_side) {
    if (cur == comm) return 0;
    cudaStreamCaptureStatus cap_cur = cudaStreamCaptureStatusNone;
    DMP_CHECK_CUDA(cudaStreamIsCapturing(cur, &cap_cur));
    cudaEvent_t ev;
    if (int rc = next_event(&ev)) return rc;
    DMP_CHECK_CUDA(cudaEventRecord(ev, cur));
    DMP_CHECK_CUDA(cudaStreamWaitEvent(comm, ev, 0));
    for (int i = 0; i < n_side; ++i) {
        if (side[i] == cur || side[i] == comm) continue;
        if (cap_cur == cudaStreamCaptureStatusActive) {
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            DMP_CHECK_CUDA(cudaStreamIsCapturing(side[i], &cs));
            if (cs != cudaStreamCaptureStatusActive) continue;
        }
        if (int rc = next_event(&ev)) return rc;
        DMP_CHECK_CUDA(cudaEventRecord(ev, side[i]));
        DMP_CHECK_CUDA(cudaStreamWaitEvent(comm, ev, 0));
    }
    return 0;
}

// `cur` waits for everything issued so far on `comm`.
DMP_API int dmp_stream_join(cudaStream_t cur, cudaStream_t comm) {
    if (cur == comm) return 0;
    cudaEvent_t ev;
    if (int rc = next_event(&ev)) return rc;
    DMP_CHECK_CUDA(cudaEventRecord(ev, comm));
    DMP_CHECK_CUDA(cudaStreamWaitEvent(cur, ev, 0));
    return 0;
}
