// Host-side gradient Reducer bookkeeping (no device code): bucket assignment by size, per-bucket pending
// counts driven from autograd hooks, strict in-order bucket launch, gradient-ready order capture for the
// one-time bucket rebuild, unused-parameter accounting.
//
// Behavioural model: torch's c10d Reducer as used by DistributedDataParallel
// (torch/nn/parallel/distributed.py:1183-1304 `_ddp_init_helper`, `_compute_bucket_assignment_by_size`,
// `Reducer.prepare_for_backward/autograd_hook/finalize`), which the reference README studies
// (/root/reference/Readme.md:145-157).  Tensors never cross this ABI: Python owns the flat bucket
// buffers and launches the collective; this object only answers "which buckets may launch now".
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#define RT_API extern "C" __attribute__((visibility("default")))

RT_API int dmp_rt_version() { return 100; }

// ---------------------------------------------------------------------------------------------------
// Bucket assignment.  Parameters are visited in the given order.  Each (dtype, device) key fills its own
// open bucket; a bucket closes when its byte size reaches the key's current limit, and the key then moves
// on to its next limit (the last limit repeats).  Buckets are finally ordered by their smallest parameter
// index.  out_bucket_of_param[i] receives the bucket id of parameter i; returns the bucket count.
// ---------------------------------------------------------------------------------------------------
RT_API int dmp_bucket_assign(int n, const int64_t* nbytes, const int32_t* keys, const int64_t* limits, int n_limits,
                             int32_t* out_bucket_of_param) {
    if (n <= 0 || n_limits <= 0) return 0;
    struct Open {
        std::vector<int> idx;
        int64_t bytes = 0;
        int limit_pos = 0;
    };
    std::map<int32_t, Open> open;
    std::vector<std::vector<int>> closed;
    for (int i = 0; i < n; ++i) {
        Open& o = open[keys ? keys[i] : 0];
        o.idx.push_back(i);
        o.bytes += nbytes[i];
        const int64_t lim = limits[std::min(o.limit_pos, n_limits - 1)];
        if (o.bytes >= lim) {
            closed.push_back(std::move(o.idx));
            o.idx.clear();
            o.bytes = 0;
            if (o.limit_pos < n_limits - 1) ++o.limit_pos;
        }
    }
    for (auto& kv : open)
        if (!kv.second.idx.empty()) closed.push_back(std::move(kv.second.idx));
    std::sort(closed.begin(), closed.end(), [](const std::vector<int>& a, const std::vector<int>& b) {
        return *std::min_element(a.begin(), a.end()) < *std::min_element(b.begin(), b.end());
    });
    for (size_t b = 0; b < closed.size(); ++b)
        for (int i : closed[b]) out_bucket_of_param[i] = (int32_t)b;
    return (int)closed.size();
}

// ---------------------------------------------------------------------------------------------------
// Reducer state machine
// ---------------------------------------------------------------------------------------------------
struct Reducer {
    int n_params = 0, n_buckets = 0;
    std::vector<int32_t> bucket_of;     // param -> bucket
    std::vector<int32_t> bucket_size;   // params per bucket
    std::vector<int32_t> pending;       // params not yet ready, per bucket
    std::vector<uint8_t> ready;         // per param: 0 = pending, 1 = ready(grad), 2 = ready(unused)
    std::vector<uint8_t> launched;      // per bucket
    std::vector<int32_t> ready_order;   // params in the order they became ready this iteration
    int next_bucket = 0;                // lowest bucket index that has not launched
    int n_launched = 0;
    bool expect_hooks = false;          // between prepare_for_backward and finalize
    long long iterations = 0;
    std::mutex mu;                      // hooks may fire from autograd worker threads
};

RT_API void* dmp_reducer_create(int n_params, int n_buckets, const int32_t* bucket_of_param) {
    if (n_params <= 0 || n_buckets <= 0) return nullptr;
    auto* r = new Reducer();
    r->n_params = n_params;
    r->n_buckets = n_buckets;
    r->bucket_of.assign(bucket_of_param, bucket_of_param + n_params);
    r->bucket_size.assign(n_buckets, 0);
    for (int i = 0; i < n_params; ++i) {
        if (r->bucket_of[i] < 0 || r->bucket_of[i] >= n_buckets) {
            delete r;
            return nullptr;
        }
        r->bucket_size[r->bucket_of[i]]++;
    }
    r->pending = r->bucket_size;
    r->ready.assign(n_params, 0);
    r->launched.assign(n_buckets, 0);
    return r;
}

RT_API void dmp_reducer_destroy(void* h) { delete (Reducer*)h; }

// Start of an iteration that will synchronise gradients.
RT_API void dmp_reducer_prepare(void* h) {
    auto* r = (Reducer*)h;
    std::lock_guard<std::mutex> g(r->mu);
    r->pending = r->bucket_size;
    std::fill(r->ready.begin(), r->ready.end(), 0);
    std::fill(r->launched.begin(), r->launched.end(), 0);
    r->ready_order.clear();
    r->next_bucket = 0;
    r->n_launched = 0;
    r->expect_hooks = true;
}

// Marks parameter `p` ready (unused != 0: it did not take part in this backward).  Writes the ids of the
// buckets that become launchable, in launch order, to out (capacity cap) and returns how many.
// Returns -1 on a double mark, -2 on a bad index, -3 if no iteration is prepared.
RT_API int dmp_reducer_mark_ready(void* h, int p, int unused, int32_t* out, int cap) {
    auto* r = (Reducer*)h;
    std::lock_guard<std::mutex> g(r->mu);
    if (!r->expect_hooks) return -3;
    if (p < 0 || p >= r->n_params) return -2;
    if (r->ready[p]) return -1;
    r->ready[p] = unused ? 2 : 1;
    r->ready_order.push_back(p);
    const int b = r->bucket_of[p];
    r->pending[b]--;
    int n = 0;
    // in-order launch: a complete bucket waits for every lower-indexed bucket (all ranks launch the same
    // sequence of collectives, which is what keeps the in-kernel barriers deadlock-free)
    while (r->next_bucket < r->n_buckets && r->pending[r->next_bucket] == 0) {
        r->launched[r->next_bucket] = 1;
        r->n_launched++;
        if (n < cap) out[n] = r->next_bucket;
        ++n;
        ++r->next_bucket;
    }
    return n;
}

// End of backward.  Returns the number of buckets that never launched (parameters that did not receive a
// gradient and were not marked unused) and writes up to cap such parameter indices to out_missing.
RT_API int dmp_reducer_finalize(void* h, int32_t* out_missing, int cap) {
    auto* r = (Reducer*)h;
    std::lock_guard<std::mutex> g(r->mu);
    r->expect_hooks = false;
    r->iterations++;
    int m = 0;
    for (int p = 0; p < r->n_params; ++p)
        if (!r->ready[p]) {
            if (m < cap) out_missing[m] = p;
            ++m;
        }
    return m;
}

RT_API int dmp_reducer_num_launched(void* h) { return ((Reducer*)h)->n_launched; }
RT_API long long dmp_reducer_iterations(void* h) { return ((Reducer*)h)->iterations; }

// Copies the gradient-ready order of the last iteration (used to rebuild buckets after iteration 1 so that
// bucket 0 holds the parameters whose gradients arrive first).  Returns the count.
RT_API int dmp_reducer_ready_order(void* h, int32_t* out, int cap) {
    auto* r = (Reducer*)h;
    std::lock_guard<std::mutex> g(r->mu);
    const int n = (int)std::min<size_t>(r->ready_order.size(), (size_t)cap);
    std::memcpy(out, r->ready_order.data(), sizeof(int32_t) * n);
    return (int)r->ready_order.size();
}

// Per-parameter unused flags of the last iteration (1 = marked unused).
RT_API void dmp_reducer_unused_flags(void* h, int32_t* out) {
    auto* r = (Reducer*)h;
    std::lock_guard<std::mutex> g(r->mu);
    for (int p = 0; p < r->n_params; ++p) out[p] = r->ready[p] == 2 ? 1 : 0;
}
