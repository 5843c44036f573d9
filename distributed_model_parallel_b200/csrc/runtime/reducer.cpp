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

RT_API void dmp_reducer_clear_plans(void* h);
RT_API void dmp_reducer_destroy(void* h) {
    dmp_reducer_clear_plans(h);
    delete (Reducer*)h;
}

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

// ---------------------------------------------------------------------------------------------------
// Native launch path: the whole gradient-ready hook behind ONE call.
//
// With only the state machine above, every autograd hook still walked Python for the launch decision, the
// stream fork and the collective's argument marshalling (torch does all of this inside c10d's C++ Reducer:
// `Reducer::autograd_hook` -> `mark_variable_ready` -> `mark_bucket_ready` -> `all_reduce_bucket`).  Here each
// bucket carries a launch plan -- which kernel-library entry point reduces it and with which arguments -- and
// dmp_reducer_grad_ready() marks the parameter, and for every bucket that becomes launchable (in bucket order)
// orders the communication stream after the stream backward runs on and the weight-gradient side streams,
// then calls the entry point.  dmp_reducer_finish() joins the communication stream back.
//
// This library stays free of CUDA: the entry points and the two stream helpers arrive as function pointers
// (the binding layer takes them from libdmp_b200.so; the CPU tests pass recording stubs).  A bucket whose plan
// is PLAN_CALLER (communication hook installed, torch.distributed back end) is handed back to the caller; one
// such bucket switches the whole reducer to caller-side launches so that the launch ORDER stays the bucket order.
// ---------------------------------------------------------------------------------------------------
typedef void* rt_stream;

struct DmpLaunchFns {
    int (*fork)(rt_stream cur, rt_stream comm, const rt_stream* side, int n_side);  // comm waits for cur and the side streams
    int (*join)(rt_stream cur, rt_stream comm);                                     // cur waits for comm
    int (*allreduce_fused)(int algo, int dtype, void* bufs, void* state, int rank, int world, long long n, float scale,
                           void* out, void* master, void* momentum, void* shadow, void* hp, int nesterov, void* master_ptrs,
                           void* shadow_ptrs, int zero_after, int max_blocks, rt_stream stream);
    int (*allreduce_nvls)(void* mc, void* state, int rank, int world, long long n, float scale, int max_blocks, rt_stream stream);
    int (*allreduce_nvls_sgd)(void* mc_grad, void* mc_master, void* mc_shadow, void* master, void* momentum, void* hp,
                              int nesterov, void* state, int rank, int world, long long n, float scale, int max_blocks,
                              rt_stream stream);
    int (*sgd_flat)(void* master, void* shadow, void* grad, int grad_dtype, void* momentum, long long n, void* hp, int nesterov,
                    int zero_grad, rt_stream stream);
};

enum { PLAN_CALLER = 0, PLAN_ALLREDUCE_FUSED = 1, PLAN_NVLS = 2, PLAN_NVLS_SGD = 3, PLAN_SGD_FLAT = 4, PLAN_NONE = 5,
       PLAN_NVLS_THEN_SGD = 6 };  // switch-side all-reduce, then the flat SGD kernel on this rank's copy (large buckets)

struct DmpBucketPlan {
    int kind;
    int algo, dtype, rank, world, nesterov, zero_after, max_blocks;
    long long numel;
    float scale;
    void *bufs, *state, *out, *master, *momentum, *shadow, *hp, *master_ptrs, *shadow_ptrs;
    void *mc_grad, *mc_master, *mc_shadow;
    void* grad;
};

struct Engine {
    DmpLaunchFns fns{};
    std::vector<DmpBucketPlan> plans;
    rt_stream comm = nullptr;
    std::vector<rt_stream> side;
    bool native = false;       // every bucket has a native plan
    bool forked = false;       // a native launch happened in this iteration: finish() must join
    long long native_launches = 0;
    int last_rc = 0;
};

static std::mutex g_engine_mu;
static std::map<void*, Engine> g_engines;  // keyed by Reducer*, so the Reducer struct (and its C ABI) stays as it was

static Engine* engine_of(void* h) {
    std::lock_guard<std::mutex> g(g_engine_mu);
    auto it = g_engines.find(h);
    return it == g_engines.end() ? nullptr : &it->second;
}

// Installs the launch plans (n_plans must equal the bucket count).  Returns 1 if the native path is active, 0 if some bucket
// is PLAN_CALLER (all launches then go back to the caller), negative on bad arguments.
RT_API int dmp_reducer_set_plans(void* h, const DmpLaunchFns* fns, const DmpBucketPlan* plans, int n_plans, rt_stream comm,
                                 const rt_stream* side, int n_side) {
    auto* r = (Reducer*)h;
    if (!r || !fns || !plans || n_plans != r->n_buckets || n_side < 0) return -1;
    Engine e;
    e.fns = *fns;
    e.plans.assign(plans, plans + n_plans);
    e.comm = comm;
    if (side) e.side.assign(side, side + n_side);
    e.native = true;
    for (const auto& p : e.plans) {
        if (p.kind == PLAN_CALLER) e.native = false;
        if (p.kind < PLAN_CALLER || p.kind > PLAN_NVLS_THEN_SGD) return -2;
        if ((p.kind == PLAN_ALLREDUCE_FUSED && !fns->allreduce_fused) || (p.kind == PLAN_NVLS && !fns->allreduce_nvls) ||
            (p.kind == PLAN_NVLS_SGD && !fns->allreduce_nvls_sgd) || (p.kind == PLAN_SGD_FLAT && !fns->sgd_flat) ||
            (p.kind == PLAN_NVLS_THEN_SGD && (!fns->allreduce_nvls || !fns->sgd_flat)))
            return -3;
    }
    if (e.native && (!fns->fork || !fns->join)) return -3;
    std::lock_guard<std::mutex> g(g_engine_mu);
    auto it = g_engines.find(h);
    if (it != g_engines.end()) {  // re-installed in the middle of an iteration (the side-stream list changed): keep its progress
        e.forked = it->second.forked;
        e.native_launches = it->second.native_launches;
    }
    g_engines[h] = std::move(e);
    return g_engines[h].native ? 1 : 0;
}

RT_API void dmp_reducer_clear_plans(void* h) {
    std::lock_guard<std::mutex> g(g_engine_mu);
    g_engines.erase(h);
}

static int launch_plan(Engine* e, const DmpBucketPlan& p, rt_stream cur) {
    if (p.kind == PLAN_NONE) return 0;
    int rc = e->fns.fork(cur, e->comm, e->side.empty() ? nullptr : e->side.data(), (int)e->side.size());
    if (rc != 0) return rc;
    e->forked = true;
    switch (p.kind) {
        case PLAN_ALLREDUCE_FUSED:
            rc = e->fns.allreduce_fused(p.algo, p.dtype, p.bufs, p.state, p.rank, p.world, p.numel, p.scale, p.out, p.master,
                                        p.momentum, p.shadow, p.hp, p.nesterov, p.master_ptrs, p.shadow_ptrs, p.zero_after,
                                        p.max_blocks, e->comm);
            break;
        case PLAN_NVLS:
            rc = e->fns.allreduce_nvls(p.mc_grad, p.state, p.rank, p.world, p.numel, p.scale, p.max_blocks, e->comm);
            break;
        case PLAN_NVLS_SGD:
            rc = e->fns.allreduce_nvls_sgd(p.mc_grad, p.mc_master, p.mc_shadow, p.master, p.momentum, p.hp, p.nesterov, p.state,
                                           p.rank, p.world, p.numel, p.scale, p.max_blocks, e->comm);
            break;
        case PLAN_SGD_FLAT:
            rc = e->fns.sgd_flat(p.master, p.shadow, p.grad, p.dtype, p.momentum, p.numel, p.hp, p.nesterov, 1, e->comm);
            break;
        case PLAN_NVLS_THEN_SGD:
            rc = e->fns.allreduce_nvls(p.mc_grad, p.state, p.rank, p.world, p.numel, p.scale, p.max_blocks, e->comm);
            if (rc == 0) {
                e->native_launches++;
                rc = e->fns.sgd_flat(p.master, p.shadow, p.grad, p.dtype, p.momentum, p.numel, p.hp, p.nesterov, 1, e->comm);
            }
            break;
        default:
            rc = -20;
    }
    if (rc == 0) e->native_launches++;
    return rc;
}

// The autograd hook.  Marks parameter p; launches (native plans) or reports (caller plans, ids written to out_caller) every
// bucket that became launchable.  Returns the number of ids written to out_caller (0 on the native path), or a negative
// code: -1 double mark, -2 bad index, -3 no iteration prepared, <= -100: entry point failed with code (-100 - rc') where
// rc' = |rc| of the failing call (its message is in the kernel library's dmp_last_error()).
RT_API int dmp_reducer_grad_ready(void* h, int p, int unused, rt_stream cur, int32_t* out_caller, int cap) {
    Engine* e = engine_of(h);
    int32_t ids[64];
    std::vector<int32_t> big;
    auto* r = (Reducer*)h;
    int32_t* buf = ids;
    int bcap = 64;
    if (r && r->n_buckets > 64) {
        big.resize(r->n_buckets);
        buf = big.data();
        bcap = r->n_buckets;
    }
    const int n = dmp_reducer_mark_ready(h, p, unused, buf, bcap);
    if (n <= 0) return n;
    if (e == nullptr || !e->native) {
        const int m = std::min(n, cap);
        if (out_caller) std::memcpy(out_caller, buf, sizeof(int32_t) * m);
        return n;
    }
    for (int i = 0; i < n; ++i) {
        const int rc = launch_plan(e, e->plans[buf[i]], cur);
        if (rc != 0) {
            e->last_rc = rc;
            return -100 - (rc < 0 ? -rc : rc);
        }
    }
    return 0;
}

// End of backward on the native path: joins the communication stream into `cur` when something was launched, then behaves
// like dmp_reducer_finalize.
RT_API int dmp_reducer_finish(void* h, rt_stream cur, int32_t* out_missing, int cap) {
    Engine* e = engine_of(h);
    if (e != nullptr && e->native && e->forked) {
        e->forked = false;
        const int rc = e->fns.join(cur, e->comm);
        if (rc != 0) {
            e->last_rc = rc;
            return -100 - (rc < 0 ? -rc : rc);
        }
    }
    return dmp_reducer_finalize(h, out_missing, cap);
}

RT_API long long dmp_reducer_native_launches(void* h) {
    Engine* e = engine_of(h);
    return e ? e->native_launches : 0;
}
RT_API int dmp_reducer_plan_bytes() { return (int)sizeof(DmpBucketPlan); }
RT_API int dmp_reducer_fns_bytes() { return (int)sizeof(DmpLaunchFns); }
