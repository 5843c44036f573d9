// Pipeline-parallel host logic: balanced layer partitioning and micro-batch schedule generation/validation.
//
// The reference hard-codes a 3/6/6/2 block split that only works for 4 ranks
// (/root/reference/code/distributed_training/model_parallel.py:102-103,129,143-144) and runs exactly one
// batch in flight (utils.py:34-210).  Here the split is computed from per-layer costs and the per-stage
// order of forward/backward micro-batch steps is generated for three schedules:
//   0 naive : M must be 1 -- the reference behaviour
//   1 gpipe : all forwards, then all backwards
//   2 1f1b  : warm-up of (W-1-s) forwards, steady 1F1B, cool-down backwards
#include <algorithm>
#include <cstdint>
#include <vector>

#define RT_API extern "C" __attribute__((visibility("default")))

// Contiguous partition of n layer costs into W stages minimising the maximum stage cost (DP, O(W n^2)).
// out_bounds has W+1 entries: stage s owns layers [out_bounds[s], out_bounds[s+1]).  Returns 0, or -1 if W > n.
RT_API int dmp_partition_balanced(int n, const double* cost, int W, int32_t* out_bounds) {
    if (W <= 0 || n <= 0 || W > n) return -1;
    std::vector<double> pre(n + 1, 0.0);
    for (int i = 0; i < n; ++i) pre[i + 1] = pre[i] + cost[i];
    const double INF = 1e300;
    std::vector<std::vector<double>> best(W + 1, std::vector<double>(n + 1, INF));
    std::vector<std::vector<int>> cut(W + 1, std::vector<int>(n + 1, -1));
    best[0][0] = 0.0;
    for (int s = 1; s <= W; ++s)
        for (int i = s; i <= n - (W - s); ++i)
            for (int j = s - 1; j < i; ++j) {
                if (best[s - 1][j] >= INF) continue;
                const double v = std::max(best[s - 1][j], pre[i] - pre[j]);
                if (v < best[s][i]) {
                    best[s][i] = v;
                    cut[s][i] = j;
                }
            }
    int i = n;
    out_bounds[W] = n;
    for (int s = W; s >= 1; --s) {
        i = cut[s][i];
        out_bounds[s - 1] = i;
    }
    return 0;
}

// Writes the step order of stage s: +(m+1) = forward of micro-batch m, -(m+1) = backward of micro-batch m.
// Returns the number of steps (2*M) or a negative error.
RT_API int dmp_schedule_generate(int kind, int W, int s, int M, int32_t* out, int cap) {
    if (W <= 0 || s < 0 || s >= W || M <= 0 || cap < 2 * M) return -1;
    int n = 0;
    if (kind == 0) {
        if (M != 1) return -2;
        out[n++] = 1;
        out[n++] = -1;
    } else if (kind == 1) {
        for (int m = 0; m < M; ++m) out[n++] = m + 1;
        for (int m = 0; m < M; ++m) out[n++] = -(m + 1);
    } else if (kind == 2) {
        const int warm = std::min(W - 1 - s, M);
        int f = 0, b = 0;
        for (; f < warm; ++f) out[n++] = f + 1;
        while (f < M) {
            out[n++] = ++f;
            out[n++] = -(++b);
        }
        while (b < M) out[n++] = -(++b);
    } else {
        return -3;
    }
    return n;
}

// Event-driven simulation of all W stages with unit forward cost tf and backward cost tb, blocking
// dependencies F(s,m) <- F(s-1,m), B(s,m) <- B(s+1,m), B(W-1,m) <- F(W-1,m).  Returns the makespan, or a
// negative value if the schedule deadlocks.  out_peak_inflight[s] = max activations stage s holds at once.
RT_API double dmp_schedule_simulate(int kind, int W, int M, double tf, double tb, int32_t* out_peak_inflight) {
    std::vector<std::vector<int32_t>> ops(W, std::vector<int32_t>(2 * M));
    for (int s = 0; s < W; ++s)
        if (dmp_schedule_generate(kind, W, s, M, ops[s].data(), 2 * M) != 2 * M) return -1.0;
    std::vector<std::vector<double>> fdone(W, std::vector<double>(M, -1.0)), bdone(W, std::vector<double>(M, -1.0));
    std::vector<int> pos(W, 0), inflight(W, 0), peak(W, 0);
    std::vector<double> t(W, 0.0);
    int remaining = W * 2 * M;
    while (remaining > 0) {
        bool progressed = false;
        for (int s = 0; s < W; ++s) {
            while (pos[s] < 2 * M) {
                const int op = ops[s][pos[s]];
                const int m = (op > 0 ? op : -op) - 1;
                double dep = 0.0;
                if (op > 0) {
                    if (s > 0) {
                        if (fdone[s - 1][m] < 0) break;
                        dep = fdone[s - 1][m];
                    }
                } else {
                    if (fdone[s][m] < 0) return -2.0;  // backward before its own forward: invalid order
                    if (s < W - 1) {
                        if (bdone[s + 1][m] < 0) break;
                        dep = bdone[s + 1][m];
                    } else {
                        dep = fdone[s][m];
                    }
                }
                const double start = std::max(t[s], dep);
                if (op > 0) {
                    t[s] = start + tf;
                    fdone[s][m] = t[s];
                    peak[s] = std::max(peak[s], ++inflight[s]);
                } else {
                    t[s] = start + tb;
                    bdone[s][m] = t[s];
                    --inflight[s];
                }
                ++pos[s];
                --remaining;
                progressed = true;
            }
        }
        if (!progressed) return -3.0;  // deadlock
    }
    if (out_peak_inflight)
        for (int s = 0; s < W; ++s) out_peak_inflight[s] = peak[s];
    return *std::max_element(t.begin(), t.end());
}
