// device_on_host.cpp — TEST HARNESS, not product code: compiles the thread-serial building blocks of
// nano-gpu-scheduler_b200/csrc/ngs_device.cuh (the very source the kernels include) with g++, with
// the handful of CUDA intrinsics they use spelled out in plain C++, so that the CPU test suite can
// compare them with the oracle on millions of rows without a GPU (tests/test_device_source_on_host.py).
// The warp-cooperative pieces (choose_group8 and everything in the .cu files) only run on the GPU
// and are covered by the -m gpu tests.  Nothing in the package links or loads this file.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

// ---- the intrinsics ngs_device.cuh uses, for a host compiler (built with -ffp-contract=off) ----
static inline int __ffs(unsigned x) { return x ? __builtin_ctz(x) + 1 : 0; }
static inline int __clz(unsigned x) { return x ? __builtin_clz(x) : 32; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __double2int_rz(double x) { return (int)x; }  // truncation toward zero, like Go's int(float64)
static inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
static inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
static inline double __ddiv_rn(double a, double b) { volatile double r = a / b; return r; }
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline int min(int a, int b) { return a < b ? a : b; }  // CUDA's global-namespace overloads
static inline int max(int a, int b) { return a > b ? a : b; }
static inline uint32_t min(uint32_t a, uint32_t b) { return a < b ? a : b; }
static inline uint32_t max(uint32_t a, uint32_t b) { return a > b ? a : b; }
[[noreturn]] static inline void warp_only() {
    fprintf(stderr, "device_on_host: a warp-cooperative function was called on the host\n");
    abort();
}
static inline unsigned __ballot_sync(unsigned, bool) { warp_only(); }
template <class T> static inline T __shfl_sync(unsigned, T, int) { warp_only(); }
template <class T> static inline T __shfl_xor_sync(unsigned, T, int) { warp_only(); }

#include "../../nano-gpu-scheduler_b200/csrc/ngs_device.cuh"

using namespace ngs;

extern "C" {

// GPUs.Choose for one row as the per-thread kernels evaluate it: Rate (fp64, no table), the exact plan search
// on packed cells, the two shortcut classifiers.  rl = RemainLoad per cell (-2..2).
//   out[0] feasible  out[1] score  out[2] n_emit  out[3] classify_row  out[4] classify_top4  out[5] top4 word
//   plan[C]: container order if feasible, else the partial emission list (INT32_MIN beyond n_emit)
void doh_eval_row(const int32_t* free, int G, const int8_t* rl, const int32_t* percent, int C, int policy, int load_int, int32_t* out,
                  int32_t* plan) {
    SortedDemand dem;
    sort_demand(percent, C, policy, &dem);
    int f[8];
    for (int g = 0; g < 8; g++) f[g] = g < G ? free[g] : 0;
    out[1] = rate_policy(policy, f, G, load_int, nullptr);
    out[3] = classify_row(f, &dem, policy != NGS_POLICY_RANDOM);
    const uint32_t t4 = top4_of(f);
    out[5] = (int32_t)t4;
    out[4] = policy == NGS_POLICY_RANDOM ? FEAS_FALLBACK : classify_top4(t4, &dem, true);
    uint64_t emit = 0;
    int n_emit = 0;
    bool ok;
    if (policy == NGS_POLICY_RANDOM) {
        ok = choose_sample<true>(f, G, &dem, &emit, &n_emit);
    } else {
        uint32_t w[8];
        bool rl0 = true;
        for (int g = 0; g < 8; g++) {
            const int rlc = (g < G ? rl[g] : 0) + 2;
            w[g] = g < G ? pack_cell(f[g], rlc, g) : W_PAD;
            rl0 = rl0 && (g >= G || rlc == 2);
        }
        ok = choose_sorted<true>(w, G, policy == NGS_POLICY_SPREAD, rl0, &dem, &emit, &n_emit);
    }
    out[0] = ok ? 1 : 0;
    out[2] = n_emit;
    write_plan(plan, &dem, policy != NGS_POLICY_RANDOM, ok, emit, n_emit);
}

// feasibility only, through the variant the streaming scan uses (no plan: early exit, no sort for the last container)
int doh_feasible_noplan(const int32_t* free, int G, const int8_t* rl, const int32_t* percent, int C, int policy) {
    SortedDemand dem;
    sort_demand(percent, C, policy, &dem);
    int f[8];
    for (int g = 0; g < 8; g++) f[g] = g < G ? free[g] : 0;
    if (policy == NGS_POLICY_RANDOM) return choose_sample<false>(f, G, &dem, nullptr, nullptr) ? 1 : 0;
    uint32_t w[8];
    bool rl0 = true;
    for (int g = 0; g < 8; g++) {
        const int rlc = (g < G ? rl[g] : 0) + 2;
        w[g] = g < G ? pack_cell(f[g], rlc, g) : W_PAD;
        rl0 = rl0 && (g >= G || rlc == 2);
    }
    return choose_sorted<false>(w, G, policy == NGS_POLICY_SPREAD, rl0, &dem, nullptr, nullptr) ? 1 : 0;
}

// sort.Sort(sortableDemand) + traversal order: d[12], order[12], nz
void doh_sort_demand(const int32_t* percent, int C, int policy, int32_t* d, int32_t* order, int32_t* nz) {
    SortedDemand dem;
    sort_demand(percent, C, policy, &dem);
    memcpy(d, dem.d, sizeof dem.d);
    memcpy(order, dem.order, sizeof dem.order);
    *nz = dem.nz;
}

// GPUResource.LoadUsage over one row.  mode 0: general loop, 1: canonical (core, memory) unrolled, 2: canonical with the
// pair table.  Returns int(loadUsage)/G; rl_out[g] = RemainLoad.
int doh_load_usage(const uint8_t* tenths /*[2][8]*/, const uint32_t* ts /*[2][8]*/, int n_periods, const int32_t* metric,
                   const uint32_t* min_ts, const uint32_t* never, int G, int mode, int32_t* rl_out) {
    LoadParams lp;
    memset(&lp, 0, sizeof lp);
    lp.enabled = 1;
    lp.n_periods = n_periods;
    for (int e = 0; e < n_periods; e++) {
        lp.metric[e] = metric[e];
        lp.min_ts[e] = min_ts[e];
        lp.never[e] = never[e];
    }
    lp.canonical = mode != 0;
    uint8_t t10[2][8];
    uint32_t tt[2][8];
    memcpy(t10, tenths, sizeof t10);
    memcpy(tt, ts, sizeof tt);
    double pair[121];
    for (int i = 0; i < 121; i++) fill_usage_pair_table(pair, i, 121 + i);  // one entry per call: pair[i]
    int rlc[8], load_int = 0;
    load_usage_row(t10, tt, lp, G, rlc, &load_int, mode == 2 ? pair : nullptr);
    for (int g = 0; g < 8; g++) rl_out[g] = rlc[g] - 2;
    return load_int;
}

uint64_t doh_splitmix64(uint64_t x) { return splitmix64(x); }
uint64_t doh_make_key(int score, uint32_t node) { return make_key(score, node); }
uint32_t doh_key_node(uint64_t k) { return key_node(k); }
int doh_key_score(uint64_t k) { return key_score(k); }

}  // extern "C"
