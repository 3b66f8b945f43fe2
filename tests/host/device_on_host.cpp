// device_on_host.cpp — TEST HARNESS, not product code: compiles the thread-serial building blocks of
// nano-gpu-scheduler_b200/csrc/ngs_device.cuh (the very source the kernels include) with g++, with
// the handful of CUDA intrinsics they use spelled out in plain C++, so that the CPU test suite can
// compare them with the oracle on millions of rows without a GPU (tests/test_device_source_on_host.py).
// The 8-lanes-per-row plan search (choose_group8) runs on a small lockstep warp simulator (32 coroutines,
// ballots and shuffles resolved when every lane has arrived), which also reports what would hang a real
// warp: lanes meeting different sync ops, or returning while others wait.  The kernels themselves (.cu)
// only run on the GPU and are covered by the -m gpu tests.  Nothing in the package links or loads this file.
#include "device_shims.h"

#include "../../nano-gpu-scheduler_b200/csrc/ngs_device.cuh"

using namespace ngs;

// the one-thread evaluation of nodes with 9..12 GPUs, as the kernels include it
#include "../../nano-gpu-scheduler_b200/csrc/ngs_tables.cuh"
#include "../../nano-gpu-scheduler_b200/csrc/ngs_wide.cuh"

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

// GPUs.Choose on one node of up to 12 GPUs through ngs_wide.cuh, on one-row tables built here.
//   tenths / ts: [2][12] usage cells per metric (nullptr: load scheduling off); rl in: the RemainLoad the row holds
//   out[0] feasible  out[1] score  out[2] n_emit; plan[C] as doh_eval_row; free_after[12] / rl_after[12]: the row after the call
//   (free_after = the plan applied when feasible)
void doh_eval_wide_row(const int32_t* free, int G, const int8_t* rl, const int32_t* percent, int C, int policy, const uint8_t* tenths,
                       const uint32_t* ts, int n_periods, const int32_t* metric, const uint32_t* min_ts, const uint32_t* never, int do_rate,
                       int32_t* out, int32_t* plan, int32_t* free_after, int8_t* rl_after) {
    SortedDemand dem;
    sort_demand(percent, C, policy, &dem);
    int32_t f_lo[8] = {0}, f_hi[4] = {0};
    int8_t r_lo[8] = {0}, r_hi[4] = {0};
    uint8_t t_lo[2 * 8], t_hi[2 * 4];
    uint32_t s_lo[2 * 8], s_hi[2 * 4];
    for (int g = 0; g < G; g++) {
        (g < 8 ? f_lo[g] : f_hi[g - 8]) = free[g];
        (g < 8 ? r_lo[g] : r_hi[g - 8]) = rl ? rl[g] : 0;
    }
    for (int m = 0; m < 2; m++)
        for (int g = 0; g < 12; g++) {
            (g < 8 ? t_lo[m * 8 + g] : t_hi[m * 4 + g - 8]) = tenths ? tenths[m * 12 + g] : 255;
            (g < 8 ? s_lo[m * 8 + g] : s_hi[m * 4 + g - 8]) = ts ? ts[m * 12 + g] : 0;
        }
    uint8_t gc = (uint8_t)G;
    WideCtx cx;
    memset(&cx, 0, sizeof(cx));
    Tables& T = cx.T;
    T.free = f_lo; T.free_hi = f_hi; T.rl = r_lo; T.rl_hi = r_hi; T.tenths = t_lo; T.tenths_hi = t_hi; T.ts = s_lo; T.ts_hi = s_hi;
    T.gcount = &gc;
    T.nstride = 1;
    T.n = 1;
    LoadParams& lp = cx.lp;
    lp.enabled = tenths != nullptr;
    lp.n_periods = n_periods;
    for (int e = 0; e < n_periods; e++) {
        lp.metric[e] = metric[e];
        lp.min_ts[e] = min_ts[e];
        lp.never[e] = never[e];
    }
    const RowEval ev = wide_row_eval(&cx, 0, &dem, (do_rate ? WIDE_RATE : 0u) | WIDE_LOAD | WIDE_APPLY);
    out[0] = ev.feasible;
    out[1] = ev.score;
    out[2] = ev.n_emit;
    write_plan(plan, &dem, policy != NGS_POLICY_RANDOM, ev.feasible, ev.emit, ev.n_emit);
    for (int g = 0; g < 12; g++) {
        free_after[g] = g < 8 ? f_lo[g] : f_hi[g - 8];
        rl_after[g] = g < 8 ? r_lo[g] : r_hi[g - 8];
    }
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

// Binpack / Spread.Choose with eight lanes per row (choose_group8), four rows per simulated warp.
//   free[4][8], G[4], rl[4][8] (RemainLoad); one demand for all four rows.
//   ok[4]; with plan != 0 also emit[4] (4 bits per entry), n_emit[4], fr_out[4][8] (cells after GPUs.Allocate)
// Returns 0, or 1 when the simulator saw what would hang a warp (message via doh_warp_error).
struct G8Args {
    const int32_t* free;
    const int32_t* G;
    const int8_t* rl;
    SortedDemand dem;
    bool spread, plan;
    int32_t* ok;
    uint64_t* emit;
    int32_t* n_emit;
    int32_t* fr_out;
};
static void g8_body(int lane, void* p) {
    G8Args& a = *(G8Args*)p;
    const int g = lane & 7, gbase = lane & 24, row = lane >> 3;
    const int G = a.G[row];
    const int fr = a.free[row * 8 + g], rlc = a.rl[row * 8 + g] + 2;
    uint64_t emit = 0;
    int n_emit = 0, fr_after = 0;
    bool ok;
    if (a.plan) ok = choose_group8<true>(g, gbase, fr, rlc, g < G, a.spread, &a.dem, &emit, &n_emit, &fr_after);
    else ok = choose_group8<false>(g, gbase, fr, rlc, g < G, a.spread, &a.dem, nullptr, nullptr, nullptr);
    if (g == 0) {
        a.ok[row] = ok ? 1 : 0;
        if (a.plan) {
            a.emit[row] = emit;
            a.n_emit[row] = n_emit;
        }
    }
    if (a.plan) a.fr_out[row * 8 + g] = g < G ? fr_after : 0;
}
static const char* g_warp_error = "";
const char* doh_warp_error() { return g_warp_error; }
int doh_choose_group8(const int32_t* free, const int32_t* G, const int8_t* rl, const int32_t* percent, int C, int spread, int plan, int32_t* ok,
                      uint64_t* emit, int32_t* n_emit, int32_t* fr_out) {
    G8Args a;
    a.free = free;
    a.G = G;
    a.rl = rl;
    sort_demand(percent, C, spread ? NGS_POLICY_SPREAD : NGS_POLICY_BINPACK, &a.dem);
    a.spread = spread != 0;
    a.plan = plan != 0;
    a.ok = ok;
    a.emit = emit;
    a.n_emit = n_emit;
    a.fr_out = fr_out;
    const char* e = warpsim::run_warp(g8_body, &a);
    g_warp_error = e ? e : "";
    return e ? 1 : 0;
}
// choose_group8_multi: four (row, pod) pairs per simulated warp, each group with its own demand and policy
// (C[j] < 0: the group has nothing to ask).  percent is [4][12].
struct G8MultiArgs {
    const int32_t* free;
    const int32_t* G;
    const int8_t* rl;
    SortedDemand dem[4];
    bool has[4];
    int32_t* ok;
};
static void g8m_body(int lane, void* p) {
    G8MultiArgs& a = *(G8MultiArgs*)p;
    const int g = lane & 7, gbase = lane & 24, row = lane >> 3;
    const bool ok = choose_group8_multi(g, gbase, a.free[row * 8 + g], a.rl[row * 8 + g] + 2, g < a.G[row], a.has[row] ? &a.dem[row] : nullptr);
    if (g == 0) a.ok[row] = ok ? 1 : 0;
}
int doh_choose_group8_multi(const int32_t* free, const int32_t* G, const int8_t* rl, const int32_t* percent, const int32_t* C, const int32_t* spread,
                            int32_t* ok) {
    G8MultiArgs a;
    a.free = free;
    a.G = G;
    a.rl = rl;
    a.ok = ok;
    for (int j = 0; j < 4; j++) {
        a.has[j] = C[j] >= 0;
        if (a.has[j]) sort_demand(percent + 12 * j, C[j], spread[j] ? NGS_POLICY_SPREAD : NGS_POLICY_BINPACK, &a.dem[j]);
    }
    const char* e = warpsim::run_warp(g8m_body, &a);
    g_warp_error = e ? e : "";
    return e ? 1 : 0;
}
// greedy_group8: the position-free greedy search, four rows per simulated warp, one demand for all.
//   state[4]: bit 0 ok, bit 1 ambiguous;  fr_out[4][8]: the cells afterwards (apply != 0)
struct GGArgs {
    const int32_t* free;
    const int32_t* G;
    const int8_t* rl;
    SortedDemand dem;
    bool spread, apply;
    int32_t* state;
    int32_t* fr_out;
};
static void gg_body(int lane, void* p) {
    GGArgs& a = *(GGArgs*)p;
    const int g = lane & 7, gbase = lane & 24, row = lane >> 3;
    int fr = a.free[row * 8 + g];
    const int rlc = a.rl[row * 8 + g] + 2;
    const int st = a.apply ? greedy_group8<true>(g, gbase, fr, rlc, g < a.G[row], a.spread, &a.dem)
                           : greedy_group8<false>(g, gbase, fr, rlc, g < a.G[row], a.spread, &a.dem);
    if (g == 0) a.state[row] = st;
    a.fr_out[row * 8 + g] = fr;
}
int doh_greedy_group8(const int32_t* free, const int32_t* G, const int8_t* rl, const int32_t* percent, int C, int spread, int apply,
                      int32_t* state, int32_t* fr_out) {
    GGArgs a;
    a.free = free;
    a.G = G;
    a.rl = rl;
    sort_demand(percent, C, spread ? NGS_POLICY_SPREAD : NGS_POLICY_BINPACK, &a.dem);
    a.spread = spread != 0;
    a.apply = apply != 0;
    a.state = state;
    a.fr_out = fr_out;
    const char* e = warpsim::run_warp(gg_body, &a);
    g_warp_error = e ? e : "";
    return e ? 1 : 0;
}
// the emission list of choose_group8 in container order, as the kernels publish it (rater.go:102-105)
void doh_unpermute(const int32_t* percent, int C, int spread, uint64_t emit, int32_t* plan) {
    SortedDemand dem;
    sort_demand(percent, C, spread ? NGS_POLICY_SPREAD : NGS_POLICY_BINPACK, &dem);
    for (int k = 0; k < C; k++) {
        const int e = (int)((emit >> (4 * k)) & 0xF);
        plan[dem.order[k]] = e == 0xF ? NGS_NOT_NEED_GPU : e;
    }
}

// two deliberately wrong warp programs: the simulator must report both (they are the bug classes that hang a GPU)
static void bad_short_circuit(int lane, void*) {
    bool ok = (lane & 8) == 0;                                      // lanes disagree ...
    ok = ok && (__ballot_sync(0xFFFFFFFFu, lane < 4) != 0u);        // ... so only some of them reach the ballot
    if (__shfl_xor_sync(0xFFFFFFFFu, ok ? 1 : 0, 1) == 12345) abort();
}
static void bad_early_return(int lane, void*) {
    if (lane == 7) return;
    if (__ballot_sync(0xFFFFFFFFu, true) == 0u) abort();
}
int doh_selftest_detects_hangs() {
    const char* e1 = warpsim::run_warp(bad_short_circuit, nullptr);
    const char* e2 = warpsim::run_warp(bad_early_return, nullptr);
    return (e1 != nullptr ? 1 : 0) | (e2 != nullptr ? 2 : 0);
}

uint64_t doh_splitmix64(uint64_t x) { return splitmix64(x); }
uint64_t doh_make_key(int score, uint32_t node) { return make_key(score, node); }
uint32_t doh_key_node(uint64_t k) { return key_node(k); }
int doh_key_score(uint64_t k) { return key_score(k); }

}  // extern "C"
