// k6_on_host.cpp — TEST HARNESS, not product code: the warp-cooperative building blocks of the batch kernel
// (nano-gpu-scheduler_b200/csrc/ngs_k6_batch.cuh: filter_pass, warp_top2, fresh_key_of_row, plan_group) compiled
// for the host and run on the lockstep warp simulator of device_shims.h, one simulated warp owning all rows of a
// small "CTA" in ordinary memory.  The kernel itself (k_schedule, a template here never instantiated) and its
// grid-wide exchange only run on the GPU.  See tests/test_device_source_on_host.py.
#include "device_shims.h"

#include <vector>

#include <cuda_runtime.h>
static uint3 threadIdx, blockIdx;
static dim3 blockDim, gridDim;
namespace cooperative_groups {}

#include "../../nano-gpu-scheduler_b200/csrc/ngs_device.cuh"
namespace cg = cooperative_groups;
using namespace ngs;
namespace {
#include "../../nano-gpu-scheduler_b200/csrc/ngs_rows.cuh"
#include "../../nano-gpu-scheduler_b200/csrc/ngs_k6_batch.cuh"
}  // namespace

namespace {
// the rows of one CTA as k_schedule's prologue lays them out in shared memory (restated here: that code is part of the kernel)
struct HostRows {
    std::vector<uint2> fA, fB;
    std::vector<uint32_t> meta, top4, feasbits;
    std::vector<int32_t> scB, scS;
    SmemRows S;
    HostRows(const int32_t* free, const int32_t* G, const int8_t* rl, const int32_t* load_int, uint32_t n) {
        fA.resize(n); fB.resize(n); meta.resize(n); top4.resize(n); scB.resize(n); scS.resize(n);
        feasbits.assign(2 * ((n + 31) / 32) + 2, 0);
        for (uint32_t r = 0; r < n; r++) {
            int f[8];
            for (int g = 0; g < 8; g++) f[g] = g < G[r] ? free[r * 8 + g] : 0;
            uint32_t m = ((uint32_t)G[r] << 24) | ((uint32_t)load_int[r] << 28);
            for (int c = 0; c < 8; c++) m |= (uint32_t)(c < G[r] ? rl[r * 8 + c] + 2 : 2) << (3 * c);
            fA[r].x = (uint32_t)f[0] | ((uint32_t)f[1] << 16); fA[r].y = (uint32_t)f[2] | ((uint32_t)f[3] << 16);
            fB[r].x = (uint32_t)f[4] | ((uint32_t)f[5] << 16); fB[r].y = (uint32_t)f[6] | ((uint32_t)f[7] << 16);
            meta[r] = m;
            scB[r] = G[r] ? rate_binpack(f, G[r], load_int[r], nullptr) : 0;
            scS[r] = G[r] ? rate_spread(f, G[r], load_int[r]) : 0;
            top4[r] = top4_of(f);
        }
        S.fA = fA.data(); S.fB = fB.data(); S.meta = meta.data(); S.scB = scB.data(); S.scS = scS.data(); S.top4 = top4.data();
        S.feasbits = feasbits.data();
    }
};

struct FilterArgs {
    HostRows* rows;
    SortedDemand dem;
    uint32_t n, node0;
    uint64_t m1, m2;
    uint32_t cnt;
};
void filter_body(int lane, void* p) {
    FilterArgs& a = *(FilterArgs*)p;
    uint64_t k1 = 0, k2 = 0;
    uint32_t cnt = 0;
    const uint32_t WT = 32, n_iter = (a.n + WT - 1) / WT;
    filter_pass<false>(a.rows->S, &a.dem, (uint32_t)lane, WT, n_iter, a.n, a.node0, (uint32_t)lane, lane & 7, lane & 24, a.rows->S.feasbits, k1, k2, cnt,
                       nullptr, nullptr, 0u);
    if (a.dem.policy != NGS_POLICY_RANDOM) {
        uint64_t m1, m2;
        warp_top2(k1, k2, m1, m2);
        if (lane == 0) { a.m1 = m1; a.m2 = m2; }
    } else if (lane == 0) {
        a.cnt = cnt;
    }
}

struct PlanArgs {
    HostRows* rows;
    BatchParams bp;
    SortedDemand dem, dem_next;
    bool has_next;
    uint32_t r, node;
    GroupCommit cm[32];
    uint64_t fresh_only[32];
    int fresh_only_ok[32];
};
void plan_body(int lane, void* p) {
    PlanArgs& a = *(PlanArgs*)p;
    GroupCommit cm;
    memset(&cm, 0, sizeof cm);
    plan_group(a.bp, a.rows->S, a.r, a.node, &a.dem, a.has_next ? &a.dem_next : nullptr, lane & 7, lane & 24, cm);
    a.cm[lane] = cm;
    bool ok = false;
    a.fresh_only[lane] = fresh_key_of_row(a.rows->S, a.r, a.node, &a.dem, lane & 7, lane & 24, &ok);  // the row as it stands, this pod
    a.fresh_only_ok[lane] = ok ? 1 : 0;
}
const char* g_err = "";
}  // namespace

extern "C" {
const char* k6h_error() { return g_err; }

// filter_pass + warp_top2 of one warp over n rows: the warp's best two keys (arg-max policies) or the feasibility
// bits and their count (random policy).  Returns 0, or 1 if the simulated warp would hang.
int k6h_filter(const int32_t* free, const int32_t* G, const int8_t* rl, const int32_t* load_int, uint32_t n, uint32_t node0,
               const int32_t* percent, int C, int policy, uint64_t* m1, uint64_t* m2, uint32_t* cnt, uint32_t* bits) {
    HostRows rows(free, G, rl, load_int, n);
    FilterArgs a;
    a.rows = &rows;
    sort_demand(percent, C, policy, &a.dem);
    a.n = n;
    a.node0 = node0;
    a.m1 = a.m2 = 0;
    a.cnt = 0;
    const char* e = warpsim::run_warp(filter_body, &a);
    g_err = e ? e : "";
    *m1 = a.m1;
    *m2 = a.m2;
    *cnt = a.cnt;
    for (uint32_t w = 0; w < (n + 31) / 32; w++) bits[w] = rows.feasbits[w];
    return e ? 1 : 0;
}

// plan_group on row r for `percent`, with the next pod's demand (C_next < 0: none), and fresh_key_of_row for `percent` on
// the row as it stands.  out per call: fr[8], emit, n_emit, top4, scB, scS, fresh, fresh_ok, key_now, ok_now; `same` = 1 when
// all 32 lanes agree on everything that is not per-cell.
int k6h_plan(const int32_t* free, const int32_t* G, const int8_t* rl, const int32_t* load_int, uint32_t n, uint32_t r, uint32_t node,
             const int32_t* percent, int C, int policy, const int32_t* percent_next, int C_next, int policy_next, int32_t* fr, uint64_t* emit,
             int32_t* n_emit, uint32_t* top4, int32_t* scB, int32_t* scS, uint64_t* fresh, int32_t* fresh_ok, uint64_t* key_now,
             int32_t* ok_now, int32_t* same) {
    HostRows rows(free, G, rl, load_int, n);
    static PlanArgs a;
    a.rows = &rows;
    memset(&a.bp, 0, sizeof a.bp);
    sort_demand(percent, C, policy, &a.dem);
    a.has_next = C_next >= 0;
    if (a.has_next) sort_demand(percent_next, C_next, policy_next, &a.dem_next);
    a.r = r;
    a.node = node;
    const char* e = warpsim::run_warp(plan_body, &a);
    g_err = e ? e : "";
    if (e) return 1;
    for (int g = 0; g < 8; g++) fr[g] = a.cm[g].fr;
    *emit = a.cm[0].emit; *n_emit = a.cm[0].n_emit; *top4 = a.cm[0].top4; *scB = a.cm[0].scB; *scS = a.cm[0].scS;
    *fresh = a.cm[0].fresh; *fresh_ok = a.cm[0].fresh_ok; *key_now = a.fresh_only[0]; *ok_now = a.fresh_only_ok[0];
    int s = 1;
    for (int l = 0; l < 32; l++) {
        const GroupCommit& c = a.cm[l];
        s &= c.fr == a.cm[l & 7].fr && c.emit == *emit && c.n_emit == *n_emit && c.top4 == *top4 && c.scB == *scB && c.scS == *scS &&
             c.fresh == *fresh && c.fresh_ok == *fresh_ok && a.fresh_only[l] == *key_now && a.fresh_only_ok[l] == *ok_now;
    }
    *same = s;
    return 0;
}
// eval_row_values (ngs_rows.cuh): the whole GPUs.Choose of the per-verb kernels for n rows, one after the other —
// Rate (with LoadUsage and the RemainLoad write-back when load != 0), shortcuts, exact search.
//   tenths / ts: [2][n][8] as the device tables hold them; min_ts[2]: the (core, memory) thresholds of make_load
//   plan != 0: PLAN instantiation (plans / partial emission lists), else the feasibility-only one
void k6h_eval_rows(const int32_t* free, const int32_t* G, int8_t* rl, const uint8_t* tenths, const uint32_t* ts, uint32_t n,
                   const int32_t* percent, int C, int policy, int load, const uint32_t* min_ts, int use_pair, int plan, int32_t* feasible,
                   int32_t* score, int32_t* plan_out, int32_t* n_emit) {
    SortedDemand dem;
    sort_demand(percent, C, policy, &dem);
    LoadParams lp;
    memset(&lp, 0, sizeof lp);
    if (load) {
        lp.enabled = 1;
        lp.n_periods = 2;
        lp.metric[0] = NGS_METRIC_CORE;
        lp.metric[1] = NGS_METRIC_MEMORY;
        lp.min_ts[0] = min_ts[0];
        lp.min_ts[1] = min_ts[1];
        lp.canonical = 1;
    }
    double pair[121];
    for (int i = 0; i < 121; i++) fill_usage_pair_table(pair, i, 121 + i);
    Tables T;
    memset(&T, 0, sizeof T);
    T.rl = rl;
    T.n = n;
    T.nstride = n;
    for (uint32_t r = 0; r < n; r++) {
        RowIn in;
        memset(&in, 0, sizeof in);
        in.G = G[r];
        for (int g = 0; g < 8; g++) in.v[g] = (uint32_t)free[r * 8 + g];
        for (int m = 0; m < 2; m++) {
            const uint8_t* t = tenths + ((size_t)m * n + r) * 8;
            in.t8[m].x = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
            in.t8[m].y = t[4] | (t[5] << 8) | (t[6] << 16) | ((uint32_t)t[7] << 24);
            for (int g = 0; g < 8; g++) in.ts[m][g] = ts[((size_t)m * n + r) * 8 + g];
        }
        const int8_t* rr = rl + (size_t)r * 8;
        in.r8.x = (uint8_t)rr[0] | ((uint8_t)rr[1] << 8) | ((uint8_t)rr[2] << 16) | ((uint32_t)(uint8_t)rr[3] << 24);
        in.r8.y = (uint8_t)rr[4] | ((uint8_t)rr[5] << 8) | ((uint8_t)rr[6] << 16) | ((uint32_t)(uint8_t)rr[7] << 24);
        RowEval ev = plan ? eval_row_values<true>(T, r, in, &dem, lp, true, use_pair ? pair : nullptr, nullptr)
                          : eval_row_values<false>(T, r, in, &dem, lp, true, use_pair ? pair : nullptr, nullptr);
        const bool ok = ev.known && ev.feasible;
        feasible[r] = ok ? 1 : 0;
        score[r] = ok ? ev.score : NGS_SCORE_MIN;
        if (plan) {
            int32_t* out = plan_out + (size_t)r * (C > 0 ? C : 1);
            if (ev.known) write_plan(out, &dem, dem.policy != NGS_POLICY_RANDOM, ok, ev.emit, ev.n_emit);
            else for (int j = 0; j < C; j++) out[j] = INT32_MIN;
            n_emit[r] = ev.known ? ev.n_emit : 0;
        }
    }
}

// the emission list in container order, as publish_group writes it (rater.go:102-105)
void k6h_unpermute(const int32_t* percent, int C, int policy, uint64_t emit, int32_t* plan) {
    SortedDemand dem;
    sort_demand(percent, C, policy, &dem);
    for (int k = 0; k < C; k++) {
        const int e = (int)((emit >> (4 * k)) & 0xF);
        plan[dem.order[k]] = e == 0xF ? NGS_NOT_NEED_GPU : e;
    }
}
uint32_t k6h_key_node(uint64_t k) { return key_node(k); }
int k6h_key_score(uint64_t k) { return key_score(k); }
}  // extern "C"
