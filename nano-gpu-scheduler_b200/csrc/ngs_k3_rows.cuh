// ngs_k3_rows.cuh — K3 / K4, the single-row kernels: bind, apply, batched release, node upsert,
// usage-table updates.  Included by ngs_dealer.cu inside its anonymous namespace.
#pragma once

struct BindResult {
    int32_t status;
    int32_t score;
    int32_t gpu_idx[NGS_MAX_CONTAINERS];
};

// K3.  NodeInfo.Bind (node.go:70-84) for one row, by one thread: cached or fresh plan, GPUs.Allocate, cleanPlan.
// Fills *res (status included) in memory the caller owns.
template <bool WIDE>
__device__ __forceinline__ void bind_row(const Tables& T, uint32_t gidx, int n_containers, const int32_t* percent, const SortedDemand* dem,
                                         const LoadParams& lp, uint32_t gen, const WideCtx* wcx, BindResult* res) {
    const uint32_t row = gidx - T.first;
    for (int j = 0; j < NGS_MAX_CONTAINERS; j++) res->gpu_idx[j] = INT32_MIN;
    res->score = NGS_SCORE_MIN;
    if (gidx < T.first || row >= T.n || T.gcount[row] == 0) {
        res->status = NGS_ERR_UNKNOWN_NODE;
        return;
    }
    const bool hit = T.tag[row] == gen;
    RowEval ev = eval_row_global<true, WIDE>(T, row, dem, lp, !hit, wcx);
    if (!ev.feasible) {  // node.go:74-76
        res->status = NGS_ERR_INFEASIBLE;
        return;
    }
    int32_t planv[NGS_MAX_CONTAINERS];
    write_plan(planv, dem, dem->policy != NGS_POLICY_RANDOM, true, ev.emit, ev.n_emit);
    for (int k = 0; k < n_containers; k++) {  // GPUs.Allocate (allocate.go:102-118); the plan fits by construction
        res->gpu_idx[k] = planv[k];
        if (planv[k] >= 0) *(WIDE ? wide_free_cell(T, row, planv[k]) : T.free + (size_t)row * 8 + planv[k]) -= percent[k];
    }
    res->score = hit ? T.cscore[row] : ev.score;
    T.tag[row] = 0;  // node.go:82
    res->status = NGS_OK;
}

// res_out may be mapped host memory: the status word is written last, behind a system-wide fence, and the host polls it
static_assert(param_block_bytes<Tables, uint32_t, PodDemand, LoadParams, uint32_t, BindResult*>() % 64 >= 20 &&
                  param_block_bytes<Tables, uint32_t, PodDemand, LoadParams, uint32_t, BindResult*>() % 64 <= 44, "see LaunchPad");
template <bool WIDE>
__global__ void k_bind(Tables T, uint32_t gidx, PodDemand pd, LoadParams lp, uint32_t gen, BindResult* res_out) {
    __shared__ SortedDemand dem;
    __shared__ BindResult s_res;
    __shared__ WideCtx s_wcx;
    if (threadIdx.x != 0) return;
    if (WIDE) wide_ctx_fill(&s_wcx, T, lp, true);
    sort_demand(pd.percent, pd.n, pd.policy, &dem);
    bind_row<WIDE>(T, gidx, pd.n, pd.percent, &dem, lp, gen, &s_wcx, &s_res);
    for (int j = 0; j < NGS_MAX_CONTAINERS; j++) res_out->gpu_idx[j] = s_res.gpu_idx[j];
    res_out->score = s_res.score;
    __threadfence_system();
    *reinterpret_cast<volatile int32_t*>(&res_out->status) = s_res.status;
}

// K3.  GPUs.Allocate (sign>0, all-or-nothing) / GPUs.Release (sign<0) with a caller-given plan, by one thread.
__device__ __forceinline__ int apply_row(const Tables& T, uint32_t gidx, int n, const int32_t* percent, const int32_t* gpu_idx_in, int sign) {
    const uint32_t row = gidx - T.first;
    if (gidx < T.first || row >= T.n || T.gcount[row] == 0) return NGS_ERR_UNKNOWN_NODE;
    const int G = min((int)T.gcount[row], WIDE_MAX);
    int f[WIDE_MAX];
    for (int g = 0; g < G; g++) f[g] = *wide_free_cell(T, row, g);
    T.tag[row] = 0;  // node.go:87,92: cleanPlan runs before the apply, whatever its outcome
    int rc = NGS_OK;
    if (sign > 0) {
        for (int k = 0; k < n; k++) {
            const int gi = gpu_idx_in[k];
            if (gi < 0) continue;  // allocate.go:105-107
            if (gi >= G || f[gi] < percent[k]) {  // allocate.go:108 (an index >= G panics in Go)
                rc = NGS_ERR_APPLY;
                break;
            }
            f[gi] -= percent[k];
        }
        if (rc == NGS_OK)
            for (int g = 0; g < G; g++) *wide_free_cell(T, row, g) = f[g];
    } else {
        for (int k = 0; k < n; k++) {
            const int gi = gpu_idx_in[k];
            if (gi < 0) continue;
            if (gi >= G) {  // allocate.go:125-127: earlier containers stay released
                rc = NGS_ERR_APPLY;
                break;
            }
            if (f[gi] + percent[k] > NGS_MAX_FREE) {
                rc = NGS_ERR_RANGE;
                break;
            }
            f[gi] += percent[k];
        }
        if (rc != NGS_ERR_RANGE)
            for (int g = 0; g < G; g++) *wide_free_cell(T, row, g) = f[g];
    }
    return rc;
}
__global__ void k_apply(Tables T, uint32_t gidx, PodDemand pd, const int32_t* __restrict__ gpu_idx_in, int sign, int32_t* status) {
    if (threadIdx.x != 0) return;
    *status = apply_row(T, gidx, pd.n, pd.percent, gpu_idx_in, sign);
}

// K3, batched: GPUs.Release (allocate.go:120-131) for many finished pods at once (the informer's
// completion path, pkg/controller/controller.go:210-243 -> dealer.go:230-255).  Releases only add,
// and additions commute, so one thread per pod with atomic adds gives the same table as releasing
// the pods one after the other.  Per pod: stops at the first GPU index >= G, as the reference does
// (earlier containers stay released), and reports NGS_ERR_APPLY; unknown node: NGS_ERR_UNKNOWN_NODE.
__global__ void k_release_batch(Tables T, uint32_t n, const uint32_t* __restrict__ node, const int32_t* __restrict__ gpu_idx,
                                const int32_t* __restrict__ percent, const int32_t* __restrict__ ncont, int max_c,
                                int32_t* __restrict__ status) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t gidx = node[i], row = gidx - T.first;
    int rc = NGS_OK;
    if (gidx < T.first || row >= T.n || T.gcount[row] == 0) {
        rc = NGS_ERR_UNKNOWN_NODE;
    } else {
        const int G = T.gcount[row];
        T.tag[row] = 0;  // node.go:92: cleanPlan
        const int C = min(max(ncont[i], 0), max_c);
        for (int k = 0; k < C; k++) {
            const int gi = gpu_idx[(size_t)i * max_c + k], pc = percent[(size_t)i * max_c + k];
            if (gi < 0) continue;
            if (gi >= G) {
                rc = NGS_ERR_APPLY;
                break;
            }
            if (pc < 0 || pc > NGS_MAX_FREE) {
                rc = NGS_ERR_INVALID;
                break;
            }
            int32_t* cell = wide_free_cell(T, row, gi);
            const int old = atomicAdd(cell, pc);
            if (old + pc > NGS_MAX_FREE) {  // keep cells inside what the kernels can hold
                atomicSub(cell, pc);
                rc = NGS_ERR_RANGE;
                break;
            }
        }
    }
    if (status) status[i] = rc;
}

// NewNodeInfo (node.go:25-42)
__global__ void k_upsert(Tables T, uint32_t row, int G) {
    if (threadIdx.x != 0) return;
    T.gcount[row] = (uint8_t)G;
    for (int g = 0; g < 8; g++) {
        T.free[(size_t)row * 8 + g] = g < G ? 100 : 0;
        T.rl[(size_t)row * 8 + g] = 0;
    }
    if (T.free_hi)
        for (int g = 8; g < WIDE_MAX; g++) {
            T.free_hi[(size_t)row * 4 + g - 8] = g < G ? 100 : 0;
            T.rl_hi[(size_t)row * 4 + g - 8] = 0;
        }
    T.tag[row] = 0;
}

// K4.  Update{Core,Memory}Usage (nodeusage.go:70-80) on one cell / Add*Usage (nodeusage.go:58-68)
__global__ void k_load_update(Tables T, uint32_t row, int metric, int card, uint8_t tenths, uint32_t ts) {
    if (threadIdx.x != 0) return;
    bool hi;
    const size_t off = wide_usage_off(T, row, metric, card, &hi);
    (hi ? T.tenths_hi : T.tenths)[off] = tenths;
    (hi ? T.ts_hi : T.ts)[off] = ts;
}
__global__ void k_load_add_node(Tables T, uint32_t row, int metric) {
    const size_t off = ((size_t)metric * T.nstride + row) * 8 + threadIdx.x;
    if (threadIdx.x < 8) {
        T.tenths[off] = 255;
        T.ts[off] = 0;
    } else if (threadIdx.x < WIDE_MAX && T.tenths_hi) {
        const size_t off_hi = ((size_t)metric * T.nstride + row) * 4 + threadIdx.x - 8;
        T.tenths_hi[off_hi] = 255;
        T.ts_hi[off_hi] = 0;
    }
}
