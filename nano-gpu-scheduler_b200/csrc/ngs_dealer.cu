// ngs_dealer.cu — CUDA kernels (sm_100a) and the C ABI of the B200-native dealer.
//
// Snapshot layout in HBM (struct-of-arrays, one row per node, rows padded to 8 GPUs):
//   free    int32 [N][8]   GPUResource.Percent per cell (32 B row = 2 x int4)   allocate.go:141-145
//   gcount  uint8 [N]      len(NodeInfo.GPUs); 0 = node unknown                 node.go:25-42
//   rl      int8  [N][8]   GPUResource.RemainLoad (written by load-aware Rate)  allocate.go:193
//   tenths  uint8 [2][N][8] ceil(10*usage) per (metric, node, card); 255 = skip  nodeusage.go:82-111
//   ts      uint32[2][N][8] update time, unix seconds                            stats.go:30-55
//   tag     uint32[N], cscore int32[N]   per-node plan cache (NodeInfo.PlanCache, node.go:22)
//
// Kernels:
//   k_assume     K1  one thread per candidate row: Rate + Choose, feasibility / score / plan out
//   k_schedule   K6  persistent cooperative kernel, rows resident in shared memory, per pod:
//                    filter+score every row (K1), block + grid arg-max (K2), bind on the owner (K3)
//   k_bind, k_apply  K3  single-row plan application
//   k_load_*     K4  usage-table updates
// All paths relative to /root/reference/pkg/dealer/.
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <pthread.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "ngs_device.cuh"

namespace cg = cooperative_groups;
using namespace ngs;

namespace {

struct Tables {
    int32_t* free;
    uint8_t* gcount;
    int8_t* rl;
    uint8_t* tenths;
    uint32_t* ts;
    uint32_t* tag;
    int32_t* cscore;
    const int16_t* blut;  // Binpack.Rate percent table [9][801] (nullptr: compute in fp64)
    uint32_t rl_zero;     // the rl table is known to hold only zeros (no load-aware pass ran yet): skip reading it
    uint32_t nstride;  // rows per metric plane of tenths / ts (>= n; padded so that every plane is 256-row aligned)
    uint32_t n;      // rows in this shard
    uint32_t first;  // global node index of row 0
};

struct RowEval {
    bool known;
    bool feasible;
    int score;
    uint64_t emit;
    int n_emit;
};

__device__ __forceinline__ int4 ld_row4(const int32_t* p) { return __ldcg(reinterpret_cast<const int4*>(p)); }
// one 256-bit load per 32-byte row (LDG.E.256 on sm_100a): every sector a warp touches is fully used
__device__ __forceinline__ void ld_row8(const void* p, uint32_t (&v)[8]) {
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "l"(p));
}
// exact plan search kept out of line: the scan's common path (shortcuts decide) stays small
// (arguments by value — four words of u16 pairs and the 3-bit RemainLoad codes — so that the caller's
// row stays in registers)
__device__ __noinline__ bool choose_sorted_slow(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3, uint32_t rlmeta, int G, bool spread,
                                                const SortedDemand* dem) {
    const uint32_t pk[4] = {p0, p1, p2, p3};
    uint32_t w[8];
#pragma unroll
    for (int g = 0; g < 8; g++) {
        const int fr = (int)((pk[g >> 1] >> (16 * (g & 1))) & 0xFFFFu);
        w[g] = g < G ? pack_cell(fr, (int)((rlmeta >> (3 * g)) & 7u), g) : W_PAD;
    }
    return choose_sorted<false>(w, G, spread, (rlmeta & 0xFFFFFFu) == RL0_META, dem, nullptr, nullptr);
}
__global__ void k_build_rate_lut(int16_t* lut) {
    const int G = blockIdx.x;  // 0..8
    for (int used = threadIdx.x; used < RATE_LUT_STRIDE; used += blockDim.x)
        lut[G * RATE_LUT_STRIDE + used] = (G > 0 && used <= 100 * G) ? (int16_t)binpack_percent_fp64(used, G) : (int16_t)0;
}

// One full GPUs.Choose (allocate.go:92-100) on a row read from the global tables.
// do_rate=false is the plan-cache hit of NodeInfo.Bind (node.go:72-78): the plan is re-derived
// from the RemainLoad the filter pass left behind, without running Rate / LoadUsage again.
// everything one evaluation reads of its row, as it arrives from memory
struct RowIn {
    int G;
    uint32_t v[8];      // free
    uint2 t8[2];        // tenths per metric (load-aware pass only)
    uint32_t ts[2][8];  // update times per metric (load-aware pass only)
    uint2 r8;           // RemainLoad (other passes; zeros when the table is known to hold only zeros)
};

template <bool PLAN>
__device__ __forceinline__ RowEval eval_row_values(const Tables& T, uint32_t row, const RowIn& in, const SortedDemand* dem,
                                                   const LoadParams& lp, bool do_rate, const double* pair) {
    RowEval ev;
    ev.known = false;
    ev.feasible = false;
    ev.score = NGS_SCORE_MIN;
    ev.emit = 0;
    ev.n_emit = 0;
    const int policy = dem->policy;
    const bool load_path = lp.enabled && do_rate && policy != NGS_POLICY_RANDOM;
    const int G = in.G;
    const uint32_t (&v)[8] = in.v;
    const uint2 (&t8)[2] = in.t8;
    const uint32_t (&ts)[2][8] = in.ts;
    const uint2 r8 = in.r8;
    if (G == 0) return ev;
    ev.known = true;
    int f[8];
#pragma unroll
    for (int g = 0; g < 8; g++) f[g] = (int)v[g];
    int rlc[8];
    int load_int = 0;
    if (load_path) {
        // rater.go:62-66 / 116-120: LoadUsage per GPU, which also rewrites RemainLoad
        uint8_t tenths[2][8];
#pragma unroll
        for (int m = 0; m < 2; m++) {
#pragma unroll
            for (int g = 0; g < 4; g++) {
                tenths[m][g] = (uint8_t)(t8[m].x >> (8 * g));
                tenths[m][g + 4] = (uint8_t)(t8[m].y >> (8 * g));
            }
        }
        load_usage_row(tenths, ts, lp, G, rlc, &load_int, pair);
        uint2 out;
        out.x = out.y = 0;
#pragma unroll
        for (int g = 0; g < 4; g++) {
            out.x |= (uint32_t)(uint8_t)(int8_t)(g < G ? rlc[g] - 2 : 0) << (8 * g);
            out.y |= (uint32_t)(uint8_t)(int8_t)(g + 4 < G ? rlc[g + 4] - 2 : 0) << (8 * g);
        }
        *reinterpret_cast<uint2*>(T.rl + (size_t)row * 8) = out;
    } else {  // (r8 is all zeros when the table is known to be)
#pragma unroll
        for (int g = 0; g < 4; g++) {
            rlc[g] = (int)(int8_t)(r8.x >> (8 * g)) + 2;
            rlc[g + 4] = (int)(int8_t)(r8.y >> (8 * g)) + 2;
        }
    }
    if (do_rate) ev.score = rate_policy(policy, f, G, load_int, T.blut);
    if (!PLAN) {
        const int cls = classify_row(f, dem, policy != NGS_POLICY_RANDOM);
        if (cls != FEAS_UNDECIDED) {
            ev.feasible = cls == FEAS_YES;
            return ev;
        }
        if (policy != NGS_POLICY_RANDOM) {
            uint32_t rlmeta = 0;
#pragma unroll
            for (int g = 0; g < 8; g++) rlmeta |= (uint32_t)(g < G ? rlc[g] : 2) << (3 * g);
            ev.feasible = choose_sorted_slow((uint32_t)f[0] | ((uint32_t)f[1] << 16), (uint32_t)f[2] | ((uint32_t)f[3] << 16),
                                             (uint32_t)f[4] | ((uint32_t)f[5] << 16), (uint32_t)f[6] | ((uint32_t)f[7] << 16), rlmeta, G,
                                             policy == NGS_POLICY_SPREAD, dem);
            return ev;
        }
    }
    if (policy == NGS_POLICY_RANDOM) {
        ev.feasible = choose_sample<PLAN>(f, G, dem, &ev.emit, &ev.n_emit);
    } else {
        uint32_t w[8];
        bool rl0 = true;
#pragma unroll
        for (int g = 0; g < 8; g++) {
            w[g] = g < G ? pack_cell(f[g], rlc[g], g) : W_PAD;
            rl0 = rl0 && (g >= G || rlc[g] == 2);
        }
        ev.feasible = choose_sorted<PLAN>(w, G, policy == NGS_POLICY_SPREAD, rl0, dem, &ev.emit, &ev.n_emit);
    }
    return ev;
}

// One full GPUs.Choose (allocate.go:92-100) on a row read from the global tables.
// do_rate=false is the plan-cache hit of NodeInfo.Bind (node.go:72-78): the plan is re-derived
// from the RemainLoad the filter pass left behind, without running Rate / LoadUsage again.
template <bool PLAN>
__device__ __forceinline__ RowEval eval_row_global(const Tables& T, uint32_t row, const SortedDemand* dem, const LoadParams& lp,
                                                   bool do_rate, const double* pair = nullptr) {
    // every load of the row is issued before the first use, so that they are all in flight together
    const bool load_path = lp.enabled && do_rate && dem->policy != NGS_POLICY_RANDOM;
    RowIn in;
    in.G = T.gcount[row];
    ld_row8(T.free + (size_t)row * 8, in.v);
    in.r8 = make_uint2(0u, 0u);
    if (load_path) {
#pragma unroll
        for (int m = 0; m < 2; m++) {
            const size_t off = ((size_t)m * T.nstride + row) * 8;
            in.t8[m] = __ldcg(reinterpret_cast<const uint2*>(T.tenths + off));
            ld_row8(T.ts + off, in.ts[m]);
        }
    } else if (!T.rl_zero) {
        in.r8 = __ldcg(reinterpret_cast<const uint2*>(T.rl + (size_t)row * 8));
    }
    return eval_row_values<PLAN>(T, row, in, dem, lp, do_rate, pair);
}

enum { MODE_ASSUME = 0, MODE_SCORE = 1 };

// K1.  DealerImpl.Assume / DealerImpl.Score fan-out (dealer.go:109-134, 138-153): one thread
// per candidate.  MODE_ASSUME recomputes every candidate (cleanPlan + Assume, dealer.go:123-124);
// MODE_SCORE serves NodeInfo.Score (node.go:59-68) from the per-node cache when it holds a
// plan for this demand generation.
#ifndef NGS_K1_BLOCKS
#define NGS_K1_BLOCKS 4  // resident 256-thread blocks per SM for the feasibility/score scan
#endif
template <bool PLAN>
__global__ void __launch_bounds__(256, PLAN ? 2 : NGS_K1_BLOCKS) k_assume(Tables T, const uint32_t* __restrict__ idx, uint32_t n, PodDemand pd, LoadParams lp,
                                                int mode, uint32_t gen, uint8_t* __restrict__ feasible,
                                                int8_t* __restrict__ reason, int32_t* __restrict__ score,
                                                int32_t* __restrict__ plan, int32_t* __restrict__ n_emitted) {
    __shared__ SortedDemand dem;
    __shared__ double s_pair[121];
    if (threadIdx.x == 0) sort_demand(pd.percent, pd.n, pd.policy, &dem);
    if (lp.enabled) fill_usage_pair_table(s_pair, threadIdx.x, blockDim.x);
    __syncthreads();
    const int C = pd.n;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint32_t gidx = idx ? idx[k] : T.first + k;
        const uint32_t row = gidx - T.first;
        RowEval ev;
        ev.known = false;
        ev.feasible = false;
        ev.score = NGS_SCORE_MIN;
        ev.emit = 0;
        ev.n_emit = 0;
        bool cached = false;
        if (gidx >= T.first && row < T.n) {
            if (mode == MODE_SCORE && T.tag[row] == gen && T.gcount[row] != 0) {
                ev.known = ev.feasible = cached = true;
                ev.score = T.cscore[row];
            } else {
                ev = eval_row_global<PLAN>(T, row, &dem, lp, true, s_pair);
            }
        }
        const bool ok = ev.known && ev.feasible;
        // The per-node plan cache only matters when load scheduling is on (the usage tables may move
        // between filter and prioritize); without it a later Score / Bind re-derives the same values
        // from the unchanged row, so the scan does not spend 8 bytes per node on the cache.
        if (ev.known && !cached && lp.enabled) {
            if (ok) {  // node.go:55
                T.tag[row] = gen;
                T.cscore[row] = ev.score;
            } else if (mode == MODE_ASSUME) {
                T.tag[row] = 0;  // dealer.go:123 cleared it, the failed Assume caches nothing
            }
        }
        if (feasible) feasible[k] = ok ? 1 : 0;
        if (reason) reason[k] = !ev.known ? NGS_REASON_UNKNOWN_NODE : (ok ? NGS_REASON_OK : NGS_REASON_INFEASIBLE);
        if (score) score[k] = ok ? ev.score : NGS_SCORE_MIN;
        if (PLAN) {
            if (plan) {
                int32_t* out = plan + (size_t)k * C;
                if (ev.known) write_plan(out, &dem, dem.policy != NGS_POLICY_RANDOM, ok, ev.emit, ev.n_emit);
                else for (int j = 0; j < C; j++) out[j] = INT32_MIN;
            }
            if (n_emitted) n_emitted[k] = ev.known ? ev.n_emit : 0;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K1, streaming form: the same evaluation as k_assume for "all rows, no per-node plans", with the
// row data staged through shared memory by the TMA engine (1-D bulk copies, cp.async.bulk +
// mbarrier transaction counts) in a 2- or 3-stage ring of 256-row tiles, so that several tiles of every
// table are in flight per CTA while the threads evaluate the tile that has landed.
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_TILE = 256;
#ifndef NGS_SCAN_STAGES_LOAD
#define NGS_SCAN_STAGES_LOAD 2
#endif
#ifndef NGS_SCAN_STAGES_PLAIN
#define NGS_SCAN_STAGES_PLAIN 3
#endif
#ifndef NGS_SCAN_BLOCKS_LOAD
#define NGS_SCAN_BLOCKS_LOAD 3
#endif
constexpr uint32_t SCAN_FREE_BYTES = SCAN_TILE * 8 * 4;    // 8 KB
constexpr uint32_t SCAN_GC_BYTES = SCAN_TILE;              // 256 B
constexpr uint32_t SCAN_RL_BYTES = SCAN_TILE * 8;          // 2 KB   (pass without load scheduling)
constexpr uint32_t SCAN_TENTHS_BYTES = SCAN_TILE * 8;      // 2 KB per metric (load-aware pass)
constexpr uint32_t SCAN_TS_BYTES = SCAN_TILE * 8 * 4;      // 8 KB per metric (load-aware pass)
template <bool LOAD>
__host__ __device__ constexpr uint32_t scan_stage_bytes() {
    return SCAN_FREE_BYTES + SCAN_GC_BYTES + (LOAD ? 2 * SCAN_TENTHS_BYTES + 2 * SCAN_TS_BYTES : SCAN_RL_BYTES);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

template <bool LOAD>
__global__ void __launch_bounds__(SCAN_TILE, LOAD ? NGS_SCAN_BLOCKS_LOAD : 4) k_scan_tma(Tables T, uint32_t n, PodDemand pd, LoadParams lp, uint32_t gen,
                                                                        uint8_t* __restrict__ feasible, int8_t* __restrict__ reason,
                                                                        int32_t* __restrict__ score) {
    extern __shared__ __align__(128) unsigned char tiles[];
    __shared__ SortedDemand dem;
    __shared__ double s_pair[121];
    constexpr int SCAN_STAGES = LOAD ? NGS_SCAN_STAGES_LOAD : NGS_SCAN_STAGES_PLAIN;
    __shared__ __align__(8) uint64_t bars[SCAN_STAGES];
    constexpr uint32_t STAGE = scan_stage_bytes<LOAD>();
    const uint32_t tid = threadIdx.x;
    const uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    const bool want_rl = !LOAD && !T.rl_zero;

    // one thread issues the copies of a tile: every table's 256-row slice is one contiguous range
    auto issue = [&](uint32_t tile, uint32_t stage) {
        unsigned char* base = tiles + (size_t)stage * STAGE;
        const size_t row = (size_t)tile * SCAN_TILE;
        uint32_t bytes = SCAN_FREE_BYTES + SCAN_GC_BYTES;
        if (LOAD) bytes += 2 * SCAN_TENTHS_BYTES + 2 * SCAN_TS_BYTES;
        else if (want_rl) bytes += SCAN_RL_BYTES;
        mbar_expect_tx(&bars[stage], bytes);
        tma_load_1d(base, T.free + row * 8, SCAN_FREE_BYTES, &bars[stage]);
        tma_load_1d(base + SCAN_FREE_BYTES, T.gcount + row, SCAN_GC_BYTES, &bars[stage]);
        unsigned char* q = base + SCAN_FREE_BYTES + SCAN_GC_BYTES;
        if (LOAD) {
#pragma unroll
            for (int m = 0; m < 2; m++) {
                const size_t off = ((size_t)m * T.nstride + row) * 8;
                tma_load_1d(q + m * SCAN_TENTHS_BYTES, T.tenths + off, SCAN_TENTHS_BYTES, &bars[stage]);
                tma_load_1d(q + 2 * SCAN_TENTHS_BYTES + m * SCAN_TS_BYTES, T.ts + off, SCAN_TS_BYTES, &bars[stage]);
            }
        } else if (want_rl) {
            tma_load_1d(q, T.rl + row * 8, SCAN_RL_BYTES, &bars[stage]);
        }
    };

    if (tid == 0) {
        for (int st = 0; st < SCAN_STAGES; st++) mbar_init(&bars[st], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        sort_demand(pd.percent, pd.n, pd.policy, &dem);
    }
    if (LOAD) fill_usage_pair_table(s_pair, tid, blockDim.x);
    __syncthreads();
    if (tid == 0)
        for (uint32_t i = 0; i < SCAN_STAGES; i++) {
            const uint32_t tile = blockIdx.x + i * gridDim.x;
            if (tile < ntiles) issue(tile, i);
        }

    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
        const uint32_t stage = it % SCAN_STAGES, parity = (it / SCAN_STAGES) & 1u;
        mbar_wait(&bars[stage], parity);
        const unsigned char* base = tiles + (size_t)stage * STAGE;
        const uint32_t row = tile * SCAN_TILE + tid;
        RowIn in;
        {
            const uint4 a = *reinterpret_cast<const uint4*>(base + (size_t)tid * 32);
            const uint4 b = *reinterpret_cast<const uint4*>(base + (size_t)tid * 32 + 16);
            in.v[0] = a.x; in.v[1] = a.y; in.v[2] = a.z; in.v[3] = a.w;
            in.v[4] = b.x; in.v[5] = b.y; in.v[6] = b.z; in.v[7] = b.w;
        }
        in.G = base[SCAN_FREE_BYTES + tid];
        in.r8 = make_uint2(0u, 0u);
        const unsigned char* q = base + SCAN_FREE_BYTES + SCAN_GC_BYTES;
        if (LOAD) {
#pragma unroll
            for (int m = 0; m < 2; m++) {
                in.t8[m] = *reinterpret_cast<const uint2*>(q + m * SCAN_TENTHS_BYTES + (size_t)tid * 8);
                const unsigned char* tsp = q + 2 * SCAN_TENTHS_BYTES + m * SCAN_TS_BYTES + (size_t)tid * 32;
                const uint4 a = *reinterpret_cast<const uint4*>(tsp), b = *reinterpret_cast<const uint4*>(tsp + 16);
                in.ts[m][0] = a.x; in.ts[m][1] = a.y; in.ts[m][2] = a.z; in.ts[m][3] = a.w;
                in.ts[m][4] = b.x; in.ts[m][5] = b.y; in.ts[m][6] = b.z; in.ts[m][7] = b.w;
            }
        } else if (want_rl) {
            in.r8 = *reinterpret_cast<const uint2*>(q + (size_t)tid * 8);
        }
        __syncthreads();  // the stage has been read by everyone: refill it with the tile 3 steps ahead
        if (tid == 0) {
            const uint32_t next = tile + SCAN_STAGES * gridDim.x;
            if (next < ntiles) issue(next, stage);
        }
        if (row < n) {
            const RowEval ev = eval_row_values<false>(T, row, in, &dem, lp, true, s_pair);
            const bool ok = ev.known && ev.feasible;
            if (ev.known && LOAD) {  // plan cache, as in k_assume (MODE_ASSUME)
                T.tag[row] = ok ? gen : 0u;
                if (ok) T.cscore[row] = ev.score;
            }
            if (feasible) feasible[row] = ok ? 1 : 0;
            if (reason) reason[row] = !ev.known ? NGS_REASON_UNKNOWN_NODE : (ok ? NGS_REASON_OK : NGS_REASON_INFEASIBLE);
            if (score) score[row] = ok ? ev.score : NGS_SCORE_MIN;
        }
    }
}

struct BindResult {
    int32_t status;
    int32_t score;
    int32_t gpu_idx[NGS_MAX_CONTAINERS];
};

// K3.  NodeInfo.Bind (node.go:70-84) for one row: cached or fresh plan, GPUs.Allocate, cleanPlan.
__global__ void k_bind(Tables T, uint32_t gidx, PodDemand pd, LoadParams lp, uint32_t gen, BindResult* res) {
    __shared__ SortedDemand dem;
    if (threadIdx.x != 0) return;
    sort_demand(pd.percent, pd.n, pd.policy, &dem);
    const uint32_t row = gidx - T.first;
    for (int j = 0; j < NGS_MAX_CONTAINERS; j++) res->gpu_idx[j] = INT32_MIN;
    res->score = NGS_SCORE_MIN;
    if (gidx < T.first || row >= T.n || T.gcount[row] == 0) {
        res->status = NGS_ERR_UNKNOWN_NODE;
        return;
    }
    const bool hit = T.tag[row] == gen;
    RowEval ev = eval_row_global<true>(T, row, &dem, lp, !hit);
    if (!ev.feasible) {  // node.go:74-76
        res->status = NGS_ERR_INFEASIBLE;
        return;
    }
    int32_t planv[NGS_MAX_CONTAINERS];
    write_plan(planv, &dem, dem.policy != NGS_POLICY_RANDOM, true, ev.emit, ev.n_emit);
    int32_t* f = T.free + (size_t)row * 8;
    for (int k = 0; k < pd.n; k++) {  // GPUs.Allocate (allocate.go:102-118); the plan fits by construction
        res->gpu_idx[k] = planv[k];
        if (planv[k] >= 0) f[planv[k]] -= pd.percent[k];
    }
    res->score = hit ? T.cscore[row] : ev.score;
    T.tag[row] = 0;  // node.go:82
    res->status = NGS_OK;
}

// K3.  GPUs.Allocate (sign>0, all-or-nothing) / GPUs.Release (sign<0) with a caller-given plan.
__global__ void k_apply(Tables T, uint32_t gidx, PodDemand pd, const int32_t* __restrict__ gpu_idx_in, int sign, int32_t* status) {
    if (threadIdx.x != 0) return;
    const uint32_t row = gidx - T.first;
    if (gidx < T.first || row >= T.n || T.gcount[row] == 0) {
        *status = NGS_ERR_UNKNOWN_NODE;
        return;
    }
    const int G = T.gcount[row];
    int32_t* fp = T.free + (size_t)row * 8;
    int f[8];
    for (int g = 0; g < 8; g++) f[g] = fp[g];
    T.tag[row] = 0;  // node.go:87,92: cleanPlan runs before the apply, whatever its outcome
    int rc = NGS_OK;
    if (sign > 0) {
        for (int k = 0; k < pd.n; k++) {
            const int gi = gpu_idx_in[k];
            if (gi < 0) continue;  // allocate.go:105-107
            if (gi >= G || f[gi] < pd.percent[k]) {  // allocate.go:108 (an index >= G panics in Go)
                rc = NGS_ERR_APPLY;
                break;
            }
            f[gi] -= pd.percent[k];
        }
        if (rc == NGS_OK)
            for (int g = 0; g < 8; g++) fp[g] = f[g];
    } else {
        for (int k = 0; k < pd.n; k++) {
            const int gi = gpu_idx_in[k];
            if (gi < 0) continue;
            if (gi >= G) {  // allocate.go:125-127: earlier containers stay released
                rc = NGS_ERR_APPLY;
                break;
            }
            if (f[gi] + pd.percent[k] > NGS_MAX_FREE) {
                rc = NGS_ERR_RANGE;
                break;
            }
            f[gi] += pd.percent[k];
        }
        if (rc != NGS_ERR_RANGE)
            for (int g = 0; g < 8; g++) fp[g] = f[g];
    }
    *status = rc;
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
            const int old = atomicAdd(&T.free[(size_t)row * 8 + gi], pc);
            if (old + pc > NGS_MAX_FREE) {  // keep cells inside what the kernels can hold
                atomicSub(&T.free[(size_t)row * 8 + gi], pc);
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
    T.tag[row] = 0;
}

// K4.  Update{Core,Memory}Usage (nodeusage.go:70-80) on one cell / Add*Usage (nodeusage.go:58-68)
__global__ void k_load_update(Tables T, uint32_t row, int metric, int card, uint8_t tenths, uint32_t ts) {
    if (threadIdx.x != 0) return;
    const size_t off = ((size_t)metric * T.nstride + row) * 8 + card;
    T.tenths[off] = tenths;
    T.ts[off] = ts;
}
__global__ void k_load_add_node(Tables T, uint32_t row, int metric) {
    const size_t off = ((size_t)metric * T.nstride + row) * 8 + threadIdx.x;
    if (threadIdx.x < 8) {
        T.tenths[off] = 255;
        T.ts[off] = 0;
    }
}

// ---------------------------------------------------------------------------------------------
// K6: persistent batch kernel
// ---------------------------------------------------------------------------------------------
__global__ void k_prep_demands(const int32_t* __restrict__ percent, const int32_t* __restrict__ ncont, int max_c,
                               const int32_t* __restrict__ policy, uint32_t P, SortedDemand* __restrict__ out, int32_t* err) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    int C = ncont[p];
    int pol = policy[p];
    bool bad = C < 0 || C > NGS_MAX_CONTAINERS || C > max_c || pol < 0 || pol > NGS_POLICY_RANDOM;
    if (bad) C = 0;
    int32_t pc[NGS_MAX_CONTAINERS];
    for (int k = 0; k < NGS_MAX_CONTAINERS; k++) {
        pc[k] = k < C ? percent[(size_t)p * max_c + k] : 0;
        if (pc[k] < 0 || pc[k] > NGS_MAX_FREE) bad = true;
    }
    if (bad) {
        atomicExch(err, NGS_ERR_INVALID);
        for (int k = 0; k < NGS_MAX_CONTAINERS; k++) pc[k] = 0;
        C = 0;
        pol = NGS_POLICY_BINPACK;
    }
    sort_demand(pc, C, pol, &out[p]);
}

struct BatchParams {
    Tables T;
    const SortedDemand* dem;  // [P]
    uint32_t P;
    LoadParams lp;
    uint64_t seed_seq;  // seed + seq0 of pod 0
    int32_t* node_out;
    int32_t* score_out;
    int32_t* gpu_idx_out;
    int max_c;
    unsigned long long* keys;  // [P][gridDim.x] one slot per CTA per pod, zero before launch
    uint32_t* done;            // [P] arrival counters of the random-policy exchange, zero before launch
    ulonglong2* xw;            // [P] arg-max exchange words {x: atomicMax key, y: arrival counter}, zero before launch
    // node-range sharding over the GPUs of one box: every rank runs this same kernel on
    // its own rows; CTA 0 of each rank trades the rank's result with the peers through
    // peer-mapped memory (plain 8-byte stores over NVLink), then tells its own CTAs
    uint32_t world, rank;
    unsigned long long* xslots[NGS_MAX_RANKS];  // [rank r] -> r's slot array [P][NGS_MAX_RANKS] (this batch's parity half)
    unsigned long long* gwin;                   // [P] this rank's copy of the global result, zero before launch
    uint32_t keys_stride, done_stride, xw_stride;  // distance between consecutive pods, in elements of each array
    int32_t* status_out;       // optional host-visible copy of the error word, written when the kernel ends
    int32_t* err;              // device error word (0 = ok)
    uint32_t rows_per_block;
#ifdef NGS_PHASE_TIMING
    unsigned long long* dbg;   // [gridDim.x][8] accumulated cycles per phase (profiling builds only)
#endif
};

__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

constexpr long long SPIN_TIMEOUT_CYCLES = 2000000000LL;  // ~1 s

__device__ __forceinline__ uint64_t warp_max_u64(uint64_t v) {
    const uint32_t hi = (uint32_t)(v >> 32);
    const uint32_t mhi = __reduce_max_sync(0xFFFFFFFFu, hi);
    const uint32_t lo = hi == mhi ? (uint32_t)v : 0u;
    const uint32_t mlo = __reduce_max_sync(0xFFFFFFFFu, lo);
    return ((uint64_t)mhi << 32) | mlo;
}

// shared-memory row: 8 x u16 free (two uint2), meta = RemainLoad codes [23:0] | G [27:24] | load_int [31:28]
struct SmemRows {
    uint2* fA;
    uint2* fB;
    uint32_t* meta;
    int32_t* scB;  // Binpack.Rate of the row's current state
    int32_t* scS;  // Spread.Rate
    uint32_t* top4;  // the row's four largest free values, one byte each (classify_top4)
    uint32_t* feasbits;
};

__device__ __forceinline__ void unpack_free(uint2 a, uint2 b, int (&f)[8]) {
    f[0] = a.x & 0xFFFF; f[1] = a.x >> 16; f[2] = a.y & 0xFFFF; f[3] = a.y >> 16;
    f[4] = b.x & 0xFFFF; f[5] = b.x >> 16; f[6] = b.y & 0xFFFF; f[7] = b.y >> 16;
}
__device__ __forceinline__ int smem_cell(const SmemRows& S, uint32_t r, int g) {
    const uint16_t* p16 = reinterpret_cast<const uint16_t*>(g < 4 ? &S.fA[r] : &S.fB[r]);
    return p16[g & 3];
}

// thread-per-row exact filter (used for the random policy, whose plan search has no sorting)
__device__ __forceinline__ bool eval_row_sample(const SmemRows& S, uint32_t r, const SortedDemand* dem) {
    const uint32_t meta = S.meta[r];
    const int G = (meta >> 24) & 15;
    if (G == 0) return false;
    int f[8];
    unpack_free(S.fA[r], S.fB[r], f);
    const int cls = classify_row(f, dem, false);
    if (cls != FEAS_UNDECIDED) return cls == FEAS_YES;
    return choose_sample<false>(f, G, dem, nullptr, nullptr);
}

// The decision for one row, worked out by a full warp (lane = GPU cell, the four 8-lane groups hold
// identical copies) while the pod's exchange is still in flight, and held in registers until the
// row is known to be the global winner: the plan, the row as GPUs.Allocate leaves it, its new
// rates and top-4, and already the row's verdict for the NEXT pod.
struct GroupCommit {
    int32_t fr;      // this lane's cell after GPUs.Allocate (0 beyond the GPU count)
    uint64_t emit;   // emission list
    int32_t n_emit;
    uint32_t top4;
    int32_t scB, scS;
    uint64_t fresh;  // key of the changed row for the next pod (0 = not feasible)
    int32_t fresh_ok;
};

__device__ __forceinline__ void plan_group(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t node, const SortedDemand* dem,
                                           const SortedDemand* dem_next, int g, int gbase, GroupCommit& cm, long long* stamps = nullptr) {
    const uint32_t meta = S.meta[r];
    const int G = (meta >> 24) & 15;
    const int load_int = (int)(meta >> 28);
    const int rlc = (int)((meta >> (3 * g)) & 7);
    const bool valid = g < G;
    choose_group8<true>(g, gbase, smem_cell(S, r, g), rlc, valid, dem->policy == NGS_POLICY_SPREAD, dem, &cm.emit, &cm.n_emit, &cm.fr);
    const int fr = valid ? cm.fr : 0;
    cm.fr = fr;
#ifdef NGS_PHASE_TIMING
    if (stamps) stamps[0] = clock64() + (fr == -12345);
#endif
    // the row after the bind: sum, fully free cells, four largest (each lane ranks its own cell:
    // descending, ties by index), both rates (Rate is a function of the row only, allocate.go:96)
    int sumf = fr;
    sumf += __shfl_xor_sync(FULL, sumf, 1);
    sumf += __shfl_xor_sync(FULL, sumf, 2);
    sumf += __shfl_xor_sync(FULL, sumf, 4);
    const int cnt100 = __popc((__ballot_sync(FULL, valid && fr == 100) >> gbase) & 0xFFu);
    const uint32_t c = ((uint32_t)fr << 3) | (uint32_t)(7 - g);
    int rank = 0;
#pragma unroll
    for (int sft = 1; sft < 8; sft++) rank += __shfl_xor_sync(FULL, c, sft) > c ? 1 : 0;
    uint32_t t4 = rank < 4 ? (uint32_t)min(fr, 255) << (8 * rank) : 0u;
    t4 |= __shfl_xor_sync(FULL, t4, 1);
    t4 |= __shfl_xor_sync(FULL, t4, 2);
    t4 |= __shfl_xor_sync(FULL, t4, 4);
    const int used = 100 * G - sumf;  // allocate.go:164-171
    const int pct = (bp.T.blut != nullptr && used >= 0 && used <= 100 * G) ? (int)__ldg(bp.T.blut + G * RATE_LUT_STRIDE + used)
                                                                            : binpack_percent_fp64(used, G);
    cm.top4 = t4;
    cm.scB = pct + load_int * 50 - G;                    // rater.go:69
    cm.scS = 100 * cnt100 + sumf / 10 - G - load_int;    // rater.go:122
    cm.fresh = 0;
    cm.fresh_ok = 0;
#ifdef NGS_PHASE_TIMING
    if (stamps) stamps[1] = clock64() + (cm.scB == -12345678);
#endif
    if (dem_next == nullptr) return;
    // the next pod on the row as it will stand after this bind
    const int pol2 = dem_next->policy;
    int cls = pol2 == NGS_POLICY_RANDOM ? FEAS_FALLBACK : classify_top4(t4, dem_next, true);
    if (cls == FEAS_FALLBACK || dem_next->nz == 0) {  // (warp-uniform: all groups look at the same row)
        int f[8];
#pragma unroll
        for (int i = 0; i < 8; i++) f[i] = __shfl_sync(FULL, fr, gbase + i);
        cls = classify_row(f, dem_next, pol2 != NGS_POLICY_RANDOM);
        if (cls == FEAS_UNDECIDED && pol2 == NGS_POLICY_RANDOM) cls = choose_sample<false>(f, G, dem_next, nullptr, nullptr) ? FEAS_YES : FEAS_NO;
    }
    if (cls == FEAS_UNDECIDED)
        cls = choose_group8<false>(g, gbase, fr, rlc, valid, pol2 == NGS_POLICY_SPREAD, dem_next, nullptr, nullptr, nullptr) ? FEAS_YES
                                                                                                                               : FEAS_NO;
    cm.fresh_ok = cls == FEAS_YES;
    if (cls == FEAS_YES && pol2 != NGS_POLICY_RANDOM) cm.fresh = make_key(pol2 == NGS_POLICY_SPREAD ? cm.scS : cm.scB, node);
}

// publish a winning decision (all 32 lanes of control warp 0): shared row, cached rates, the global
// table and the pod's outputs -- stores only, everything was computed by plan_group
__device__ __forceinline__ void publish_group(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t row,
                                              const SortedDemand* dem, uint32_t p, int score, const GroupCommit& cm, uint32_t lane) {
    if (lane < 8) {
        reinterpret_cast<uint16_t*>(lane < 4 ? &S.fA[r] : &S.fB[r])[lane & 3] = (uint16_t)cm.fr;
        bp.T.free[(size_t)row * 8 + lane] = cm.fr;  // 8 lanes: one 32 B row store
        int32_t* out = bp.gpu_idx_out + (size_t)p * bp.max_c;
        const int C = dem->C;
        for (int k = (int)lane; k < bp.max_c; k += 8)
            if (k >= C) out[k] = INT32_MIN;
        for (int k = (int)lane; k < C; k += 8) {  // rater.go:102-105: un-permute the emission order
            const int e = (int)((cm.emit >> (4 * k)) & 0xF);
            out[dem->order[k]] = e == 0xF ? NGS_NOT_NEED_GPU : e;
        }
    }
    if (lane == 0) {
        S.scB[r] = cm.scB;
        S.scS[r] = cm.scS;
        S.top4[r] = cm.top4;
        bp.T.tag[row] = 0;  // node.go:82
        bp.node_out[p] = (int32_t)(bp.T.first + row);
        bp.score_out[p] = score;
    }
}

// thread-serial publish for the random policy (plan search without sorting)
__device__ __forceinline__ void publish_sample(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t row,
                                               const SortedDemand* dem, uint32_t p) {
    const uint32_t meta = S.meta[r];
    const int G = (meta >> 24) & 15;
    const int load_int = (int)(meta >> 28);
    int f[8];
    unpack_free(S.fA[r], S.fB[r], f);
    uint64_t emit = 0;
    int n_emit = 0;
    choose_sample<true>(f, G, dem, &emit, &n_emit);  // places on f directly
    uint2 a, b;
    a.x = (uint32_t)f[0] | ((uint32_t)f[1] << 16); a.y = (uint32_t)f[2] | ((uint32_t)f[3] << 16);
    b.x = (uint32_t)f[4] | ((uint32_t)f[5] << 16); b.y = (uint32_t)f[6] | ((uint32_t)f[7] << 16);
    S.fA[r] = a;
    S.fB[r] = b;
    S.scB[r] = rate_binpack(f, G, load_int, bp.T.blut);
    S.scS[r] = rate_spread(f, G, load_int);
    S.top4[r] = top4_of(f);
    int4* gp = reinterpret_cast<int4*>(bp.T.free + (size_t)row * 8);
    gp[0] = make_int4(f[0], f[1], f[2], f[3]);
    gp[1] = make_int4(f[4], f[5], f[6], f[7]);
    bp.T.tag[row] = 0;
    bp.node_out[p] = (int32_t)(bp.T.first + row);
    bp.score_out[p] = 100;  // SampleRater.Rate (rater.go:24-26)
    int32_t* out = bp.gpu_idx_out + (size_t)p * bp.max_c;
    for (int k = 0; k < bp.max_c; k++) {
        const int e = (int)((emit >> (4 * k)) & 0xF);
        out[k] = k < dem->C ? (e == 0xF ? NGS_NOT_NEED_GPU : e) : INT32_MIN;
    }
}

__device__ __forceinline__ void publish_none(const BatchParams& bp, uint32_t p) {
    bp.node_out[p] = -1;
    bp.score_out[p] = NGS_SCORE_MIN;
    for (int k = 0; k < bp.max_c; k++) bp.gpu_idx_out[(size_t)p * bp.max_c + k] = INT32_MIN;
}

// grid-wide exchange: every CTA stores its 64-bit slot, then bumps the pod's arrival counter
// with release semantics; one lane per CTA spins (acquire) on that single word, after which the
// slots are read once.  Bounded: gives up and raises the batch's error word after ~1 s.
__device__ __forceinline__ void post_slot(unsigned long long* slot, unsigned long long v, uint32_t* done) {
    st_relaxed_u64(slot, v | KEY_VALID);
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(done) : "memory");
}
// The arrival counter is only a cheap thing to spin on (one word, one lane per CTA): it is not
// ordered against the slot stores, so readers still check each slot's VALID bit (read_slot).
__device__ __forceinline__ bool wait_done(const uint32_t* done, uint32_t target, int32_t* err) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(done) : "memory");
    if (v >= target) return true;
    const long long t0 = clock64();
    uint32_t spins = 0;
    for (;;) {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(done) : "memory");
        if (v >= target) return true;
        if ((++spins & 255u) == 0) {
            if (*reinterpret_cast<volatile int32_t*>(err) != 0 || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                atomicCAS(err, 0, NGS_ERR_TIMEOUT);
                return false;
            }
        }
    }
}

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// bounded wait for a VALID word (written by a peer GPU or by this rank's CTA 0)
__device__ __forceinline__ unsigned long long wait_valid(const unsigned long long* p, int32_t* err, bool& dead) {
    unsigned long long v = ld_relaxed_sys_u64(p);
    if (v & KEY_VALID) return v & ~KEY_VALID;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!((v = ld_relaxed_sys_u64(p)) & KEY_VALID)) {
        if ((++spins & 255u) == 0) {
            if (*reinterpret_cast<volatile int32_t*>(err) != 0 || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                atomicCAS(err, 0, NGS_ERR_TIMEOUT);
                dead = true;
                return 0;
            }
        }
    }
    return v & ~KEY_VALID;
}
// CTA 0, one full warp: post this rank's word to every rank's slot array, collect all ranks'
// words from our own array.  Returns lane j's word of rank j (0 for lanes >= world).
__device__ __forceinline__ unsigned long long cross_rank_exchange(const BatchParams& bp, uint32_t p, unsigned long long mine,
                                                                  uint32_t lane, bool& dead) {
    unsigned long long got = 0;
    if (lane < bp.world) {
        st_relaxed_sys_u64(bp.xslots[lane] + (size_t)p * NGS_MAX_RANKS + bp.rank, mine | KEY_VALID);
        got = wait_valid(bp.xslots[bp.rank] + (size_t)p * NGS_MAX_RANKS + lane, bp.err, dead);
    }
    dead = __any_sync(0xFFFFFFFFu, dead);
    return got;
}

// arg-max exchange word: the key (atomicMax) and the arrival counter share one 16-byte word, so one
// 16-byte load returns both.  A CTA bumps the counter only after its atomicMax has returned, hence
// a load that sees the counter full also sees the final key.  Returns false on time-out.
__device__ __forceinline__ bool wait_exchange(const ulonglong2* w, uint32_t target, int32_t* err, unsigned long long* key) {
    unsigned long long a, b;
    const long long t0 = clock64();
    uint32_t spins = 0;
    for (;;) {
        asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(w) : "memory");
        if ((uint32_t)b >= target) {
            *key = a;
            return true;
        }
        if ((++spins & 255u) == 0) {
            if (*reinterpret_cast<volatile int32_t*>(err) != 0 || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                atomicCAS(err, 0, NGS_ERR_TIMEOUT);
                *key = 0;
                return false;
            }
        }
    }
}

__device__ __forceinline__ unsigned long long read_slot(const unsigned long long* slot, bool dead) {
    unsigned long long v = ld_relaxed_u64(slot);
    while (!(v & KEY_VALID) && !dead) v = ld_relaxed_u64(slot);  // normally valid at the first read
    return v & ~KEY_VALID;
}

// filter (and rate) this worker warp's rows for one pod, on the rows as they stand.  Argmax
// policies: every thread keeps the best two keys of its rows (k1 >= k2).  Random policy: one
// feasibility bit per row in the pod's feasbits words, the warp's count in cnt (lane 0).
// Must be called by all 32 lanes.
__device__ __forceinline__ void filter_pass(const SmemRows& S, const SortedDemand* dem, uint32_t wt, uint32_t WT, uint32_t n_iter,
                                            uint32_t my_rows, uint32_t node0, uint32_t lane, int g, int gbase, uint32_t* fb,
                                            uint64_t& k1, uint64_t& k2, uint32_t& cnt) {
    const int policy = dem->policy;
    const bool spread = policy == NGS_POLICY_SPREAD;
    if (policy == NGS_POLICY_RANDOM) {
        uint32_t c = 0;
        for (uint32_t it = 0; it < n_iter; it++) {
            const uint32_t r = it * WT + wt;
            const bool ok = r < my_rows && eval_row_sample(S, r, dem);
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, ok);
            if (lane == 0 && r < my_rows) {
                fb[r >> 5] = bal;
                c += __popc(bal);
            }
        }
        cnt = c;
        return;
    }
    k1 = 0;
    k2 = 0;
    for (uint32_t it = 0; it < n_iter; it++) {
        const uint32_t r = it * WT + wt;
        int cls = FEAS_NO;
        if (r < my_rows) {
            cls = classify_top4(S.top4[r], dem, true);  // unknown nodes hold all zeros: FEAS_NO (or YES for an empty demand,
            if (cls == FEAS_FALLBACK || dem->nz == 0) {  //  which the full-row path below sorts out)
                cls = FEAS_NO;
                const uint32_t meta = S.meta[r];
                if (((meta >> 24) & 15) != 0) {
                    int f[8];
                    unpack_free(S.fA[r], S.fB[r], f);
                    cls = classify_row(f, dem, true);
                }
            }
        }
        // rows the shortcuts leave open are searched exactly right away, 8 lanes per row, 4 rows a time;
        // the verdict goes back to the lane that owns the row
        uint32_t und = __ballot_sync(0xFFFFFFFFu, cls == FEAS_UNDECIDED);
        while (und) {  // warp-uniform
            uint32_t t = und;
            for (uint32_t q = 0; q < (lane >> 3); q++) t &= t - 1;  // group j takes the j-th open row
            const bool has = t != 0;
            const uint32_t rr = __shfl_sync(0xFFFFFFFFu, r, has ? __ffs(t) - 1 : 0);
            const uint32_t meta = S.meta[rr];
            const int G = has ? (int)((meta >> 24) & 15) : 0;
            const bool ok = choose_group8<false>(g, gbase, smem_cell(S, rr, g), (int)((meta >> (3 * g)) & 7), g < G, spread, dem,
                                                 nullptr, nullptr, nullptr);
            const uint32_t okb = __ballot_sync(0xFFFFFFFFu, has && ok && g == 0);  // group j's verdict at bit 8j
            if ((und >> lane) & 1u) {
                const uint32_t j = __popc(und & ((1u << lane) - 1u));
                if (j < 4) cls = ((okb >> (8 * j)) & 1u) ? FEAS_YES : FEAS_NO;
            }
            und &= und - 1;
            und &= und - 1;
            und &= und - 1;
            und &= und - 1;
        }
        if (cls == FEAS_YES) {
            const uint64_t key = make_key(spread ? S.scS[r] : S.scB[r], node0 + r);
            if (key > k1) {
                k2 = k1;
                k1 = key;
            } else if (key > k2) {
                k2 = key;
            }
        }
    }
}

// the warp's best two keys from every lane's best two: m1 = max of the k1; m2 = the best of what is
// left (the holder of m1 offers its k2, everyone else their k1).  Keys are distinct (one per row).
__device__ __forceinline__ void warp_top2(uint64_t k1, uint64_t k2, uint64_t& m1, uint64_t& m2) {
    m1 = warp_max_u64(k1);
    m2 = warp_max_u64(k1 == m1 ? k2 : k1);
}

// feasibility and key of ONE row for one pod, computed by a full warp (control warp 0 right after it
// has bound the previous pod to that row): shortcuts first, the exact 8-lane search if they do not
// decide.  All 32 lanes get the same answer.
__device__ __forceinline__ uint64_t fresh_key_of_row(const SmemRows& S, uint32_t r, uint32_t node, const SortedDemand* dem, int g,
                                                     int gbase, bool* feasible) {
    const int policy = dem->policy;
    if (policy == NGS_POLICY_RANDOM) {
        *feasible = eval_row_sample(S, r, dem);
        return 0;
    }
    const bool spread = policy == NGS_POLICY_SPREAD;
    int cls = classify_top4(S.top4[r], dem, true);
    if (cls == FEAS_FALLBACK || dem->nz == 0) {
        int f[8];
        unpack_free(S.fA[r], S.fB[r], f);
        cls = classify_row(f, dem, true);
    }
    if (cls == FEAS_UNDECIDED) {  // warp-uniform: every lane looked at the same row
        const uint32_t meta = S.meta[r];
        const int G = (meta >> 24) & 15;
        cls = choose_group8<false>(g, gbase, smem_cell(S, r, g), (int)((meta >> (3 * g)) & 7), g < G, spread, dem, nullptr, nullptr,
                                   nullptr)
                  ? FEAS_YES
                  : FEAS_NO;
    }
    *feasible = cls == FEAS_YES;
    return cls == FEAS_YES ? make_key(spread ? S.scS[r] : S.scB[r], node) : 0ull;
}

// K6.  One CTA per SM (cooperative launch, so every CTA is resident), each owning a contiguous
// node range held in shared memory for the whole batch.  Warps 0 and 1 are control warps, the
// others ("workers") own the rows.  Pods are decided strictly in order, but the work is pipelined:
// while the control warps settle pod p, the workers already filter pod p+1 on the rows as they
// stand; binding pod p changes exactly one row, so afterwards only that row is filtered again.
//   filter (workers): one thread per row settles feasibility with the exact shortcuts of
//      classify_row; the ~1% of rows they leave open are searched exactly right away, 8 lanes per
//      row (choose_group8).  Rate is a function of the row only (allocate.go:96), so both
//      policies' rates sit beside the row and are refreshed when the row changes.
//   arg-max (warp 1): per-warp maxima -> one 64-bit key per CTA -> atomicMax into the pod's word in
//      L2, arrival counter, spin, read (K5: plus the cross-GPU trade).  Meanwhile warp 0 works out the
//      exact plan of this CTA's best row, 8 lanes per row.
//      (random policy: slots carry feasible counts; seeded k-th pick over the feasible set)
//   bind (warp 0 of the CTA that owns the winner): publish the plan, update the row.
__global__ void __launch_bounds__(1024, 1) k_schedule(BatchParams bp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ SortedDemand s_dem[4];
    __shared__ uint64_t s_red[2][32][2];  // per pod parity, per worker warp: best two keys (or {count, -})
    __shared__ uint64_t s_win;     // the pod's winning key, handed from the exchange warp to the plan warp
    // what binding pod p-1 leaves for pod p, indexed by p & 1 (written during iteration p-1, read in p):
    __shared__ int s_dirty[2];        // row of this CTA pod p-1 was bound to, -1 if none
    __shared__ uint64_t s_fresh[2];   // that row's key for pod p, evaluated on the row after the bind (0 = not feasible)
    __shared__ int s_fresh_ok[2];     // ... and its feasibility (what the random policy needs)
    __shared__ int s_abort;
    const uint32_t rpb = bp.rows_per_block;
    if (threadIdx.x == 0) {
        s_abort = 0;
        s_dirty[0] = s_dirty[1] = -1;
        s_fresh[0] = s_fresh[1] = 0;
        s_fresh_ok[0] = s_fresh_ok[1] = 0;
    }
    const uint32_t fwords = (rpb + 31) / 32;
    SmemRows S;
    S.fA = reinterpret_cast<uint2*>(smem_raw);
    S.fB = S.fA + rpb;
    S.meta = reinterpret_cast<uint32_t*>(S.fB + rpb);
    S.scB = reinterpret_cast<int32_t*>(S.meta + rpb);
    S.scS = S.scB + rpb;
    S.top4 = reinterpret_cast<uint32_t*>(S.scS + rpb);
    S.feasbits = S.top4 + rpb;  // two buffers of fwords words (pod parity)

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    // the two control warps are the LAST two: the SM's issue arbiter favours high warp ids, and the
    // control path (exchange, plan) is the latency-critical one
    // Warps map to the SM's four schedulers by warp id % 4.  Control warp 0 runs the latency-critical plan
    // search; the worker warps that would share its scheduler stay idle (they own no rows), so that its
    // instructions never queue behind the filter's.
    const uint32_t NWARP = nwarps - 2;                     // warps below the control warps
    const uint32_t csm = NWARP & 3u;                       // scheduler of control warp 0
    const uint32_t idle_below = (warp + 3u - csm) >> 2;    // idle warps among [0, warp)
    const uint32_t NW = NWARP - ((NWARP + 3u - csm) >> 2); // worker warps that own rows
    const uint32_t WT = NW * 32u;
    const bool is_worker = warp < NWARP && (warp & 3u) != csm;
    const bool is_control = warp >= NWARP;
    const uint32_t ww = warp - idle_below;                 // worker index (meaningful for workers)
    const uint32_t wt = ww * 32u + lane;
    const uint32_t cw = warp - NWARP;                      // control warp 0 (plan, bind) / 1 (exchange)
    const uint32_t row0 = blockIdx.x * rpb;
    const uint32_t my_rows = row0 < bp.T.n ? min(rpb, bp.T.n - row0) : 0u;
    const Tables& T = bp.T;
    const int g = lane & 7, gbase = lane & 24;

    // ---- prologue: rows -> shared memory (and the load-aware quantities, frozen for the batch) ----
    for (uint32_t r = tid; r < my_rows; r += blockDim.x) {
        const uint32_t row = row0 + r;
        const int G = T.gcount[row];
        int f[8];
        const int4 a = ld_row4(T.free + (size_t)row * 8), b = ld_row4(T.free + (size_t)row * 8 + 4);
        f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
        f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
        int rlc[8];
        int load_int = 0;
        if (bp.lp.enabled && G > 0) {
            uint8_t tenths[2][8];
            uint32_t ts[2][8];
#pragma unroll
            for (int m = 0; m < 2; m++) {
                const size_t off = ((size_t)m * T.nstride + row) * 8;
                const uint2 t8 = __ldcg(reinterpret_cast<const uint2*>(T.tenths + off));
                const uint4 t0 = __ldcg(reinterpret_cast<const uint4*>(T.ts + off));
                const uint4 t1 = __ldcg(reinterpret_cast<const uint4*>(T.ts + off + 4));
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    tenths[m][c] = (uint8_t)(t8.x >> (8 * c));
                    tenths[m][c + 4] = (uint8_t)(t8.y >> (8 * c));
                }
                ts[m][0] = t0.x; ts[m][1] = t0.y; ts[m][2] = t0.z; ts[m][3] = t0.w;
                ts[m][4] = t1.x; ts[m][5] = t1.y; ts[m][6] = t1.z; ts[m][7] = t1.w;
            }
            load_usage_row(tenths, ts, bp.lp, G, rlc, &load_int);
        } else {
            const uint2 r8 = __ldcg(reinterpret_cast<const uint2*>(T.rl + (size_t)row * 8));
#pragma unroll
            for (int c = 0; c < 4; c++) {
                rlc[c] = (int)(int8_t)(r8.x >> (8 * c)) + 2;
                rlc[c + 4] = (int)(int8_t)(r8.y >> (8 * c)) + 2;
            }
        }
        uint32_t meta = ((uint32_t)G << 24) | ((uint32_t)load_int << 28);
#pragma unroll
        for (int c = 0; c < 8; c++) meta |= (uint32_t)(c < G ? rlc[c] : 2) << (3 * c);
        uint2 pa, pb;
        pa.x = (uint32_t)f[0] | ((uint32_t)f[1] << 16); pa.y = (uint32_t)f[2] | ((uint32_t)f[3] << 16);
        pb.x = (uint32_t)f[4] | ((uint32_t)f[5] << 16); pb.y = (uint32_t)f[6] | ((uint32_t)f[7] << 16);
        S.fA[r] = pa;
        S.fB[r] = pb;
        S.meta[r] = meta;
        S.scB[r] = G ? rate_binpack(f, G, load_int, T.blut) : 0;
        S.scS[r] = G ? rate_spread(f, G, load_int) : 0;
        S.top4[r] = top4_of(f);  // cells beyond G hold 0
    }
    for (uint32_t i = tid; i < 2 * sizeof(SortedDemand) / 4; i += blockDim.x)
        if (i / (sizeof(SortedDemand) / 4) < bp.P) reinterpret_cast<uint32_t*>(&s_dem[0])[i] = reinterpret_cast<const uint32_t*>(&bp.dem[0])[i];
    __syncthreads();

#ifdef NGS_PHASE_TIMING
    long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tq = clock64();
#define PHASE(i) do { const long long t_ = clock64(); ph[i] += t_ - tq; tq = t_; } while (0)
// BAR.SYNC blocks lazily (at the first instruction that needs the barrier): consume a shared word
// before reading the clock so that the wait lands in the phase it belongs to
#define PHASE_AFTER_BARRIER(i) do { if (*reinterpret_cast<volatile int*>(&s_abort) == 0x7fffffff) ph[7]++; PHASE(i); } while (0)
#else
#define PHASE(i) do { } while (0)
#define PHASE_AFTER_BARRIER(i) do { } while (0)
#endif
    const uint32_t n_iter = (my_rows + WT - 1) / WT;
    bool rl_pending = bp.lp.enabled != 0;  // RemainLoad write-back of the first load-aware Rate
    uint64_t k1 = 0, k2 = 0;  // this worker's best two keys for the pod being filtered
    uint32_t cnt = 0;

    // allocate.go:193: LoadUsage leaves RemainLoad in the live node state.  SampleRater.Rate never
    // calls it, so the table is only touched once a binpack/spread pod comes by.
    auto write_back_rl = [&]() {
        for (uint32_t r = tid; r < my_rows; r += blockDim.x) {
            const uint32_t meta = S.meta[r];
            const int G = (meta >> 24) & 15;
            if (G == 0) continue;
            uint2 out;
            out.x = out.y = 0;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                out.x |= (uint32_t)(uint8_t)(int8_t)(c < G ? (int)((meta >> (3 * c)) & 7) - 2 : 0) << (8 * c);
                out.y |= (uint32_t)(uint8_t)(int8_t)(c + 4 < G ? (int)((meta >> (3 * (c + 4))) & 7) - 2 : 0) << (8 * c);
            }
            *reinterpret_cast<uint2*>(T.rl + (size_t)(row0 + r) * 8) = out;
        }
    };

    auto publish_filter = [&](uint32_t buf, int policy) {  // hand the warp's result for a pod to the control warps
        if (policy != NGS_POLICY_RANDOM) {
            uint64_t m1, m2;
            warp_top2(k1, k2, m1, m2);
            if (lane == 0) {
                s_red[buf][ww][0] = m1;
                s_red[buf][ww][1] = m2;
            }
        } else if (lane == 0) {
            s_red[buf][ww][0] = cnt;
        }
    };

    // pod 0 is filtered before the loop; inside the loop the workers are always one pod ahead
    if (bp.P > 0) {
        if (rl_pending && s_dem[0].policy != NGS_POLICY_RANDOM) {
            rl_pending = false;
            write_back_rl();
        }
        if (is_worker) {
            filter_pass(S, &s_dem[0], wt, WT, n_iter, my_rows, T.first + row0, lane, g, gbase, S.feasbits, k1, k2, cnt);
            publish_filter(0, s_dem[0].policy);
        }
    }
    __syncthreads();

    // ONE CTA-wide barrier per pod.  In iteration p the workers filter pod p+1 on the rows as they stand
    // and leave each warp's best TWO keys; the control warps settle pod p, and if this CTA binds it, the
    // row it was bound to is evaluated once more for pod p+1 by the binding warp itself.  At the start of
    // iteration p+1 the control warps put the two together: a warp whose best key belongs to the changed
    // row contributes its second best, and the changed row competes with its fresh key.
    for (uint32_t p = 0; p < bp.P; p++) {
        const SortedDemand* dem = &s_dem[p & 3];
        const int policy = dem->policy;
        const bool spread = policy == NGS_POLICY_SPREAD;
        const bool more = p + 1 < bp.P;
        const SortedDemand* dem_next = &s_dem[(p + 1) & 3];
        if (more && rl_pending && dem_next->policy != NGS_POLICY_RANDOM) {
            rl_pending = false;
            write_back_rl();
        }
        if (is_worker) {
            // ---- workers: filter pod p+1 (pod p's bind may still be in flight: its row is dealt with above) ----
            if (more) {
                filter_pass(S, dem_next, wt, WT, n_iter, my_rows, T.first + row0, lane, g, gbase, S.feasbits + ((p + 1) & 1) * fwords, k1,
                            k2, cnt);
                publish_filter((p + 1) & 1, dem_next->policy);
            }
            PHASE(0);
        } else if (is_control) {
            // what the previous iteration left: the row pod p-1 was bound to (if it is ours) and its fresh verdict
            const int dirty = s_dirty[p & 1];
            const uint64_t fresh = s_fresh[p & 1];
            const int fresh_ok = s_fresh_ok[p & 1];
            const uint32_t dirty_node = dirty >= 0 ? T.first + row0 + (uint32_t)dirty : 0xFFFFFFFFu;
            const uint32_t nx = (p + 1) & 1;  // slot control warp 0 fills for the next pod
            PHASE(3);  // (profiling) control warps: start of iteration
            if (policy == NGS_POLICY_RANDOM) {
              if (cw == 0) {
                // ---- control, random policy: counts -> seeded pick -> bind ----
                bool dead = false;
                unsigned long long* krow = bp.keys + (size_t)p * bp.keys_stride;
                uint32_t* done = bp.done + (size_t)p * bp.done_stride;
                uint32_t* fb = S.feasbits + (p & 1) * fwords;
                uint64_t v = lane < NW ? s_red[p & 1][lane][0] : 0;
                if (dirty >= 0) {  // the changed row: replace its (possibly stale) bit by the fresh verdict
                    const uint32_t wi = (uint32_t)dirty >> 5, bit = 1u << (dirty & 31);
                    const uint32_t old = fb[wi];
                    __syncwarp();
                    if (lane == 0) {
                        fb[wi] = fresh_ok ? (old | bit) : (old & ~bit);
                        v += (uint64_t)(fresh_ok ? 1 : 0) - (uint64_t)((old & bit) ? 1 : 0);  // lane 0 holds warp 0's count
                    }
                    __syncwarp();
                }
                if (lane == 0) {
                    s_dirty[nx] = -1;
                    s_fresh[nx] = 0;
                    s_fresh_ok[nx] = 0;
                }
                if (p + 2 < bp.P)  // stage the demand the workers will need next
                    for (uint32_t i = lane; i < sizeof(SortedDemand) / 4; i += 32)
                        reinterpret_cast<uint32_t*>(&s_dem[(p + 2) & 3])[i] = reinterpret_cast<const uint32_t*>(&bp.dem[p + 2])[i];
                int bound_row = -1;
                {
                uint32_t c = (uint32_t)v;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, o);
                if (lane == 0) {
                    post_slot(krow + blockIdx.x, c, done);
                    dead = !wait_done(done, gridDim.x, bp.err);
                }
                __syncwarp();
                // contiguous chunk of CTAs per lane so that a prefix over lanes is a prefix over CTAs
                const uint32_t chunk = (gridDim.x + 31) / 32;
                const uint32_t j0 = min(lane * chunk, gridDim.x), j1 = min(j0 + chunk, gridDim.x);
                uint64_t sum = 0;
                dead = __any_sync(0xFFFFFFFFu, dead);
                for (uint32_t j = j0; j < j1; j++) sum += read_slot(krow + j, dead);
                uint64_t incl = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint64_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                    if (lane >= (uint32_t)o) incl += t;
                }
                const uint64_t ltotal = __shfl_sync(0xFFFFFFFFu, incl, 31);  // feasible rows on this rank
                uint64_t total = ltotal, before_rank = 0;
                if (bp.world > 1) {
                    // K5: ranks trade their totals; rank r's feasible rows follow those of ranks < r
                    unsigned long long packed = 0;
                    if (blockIdx.x == 0) {
                        const unsigned long long cj = cross_rank_exchange(bp, p, ltotal, lane, dead);
                        unsigned long long inc2 = cj;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc2, o);
                            if (lane >= (uint32_t)o) inc2 += t;
                        }
                        const unsigned long long gt = __shfl_sync(0xFFFFFFFFu, inc2, 31);
                        const unsigned long long bf = __shfl_sync(0xFFFFFFFFu, inc2 - cj, bp.rank);
                        packed = (gt << 31) | bf;
                        if (lane == 0) st_relaxed_u64(bp.gwin + p, packed | KEY_VALID);
                    } else {
                        if (lane == 0) packed = wait_valid(bp.gwin + p, bp.err, dead);
                        packed = __shfl_sync(0xFFFFFFFFu, packed, 0);
                        dead = __any_sync(0xFFFFFFFFu, dead);
                    }
                    total = packed >> 31;
                    before_rank = packed & 0x7FFFFFFFull;
                }
                if (total == 0) {
                    if (blockIdx.x == 0 && lane == 0) publish_none(bp, p);
                } else {
                    const uint64_t gpick = splitmix64(bp.seed_seq + p) % total;
                    const bool mine = gpick >= before_rank && gpick < before_rank + ltotal;
                    if (!mine && blockIdx.x == 0 && lane == 0) {
                        publish_none(bp, p);  // another rank owns the pick; it reports node and plan
                        bp.score_out[p] = 100;
                    }
                    const uint64_t pick = mine ? gpick - before_rank : ~0ull;  // ~0: matches no CTA
                    // rank of the pick inside this CTA, if it falls here
                    uint64_t run = incl - sum;
                    long long local_rank = -1;
                    for (uint32_t j = j0; j < j1; j++) {
                        const uint64_t cj = read_slot(krow + j, dead);
                        if (j == blockIdx.x && pick >= run && pick < run + cj) local_rank = (long long)(pick - run);
                        run += cj;
                    }
                    // exactly one lane of exactly one CTA sees local_rank >= 0
                    const uint32_t who = __ballot_sync(0xFFFFFFFFu, local_rank >= 0);
                    if (who) {
                        const uint32_t rank = (uint32_t)__shfl_sync(0xFFFFFFFFu, (int)local_rank, __ffs(who) - 1);
                        // locate the rank-th set bit of feasbits (rows of this CTA in index order)
                        const uint32_t nwords = (my_rows + 31) / 32;
                        const uint32_t wchunk = (nwords + 31) / 32;
                        const uint32_t w0 = min(lane * wchunk, nwords), w1 = min(w0 + wchunk, nwords);
                        uint32_t sb = 0;
                        for (uint32_t wi = w0; wi < w1; wi++) sb += __popc(fb[wi]);
                        uint32_t inc = sb;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                            if (lane >= (uint32_t)o) inc += t;
                        }
                        const uint32_t before = inc - sb;
                        if (rank >= before && rank < before + sb) {
                            uint32_t need = rank - before;
                            for (uint32_t wi = w0; wi < w1; wi++) {
                                uint32_t bits = fb[wi];
                                const uint32_t pc = __popc(bits);
                                if (need < pc) {
                                    for (uint32_t q = 0; q < need; q++) bits &= bits - 1;
                                    const uint32_t r = wi * 32 + (__ffs(bits) - 1);
                                    publish_sample(bp, S, r, row0 + r, dem, p);
                                    bound_row = (int)r;
                                    break;
                                }
                                need -= pc;
                            }
                        }
                    }
                }
                }
                // if this CTA bound the pod, evaluate that row for the next pod right away
                const uint32_t bmask = __ballot_sync(0xFFFFFFFFu, bound_row >= 0);
                if (bmask && more) {
                    const int br = __shfl_sync(0xFFFFFFFFu, bound_row, __ffs(bmask) - 1);
                    bool ok = false;
                    const uint64_t fk = fresh_key_of_row(S, (uint32_t)br, T.first + row0 + (uint32_t)br, dem_next, g, gbase, &ok);
                    if (lane == 0) {
                        s_dirty[nx] = br;
                        s_fresh[nx] = fk;
                        s_fresh_ok[nx] = ok ? 1 : 0;
                    }
                }
                if (dead && lane == 0) s_abort = 1;
              }
            } else {
                // ---- control, arg-max policies: warp 1 runs the exchange while warp 0 works out the exact
                // plan of this CTA's best row (published only if that row is the global winner) ----
                uint64_t v = 0;
                if (lane < NW) {
                    const uint64_t m1 = s_red[p & 1][lane][0], m2 = s_red[p & 1][lane][1];
                    v = ((m1 & KEY_FEASIBLE) && key_node(m1) == dirty_node) ? m2 : m1;  // drop the changed row's old key
                }
                v = warp_max_u64(v);
                v = fresh > v ? fresh : v;  // ... and let it compete with its fresh one
#ifdef NGS_PHASE_TIMING
                if (v == 12345) ph[5]++;
                PHASE(7);  // (profiling) combine: per-warp best two -> this CTA's key
#endif
                const bool have = (v & KEY_FEASIBLE) != 0;
                const uint32_t my_r = have ? key_node(v) - T.first - row0 : 0u;
                GroupCommit cm;
                cm.fr = 0;
                cm.emit = 0;
                cm.n_emit = 0;
                cm.top4 = 0;
                cm.scB = cm.scS = 0;
                cm.fresh = 0;
                cm.fresh_ok = 0;
                if (cw == 1) {
                    // exchange (fastest variant of profiles/tools/xchg_bench.cu): atomicMax into the pod's
                    // word; the arrival counter is bumped only after that atomic has returned (data
                    // dependence), so a full counter implies every CTA's key is in the word
                    bool dead = false;
                    ulonglong2* xw = bp.xw + (size_t)p * bp.xw_stride;
                    // the demand the workers will need next is fetched into registers now and stored to shared
                    // memory after the exchange, so that its global-load latency hides behind the wait
                    uint32_t stg0 = 0, stg1 = 0;
                    const bool stage = p + 2 < bp.P;
                    if (stage) {
                        const uint32_t* src = reinterpret_cast<const uint32_t*>(&bp.dem[p + 2]);
                        stg0 = __ldcg(src + lane);
                        if (lane + 32 < sizeof(SortedDemand) / 4) stg1 = __ldcg(src + lane + 32);
                    }
                    if (lane == 0) {
                        const unsigned long long old = atomicMax(&xw->x, (unsigned long long)v);
                        const unsigned long long one = 1ull + (old >> 63);  // bit 63 is never set: always 1
                        asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(&xw->y), "l"(one) : "memory");
                    }
                    PHASE(2);
                    if (bp.world == 1) {
                        if (lane == 0) {
                            unsigned long long m;
                            dead = !wait_exchange(xw, gridDim.x, bp.err, &m);
                            s_win = m;
                        }
                    } else if (blockIdx.x == 0) {
                        // K5: this rank's best key goes to every peer; the best over all ranks comes back
                        unsigned long long mloc = 0;
                        if (lane == 0) dead = !wait_exchange(xw, gridDim.x, bp.err, &mloc);
                        mloc = __shfl_sync(0xFFFFFFFFu, mloc, 0);
                        dead = __any_sync(0xFFFFFFFFu, dead);
                        const uint64_t mg = warp_max_u64(cross_rank_exchange(bp, p, mloc, lane, dead));
                        if (lane == 0) {
                            st_relaxed_u64(bp.gwin + p, mg | KEY_VALID);
                            s_win = mg;
                        }
                    } else if (lane == 0) {
                        s_win = wait_valid(bp.gwin + p, bp.err, dead);
                    }
                    if (dead && lane == 0) s_abort = 1;
                    if (stage) {
                        uint32_t* dst = reinterpret_cast<uint32_t*>(&s_dem[(p + 2) & 3]);
                        dst[lane] = stg0;
                        if (lane + 32 < sizeof(SortedDemand) / 4) dst[lane + 32] = stg1;
                    }
                } else {
                    if (lane == 0) {
                        s_dirty[nx] = -1;
                        s_fresh[nx] = 0;
                        s_fresh_ok[nx] = 0;
                    }
#ifdef NGS_PHASE_TIMING
                    long long st[2] = {0, 0};
                    const long long t_in = clock64();
                    if (have) plan_group(bp, S, my_r, T.first + row0 + my_r, dem, more ? dem_next : nullptr, g, gbase, cm, st);
                    if (have && cm.fresh != 12345) {
                        ph[0] += st[0] - t_in;   // exact plan (choose_group8<PLAN>)
                        ph[2] += st[1] - st[0];  // post-bind statistics and rates
                        ph[6] -= 0;
                    }
#else
                    if (have) plan_group(bp, S, my_r, T.first + row0 + my_r, dem, more ? dem_next : nullptr, g, gbase, cm);
#endif
                }
                PHASE(4);
                asm volatile("bar.sync 1, 64;" ::: "memory");  // the two control warps only
                PHASE_AFTER_BARRIER(5);
                if (cw == 0) {
                    const uint64_t m = s_win;
                    if (!(m & KEY_FEASIBLE)) {
                        if (blockIdx.x == 0 && lane == 0) publish_none(bp, p);
                    } else if (blockIdx.x == 0 && lane == 0 && key_node(m) - T.first >= T.n) {
                        // the winner lives on another rank: node and score are known from the key, the
                        // plan is filled in by the owner (merged on the host side by an allreduce-max)
                        publish_none(bp, p);
                        bp.node_out[p] = (int32_t)key_node(m);
                        bp.score_out[p] = key_score(m);
                    }
                    if (have && m == v) {  // warp-uniform
                        // ---- bind on the owner: stores only, everything was worked out during the exchange ----
                        publish_group(bp, S, my_r, row0 + my_r, dem, p, key_score(m), cm, lane);
                        if (lane == 0) {
                            s_dirty[nx] = (int)my_r;
                            s_fresh[nx] = cm.fresh;
                            s_fresh_ok[nx] = cm.fresh_ok;
                        }
                    }
                    PHASE(6);
                }
            }
        }
        __syncthreads();  // pod p is settled; the workers' results for pod p+1 are in s_red / feasbits
        PHASE_AFTER_BARRIER(1);
        if (s_abort) break;  // a peer never posted: give up instead of hanging the GPU
    }
    if (blockIdx.x == 0 && tid == 0 && bp.status_out) *bp.status_out = *reinterpret_cast<volatile int32_t*>(bp.err);
#ifdef NGS_PHASE_TIMING
    if (tid == NWARP * 32 && bp.dbg)
        for (int i = 0; i < 8; i++) bp.dbg[blockIdx.x * 16 + i] = (unsigned long long)ph[i];
    if (tid == NWARP * 32 + 32 && bp.dbg)
        for (int i = 0; i < 8; i++) bp.dbg[blockIdx.x * 16 + 8 + i] = (unsigned long long)ph[i];
#endif
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
}  // namespace

struct ngs_dealer {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    Tables T{};
    int num_sms = 0;
    int smem_optin = 0;
    int coop = 0;
    // plan-cache generation (NodeInfo.PlanCache keyed by Demand.Hash, allocate.go:72-75)
    uint32_t gen = 1;
    std::vector<int32_t> cur_demand;
    // device scratch
    uint32_t* d_idx = nullptr;
    uint8_t* d_feas = nullptr;
    int8_t* d_reason = nullptr;
    int32_t* d_score = nullptr;
    int32_t* d_plan = nullptr;
    int32_t* d_nem = nullptr;
    size_t verb_cap = 0;   // candidates the verb scratch holds
    size_t plan_cap = 0;   // int32 entries d_plan holds
    BindResult* d_bind = nullptr;
    BindResult* h_bind = nullptr;  // pinned
    int32_t* d_small = nullptr;    // gpu_idx in / status out for k_apply
    int32_t* h_small = nullptr;    // pinned
    // batch
    uint32_t batch_cap = 0;  // pods per launch chunk
    SortedDemand* d_dem = nullptr;
    unsigned long long* d_keys = nullptr;  // per-pod exchange records: {xw 16 B, done 4 B, pad, keys[num_sms]}
    size_t pod_rec_bytes = 0;
    int32_t* d_pod_in_dev = nullptr;       // device aliases of the mapped staging buffers
    int32_t* d_pod_out_dev = nullptr;
    int32_t* h_pod_in = nullptr;   // pinned mirrors
    int32_t* h_pod_out = nullptr;
    int32_t* d_err = nullptr;
    int32_t* h_err = nullptr;
    int32_t* d_rel = nullptr;  // scratch of ngs_release_batch
    size_t rel_cap = 0;
    // cross-GPU exchange (K5)
    int world = 1, rank = 0;
    unsigned long long* d_xslots = nullptr;               // own buffer: [2 parities][BATCH_CHUNK][NGS_MAX_RANKS]
    unsigned long long* peer_xslots[NGS_MAX_RANKS] = {};  // peer-mapped bases (own entry = d_xslots)
    unsigned long long* d_gwin = nullptr;                 // [BATCH_CHUNK]
    uint64_t xbatch = 0;                                  // launches so far: selects the parity half
    ngs_stats_t stats{};
    unsigned long long* d_dbg = nullptr;  // profiling builds only
    uint32_t dbg_grid = 0;
    char err[512] = {0};
};

namespace {

constexpr uint32_t BATCH_CHUNK = 4096;
constexpr size_t XSLOT_HALF = (size_t)BATCH_CHUNK * NGS_MAX_RANKS;  // u64 words per parity half

int fail(ngs_dealer* h, int code, const char* fmt, ...) {
    if (h) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(h->err, sizeof(h->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

#define CU(h, call)                                                                                   \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return fail(h, NGS_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

int make_demand(ngs_dealer* h, const int32_t* percent, int n, int policy, PodDemand* pd) {
    if (n < 0 || (n > 0 && !percent)) return fail(h, NGS_ERR_INVALID, "demand: bad container count %d", n);
    if (n > NGS_MAX_CONTAINERS) return fail(h, NGS_ERR_UNSUPPORTED, "demand: %d containers > NGS_MAX_CONTAINERS", n);
    if (policy < 0 || policy > NGS_POLICY_RANDOM) return fail(h, NGS_ERR_INVALID, "unknown policy %d", policy);
    memset(pd, 0, sizeof(*pd));
    pd->n = n;
    pd->policy = policy;
    for (int i = 0; i < n; i++) {
        if (percent[i] < 0 || percent[i] > NGS_MAX_FREE) return fail(h, NGS_ERR_INVALID, "demand: percent[%d]=%d out of range", i, percent[i]);
        pd->percent[i] = percent[i];
    }
    return NGS_OK;
}

// floor division for possibly negative numerators
int64_t floor_div(int64_t a, int64_t b) {
    int64_t q = a / b, r = a % b;
    return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q;
}

int make_load(ngs_dealer* h, const ngs_load_ctx* L, LoadParams* lp) {
    memset(lp, 0, sizeof(*lp));
    if (!L) return NGS_OK;
    if (L->n_periods < 0) return fail(h, NGS_ERR_INVALID, "load ctx: n_periods < 0");
    if (L->n_periods > NGS_MAX_PERIODS) return fail(h, NGS_ERR_UNSUPPORTED, "load ctx: %d periods > NGS_MAX_PERIODS", L->n_periods);
    lp->enabled = 1;
    lp->n_periods = L->n_periods;
    for (int e = 0; e < L->n_periods; e++) {
        if (L->metric[e] != NGS_METRIC_CORE && L->metric[e] != NGS_METRIC_MEMORY) return fail(h, NGS_ERR_INVALID, "load ctx: metric[%d]", e);
        lp->metric[e] = L->metric[e];
        // now < ts*1e9 + window  <=>  ts >= floor((now - window) / 1e9) + 1
        const int64_t min_ts = floor_div(L->now_ns - L->window_ns[e], 1000000000LL) + 1;
        if (min_ts <= 0) lp->min_ts[e] = 0;
        else if (min_ts > 0xFFFFFFFFLL) lp->never[e] = 1;
        else lp->min_ts[e] = (uint32_t)min_ts;
    }
    lp->canonical = lp->n_periods == 2 && lp->metric[0] == NGS_METRIC_CORE && lp->metric[1] == NGS_METRIC_MEMORY && !lp->never[0] &&
                    !lp->never[1];
    return NGS_OK;
}

// Demand.Hash keyed plan cache (node.go:45-47): a new demand opens a new generation
uint32_t demand_generation(ngs_dealer* h, const PodDemand& pd, bool load_on) {
    std::vector<int32_t> d(pd.percent, pd.percent + pd.n);
    d.push_back(load_on ? 1 : 0);  // a switch of isLoadSchedule also starts a new generation
    if (load_on) {
        h->T.rl_zero = 0;  // a load-aware Rate writes RemainLoad (allocate.go:193)
    }
    if (d != h->cur_demand) {
        h->cur_demand.swap(d);
        h->gen++;
        if (h->gen == 0) h->gen = 1;
    }
    return h->gen;
}

int ensure_verb_scratch(ngs_dealer* h, size_t n, size_t plan_entries) {
    if (n > h->verb_cap) {
        size_t cap = n < 1024 ? 1024 : n;
        cudaFree(h->d_idx); cudaFree(h->d_feas); cudaFree(h->d_reason); cudaFree(h->d_score); cudaFree(h->d_nem);
        h->d_idx = nullptr; h->d_feas = nullptr; h->d_reason = nullptr; h->d_score = nullptr; h->d_nem = nullptr;
        h->verb_cap = 0;
        CU(h, cudaMalloc(&h->d_idx, cap * 4));
        CU(h, cudaMalloc(&h->d_feas, cap));
        CU(h, cudaMalloc(&h->d_reason, cap));
        CU(h, cudaMalloc(&h->d_score, cap * 4));
        CU(h, cudaMalloc(&h->d_nem, cap * 4));
        h->verb_cap = cap;
    }
    if (plan_entries > h->plan_cap) {
        cudaFree(h->d_plan);
        h->d_plan = nullptr;
        h->plan_cap = 0;
        CU(h, cudaMalloc(&h->d_plan, plan_entries * 4));
        h->plan_cap = plan_entries;
    }
    return NGS_OK;
}

int ensure_batch(ngs_dealer* h) {
    if (h->batch_cap) return NGS_OK;
    const uint32_t cap = BATCH_CHUNK;
    const size_t in_words = (size_t)cap * (NGS_MAX_CONTAINERS + 2), out_words = (size_t)cap * (NGS_MAX_CONTAINERS + 2);
    CU(h, cudaMalloc(&h->d_dem, sizeof(SortedDemand) * cap));
    // one exchange record per pod, pod-major, so that a batch of P pods clears ONE contiguous range:
    //   [0,16) arg-max word {key, counter}   [16,20) random-policy arrival counter   [32, 32+8*SMs) per-CTA slots
    h->pod_rec_bytes = (32 + sizeof(unsigned long long) * (size_t)h->num_sms + 31) / 32 * 32;
    CU(h, cudaMalloc(&h->d_keys, h->pod_rec_bytes * cap));
    // host-API staging lives in mapped pinned memory: the kernels read the pods and write the
    // decisions straight through PCIe, so a call needs no copy commands at all
    CU(h, cudaHostAlloc(&h->h_pod_in, (in_words + 8) * 4, cudaHostAllocMapped));
    CU(h, cudaHostAlloc(&h->h_pod_out, (out_words + 8) * 4, cudaHostAllocMapped));
    CU(h, cudaHostGetDevicePointer(&h->d_pod_in_dev, h->h_pod_in, 0));
    CU(h, cudaHostGetDevicePointer(&h->d_pod_out_dev, h->h_pod_out, 0));
    h->batch_cap = cap;
    return NGS_OK;
}

// Block size of k_schedule: two control warps on top of enough warps that, after leaving idle the
// ones sharing control warp 0's scheduler, one worker thread per row remains (at most 32 warps).
uint32_t schedule_block_size(uint32_t rpb) {
    const uint32_t need = (rpb + 31) / 32;  // worker warps wanted
    for (uint32_t nw = 1; nw <= 30; nw++) {
        const uint32_t csm = nw & 3u;
        const uint32_t workers = nw - ((nw + 3u - csm) >> 2);
        if (workers >= need) return (nw + 2) * 32;
    }
    return 1024;
}

// launch K6 for one chunk of <= BATCH_CHUNK pods whose inputs are in device memory
int launch_schedule(ngs_dealer* h, uint32_t P, const int32_t* d_percent, const int32_t* d_ncont, int max_c, const int32_t* d_policy,
                    const LoadParams& lp, uint64_t seed_seq, int32_t* d_node, int32_t* d_score, int32_t* d_gidx,
                    int32_t* status_out = nullptr) {
    if (!h->coop) return fail(h, NGS_ERR_CUDA, "device lacks cooperative launch");
    const uint32_t N = h->T.n;
    uint32_t grid = (N + 63) / 64;
    if (grid < 1) grid = 1;
    if (grid > (uint32_t)h->num_sms) grid = (uint32_t)h->num_sms;
    const uint32_t rpb = (N + grid - 1) / grid;
    const uint32_t block = schedule_block_size(rpb);
    const size_t smem = (size_t)rpb * 32 + (size_t)((rpb + 31) / 32) * 8 + 16;  // rows + two feasbits buffers
    if (smem + 2048 > (size_t)h->smem_optin)  // (2 KB: the kernel's static shared memory)
        return fail(h, NGS_ERR_UNSUPPORTED, "%u rows per SM need %zu B shared memory (> %d): shard the cluster over more GPUs", rpb, smem, h->smem_optin);
    {
        // the attribute is a per-device upper bound shared by every handle: allow the device maximum, once
        static bool s_done[64] = {false};
        static pthread_mutex_t s_mu = PTHREAD_MUTEX_INITIALIZER;
        const int dev = h->device & 63;
        pthread_mutex_lock(&s_mu);
        cudaError_t e = cudaSuccess;
        if (!s_done[dev]) {
            cudaFuncAttributes fa;
            e = cudaFuncGetAttributes(&fa, k_schedule);  // the opt-in limit covers static + dynamic shared memory
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(k_schedule, cudaFuncAttributeMaxDynamicSharedMemorySize, h->smem_optin - (int)fa.sharedSizeBytes);
            s_done[dev] = e == cudaSuccess;
        }
        pthread_mutex_unlock(&s_mu);
        CU(h, e);
    }
    k_prep_demands<<<(P + 127) / 128, 128, 0, h->stream>>>(d_percent, d_ncont, max_c, d_policy, P, h->d_dem, h->d_err);
    CU(h, cudaGetLastError());
    unsigned char* rec = reinterpret_cast<unsigned char*>(h->d_keys);
    ulonglong2* d_xw = reinterpret_cast<ulonglong2*>(rec);
    uint32_t* d_done = reinterpret_cast<uint32_t*>(rec + 16);
    CU(h, cudaMemsetAsync(rec, 0, h->pod_rec_bytes * (size_t)P, h->stream));
    if (lp.enabled) h->T.rl_zero = 0;  // the batch's load-aware Rate writes RemainLoad
    BatchParams bp;
    bp.T = h->T;
    bp.dem = h->d_dem;
    bp.P = P;
    bp.lp = lp;
    bp.seed_seq = seed_seq;
    bp.node_out = d_node;
    bp.score_out = d_score;
    bp.gpu_idx_out = d_gidx;
    bp.max_c = max_c;
    bp.keys = reinterpret_cast<unsigned long long*>(rec + 32);
    bp.done = d_done;
    bp.xw = d_xw;
    bp.keys_stride = (uint32_t)(h->pod_rec_bytes / 8);
    bp.done_stride = (uint32_t)(h->pod_rec_bytes / 4);
    bp.xw_stride = (uint32_t)(h->pod_rec_bytes / 16);
    bp.status_out = status_out;
    bp.world = (uint32_t)h->world;
    bp.rank = (uint32_t)h->rank;
    bp.gwin = h->d_gwin;
    for (int r = 0; r < NGS_MAX_RANKS; r++) bp.xslots[r] = nullptr;
    if (h->world > 1) {
        // Parity halves instead of a cross-rank barrier per batch: batch i uses half i&1 on every
        // rank; the other half is cleared now, before this batch starts.  A peer can only write into
        // that half once it has finished batch i, which needs this rank's batch-i posts first.
        const size_t cur = (size_t)(h->xbatch & 1), nxt = cur ^ 1;
        CU(h, cudaMemsetAsync(h->d_xslots + nxt * XSLOT_HALF, 0, sizeof(unsigned long long) * XSLOT_HALF, h->stream));
        CU(h, cudaMemsetAsync(h->d_gwin, 0, sizeof(unsigned long long) * (size_t)P, h->stream));
        for (int r = 0; r < h->world; r++) bp.xslots[r] = h->peer_xslots[r] + cur * XSLOT_HALF;
        h->xbatch++;
    }
    bp.err = h->d_err;
    bp.rows_per_block = rpb;
#ifdef NGS_PHASE_TIMING
    if (!h->d_dbg) CU(h, cudaMalloc(&h->d_dbg, sizeof(unsigned long long) * (4096 + 256 * 256 * 4)));
    CU(h, cudaMemsetAsync(h->d_dbg, 0, sizeof(unsigned long long) * (4096 + 256 * 256 * 4), h->stream));
    bp.dbg = h->d_dbg;
    h->dbg_grid = grid;
#endif
    void* args[] = {&bp};
    CU(h, cudaLaunchCooperativeKernel((void*)k_schedule, dim3(grid), dim3(block), args, smem, h->stream));
    h->stats.kernel_launches += 2;
    h->stats.node_evals += (uint64_t)P * N;
    h->stats.pods_scheduled += P;
    return NGS_OK;
}

int check_dev_err(ngs_dealer* h) {
    if (*h->h_err != NGS_OK) {
        int e = *h->h_err;
        *h->h_err = NGS_OK;
        cudaMemsetAsync(h->d_err, 0, 4, h->stream);
        if (e == NGS_ERR_TIMEOUT) return fail(h, e, "batch kernel gave up waiting for a peer CTA (results of this batch are incomplete)");
        return fail(h, e, "device-side validation failed (bad container count, percent or policy in a batch)");
    }
    return NGS_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
extern "C" {

uint32_t ngs_abi_version(void) { return NGS_ABI_VERSION; }

const char* ngs_strerror(int s) {
    switch (s) {
        case NGS_OK: return "ok";
        case NGS_ERR_INVALID: return "invalid argument";
        case NGS_ERR_CUDA: return "CUDA error";
        case NGS_ERR_UNKNOWN_NODE: return "unknown node";
        case NGS_ERR_INFEASIBLE: return "assume failed: demand does not fit the node";
        case NGS_ERR_APPLY: return "can't apply plan";
        case NGS_ERR_RANGE: return "capacity out of range";
        case NGS_ERR_UNSUPPORTED: return "unsupported size";
        case NGS_ERR_TIMEOUT: return "device-side wait timed out";
        case NGS_ERR_NOALLOC_LIMIT:
            return "ngs_schedule_batch_noalloc: more than NGS_NOALLOC_MAX_PODS pods or NGS_NOALLOC_MAX_CTAS CTAs (fixed scratch)";
        default: return "unknown status";
    }
}

const char* ngs_last_error(const ngs_dealer* h) { return h ? h->err : ""; }

int ngs_create(ngs_dealer** out, int device, uint32_t max_nodes, uint32_t first_node, void* stream) {
    if (!out || max_nodes == 0) return NGS_ERR_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return NGS_ERR_CUDA;
    ngs_dealer* h = new (std::nothrow) ngs_dealer();
    if (!h) return NGS_ERR_INVALID;
    h->device = device;
    *out = h;  // handed out even on failure so that ngs_last_error() can be read; destroy it either way
    CU(h, cudaSetDevice(device));
    CU(h, cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device));
    CU(h, cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    CU(h, cudaDeviceGetAttribute(&h->coop, cudaDevAttrCooperativeLaunch, device));
    if (stream) {
        h->stream = (cudaStream_t)stream;
    } else {
        CU(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
    }
    const size_t N = max_nodes;
    const size_t NP = (N + 255) / 256 * 256;  // tables are padded to whole 256-row tiles (the scan stages tiles with bulk copies)
    h->T.n = max_nodes;
    h->T.nstride = (uint32_t)NP;
    h->T.first = first_node;
    h->T.rl_zero = 1;
    CU(h, cudaMalloc(&h->T.free, NP * 8 * 4));
    CU(h, cudaMalloc(&h->T.gcount, NP));
    CU(h, cudaMalloc(&h->T.rl, NP * 8));
    CU(h, cudaMalloc(&h->T.tenths, 2 * NP * 8));
    CU(h, cudaMalloc(&h->T.ts, 2 * NP * 8 * 4));
    CU(h, cudaMalloc(&h->T.tag, NP * 4));
    CU(h, cudaMalloc(&h->T.cscore, NP * 4));
    CU(h, cudaMemsetAsync(h->T.free, 0, NP * 8 * 4, h->stream));
    CU(h, cudaMemsetAsync(h->T.gcount, 0, NP, h->stream));
    CU(h, cudaMemsetAsync(h->T.rl, 0, NP * 8, h->stream));
    CU(h, cudaMemsetAsync(h->T.tenths, 0xFF, 2 * NP * 8, h->stream));
    CU(h, cudaMemsetAsync(h->T.ts, 0, 2 * NP * 8 * 4, h->stream));
    CU(h, cudaMemsetAsync(h->T.tag, 0, NP * 4, h->stream));
    CU(h, cudaMemsetAsync(h->T.cscore, 0, NP * 4, h->stream));
    {
        int16_t* lut = nullptr;
        CU(h, cudaMalloc(&lut, sizeof(int16_t) * 9 * RATE_LUT_STRIDE));
        k_build_rate_lut<<<9, 256, 0, h->stream>>>(lut);
        CU(h, cudaGetLastError());
        h->T.blut = lut;
    }
    CU(h, cudaMemsetAsync(h->T.free, 0, N * 8 * 4, h->stream));
    CU(h, cudaMemsetAsync(h->T.gcount, 0, N, h->stream));
    CU(h, cudaMemsetAsync(h->T.rl, 0, N * 8, h->stream));
    CU(h, cudaMemsetAsync(h->T.tenths, 0xFF, 2 * N * 8, h->stream));
    CU(h, cudaMemsetAsync(h->T.ts, 0, 2 * N * 8 * 4, h->stream));
    CU(h, cudaMemsetAsync(h->T.tag, 0, N * 4, h->stream));
    CU(h, cudaMemsetAsync(h->T.cscore, 0, N * 4, h->stream));
    CU(h, cudaMalloc(&h->d_bind, sizeof(BindResult)));
    CU(h, cudaMallocHost(&h->h_bind, sizeof(BindResult)));
    CU(h, cudaMalloc(&h->d_small, 64 * 4));
    CU(h, cudaMallocHost(&h->h_small, 64 * 4));
    CU(h, cudaMalloc(&h->d_err, 4));
    CU(h, cudaMallocHost(&h->h_err, 4));
    *h->h_err = 0;
    CU(h, cudaMemsetAsync(h->d_err, 0, 4, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return NGS_OK;
}

int ngs_destroy(ngs_dealer* h) {
    if (!h) return NGS_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->T.free); cudaFree(h->T.gcount); cudaFree(h->T.rl); cudaFree(h->T.tenths); cudaFree(h->T.ts);
    cudaFree(h->T.tag); cudaFree(h->T.cscore); cudaFree(const_cast<int16_t*>(h->T.blut));
    cudaFree(h->d_idx); cudaFree(h->d_feas); cudaFree(h->d_reason); cudaFree(h->d_score); cudaFree(h->d_plan); cudaFree(h->d_nem);
    cudaFree(h->d_bind); cudaFreeHost(h->h_bind); cudaFree(h->d_small); cudaFreeHost(h->h_small);
    cudaFree(h->d_dem); cudaFree(h->d_keys);
    cudaFreeHost(h->h_pod_in); cudaFreeHost(h->h_pod_out);
    cudaFree(h->d_err); cudaFreeHost(h->h_err);
    for (int r = 0; r < NGS_MAX_RANKS; r++)
        if (h->peer_xslots[r] && r != h->rank) cudaIpcCloseMemHandle(h->peer_xslots[r]);
    cudaFree(h->d_xslots); cudaFree(h->d_gwin); cudaFree(h->d_rel);
    cudaFree(h->d_dbg);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return NGS_OK;
}

void* ngs_stream(ngs_dealer* h) { return h ? (void*)h->stream : nullptr; }

int ngs_sync(ngs_dealer* h) {
    if (!h) return NGS_ERR_INVALID;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaMemcpyAsync(h->h_err, h->d_err, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return check_dev_err(h);
}

#ifdef NGS_PHASE_TIMING
// profiling builds only: per-CTA cycles per phase of the last k_schedule launch ([grid][16])
int ngs_debug_phases(ngs_dealer* h, unsigned long long* out, uint32_t* grid) {
    if (!h || !h->d_dbg) return NGS_ERR_INVALID;
    CU(h, cudaMemcpyAsync(out, h->d_dbg, sizeof(unsigned long long) * (4096 + 256 * 256 * 4), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    *grid = h->dbg_grid;
    return NGS_OK;
}
#endif

int ngs_stats(ngs_dealer* h, ngs_stats_t* out) {
    if (!h || !out) return NGS_ERR_INVALID;
    *out = h->stats;
    return NGS_OK;
}

int ngs_node_upsert(ngs_dealer* h, uint32_t idx, int gpu_count) {
    if (!h) return NGS_ERR_INVALID;
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    if (gpu_count < 0) return fail(h, NGS_ERR_INVALID, "gpu_count %d", gpu_count);
    if (gpu_count > NGS_MAX_GPUS) return fail(h, NGS_ERR_UNSUPPORTED, "gpu_count %d > NGS_MAX_GPUS", gpu_count);
    CU(h, cudaSetDevice(h->device));
    k_upsert<<<1, 32, 0, h->stream>>>(h->T, idx - h->T.first, gpu_count);
    CU(h, cudaGetLastError());
    h->stats.kernel_launches++;
    return NGS_OK;
}

int ngs_snapshot_load(ngs_dealer* h, uint32_t first, uint32_t n, const uint8_t* gpu_count, const int32_t* free, const int8_t* remain_load) {
    if (!h || !gpu_count) return NGS_ERR_INVALID;
    if (first < h->T.first || (uint64_t)(first - h->T.first) + n > h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "rows [%u,%u) outside this shard", first, first + n);
    if (n == 0) return NGS_OK;
    const uint32_t r0 = first - h->T.first;
    std::vector<int32_t> fr((size_t)n * 8);
    std::vector<int8_t> rl((size_t)n * 8, 0);
    for (uint32_t i = 0; i < n; i++) {
        const int G = gpu_count[i];
        if (G > NGS_MAX_GPUS) return fail(h, NGS_ERR_UNSUPPORTED, "row %u: gpu_count %d > NGS_MAX_GPUS", first + i, G);
        for (int g = 0; g < 8; g++) {
            int32_t v = g < G ? (free ? free[(size_t)i * 8 + g] : 100) : 0;
            if (v < 0 || v > NGS_MAX_FREE) return fail(h, NGS_ERR_RANGE, "row %u cell %d: free %d out of range", first + i, g, v);
            fr[(size_t)i * 8 + g] = v;
            int r = (g < G && remain_load) ? remain_load[(size_t)i * 8 + g] : 0;
            if (r < -2 || r > 2) return fail(h, NGS_ERR_RANGE, "row %u cell %d: remain_load %d out of range", first + i, g, r);
            rl[(size_t)i * 8 + g] = (int8_t)r;
        }
    }
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaMemcpyAsync(h->T.free + (size_t)r0 * 8, fr.data(), fr.size() * 4, cudaMemcpyHostToDevice, h->stream));
    if (remain_load) h->T.rl_zero = 0;
    CU(h, cudaMemcpyAsync(h->T.rl + (size_t)r0 * 8, rl.data(), rl.size(), cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(h->T.gcount + r0, gpu_count, n, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemsetAsync(h->T.tag + r0, 0, (size_t)n * 4, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.h2d_bytes += fr.size() * 4 + rl.size() + n;
    return NGS_OK;
}

int ngs_snapshot_read(ngs_dealer* h, uint32_t first, uint32_t n, uint8_t* gpu_count, int32_t* free, int8_t* remain_load) {
    if (!h) return NGS_ERR_INVALID;
    if (first < h->T.first || (uint64_t)(first - h->T.first) + n > h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "rows [%u,%u) outside this shard", first, first + n);
    if (n == 0) return NGS_OK;
    const uint32_t r0 = first - h->T.first;
    CU(h, cudaSetDevice(h->device));
    if (gpu_count) CU(h, cudaMemcpyAsync(gpu_count, h->T.gcount + r0, n, cudaMemcpyDeviceToHost, h->stream));
    if (free) CU(h, cudaMemcpyAsync(free, h->T.free + (size_t)r0 * 8, (size_t)n * 32, cudaMemcpyDeviceToHost, h->stream));
    if (remain_load) CU(h, cudaMemcpyAsync(remain_load, h->T.rl + (size_t)r0 * 8, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.d2h_bytes += (gpu_count ? n : 0) + (free ? (size_t)n * 32 : 0) + (remain_load ? (size_t)n * 8 : 0);
    return NGS_OK;
}

int ngs_node_apply(ngs_dealer* h, uint32_t idx, const int32_t* gpu_idx, const int32_t* percent, int n_containers, int sign) {
    if (!h || sign == 0) return NGS_ERR_INVALID;
    PodDemand pd;
    int rc = make_demand(h, percent, n_containers, NGS_POLICY_BINPACK, &pd);
    if (rc) return rc;
    if (n_containers > 0 && !gpu_idx) return fail(h, NGS_ERR_INVALID, "apply: gpu_idx is NULL");
    CU(h, cudaSetDevice(h->device));
    for (int i = 0; i < n_containers; i++) h->h_small[i] = gpu_idx[i];
    CU(h, cudaMemcpyAsync(h->d_small, h->h_small, NGS_MAX_CONTAINERS * 4, cudaMemcpyHostToDevice, h->stream));
    k_apply<<<1, 32, 0, h->stream>>>(h->T, idx, pd, h->d_small, sign, h->d_small + 32);
    CU(h, cudaGetLastError());
    CU(h, cudaMemcpyAsync(h->h_small + 32, h->d_small + 32, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.kernel_launches++;
    rc = h->h_small[32];
    if (rc) return fail(h, rc, "apply on node %u: %s", idx, ngs_strerror(rc));
    return NGS_OK;
}

int ngs_release_batch(ngs_dealer* h, uint32_t n, const uint32_t* node, const int32_t* gpu_idx, const int32_t* percent,
                      const int32_t* n_containers, int max_c, int32_t* status_out) {
    if (!h || (n > 0 && (!node || !gpu_idx || !percent || !n_containers))) return NGS_ERR_INVALID;
    if (max_c < 1 || max_c > NGS_MAX_CONTAINERS) return fail(h, NGS_ERR_UNSUPPORTED, "max_c %d outside [1, NGS_MAX_CONTAINERS]", max_c);
    if (n == 0) return NGS_OK;
    CU(h, cudaSetDevice(h->device));
    const size_t cells = (size_t)n * max_c;
    const size_t words = 3 * (size_t)n + 2 * cells;  // node | ncont | status | gpu_idx | percent
    if (words > h->rel_cap) {
        cudaFree(h->d_rel);
        h->d_rel = nullptr;
        h->rel_cap = 0;
        CU(h, cudaMalloc(&h->d_rel, words * 4));
        h->rel_cap = words;
    }
    uint32_t* d_node = reinterpret_cast<uint32_t*>(h->d_rel);
    int32_t* d_nc = h->d_rel + n;
    int32_t* d_st = h->d_rel + 2 * (size_t)n;
    int32_t* d_gi = h->d_rel + 3 * (size_t)n;
    int32_t* d_pc = d_gi + cells;
    CU(h, cudaMemcpyAsync(d_node, node, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(d_nc, n_containers, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(d_gi, gpu_idx, cells * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(d_pc, percent, cells * 4, cudaMemcpyHostToDevice, h->stream));
    k_release_batch<<<(n + 127) / 128, 128, 0, h->stream>>>(h->T, n, d_node, d_gi, d_pc, d_nc, max_c, d_st);
    CU(h, cudaGetLastError());
    std::vector<int32_t> st(n);
    CU(h, cudaMemcpyAsync(st.data(), d_st, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.kernel_launches++;
    h->stats.h2d_bytes += (2 * (size_t)n + 2 * cells) * 4;
    h->stats.d2h_bytes += (size_t)n * 4;
    int first_bad = NGS_OK;
    for (uint32_t i = 0; i < n; i++) {
        if (status_out) status_out[i] = st[i];
        if (st[i] != NGS_OK && first_bad == NGS_OK) first_bad = st[i];
    }
    if (first_bad != NGS_OK) return fail(h, first_bad, "release batch: at least one pod failed (%s); see the per-pod status", ngs_strerror(first_bad));
    return NGS_OK;
}

static uint8_t encode_tenths(double value, int value_ok, int64_t ts_sec, int ts_ok) {
    // nodeusage.go:99-110 + allocate.go:190: every early-out means "skip this entry"
    if (!value_ok || !ts_ok || std::isnan(value) || value < 0 || value > 1) return 255;
    if (ts_sec < 0 || ts_sec > 0xFFFFFFFFLL) return 255;
    return (uint8_t)std::ceil(10 * value);
}

int ngs_load_add_node(ngs_dealer* h, uint32_t idx, int metric) {
    if (!h || metric < 0 || metric > 1) return NGS_ERR_INVALID;
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    CU(h, cudaSetDevice(h->device));
    k_load_add_node<<<1, 32, 0, h->stream>>>(h->T, idx - h->T.first, metric);
    CU(h, cudaGetLastError());
    h->stats.kernel_launches++;
    return NGS_OK;
}

int ngs_load_update(ngs_dealer* h, uint32_t idx, int metric, int card, double value, int value_ok, int64_t ts_sec, int ts_ok) {
    if (!h || metric < 0 || metric > 1) return NGS_ERR_INVALID;
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    if (card < 0) return fail(h, NGS_ERR_INVALID, "card %d", card);
    if (card >= NGS_MAX_GPUS) return NGS_OK;  // a card the kernels never look at (LoadUsage walks gpus only)
    CU(h, cudaSetDevice(h->device));
    const uint8_t t10 = encode_tenths(value, value_ok, ts_sec, ts_ok);
    k_load_update<<<1, 32, 0, h->stream>>>(h->T, idx - h->T.first, metric, card, t10, t10 == 255 ? 0u : (uint32_t)ts_sec);
    CU(h, cudaGetLastError());
    h->stats.kernel_launches++;
    return NGS_OK;
}

int ngs_load_update_bulk(ngs_dealer* h, int metric, uint32_t first, uint32_t n, const double* value, const uint8_t* value_ok,
                         const int64_t* ts_sec, const uint8_t* ts_ok) {
    if (!h || metric < 0 || metric > 1 || !value || !ts_sec) return NGS_ERR_INVALID;
    if (first < h->T.first || (uint64_t)(first - h->T.first) + n > h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "rows [%u,%u) outside this shard", first, first + n);
    if (n == 0) return NGS_OK;
    const size_t cells = (size_t)n * 8;
    std::vector<uint8_t> t10(cells);
    std::vector<uint32_t> ts(cells);
    for (size_t i = 0; i < cells; i++) {
        t10[i] = encode_tenths(value[i], value_ok ? value_ok[i] : 1, ts_sec[i], ts_ok ? ts_ok[i] : 1);
        ts[i] = t10[i] == 255 ? 0u : (uint32_t)ts_sec[i];
    }
    const size_t off = ((size_t)metric * h->T.nstride + (first - h->T.first)) * 8;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaMemcpyAsync(h->T.tenths + off, t10.data(), cells, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(h->T.ts + off, ts.data(), cells * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.h2d_bytes += cells * 5;
    return NGS_OK;
}

// launch the streaming scan over rows [0, n): TMA-staged tiles, feasibility / reason / score only
static int launch_scan(ngs_dealer* h, uint32_t n, const PodDemand& pd, const LoadParams& lp, uint32_t gen, uint8_t* d_feas, int8_t* d_reason,
                       int32_t* d_score) {
    const uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (lp.enabled) {
        // load-aware pass: instruction-heavier per row; measured best with direct 256-bit loads at full
        // occupancy (k_assume, 73 % of HBM peak) rather than TMA staging at 2-3 CTAs per SM (66-69 %)
        int grid = (int)((n + 255) / 256);
        const int max_grid = h->num_sms * NGS_K1_BLOCKS;
        if (grid > max_grid) grid = max_grid;
        k_assume<false><<<grid, 256, 0, h->stream>>>(h->T, nullptr, n, pd, lp, MODE_ASSUME, gen, d_feas, d_reason, d_score, nullptr, nullptr);
    } else {
        const size_t smem = (size_t)NGS_SCAN_STAGES_PLAIN * scan_stage_bytes<false>();
        uint32_t grid = (uint32_t)h->num_sms * 4;
        if (grid > ntiles) grid = ntiles;
        k_scan_tma<false><<<grid, SCAN_TILE, smem, h->stream>>>(h->T, n, pd, lp, gen, d_feas, d_reason, d_score);
    }
    CU(h, cudaGetLastError());
    return NGS_OK;
}

static int run_assume(ngs_dealer* h, int mode, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
                      const ngs_load_ctx* load, uint8_t* feasible, int8_t* reason, int32_t* score, int32_t* plan, int32_t* n_emitted) {
    if (!h) return NGS_ERR_INVALID;
    PodDemand pd;
    LoadParams lp;
    int rc = make_demand(h, percent, n_containers, policy, &pd);
    if (rc) return rc;
    rc = make_load(h, load, &lp);
    if (rc) return rc;
    if (!idx && n > h->T.n) return fail(h, NGS_ERR_INVALID, "n=%u exceeds the %u rows of this shard", n, h->T.n);
    if (n == 0) return NGS_OK;
    CU(h, cudaSetDevice(h->device));
    const bool want_plan = plan || n_emitted;
    rc = ensure_verb_scratch(h, n, want_plan ? (size_t)n * (n_containers > 0 ? n_containers : 1) : 0);
    if (rc) return rc;
    const uint32_t gen = demand_generation(h, pd, lp.enabled != 0);
    if (idx) {
        CU(h, cudaMemcpyAsync(h->d_idx, idx, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
        h->stats.h2d_bytes += (size_t)n * 4;
    }
    const int block = 256;
    int grid = (int)((n + block - 1) / block);
    const int max_grid = h->num_sms * (want_plan ? 2 : NGS_K1_BLOCKS);  // exactly the resident blocks: measured best
    if (grid > max_grid) grid = max_grid;
    if (!idx && mode == MODE_ASSUME && !want_plan) {
        rc = launch_scan(h, n, pd, lp, gen, feasible ? h->d_feas : nullptr, reason ? h->d_reason : nullptr, score ? h->d_score : nullptr);
        if (rc) return rc;
    } else if (want_plan)
        k_assume<true><<<grid, block, 0, h->stream>>>(h->T, idx ? h->d_idx : nullptr, n, pd, lp, mode, gen, feasible ? h->d_feas : nullptr,
                                                     reason ? h->d_reason : nullptr, score ? h->d_score : nullptr,
                                                     plan ? h->d_plan : nullptr, n_emitted ? h->d_nem : nullptr);
    else
        k_assume<false><<<grid, block, 0, h->stream>>>(h->T, idx ? h->d_idx : nullptr, n, pd, lp, mode, gen, feasible ? h->d_feas : nullptr,
                                                      reason ? h->d_reason : nullptr, score ? h->d_score : nullptr, nullptr, nullptr);
    CU(h, cudaGetLastError());
    if (feasible) CU(h, cudaMemcpyAsync(feasible, h->d_feas, n, cudaMemcpyDeviceToHost, h->stream));
    if (reason) CU(h, cudaMemcpyAsync(reason, h->d_reason, n, cudaMemcpyDeviceToHost, h->stream));
    if (score) CU(h, cudaMemcpyAsync(score, h->d_score, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    if (plan && n_containers > 0) CU(h, cudaMemcpyAsync(plan, h->d_plan, (size_t)n * n_containers * 4, cudaMemcpyDeviceToHost, h->stream));
    if (n_emitted) CU(h, cudaMemcpyAsync(n_emitted, h->d_nem, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.kernel_launches++;
    h->stats.node_evals += n;
    h->stats.d2h_bytes += (feasible ? n : 0) + (reason ? n : 0) + (score ? (size_t)n * 4 : 0) +
                          (plan ? (size_t)n * n_containers * 4 : 0) + (n_emitted ? (size_t)n * 4 : 0);
    return NGS_OK;
}

int ngs_assume(ngs_dealer* h, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
               const ngs_load_ctx* load, uint8_t* feasible, int8_t* reason, int32_t* score, int32_t* plan, int32_t* n_emitted) {
    return run_assume(h, MODE_ASSUME, idx, n, percent, n_containers, policy, load, feasible, reason, score, plan, n_emitted);
}

// DealerImpl.Assume + Score over rows [0, n) with the per-node outputs left in device memory
// (no copies, no synchronisation): what a caller that keeps the arg-max on the GPU uses, and what
// bench.py times for the scan kernel's own roofline.
int ngs_assume_dev(ngs_dealer* h, uint32_t n, const int32_t* percent, int n_containers, int policy, const ngs_load_ctx* load,
                   uint8_t* d_feasible, int32_t* d_score) {
    if (!h) return NGS_ERR_INVALID;
    PodDemand pd;
    LoadParams lp;
    int rc = make_demand(h, percent, n_containers, policy, &pd);
    if (rc) return rc;
    rc = make_load(h, load, &lp);
    if (rc) return rc;
    if (n > h->T.n) return fail(h, NGS_ERR_INVALID, "n=%u exceeds the %u rows of this shard", n, h->T.n);
    if (n == 0) return NGS_OK;
    CU(h, cudaSetDevice(h->device));
    const uint32_t gen = demand_generation(h, pd, lp.enabled != 0);
    rc = launch_scan(h, n, pd, lp, gen, d_feasible, nullptr, d_score);
    if (rc) return rc;
    h->stats.kernel_launches++;
    h->stats.node_evals += n;
    return NGS_OK;
}

int ngs_score(ngs_dealer* h, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
              const ngs_load_ctx* load, int32_t* score) {
    if (!score) return NGS_ERR_INVALID;
    return run_assume(h, MODE_SCORE, idx, n, percent, n_containers, policy, load, nullptr, nullptr, score, nullptr, nullptr);
}

int ngs_bind(ngs_dealer* h, uint32_t idx, const int32_t* percent, int n_containers, int policy, const ngs_load_ctx* load,
             int32_t* gpu_idx_out, int32_t* score_out) {
    if (!h) return NGS_ERR_INVALID;
    PodDemand pd;
    LoadParams lp;
    int rc = make_demand(h, percent, n_containers, policy, &pd);
    if (rc) return rc;
    rc = make_load(h, load, &lp);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    const uint32_t gen = demand_generation(h, pd, lp.enabled != 0);
    k_bind<<<1, 32, 0, h->stream>>>(h->T, idx, pd, lp, gen, h->d_bind);
    CU(h, cudaGetLastError());
    CU(h, cudaMemcpyAsync(h->h_bind, h->d_bind, sizeof(BindResult), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.kernel_launches++;
    h->stats.node_evals++;
    h->stats.d2h_bytes += sizeof(BindResult);
    rc = h->h_bind->status;
    if (rc) return fail(h, rc, "bind on node %u: %s", idx, ngs_strerror(rc));
    for (int i = 0; i < n_containers; i++)
        if (gpu_idx_out) gpu_idx_out[i] = h->h_bind->gpu_idx[i];
    if (score_out) *score_out = h->h_bind->score;
    h->stats.pods_scheduled++;
    return NGS_OK;
}

int ngs_schedule_batch_dev(ngs_dealer* h, uint32_t n_pods, const int32_t* d_percent, const int32_t* d_n_containers, int max_c,
                           const int32_t* d_policy, const ngs_load_ctx* load, uint64_t seed, uint64_t seq0, int32_t* d_node_out,
                           int32_t* d_score_out, int32_t* d_gpu_idx_out) {
    if (!h || !d_percent || !d_n_containers || !d_policy || !d_node_out || !d_score_out || !d_gpu_idx_out) return NGS_ERR_INVALID;
    if (max_c < 1 || max_c > NGS_MAX_CONTAINERS) return fail(h, NGS_ERR_UNSUPPORTED, "max_c %d outside [1, NGS_MAX_CONTAINERS]", max_c);
    if (n_pods > NGS_MAX_BATCH) return fail(h, NGS_ERR_UNSUPPORTED, "%u pods > NGS_MAX_BATCH", n_pods);
    LoadParams lp;
    int rc = make_load(h, load, &lp);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    rc = ensure_batch(h);
    if (rc) return rc;
    for (uint32_t p0 = 0; p0 < n_pods; p0 += BATCH_CHUNK) {
        const uint32_t P = n_pods - p0 < BATCH_CHUNK ? n_pods - p0 : BATCH_CHUNK;
        rc = launch_schedule(h, P, d_percent + (size_t)p0 * max_c, d_n_containers + p0, max_c, d_policy + p0, lp, seed + seq0 + p0,
                             d_node_out + p0, d_score_out + p0, d_gpu_idx_out + (size_t)p0 * max_c);
        if (rc) return rc;
    }
    return NGS_OK;
}

int ngs_schedule_batch(ngs_dealer* h, uint32_t n_pods, const int32_t* percent, const int32_t* n_containers, int max_c,
                       const int32_t* policy, const ngs_load_ctx* load, uint64_t seed, uint64_t seq0, int32_t* node_out,
                       int32_t* score_out, int32_t* gpu_idx_out) {
    if (!h || !percent || !n_containers || !policy || !node_out || !score_out || !gpu_idx_out) return NGS_ERR_INVALID;
    if (max_c < 1 || max_c > NGS_MAX_CONTAINERS) return fail(h, NGS_ERR_UNSUPPORTED, "max_c %d outside [1, NGS_MAX_CONTAINERS]", max_c);
    if (n_pods > NGS_MAX_BATCH) return fail(h, NGS_ERR_UNSUPPORTED, "%u pods > NGS_MAX_BATCH", n_pods);
    for (uint32_t p = 0; p < n_pods; p++) {
        if (n_containers[p] < 0 || n_containers[p] > max_c) return fail(h, NGS_ERR_INVALID, "pod %u: %d containers", p, n_containers[p]);
        if (policy[p] < 0 || policy[p] > NGS_POLICY_RANDOM) return fail(h, NGS_ERR_INVALID, "pod %u: policy %d", p, policy[p]);
        for (int k = 0; k < n_containers[p]; k++) {
            const int32_t v = percent[(size_t)p * max_c + k];
            if (v < 0 || v > NGS_MAX_FREE) return fail(h, NGS_ERR_INVALID, "pod %u: percent[%d]=%d out of range", p, k, v);
        }
    }
    LoadParams lp;
    int rc = make_load(h, load, &lp);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    rc = ensure_batch(h);
    if (rc) return rc;
    for (uint32_t p0 = 0; p0 < n_pods; p0 += BATCH_CHUNK) {
        const uint32_t P = n_pods - p0 < BATCH_CHUNK ? n_pods - p0 : BATCH_CHUNK;
        // mapped pinned staging: percent [P][max_c] | ncont [P] | policy [P]; outputs: status | node | score | gpu_idx
        int32_t* hin = h->h_pod_in;
        memcpy(hin, percent + (size_t)p0 * max_c, (size_t)P * max_c * 4);
        memcpy(hin + (size_t)P * max_c, n_containers + p0, (size_t)P * 4);
        memcpy(hin + (size_t)P * max_c + P, policy + p0, (size_t)P * 4);
        const size_t in_bytes = ((size_t)P * max_c + 2 * (size_t)P) * 4;
        const int32_t* din = h->d_pod_in_dev;
        int32_t* d_status = h->d_pod_out_dev;  // word 0..7: status (padded so that the arrays stay 32-byte aligned)
        int32_t* d_node = h->d_pod_out_dev + 8;
        int32_t* d_score = d_node + P;
        int32_t* d_gidx = d_score + P;
        h->h_pod_out[0] = NGS_ERR_CUDA;  // overwritten by the kernel's last act
        rc = launch_schedule(h, P, din, din + (size_t)P * max_c, max_c, din + (size_t)P * max_c + P, lp, seed + seq0 + p0, d_node,
                             d_score, d_gidx, d_status);
        if (rc) return rc;
        const size_t out_bytes = (2 * (size_t)P + (size_t)P * max_c) * 4;
        CU(h, cudaStreamSynchronize(h->stream));
        *h->h_err = h->h_pod_out[0];
        rc = check_dev_err(h);
        if (rc) return rc;
        const int32_t* hout = h->h_pod_out + 8;
        memcpy(node_out + p0, hout, (size_t)P * 4);
        memcpy(score_out + p0, hout + P, (size_t)P * 4);
        memcpy(gpu_idx_out + (size_t)p0 * max_c, hout + 2 * (size_t)P, (size_t)P * max_c * 4);
        h->stats.h2d_bytes += in_bytes;
        h->stats.d2h_bytes += out_bytes;
    }
    return NGS_OK;
}

int ngs_xchg_export(ngs_dealer* h, void* handle_out) {
    if (!h || !handle_out) return NGS_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == NGS_IPC_HANDLE_BYTES, "IPC handle size");
    CU(h, cudaSetDevice(h->device));
    if (!h->d_xslots) {
        CU(h, cudaMalloc(&h->d_xslots, sizeof(unsigned long long) * 2 * XSLOT_HALF));
        CU(h, cudaMalloc(&h->d_gwin, sizeof(unsigned long long) * BATCH_CHUNK));
        CU(h, cudaMemset(h->d_xslots, 0, sizeof(unsigned long long) * 2 * XSLOT_HALF));
        CU(h, cudaMemset(h->d_gwin, 0, sizeof(unsigned long long) * BATCH_CHUNK));
        CU(h, cudaDeviceSynchronize());  // cleared before any peer can learn the handle
    }
    cudaIpcMemHandle_t mh;
    CU(h, cudaIpcGetMemHandle(&mh, h->d_xslots));
    memcpy(handle_out, &mh, sizeof(mh));
    return NGS_OK;
}

int ngs_xchg_attach(ngs_dealer* h, int rank, int world, const void* handles) {
    if (!h || !handles || world < 1 || world > NGS_MAX_RANKS || rank < 0 || rank >= world) return NGS_ERR_INVALID;
    if (!h->d_xslots) return fail(h, NGS_ERR_INVALID, "ngs_xchg_attach before ngs_xchg_export");
    CU(h, cudaSetDevice(h->device));
    const unsigned char* hb = static_cast<const unsigned char*>(handles);
    for (int r = 0; r < world; r++) {
        if (r == rank) {
            h->peer_xslots[r] = h->d_xslots;
            continue;
        }
        cudaIpcMemHandle_t mh;
        memcpy(&mh, hb + (size_t)r * NGS_IPC_HANDLE_BYTES, sizeof(mh));
        void* ptr = nullptr;
        CU(h, cudaIpcOpenMemHandle(&ptr, mh, cudaIpcMemLazyEnablePeerAccess));
        h->peer_xslots[r] = static_cast<unsigned long long*>(ptr);
    }
    h->rank = rank;
    h->world = world;
    h->xbatch = 0;
    return NGS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// ngs_schedule_batch_noalloc: the batch kernel for hosts that may not allocate device memory
// (see include/ngs_dealer.h for who they are).  Same kernels as ngs_schedule_batch_dev; the only
// difference is where the scratch comes from: a fixed array in this library's device image
// instead of buffers the handle cudaMalloc'ed.
// ---------------------------------------------------------------------------------------------
namespace {

constexpr size_t NA_DEM_BYTES = sizeof(SortedDemand) * NGS_NOALLOC_MAX_PODS;                                      // sorted demands
constexpr size_t NA_KEYS_BYTES = sizeof(unsigned long long) * (size_t)NGS_NOALLOC_MAX_PODS * NGS_NOALLOC_MAX_CTAS;  // per-CTA slots
constexpr size_t NA_MX_BYTES = sizeof(ulonglong2) * NGS_NOALLOC_MAX_PODS;                                          // arg-max exchange words
constexpr size_t NA_DONE_BYTES = sizeof(uint32_t) * NGS_NOALLOC_MAX_PODS;                                          // arrival counters
constexpr size_t NA_DEM_OFF = 0;
constexpr size_t NA_KEYS_OFF = NA_DEM_OFF + NA_DEM_BYTES;
constexpr size_t NA_MX_OFF = NA_KEYS_OFF + NA_KEYS_BYTES;
constexpr size_t NA_DONE_OFF = NA_MX_OFF + NA_MX_BYTES;
constexpr size_t NA_ERR_OFF = NA_DONE_OFF + NA_DONE_BYTES;
constexpr size_t NA_TOTAL_BYTES = NA_ERR_OFF + 256;
static_assert(NA_KEYS_OFF % 256 == 0 && NA_MX_OFF % 256 == 0 && NA_DONE_OFF % 256 == 0 && NA_ERR_OFF % 256 == 0, "scratch layout");

// The scratch itself.  In host code this name is only a handle for the runtime, not a device
// address: the address the kernels get comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_TOTAL_BYTES];

struct NoAllocDevice {
    bool ready = false;
    unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
    int num_sms = 0, smem_optin = 0, coop = 0;
};
NoAllocDevice g_na_dev[64];
pthread_mutex_t g_na_mu = PTHREAD_MUTEX_INITIALIZER;

// once per device, with that device current: resolve the scratch array's address there
int noalloc_setup(int device, NoAllocDevice** out) {
    if (device < 0 || device >= 64) return NGS_ERR_INVALID;
    NoAllocDevice* d = &g_na_dev[device];
    int rc = NGS_OK;
    pthread_mutex_lock(&g_na_mu);
    if (!d->ready) {
        void* ptr = nullptr;
        if (cudaGetSymbolAddress(&ptr, g_noalloc_scratch) != cudaSuccess || ptr == nullptr ||
            cudaDeviceGetAttribute(&d->num_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess ||
            cudaDeviceGetAttribute(&d->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device) != cudaSuccess ||
            cudaDeviceGetAttribute(&d->coop, cudaDevAttrCooperativeLaunch, device) != cudaSuccess) {
            rc = NGS_ERR_CUDA;
        } else {
            d->scratch = static_cast<unsigned char*>(ptr);
            d->ready = true;
        }
    }
    pthread_mutex_unlock(&g_na_mu);
    *out = d;
    return rc;
}

}  // namespace

extern "C" int ngs_schedule_batch_noalloc(int device, void* stream, const ngs_snapshot_view* snap, uint32_t n_pods,
                                          const int32_t* d_percent, const int32_t* d_n_containers, int max_c, const int32_t* d_policy,
                                          const ngs_load_ctx* load, uint64_t seed, uint64_t seq0, int32_t* d_node_out,
                                          int32_t* d_score_out, int32_t* d_gpu_idx_out) {
    if (!snap || !snap->free || !snap->gpu_count || !snap->remain_load || !snap->tenths || !snap->ts || !snap->plan_tag ||
        !snap->plan_score || snap->n_nodes == 0)
        return NGS_ERR_INVALID;
    if (!d_percent || !d_n_containers || !d_policy || !d_node_out || !d_score_out || !d_gpu_idx_out) return NGS_ERR_INVALID;
    if (max_c < 1 || max_c > NGS_MAX_CONTAINERS) return NGS_ERR_UNSUPPORTED;
    if (n_pods > NGS_NOALLOC_MAX_PODS) return NGS_ERR_NOALLOC_LIMIT;  // the sorted demands would not fit
    if (n_pods == 0) return NGS_OK;
    LoadParams lp;
    int rc = make_load(nullptr, load, &lp);
    if (rc) return rc;
    if (cudaSetDevice(device) != cudaSuccess) return NGS_ERR_CUDA;
    NoAllocDevice* dv = nullptr;
    rc = noalloc_setup(device, &dv);  // first call on this device: cudaGetSymbolAddress
    if (rc) return rc;
    if (!dv->coop) return NGS_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // same launch shape as the handle path
    const uint32_t N = snap->n_nodes;
    uint32_t grid = (N + 63) / 64;
    if (grid < 1) grid = 1;
    if (grid > (uint32_t)dv->num_sms) grid = (uint32_t)dv->num_sms;
    if (grid > NGS_NOALLOC_MAX_CTAS) return NGS_ERR_NOALLOC_LIMIT;  // the per-CTA slots would not fit
    const uint32_t rpb = (N + grid - 1) / grid;
    const uint32_t block = schedule_block_size(rpb);
    const size_t smem = (size_t)rpb * 32 + (size_t)((rpb + 31) / 32) * 8 + 16;
    if (smem + 2048 > (size_t)dv->smem_optin) return NGS_ERR_UNSUPPORTED;

    unsigned char* sc = dv->scratch;
    SortedDemand* dem = reinterpret_cast<SortedDemand*>(sc + NA_DEM_OFF);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(sc + NA_KEYS_OFF);
    ulonglong2* mx = reinterpret_cast<ulonglong2*>(sc + NA_MX_OFF);
    uint32_t* done = reinterpret_cast<uint32_t*>(sc + NA_DONE_OFF);
    int32_t* err = reinterpret_cast<int32_t*>(sc + NA_ERR_OFF);
    // the array still holds the previous call's words: everything the kernel expects to find at
    // zero is cleared here, in stream order, before the launch
    if (cudaMemsetAsync(keys, 0, sizeof(unsigned long long) * (size_t)n_pods * grid, st) != cudaSuccess ||
        cudaMemsetAsync(mx, 0, sizeof(ulonglong2) * (size_t)n_pods, st) != cudaSuccess ||
        cudaMemsetAsync(done, 0, sizeof(uint32_t) * (size_t)n_pods, st) != cudaSuccess ||
        cudaMemsetAsync(err, 0, sizeof(int32_t), st) != cudaSuccess)
        return NGS_ERR_CUDA;
    // (an upper bound per device: asking for the device maximum once keeps it valid for every later launch, also of handles)
    {
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, k_schedule) != cudaSuccess ||
            cudaFuncSetAttribute(k_schedule, cudaFuncAttributeMaxDynamicSharedMemorySize, dv->smem_optin - (int)fa.sharedSizeBytes) != cudaSuccess)
            return NGS_ERR_CUDA;
    }
    k_prep_demands<<<(n_pods + 127) / 128, 128, 0, st>>>(d_percent, d_n_containers, max_c, d_policy, n_pods, dem, err);
    if (cudaGetLastError() != cudaSuccess) return NGS_ERR_CUDA;

    BatchParams bp;
    memset(&bp, 0, sizeof(bp));
    bp.T.free = snap->free;
    bp.T.gcount = snap->gpu_count;
    bp.T.rl = snap->remain_load;
    bp.T.tenths = snap->tenths;
    bp.T.ts = snap->ts;
    bp.T.tag = snap->plan_tag;
    bp.T.cscore = snap->plan_score;
    bp.T.blut = nullptr;  // no table to borrow without a handle: Binpack.Rate runs in fp64
    bp.T.rl_zero = 0;     // caller-owned table: always read
    bp.T.n = snap->n_nodes;
    bp.T.nstride = snap->n_nodes;  // caller-owned tables: planes are exactly n_nodes rows apart
    bp.T.first = snap->first_node;
    bp.dem = dem;
    bp.P = n_pods;
    bp.lp = lp;
    bp.seed_seq = seed + seq0;
    bp.node_out = d_node_out;
    bp.score_out = d_score_out;
    bp.gpu_idx_out = d_gpu_idx_out;
    bp.max_c = max_c;
    bp.keys = keys;
    bp.done = done;
    bp.xw = mx;
    bp.keys_stride = grid;
    bp.done_stride = 1;
    bp.xw_stride = 1;
    bp.status_out = nullptr;
    bp.world = 1;
    bp.rank = 0;
    bp.gwin = nullptr;
    bp.err = err;
    bp.rows_per_block = rpb;
    void* args[] = {&bp};
    if (cudaLaunchCooperativeKernel((void*)k_schedule, dim3(grid), dim3(block), args, smem, st) != cudaSuccess) return NGS_ERR_CUDA;
    int32_t dev_status = NGS_OK;  // device-side validation / timeout word of this batch
    if (cudaMemcpyAsync(&dev_status, err, sizeof(dev_status), cudaMemcpyDeviceToHost, st) != cudaSuccess) return NGS_ERR_CUDA;
    if (cudaStreamSynchronize(st) != cudaSuccess) return NGS_ERR_CUDA;
    return dev_status;
}
