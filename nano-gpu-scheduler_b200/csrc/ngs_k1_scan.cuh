// ngs_k1_scan.cuh — K1, the filter / prioritize scans: k_assume (one thread per candidate, direct
// 256-bit loads) and k_scan_tma (all rows, tiles staged by TMA bulk copies).
// Included by ngs_dealer.cu inside its anonymous namespace, after ngs_rows.cuh.
#pragma once

enum { MODE_ASSUME = 0, MODE_SCORE = 1 };
// or-ed into the mode of a PLAN call: only candidates that turn out infeasible get their (partial) plan worked out - what the
// FailedNodes texts need (rater.go:99,152); everyone else costs what a call without plans costs
constexpr int MODE_FAILED_PLANS_ONLY = 4;

// How a verb kernel tells the host that its outputs are complete when they go straight into mapped host memory (no copy
// command, no stream synchronisation): after the CTA's barrier one thread makes the CTA's stores visible system-wide (a
// fence is cumulative over what the barrier ordered before it - the pattern of a cooperative-groups grid sync, at system
// scope; a fence in every thread cost 1.5 us more, profiles/r02/mapped_write_bench.txt) and draws a ticket; the CTA that
// draws the last one writes the call's sequence number into the host's flag word.  flag == nullptr: the caller
// synchronises the stream.
struct VerbDone {
    uint32_t* ticket;  // device counter, zero between calls
    int32_t* flag;     // mapped host word
    int32_t seq;
};
__device__ __forceinline__ void verb_done(const VerbDone& vd) {
    if (vd.flag == nullptr) return;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        if (atomicAdd(vd.ticket, 1u) == gridDim.x - 1) {
            *vd.ticket = 0;  // ready for the next call (ordered by the kernel boundary)
            __threadfence_system();
            *reinterpret_cast<volatile int32_t*>(vd.flag) = vd.seq;
        }
    }
}

// One status byte per row from every thread of a 256-thread CTA (rows base + threadIdx.x, base a multiple of 256), written
// as two 128-byte warp stores.  The verbs' outputs go straight into mapped host memory, where what counts is the bytes one
// warp instruction writes: 32 (a warp of byte stores) drain at a third of the rate of 128 (profiles/r02/mapped_write_bench.txt).
// Every thread of the CTA must call; `in` = the thread's row exists.
__device__ __forceinline__ void store_codes_cta(int8_t* __restrict__ out, uint32_t base, uint32_t n, bool in, int code, uint8_t* s_code) {
    s_code[threadIdx.x] = in ? (uint8_t)code : (uint8_t)0;
    __syncthreads();
    if (threadIdx.x < 64) {
        const uint32_t k = base + threadIdx.x * 4;
        if (k + 3 < n) {
            *reinterpret_cast<uint32_t*>(out + k) = *reinterpret_cast<const uint32_t*>(s_code + threadIdx.x * 4);
        } else {
            for (uint32_t j = 0; j < 4 && k + j < n; j++) out[k + j] = (int8_t)s_code[threadIdx.x * 4 + j];
        }
    }
    __syncthreads();  // (s_code is rewritten in the caller's next round)
}

// K1.  DealerImpl.Assume / DealerImpl.Score fan-out (dealer.go:109-134, 138-153): one thread
// per candidate.  MODE_ASSUME recomputes every candidate (cleanPlan + Assume, dealer.go:123-124);
// MODE_SCORE serves NodeInfo.Score (node.go:59-68) from the per-node cache when it holds a
// plan for this demand generation.
#ifndef NGS_K1_BLOCKS
#define NGS_K1_BLOCKS 4  // resident 256-thread blocks per SM for the feasibility/score scan
#endif
// The candidates of one call, 256 per CTA round (every thread of a 256-thread CTA must call; the status bytes are stored
// by the CTA as a whole).  `dem`, `s_pair` (load-aware calls), `wcx` (WIDE), `s_code` and `s_plan8` (PLAN: 256 x 12 bytes) live
// in the caller's shared memory.  PLAN outputs travel packed, one byte per entry (PLAN_NONE = the INT32_MIN of the C ABI): they
// go into mapped host memory, where bytes over PCIe are what the call costs; the host widens them.
constexpr int PLAN_NONE = -128;
template <bool PLAN, bool WIDE>
__device__ __forceinline__ void assume_rows(const Tables& T, const uint32_t* __restrict__ idx, uint32_t n, const SortedDemand* dem,
                                            const LoadParams& lp, int mode, uint32_t gen, uint8_t* __restrict__ feasible,
                                            int8_t* __restrict__ reason, int32_t* __restrict__ score, int8_t* __restrict__ plan,
                                            uint8_t* __restrict__ n_emitted, const double* s_pair, const WideCtx* wcx, uint8_t* s_code,
                                            int8_t* s_plan8 = nullptr) {
    const bool failed_only = PLAN && (mode & MODE_FAILED_PLANS_ONLY) != 0;
    mode &= 3;
    const int C = dem->C;
    // (CTA-uniform trip count)
    for (uint32_t k0 = blockIdx.x * blockDim.x; k0 < n; k0 += gridDim.x * blockDim.x) {
        const uint32_t k = k0 + threadIdx.x;
        const bool in = k < n;
        RowEval ev;
        ev.known = false;
        ev.feasible = false;
        ev.score = NGS_SCORE_MIN;
        ev.emit = 0;
        ev.n_emit = 0;
        bool cached = false;
        uint32_t row = 0;
        if (in) {
            const uint32_t gidx = idx ? idx[k] : T.first + k;
            row = gidx - T.first;
            if (gidx >= T.first && row < T.n) {
                if (mode == MODE_SCORE && T.tag[row] == gen && T.gcount[row] != 0) {
                    ev.known = ev.feasible = cached = true;
                    ev.score = T.cscore[row];
                } else if (failed_only) {
                    ev = eval_row_global<false, WIDE>(T, row, dem, lp, true, wcx, s_pair);
                    if (ev.known && !ev.feasible) ev = eval_row_global<PLAN, WIDE>(T, row, dem, lp, true, wcx, s_pair);  // (same verdict, plus the partial plan)
                } else {
                    ev = eval_row_global<PLAN, WIDE>(T, row, dem, lp, true, wcx, s_pair);
                }
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
        if (feasible && in) feasible[k] = ok ? 1 : 0;
        if (reason) store_codes_cta(reason, k0, n, in, !ev.known ? NGS_REASON_UNKNOWN_NODE : (ok ? NGS_REASON_OK : NGS_REASON_INFEASIBLE), s_code);
        if (score && in) score[k] = ok ? ev.score : NGS_SCORE_MIN;
        if (PLAN) {
            if (plan && C > 0) {
                // C bytes per row: the CTA's 256 rows are one contiguous range of the output, stored as 32-bit words
                int8_t* sp = s_plan8 + threadIdx.x * C;
                int32_t tmp[NGS_MAX_CONTAINERS];
                const bool have = in && ev.known && !(failed_only && ok);
                if (have) write_plan(tmp, dem, dem->policy != NGS_POLICY_RANDOM, ok, ev.emit, ev.n_emit);
                for (int j = 0; j < C; j++) sp[j] = (have && tmp[j] != INT32_MIN) ? (int8_t)tmp[j] : (int8_t)PLAN_NONE;
                __syncthreads();
                const size_t first = (size_t)k0 * C, end = ((size_t)n * C + 3) & ~(size_t)3;  // (the staging array is padded to whole words)
                for (uint32_t w = threadIdx.x; w < 64u * (uint32_t)C; w += blockDim.x)
                    if (first + 4 * (size_t)w < end)
                        *reinterpret_cast<uint32_t*>(plan + first + 4 * (size_t)w) = *reinterpret_cast<const uint32_t*>(s_plan8 + 4 * w);
                __syncthreads();
            }
            if (n_emitted) store_codes_cta(reinterpret_cast<int8_t*>(n_emitted), k0, n, in, (in && ev.known && !(failed_only && ok)) ? ev.n_emit : 0, s_code);
        }
    }
}

// WIDE: the snapshot holds nodes of 9..12 GPUs (ngs_wide.cuh); the host launches those instantiations only then.
using AssumePad = LaunchPad<56>;
static_assert(param_block_bytes<Tables, const uint32_t*, uint32_t, PodDemand, LoadParams, int, uint32_t, uint8_t*, int8_t*, int32_t*, int8_t*, uint8_t*,
                                VerbDone, AssumePad>() % 64 == 32, "see LaunchPad");
template <bool PLAN, bool WIDE>
__global__ void __launch_bounds__(256, PLAN ? 2 : NGS_K1_BLOCKS) k_assume(Tables T, const uint32_t* __restrict__ idx, uint32_t n, PodDemand pd, LoadParams lp,
                                                int mode, uint32_t gen, uint8_t* __restrict__ feasible,
                                                int8_t* __restrict__ reason, int32_t* __restrict__ score,
                                                int8_t* __restrict__ plan, uint8_t* __restrict__ n_emitted, VerbDone vd, AssumePad) {
    __shared__ SortedDemand dem;
    __shared__ double s_pair[121];
    __shared__ WideCtx s_wcx;
    __shared__ __align__(16) uint8_t s_code[256];
    __shared__ __align__(16) int8_t s_plan8[PLAN ? 256 * NGS_MAX_CONTAINERS : 16];
    if (threadIdx.x == 0) sort_demand(pd.percent, pd.n, pd.policy, &dem);
    if (WIDE) wide_ctx_fill(&s_wcx, T, lp, threadIdx.x == 32);
    if (lp.enabled) fill_usage_pair_table(s_pair, threadIdx.x, blockDim.x);
    __syncthreads();
    assume_rows<PLAN, WIDE>(T, idx, n, &dem, lp, mode, gen, feasible, reason, score, plan, n_emitted, s_pair, &s_wcx, s_code, s_plan8);
    verb_done(vd);
}

// ---------------------------------------------------------------------------------------------
// K1, streaming form: the same evaluation as k_assume for "all rows, no per-node plans, load
// scheduling off", with the row data staged through shared memory by the TMA engine (1-D bulk
// copies, cp.async.bulk + mbarrier transaction counts) in a ring of 256-row tiles, so that several
// tiles of every table are in flight per CTA while the threads evaluate the tile that has landed.
// (The load-aware pass does more arithmetic per row and measured faster with k_assume's direct
// 256-bit loads at full occupancy than with this staging at 2-3 CTAs per SM: DESIGN.md, K1.)
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_TILE = 256;
#ifndef NGS_SCAN_STAGES
#define NGS_SCAN_STAGES 2
#endif
#ifndef NGS_SCAN_BLOCKS
#define NGS_SCAN_BLOCKS 6  // resident CTAs per SM
#endif
constexpr uint32_t SCAN_FREE_BYTES = SCAN_TILE * 8 * 4;  // 8 KB
constexpr uint32_t SCAN_GC_BYTES = SCAN_TILE;            // 256 B
constexpr uint32_t SCAN_RL_BYTES = SCAN_TILE * 8;        // 2 KB
constexpr uint32_t SCAN_STAGE_BYTES = SCAN_FREE_BYTES + SCAN_GC_BYTES + SCAN_RL_BYTES;

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

using ScanPad = LaunchPad<24>;
static_assert(param_block_bytes<Tables, uint32_t, PodDemand, LoadParams, uint8_t*, int8_t*, int32_t*, VerbDone, ScanPad>() % 64 == 32, "see LaunchPad");
template <bool WIDE>
__global__ void __launch_bounds__(SCAN_TILE, NGS_SCAN_BLOCKS) k_scan_tma(Tables T, uint32_t n, PodDemand pd, LoadParams lp_in,
                                                                       uint8_t* __restrict__ feasible, int8_t* __restrict__ reason,
                                                                       int32_t* __restrict__ score, VerbDone vd, ScanPad) {
    extern __shared__ __align__(128) unsigned char tiles[];
    __shared__ SortedDemand dem;
    __shared__ __align__(8) uint64_t bars[NGS_SCAN_STAGES];
    __shared__ WideCtx s_wcx;
    __shared__ __align__(16) uint8_t s_code[SCAN_TILE];
    const uint32_t tid = threadIdx.x;
    const uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    const bool want_rl = !T.rl_zero;
    LoadParams lp = lp_in;
    lp.enabled = 0;  // a compile-time fact for this kernel: the LoadUsage code drops out of the row evaluation

    // one thread issues the copies of a tile: every table's 256-row slice is one contiguous range
    auto issue = [&](uint32_t tile, uint32_t stage) {
        unsigned char* base = tiles + (size_t)stage * SCAN_STAGE_BYTES;
        const size_t row = (size_t)tile * SCAN_TILE;
        mbar_expect_tx(&bars[stage], SCAN_FREE_BYTES + SCAN_GC_BYTES + (want_rl ? SCAN_RL_BYTES : 0u));
        tma_load_1d(base, T.free + row * 8, SCAN_FREE_BYTES, &bars[stage]);
        tma_load_1d(base + SCAN_FREE_BYTES, T.gcount + row, SCAN_GC_BYTES, &bars[stage]);
        if (want_rl) tma_load_1d(base + SCAN_FREE_BYTES + SCAN_GC_BYTES, T.rl + row * 8, SCAN_RL_BYTES, &bars[stage]);
    };

    // thread 0 gets the first tiles moving before anything else happens; the demand is sorted by another
    // warp meanwhile, so the prologue costs no memory time
    if (tid == 0) {
        for (int st = 0; st < NGS_SCAN_STAGES; st++) mbar_init(&bars[st], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (uint32_t i = 0; i < NGS_SCAN_STAGES; i++) {
            const uint32_t tile = blockIdx.x + i * gridDim.x;
            if (tile < ntiles) issue(tile, i);
        }
    }
    if (tid == 32) sort_demand(pd.percent, pd.n, pd.policy, &dem);
    if (WIDE) wide_ctx_fill(&s_wcx, T, lp, tid == 64);
    __syncthreads();

    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, it++) {
        const uint32_t stage = it % NGS_SCAN_STAGES, parity = (it / NGS_SCAN_STAGES) & 1u;
        mbar_wait(&bars[stage], parity);
        const unsigned char* base = tiles + (size_t)stage * SCAN_STAGE_BYTES;
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
        if (want_rl) in.r8 = *reinterpret_cast<const uint2*>(base + SCAN_FREE_BYTES + SCAN_GC_BYTES + (size_t)tid * 8);
        __syncthreads();  // the stage has been read by everyone: refill it with the tile NGS_SCAN_STAGES steps ahead
        if (tid == 0) {
            const uint32_t next = tile + NGS_SCAN_STAGES * gridDim.x;
            if (next < ntiles) issue(next, stage);
        }
        RowEval ev;
        ev.known = ev.feasible = false;
        ev.score = NGS_SCORE_MIN;
        if (row < n) ev = eval_row_values<false, WIDE>(T, row, in, &dem, lp, true, nullptr, &s_wcx);
        const bool ok = ev.known && ev.feasible;
        if (feasible && row < n) feasible[row] = ok ? 1 : 0;
        if (reason) store_codes_cta(reason, tile * SCAN_TILE, n, row < n, !ev.known ? NGS_REASON_UNKNOWN_NODE : (ok ? NGS_REASON_OK : NGS_REASON_INFEASIBLE), s_code);
        if (score && row < n) score[row] = ok ? ev.score : NGS_SCORE_MIN;
    }
    verb_done(vd);
}
