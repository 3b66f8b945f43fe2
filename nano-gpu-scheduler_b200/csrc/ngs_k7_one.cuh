// ngs_k7_one.cuh — K7, one pod's whole cycle (filter + prioritize + arg-max + bind) in ONE ordinary
// launch, for the decision-latency half of the metric.  K6 is built for batches: a demand-prep
// launch, a memset and a cooperative launch that first moves every row into shared memory.  For a
// single pod none of that pays, so this kernel scans the tables where they lie (K1's row
// evaluation), folds the keys with one atomicMax per CTA, and the CTA that arrives last — it can
// tell from a ticket counter, nobody waits for anybody — binds the winner and leaves the exchange
// record zeroed for the next call.  The random policy goes the same way with counts instead of keys:
// every CTA owns a contiguous node range and leaves the number of feasible nodes in it; the last CTA
// to arrive finds the range holding the seeded k-th feasible node and evaluates that range once more.
// On a sharded handle (binpack / spread) the last CTA of every rank trades the rank's best key with the
// peers through the same peer-mapped slot arrays K6 uses; the rank that owns the winner binds it and
// drops the plan into every peer's mailbox, so all ranks return the full decision without a collective.
// Same results as K6 with P = 1 (tests/test_gpu_parity.py, tests/sharded_worker.py).
// Included by ngs_dealer.cu inside its anonymous namespace, after ngs_k6_batch.cuh.
#pragma once

struct OneRec {
    unsigned long long key;  // arg-max over the CTAs' best keys (0 = nothing feasible)
    unsigned int arrived;    // CTAs that have folded their key in
    unsigned int pad;
};

// node-range sharding: this call's parity half of every rank's exchange buffer (K6's layout: pod-major
// slots [BATCH_CHUNK][NGS_MAX_RANKS][2] - this kernel uses the first NGS_MAX_RANKS words as one word per rank - then
// XMAIL_WORDS words of mailbox for the owner's plan)
struct OneShard {
    uint32_t world, rank;
    unsigned long long* xs[NGS_MAX_RANKS];
    int32_t* err;  // the handle's device error word (bounded waits raise it)
};
constexpr size_t XMAIL_WORDS = 16;
constexpr size_t XMAIL_OFFSET = (size_t)4096 * NGS_MAX_RANKS * 2;  // == BATCH_CHUNK * NGS_MAX_RANKS * 2 (static_assert in ngs_dealer.cu)

constexpr int32_t ONE_PENDING = 0x7FFFFFFF;  // status word while the kernel has not answered yet
constexpr uint32_t ONE_MAX_CTAS = 1024;      // entries of the random policy's per-CTA count array

// exclusive prefix of v over the 256 threads of the block (in thread order) and the block total
__device__ __forceinline__ uint32_t block_scan_256(uint32_t v, uint32_t* s_warp /*[8]*/, uint32_t* total) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(FULL, inc, d);
        if (lane >= (uint32_t)d) inc += up;
    }
    __syncthreads();  // (s_warp may still be read from a previous scan)
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (uint32_t w = 0; w < 8; w++) {
        const uint32_t x = s_warp[w];
        if (w < warp) base += x;
        tot += x;
    }
    *total = tot;
    return base + inc - v;
}

// SampleRater on one row: feasibility only
template <bool WIDE>
__device__ __forceinline__ bool sample_row_feasible(const Tables& T, uint32_t row, const SortedDemand* dem, const WideCtx* wcx) {
    const int G = T.gcount[row];
    if (G == 0) return false;
    if (G > 8) return WIDE ? (eval_wide_row(wcx, row, dem, 0u).x & 1u) != 0 : false;  // ngs_wide.cuh
    uint32_t v[8];
    ld_row8(T.free + (size_t)row * 8, v);
    int f[8];
#pragma unroll
    for (int g = 0; g < 8; g++) f[g] = (int)v[g];
    return choose_sample<false>(f, G, dem, nullptr, nullptr);
}

// K7, random policy (see the file header).  cnt: [gridDim.x] feasible nodes per CTA range.
// On a sharded handle the last CTA of every rank trades the rank's number of feasible nodes with the peers (the slot arrays
// of K5); rank r's feasible nodes follow those of the ranks before it, so every rank knows which rank holds the seeded pick.
// That rank binds it and drops node and plan into every peer's mailbox; the others wait for the mailbox.
// WIDE: the snapshot holds nodes of 9..12 GPUs (ngs_wide.cuh); the host launches that instantiation only then, so the
// usual one carries none of that code (it cost the latency-critical path 0.65 us per pod when it was a run-time branch)
static_assert(param_block_bytes<Tables, PodDemand, uint64_t, OneRec*, uint32_t*, int32_t*, int, OneShard>() % 64 == 32, "see LaunchPad");
template <bool WIDE>
__global__ void __launch_bounds__(256, 4) k_schedule_one_random(Tables T, PodDemand pd, uint64_t seed_seq, OneRec* rec, uint32_t* cnt, int32_t* out,
                                                                int max_c, OneShard sh) {
    __shared__ SortedDemand dem;
    __shared__ uint32_t s_warp[8];
    __shared__ int s_last;
    __shared__ uint32_t s_pick_cta, s_pick_off;
    __shared__ unsigned long long s_total, s_before_rank;
    __shared__ int s_status;
    __shared__ WideCtx s_wcx;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) sort_demand(pd.percent, pd.n, pd.policy, &dem);
    if (WIDE && tid == 32) {
        LoadParams off;
        off.enabled = 0;  // SampleRater never looks at the load tables
        wide_ctx_fill(&s_wcx, T, off, true);
    }
    __syncthreads();
    // CTA b owns rows [b * per_cta, (b+1) * per_cta), thread t the rpt consecutive rows from b * per_cta + t * rpt
    const uint32_t rpt = (T.n + gridDim.x * 256u - 1) / (gridDim.x * 256u);
    const uint32_t per_cta = rpt * 256u;
    auto count_range = [&](uint32_t cta) {
        uint32_t c = 0;
        const uint32_t r0 = cta * per_cta + tid * rpt;
        for (uint32_t j = 0; j < rpt; j++)
            if (r0 + j < T.n && sample_row_feasible<WIDE>(T, r0 + j, &dem, &s_wcx)) c++;
        return c;
    };
    uint32_t total;
    const uint32_t mine = count_range(blockIdx.x);
    block_scan_256(mine, s_warp, &total);
    if (tid == 0) {
        cnt[blockIdx.x] = total;
        __threadfence();
        s_last = atomicAdd(&rec->arrived, 1u) == gridDim.x - 1 ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;

    // ---- the last CTA to arrive ----
    __threadfence();
    if (tid == 0) rec->arrived = 0;
    const uint32_t gpt = (gridDim.x + 255u) / 256u;  // count entries per thread
    uint32_t part = 0;
    for (uint32_t j = 0; j < gpt; j++) {
        const uint32_t b = tid * gpt + j;
        if (b < gridDim.x) part += __ldcg(cnt + b);
    }
    uint32_t feasible;
    const uint32_t before = block_scan_256(part, s_warp, &feasible);
    int32_t* gout = out + 10;
    for (int k = (int)tid; k < max_c; k += 256) gout[k] = INT32_MIN;
    if (tid == 0) {
        s_total = feasible;
        s_before_rank = 0;
        s_status = NGS_OK;
    }
    __syncthreads();
    if (sh.world > 1) {
        if (tid < 32) {  // warp 0: this rank's count to every rank (its own array included), all ranks' counts back
            bool dead = false;
            unsigned long long got = 0;
            if (tid < sh.world) {
                st_relaxed_sys_u64(sh.xs[tid] + sh.rank, (unsigned long long)feasible | KEY_VALID);
                got = wait_valid(sh.xs[sh.rank] + tid, sh.err, dead);
            }
            dead = __any_sync(FULL, dead);
            unsigned long long inc = got;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long t = __shfl_up_sync(FULL, inc, o);
                if (tid >= (uint32_t)o) inc += t;
            }
            const unsigned long long tot = __shfl_sync(FULL, inc, 31);
            const unsigned long long bef = __shfl_sync(FULL, inc - got, sh.rank);
            if (tid == 0) {
                s_total = dead ? 0ull : tot;
                s_before_rank = bef;
                if (dead) s_status = NGS_ERR_TIMEOUT;
            }
        }
        __syncthreads();
    }
    const unsigned long long gtotal = s_total, before_rank = s_before_rank;
    const unsigned long long gpick = gtotal ? splitmix64(seed_seq) % gtotal : 0ull;  // the gpick-th feasible node of the whole cluster
    const bool pick_here = gtotal != 0 && gpick >= before_rank && gpick < before_rank + feasible;
    if (gtotal == 0) {
        if (tid == 0) {
            out[8] = -1;
            out[9] = NGS_SCORE_MIN;
        }
    } else if (!pick_here) {
        // another rank holds the pick: node and plan arrive in the mailbox
        if (tid == 0) {
            bool dead = false;
            const unsigned long long* mail = sh.xs[sh.rank] + XMAIL_OFFSET;
            wait_valid(mail, sh.err, dead);
            if (dead) {
                s_status = NGS_ERR_TIMEOUT;
                out[8] = -1;
                out[9] = NGS_SCORE_MIN;
            } else {
                __threadfence_system();
                for (int k = 0; k < max_c && k < NGS_MAX_CONTAINERS; k++) {
                    const unsigned long long w = ld_relaxed_sys_u64(mail + 1 + k / 2);
                    gout[k] = (int32_t)(uint32_t)((k & 1) ? (w >> 32) : w);
                }
                out[8] = (int32_t)(uint32_t)ld_relaxed_sys_u64(mail + 1 + NGS_MAX_CONTAINERS / 2);
                out[9] = 100;
            }
        }
    } else {
        const uint32_t pick = (uint32_t)(gpick - before_rank);  // the pick-th feasible node of this rank, in index order
        if (pick >= before && pick < before + part) {
            uint32_t acc = before;
            for (uint32_t j = 0; j < gpt; j++) {
                const uint32_t b = tid * gpt + j;
                if (b >= gridDim.x) break;
                const uint32_t c = __ldcg(cnt + b);
                if (pick < acc + c) {
                    s_pick_cta = b;
                    s_pick_off = pick - acc;
                    break;
                }
                acc += c;
            }
        }
        __syncthreads();
        const uint32_t cta = s_pick_cta, off = s_pick_off;
        const uint32_t c2 = count_range(cta);
        uint32_t tot2;
        const uint32_t before2 = block_scan_256(c2, s_warp, &tot2);
        if (off >= before2 && off < before2 + c2) {  // exactly one thread: the node is among its rows
            uint32_t left = off - before2;
            const uint32_t r0 = cta * per_cta + tid * rpt;
            for (uint32_t j = 0; j < rpt; j++) {
                const uint32_t row = r0 + j;
                if (row >= T.n || !sample_row_feasible<WIDE>(T, row, &dem, &s_wcx)) continue;
                if (left-- != 0) continue;
                // NodeInfo.Bind (node.go:70-84) with SampleRater.Choose (rater.go:28-50)
                const int G = T.gcount[row];
                uint64_t emit = 0;
                int n_emit = 0;
                if (WIDE && G > 8) {  // ngs_wide.cuh
                    const uint4 r = eval_wide_row(&s_wcx, row, &dem, WIDE_APPLY);
                    emit = (uint64_t)r.z | ((uint64_t)r.w << 32);
                    n_emit = (int)(r.x >> 8);
                } else {
                    uint32_t v[8];
                    ld_row8(T.free + (size_t)row * 8, v);
                    int f[8];
#pragma unroll
                    for (int g = 0; g < 8; g++) f[g] = (int)v[g];
                    choose_sample<true>(f, G, &dem, &emit, &n_emit);  // places on f
                    int4* gp = reinterpret_cast<int4*>(T.free + (size_t)row * 8);
                    gp[0] = make_int4(f[0], f[1], f[2], f[3]);
                    gp[1] = make_int4(f[4], f[5], f[6], f[7]);
                }
                T.tag[row] = 0;
                for (int k = 0; k < dem.C; k++) {
                    const int e = (int)((emit >> (4 * k)) & 0xF);
                    gout[k] = e == 0xF ? NGS_NOT_NEED_GPU : e;
                }
                out[8] = (int32_t)(T.first + row);
                out[9] = 100;  // SampleRater.Rate (rater.go:24-26)
                if (sh.world > 1) {  // node and plan into every peer's mailbox, then (after a fence) the word that says so
                    int32_t planv[NGS_MAX_CONTAINERS];
                    for (int k = 0; k < NGS_MAX_CONTAINERS; k++) {
                        const int e = (int)((emit >> (4 * k)) & 0xF);
                        planv[k] = k < dem.C ? (e == 0xF ? NGS_NOT_NEED_GPU : e) : INT32_MIN;
                    }
                    for (uint32_t r = 0; r < sh.world; r++) {
                        if (r == sh.rank) continue;
                        unsigned long long* mail = sh.xs[r] + XMAIL_OFFSET;
                        for (int w = 0; w < NGS_MAX_CONTAINERS / 2; w++)
                            st_relaxed_sys_u64(mail + 1 + w, ((unsigned long long)(uint32_t)planv[2 * w + 1] << 32) | (uint32_t)planv[2 * w]);
                        st_relaxed_sys_u64(mail + 1 + NGS_MAX_CONTAINERS / 2, (unsigned long long)(T.first + row));
                    }
                    __threadfence_system();
                    for (uint32_t r = 0; r < sh.world; r++)
                        if (r != sh.rank) st_relaxed_sys_u64(sh.xs[r] + XMAIL_OFFSET, KEY_VALID | 1ull);
                }
                break;
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        *reinterpret_cast<volatile int32_t*>(out) = s_status;
    }
}

// out (mapped host memory): [0] status, [8] node, [9] score, [10 .. 10+max_c) gpu indexes
using OnePad = LaunchPad<32>;
static_assert(param_block_bytes<Tables, PodDemand, LoadParams, OneRec*, int32_t*, int, OneShard, OnePad>() % 64 == 32, "see LaunchPad");
template <bool WIDE>
__global__ void __launch_bounds__(256, 4) k_schedule_one(Tables T, PodDemand pd, LoadParams lp, OneRec* rec, int32_t* out, int max_c, OneShard sh,
                                                         OnePad) {
    __shared__ SortedDemand dem;
    __shared__ double s_pair[121];
    __shared__ uint64_t s_wmax[8];
    __shared__ int s_last;
    __shared__ int32_t s_plan[NGS_MAX_CONTAINERS];  // the winner's plan in container order (last CTA only)
    __shared__ WideCtx s_wcx;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
#ifdef NGS_K7_TIMING
    long long tk[8];
    tk[0] = clock64();
#define K7_STAMP(i) tk[i] = clock64()
#else
#define K7_STAMP(i) do { } while (0)
#endif
    // the first row of every thread is requested before the demand is sorted: the loads do not depend on it
    const bool load_path = lp.enabled != 0;  // (binpack / spread only)
    uint32_t row = blockIdx.x * blockDim.x + tid;
    RowIn in;
    if (row < T.n) load_row(T, row, load_path, in);
    if (tid == 32) sort_demand(pd.percent, pd.n, pd.policy, &dem);
    if (WIDE) wide_ctx_fill(&s_wcx, T, lp, tid == 64);
    if (lp.enabled) fill_usage_pair_table(s_pair, tid, blockDim.x);
    __syncthreads();
    K7_STAMP(1);
    uint64_t best = 0;
    for (; row < T.n; row += gridDim.x * blockDim.x) {
        // Rate + Choose on the row as it stands; in load-aware mode this also leaves RemainLoad behind (allocate.go:193)
        const RowEval ev = eval_row_values<false, false>(T, row, in, &dem, lp, true, s_pair, nullptr);
        if (ev.known && ev.feasible) {
            const uint64_t k = make_key(ev.score, T.first + row);
            best = k > best ? k : best;
        }
        const uint32_t nxt = row + gridDim.x * blockDim.x;
        if (nxt < T.n) load_row(T, nxt, load_path, in);
    }
    if (WIDE) {  // nodes of 9..12 GPUs (ngs_wide.cuh): a pass of their own over the GPU counts
        for (row = blockIdx.x * blockDim.x + tid; row < T.n; row += gridDim.x * blockDim.x) {
            if (T.gcount[row] <= 8) continue;
            const uint4 w = eval_wide_row(&s_wcx, row, &dem, WIDE_RATE | (lp.enabled ? WIDE_LOAD : 0u));
            if (w.x & 1u) {
                const uint64_t k = make_key((int)w.y, T.first + row);
                best = k > best ? k : best;
            }
        }
    }
    K7_STAMP(2);
    best = warp_max_u64(best);
    if (lane == 0) s_wmax[warp] = best;
    __syncthreads();
    if (tid == 0) {
        uint64_t m = 0;
        for (uint32_t w = 0; w < blockDim.x / 32; w++) m = s_wmax[w] > m ? s_wmax[w] : m;
        if (m) atomicMax(&rec->key, (unsigned long long)m);
        __threadfence();  // this CTA's RemainLoad stores and its key, before the ticket
        s_last = atomicAdd(&rec->arrived, 1u) == gridDim.x - 1 ? 1 : 0;
    }
    __syncthreads();
    if (!s_last || warp != 0) return;
    K7_STAMP(3);

    // ---- warp 0 of the last CTA to arrive: every other CTA's key and RemainLoad rows are visible now ----
    __threadfence();
    unsigned long long key = 0;
    if (lane == 0) {
        key = atomicAdd(&rec->key, 0ull);
        rec->key = 0;  // ready for the next call (ordered by the kernel boundary)
        rec->arrived = 0;
    }
    key = __shfl_sync(FULL, key, 0);
    int32_t* gout = out + 10;
    if (lane < NGS_MAX_CONTAINERS) s_plan[lane] = INT32_MIN;
    __syncwarp();
    int32_t status = NGS_OK;
    bool mine = key != 0;  // the winner is one of this rank's rows
    if (sh.world > 1) {
        // K5 for one pod: this rank's best key to every rank (its own array included), all ranks' keys back
        bool dead = false;
        unsigned long long got = 0;
        if (lane < sh.world) {
            st_relaxed_sys_u64(sh.xs[lane] + sh.rank, key | KEY_VALID);
            got = wait_valid(sh.xs[sh.rank] + lane, sh.err, dead);
        }
        dead = __any_sync(FULL, dead);
        const unsigned long long gmax = warp_max_u64(got);
        if (dead) status = NGS_ERR_TIMEOUT;
        mine = !dead && gmax != 0 && gmax == key;
        key = dead ? 0ull : gmax;
    }
    if (key == 0) {  // no node takes the pod (or the exchange gave up)
        if (lane == 0) {
            out[8] = -1;
            out[9] = NGS_SCORE_MIN;
        }
    } else if (mine) {
        // NodeInfo.Bind (node.go:70-84) on the winner, eight lanes on the row's eight cells (the four
        // groups of the warp hold identical copies): the plan from the RemainLoad the scan left behind
        const uint32_t wrow = key_node(key) - T.first;
        const int g = (int)(lane & 7u), gbase = (int)(lane & 24u);
        const int G = T.gcount[wrow];
        const int cell = __ldcg(T.free + (size_t)wrow * 8 + g);
        const int rlc = (int)__ldcg(T.rl + (size_t)wrow * 8 + g) + 2;
        uint64_t emit = 0;
        int n_emit = 0, fr = 0;
        if (WIDE && G > 8) {  // ngs_wide.cuh: lane 0 plans and places, the plan goes to the other lanes
            if (lane == 0) {  // (no WIDE_LOAD: the plan is derived from the RemainLoad the scan left behind, as below)
                const uint4 r = eval_wide_row(&s_wcx, wrow, &dem, WIDE_APPLY);
                emit = (uint64_t)r.z | ((uint64_t)r.w << 32);
            }
            emit = __shfl_sync(FULL, emit, 0);
        } else {
            choose_group8<true>(g, gbase, cell, rlc, g < G, dem.policy == NGS_POLICY_SPREAD, &dem, &emit, &n_emit, &fr);
        }
        if (lane < 8) {
            if (g < G && G <= 8) T.free[(size_t)wrow * 8 + lane] = fr;  // GPUs.Allocate (allocate.go:102-118)
            for (int k = (int)lane; k < dem.C; k += 8) {      // rater.go:102-105: un-permute the emission order
                const int e = (int)((emit >> (4 * k)) & 0xF);
                s_plan[dem.order[k]] = e == 0xF ? NGS_NOT_NEED_GPU : e;
            }
        }
        __syncwarp();
        if (lane == 0) {
            T.tag[wrow] = 0;  // node.go:82
            out[8] = (int32_t)(T.first + wrow);
            out[9] = key_score(key);
        }
        if (sh.world > 1) {
            // the plan goes into every peer's mailbox: 6 words of indexes, then (after a fence) the word that says so
            if (lane < sh.world && lane != sh.rank) {
                unsigned long long* mail = sh.xs[lane] + XMAIL_OFFSET;
                for (int w = 0; w < NGS_MAX_CONTAINERS / 2; w++)
                    st_relaxed_sys_u64(mail + 1 + w, ((unsigned long long)(uint32_t)s_plan[2 * w + 1] << 32) | (uint32_t)s_plan[2 * w]);
                __threadfence_system();
                st_relaxed_sys_u64(mail, KEY_VALID | 1ull);
            }
        }
    } else {
        // another rank owns the winner: node and score are in the key, the plan arrives in the mailbox
        bool dead = false;
        const unsigned long long* mail = sh.xs[sh.rank] + XMAIL_OFFSET;
        if (lane == 0) wait_valid(mail, sh.err, dead);
        dead = __any_sync(FULL, dead);
        if (dead) {
            status = NGS_ERR_TIMEOUT;
            if (lane == 0) {
                out[8] = -1;
                out[9] = NGS_SCORE_MIN;
            }
        } else {
            __threadfence_system();
            if (lane < NGS_MAX_CONTAINERS) {
                const unsigned long long w = ld_relaxed_sys_u64(mail + 1 + lane / 2);
                s_plan[lane] = (int32_t)(uint32_t)((lane & 1u) ? (w >> 32) : w);
            }
            if (lane == 0) {
                out[8] = (int32_t)key_node(key);
                out[9] = key_score(key);
            }
        }
    }
    __syncwarp();
    K7_STAMP(4);
    if ((int)lane < max_c) gout[lane] = lane < NGS_MAX_CONTAINERS ? s_plan[lane] : INT32_MIN;
    __syncwarp();
    if (lane == 0) {
#ifdef NGS_K7_TIMING
        for (int i = 0; i < 5; i++) out[32 + i] = (int32_t)(tk[i] - tk[0]);
#endif
        __threadfence_system();  // the answer is in host memory before the status word says so
        *reinterpret_cast<volatile int32_t*>(out) = status;
    }
}
