// ngs_k8_serve.cuh — K8, the resident form of K7: one pod's whole cycle (filter + prioritize + arg-max + bind) without
// a kernel launch per pod.
//
// A launch whose completion the host can see costs 8-9 us on this box even for an empty kernel
// (profiles/r02/param_size_bench.txt); K7's own work is about 7 us (profiles/r02/k7_phases.txt).  K8 is launched once
// (cooperatively: every CTA must be resident, they wait for each other's requests) and then serves requests the host
// drops into a 512-byte block of mapped pinned memory.  The block is 32 quads of {3 payload words, sequence number}:
//   host   : all payload words (the demand already sorted), store fence, then the request's sequence number into every quad
//   CTA 0  : warp 0 polls the block with one coalesced 512-byte read, a quad per lane; the request is complete when the 32
//            sequence words agree (a 64-byte line is read at one point in time, and the host wrote its payload before its
//            sequence words); each lane then stores its quad - one 16-byte store - into a copy of the block in device memory
//   all    : warp 0 of every other CTA polls that copy the same way (a 16-byte quad is stored and loaded as one access, so a
//            lane sees a quad entirely old or entirely new: the request arrives with the read that finds it, no second trip
//            for its body); the CTA then runs K7's body: scan of its rows, atomicMax of the best key, ticket; the CTA that
//            draws the last ticket binds the winner and writes (node, score, plan) and then the sequence number into mapped
//            host memory, where the host polls it
// The kernel leaves when asked to (a request with quit set) or when no request arrived for idle_cycles; the host restarts
// it on the next request.  Every wait is bounded.  Arg-max policies on unsharded handles; everything else goes through K7.
// The VERBS instantiation also takes ngs_assume / ngs_score (SERVE_OP_SCAN: k_assume's body, every CTA, ticket at the end)
// and ngs_bind (SERVE_OP_BIND: k_bind's body, one thread of CTA 0) as requests.
// Same results as the launched kernels and the oracle (tests/test_gpu_serve.py).  Included by ngs_dealer.cu after ngs_k7_one.cuh.
#pragma once

constexpr int SERVE_REQ_WORDS = 128;  // the request block: 32 quads {3 payload words, sequence number}
constexpr int SERVE_BODY_WORDS = 96;
constexpr uint32_t SERVE_QUIT = 0xFFFFFFFFu;

enum { SERVE_OP_POD = 0, SERVE_OP_SCAN = 1, SERVE_OP_BIND = 2, SERVE_OP_APPLY = 3, SERVE_OP_LOAD = 4 };
struct ServeBody {
    SortedDemand dem;  // sorted by the host (sort_demand): every CTA would otherwise spend 0.5 us on it per request
    LoadParams lp;
    uint32_t rl_zero;  // Tables::rl_zero for this request
    uint32_t max_c;
    uint32_t quit;
    uint32_t prev_row;  // the row the previous pod was bound to (0xFFFFFFFF: none, or not a pod): the one row that changed since
    uint32_t op;       // SERVE_OP_POD: one pod's whole cycle; _SCAN: ngs_assume / ngs_score; _BIND: ngs_bind
    // _SCAN (k_assume's arguments; the outputs are the handle's mapped staging arrays)
    uint32_t mode, gen, n;
    uint32_t idx_lo, idx_hi;        // device pointer to the candidate list, 0 = rows 0..n-1
    uint32_t reason_lo, reason_hi;  // status bytes (0 = not wanted)
    uint32_t score_lo, score_hi;
    uint32_t plan_lo, plan_hi;      // packed plans / emitted counts (0 = no plans: the cheaper instantiation of the scan)
    uint32_t nem_lo, nem_hi;
    // _BIND
    uint32_t node;
    // _APPLY (ngs_node_apply): node; dem.C containers with dem.raw[] = percent and dem.d[] = the GPU indexes; mode = sign (as int32)
    // _LOAD (ngs_load_update / ngs_load_add_node): node; mode = metric, gen = card (0xFFFFFFFF: fresh empty maps for the node),
    //       n = tenths code, idx_lo = update time in seconds
};
template <class P>
__device__ __forceinline__ P* serve_ptr(uint32_t lo, uint32_t hi) {
    return reinterpret_cast<P*>(((unsigned long long)hi << 32) | lo);
}
static_assert(sizeof(ServeBody) <= SERVE_BODY_WORDS * 4 && sizeof(ServeBody) % 4 == 0, "the request body has to fit the block");
constexpr int SERVE_BODY_USED = (int)(sizeof(ServeBody) / 4);

struct __align__(16) ServeDev {  // device memory: the request block as CTA 0 forwards it to the other CTAs
    uint32_t blk[SERVE_REQ_WORDS];
    unsigned long long t_seen;  // (profiling builds)
};
// body word i sits at block word 4 * (i / 3) + i % 3
__host__ __device__ inline int serve_block_word(int i) { return 4 * (i / 3) + i % 3; }

// The verbs' bodies.  cx: the request's tables and load parameters in shared memory.
// (the scan with plans is the rarer request and the hungrier code: out of line, so that it does not set the register
// allocation of everything else in the kernel)
template <bool WIDE>
__device__ __noinline__ void serve_scan_plans(const WideCtx* cx, const ServeBody* body, const double* s_pair, uint8_t* s_code, int8_t* s_plan8) {
    assume_rows<true, WIDE>(cx->T, serve_ptr<const uint32_t>(body->idx_lo, body->idx_hi), body->n, &body->dem, cx->lp, (int)body->mode, body->gen, nullptr,
                            serve_ptr<int8_t>(body->reason_lo, body->reason_hi), serve_ptr<int32_t>(body->score_lo, body->score_hi),
                            serve_ptr<int8_t>(body->plan_lo, body->plan_hi), serve_ptr<uint8_t>(body->nem_lo, body->nem_hi), s_pair, cx, s_code, s_plan8);
}
template <bool WIDE>
__device__ __forceinline__ void serve_scan(const WideCtx* cx, const ServeBody* body, const double* s_pair, uint8_t* s_code, int8_t* s_plan8) {
    if ((body->plan_lo | body->plan_hi) != 0)  // (CTA-uniform)
        serve_scan_plans<WIDE>(cx, body, s_pair, s_code, s_plan8);
    else
        assume_rows<false, WIDE>(cx->T, serve_ptr<const uint32_t>(body->idx_lo, body->idx_hi), body->n, &body->dem, cx->lp, (int)body->mode, body->gen,
                                 nullptr, serve_ptr<int8_t>(body->reason_lo, body->reason_hi), serve_ptr<int32_t>(body->score_lo, body->score_hi), nullptr,
                                 nullptr, s_pair, cx, s_code);
}
template <bool WIDE>
__device__ __forceinline__ void serve_bind(const WideCtx* cx, const ServeBody* body, BindResult* res) {
    bind_row<WIDE>(cx->T, body->node, body->dem.C, body->dem.raw, &body->dem, cx->lp, body->gen, cx, res);
}

// out (mapped host memory): [0] sequence number of the request answered, [1] status, [8] node, [9] score, [10 ..) gpu indexes
// (_SCAN answers through the staging arrays and writes [0], [1] only; _BIND: [1] status, [9] score, [10 ..) gpu indexes)
// VERBS: the instantiation that also serves ngs_assume / ngs_score / ngs_bind requests (ngs_serve_verbs).  Their bodies
// next to the pod cycle's want more registers than three CTAs per SM leave (500 bytes of spills and 2 us per pod when they
// were squeezed in), so that instantiation runs two CTAs per SM and the pod-only one stays as lean as it was.
template <bool WIDE, bool VERBS>
__global__ void __launch_bounds__(256, VERBS ? 2 : 3) k_serve(Tables T, OneRec* rec, ServeDev* dev, const uint32_t* req, int32_t* out, uint32_t* exited,
                                                 uint32_t last_seq, uint32_t launch_id, long long idle_cycles) {
    __shared__ ServeBody body;
    __shared__ double s_pair[121];
    __shared__ uint64_t s_wmax[8];
    __shared__ int s_last;
    __shared__ int32_t s_plan[NGS_MAX_CONTAINERS];
    __shared__ WideCtx s_wcx;
    __shared__ uint32_t s_seq;
    __shared__ __align__(16) uint8_t s_code[256];
    __shared__ BindResult s_bind;
    __shared__ __align__(16) int8_t s_plan8[VERBS ? 256 * NGS_MAX_CONTAINERS : 16];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    fill_usage_pair_table(s_pair, tid, blockDim.x);  // (constant: once per launch)
    uint32_t last = last_seq;
    const long long wait_limit = idle_cycles * 2 + SPIN_TIMEOUT_CYCLES;
    // Between two requests the tables change in one row only (the one the last pod was bound to; everything else that
    // writes them stops this kernel first), so every thread asks for its first row of the NEXT request before it starts
    // waiting for it, in the load mode of the request just served; the loads land while the request travels.
    RowIn in;
    int pref_mode = -1;  // what `in` was requested for: -1 nothing, else (load-aware ? 2 : 0) | rl_zero
    const uint32_t row_first = blockIdx.x * blockDim.x + tid;
#ifdef NGS_K8_TIMING
    unsigned long long tg[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define K8_STAMP(i) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tg[i]))
#else
#define K8_STAMP(i) do { } while (0)
#endif
    for (;;) {
        // ---- the next request: warp 0 of every CTA polls, CTA 0 the host's block, the others CTA 0's copy of it ----
        if (warp == 0) {
            const long long t0 = clock64();
            uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = last;
            const bool first = blockIdx.x == 0;
            for (uint32_t spins = 0;; spins++) {
                if (first)
                    asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "l"(req + 4 * lane) : "memory");
                else
                    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "l"(dev->blk + 4 * lane) : "memory");
                const uint32_t s0 = __shfl_sync(FULL, w3, 0);
                if (__all_sync(FULL, w3 == s0) && s0 != last) break;
                if ((spins & 15u) == 15u) {  // (lane 0's clock decides for the warp)
                    const bool over = __shfl_sync(FULL, clock64() - t0 > (first ? idle_cycles : wait_limit) ? 1 : 0, 0) != 0;
                    if (over) {  // CTA 0: nothing came for idle_cycles, everybody leaves; the others: CTA 0 is gone (never expected)
                        w3 = SERVE_QUIT;
                        break;
                    }
                }
            }
            if (first) {
#ifdef NGS_K8_TIMING
                if (lane == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dev->t_seen));
#endif
                *reinterpret_cast<uint4*>(dev->blk + 4 * lane) = make_uint4(w0, w1, w2, w3);  // one 16-byte store per quad
            }
            uint32_t* bw = reinterpret_cast<uint32_t*>(&body);
            if (3 * lane + 0 < (uint32_t)SERVE_BODY_USED) bw[3 * lane + 0] = w0;
            if (3 * lane + 1 < (uint32_t)SERVE_BODY_USED) bw[3 * lane + 1] = w1;
            if (3 * lane + 2 < (uint32_t)SERVE_BODY_USED) bw[3 * lane + 2] = w2;
            if (lane == 0) s_seq = w3;
        }
        __syncthreads();
        K8_STAMP(0);  // this CTA holds the request
        const uint32_t cur = s_seq;
        if (cur == SERVE_QUIT) break;
        last = cur;
        K8_STAMP(1);
        if (body.quit) break;
        Tables Tq = T;
        Tq.rl_zero = body.rl_zero;
        const LoadParams& lp = body.lp;
        const SortedDemand& dem = body.dem;
        if (VERBS && body.op != SERVE_OP_POD) {
            // ---- the verbs: k_assume's and k_bind's bodies (ngs_k1_scan.cuh, ngs_k3_rows.cuh) ----
            if (tid == 64) {
                s_wcx.T = Tq;
                s_wcx.lp = lp;
            }
            __syncthreads();
            if (body.op == SERVE_OP_SCAN) {
                serve_scan<WIDE>(&s_wcx, &body, s_pair, s_code, s_plan8);
                __syncthreads();
                if (tid == 0) {
                    __threadfence_system();  // the CTA's outputs (cumulative over the barrier), before the ticket
                    if (atomicAdd(&rec->arrived, 1u) == gridDim.x - 1) {
                        rec->arrived = 0;
                        out[1] = NGS_OK;
                        __threadfence_system();
                        *reinterpret_cast<volatile int32_t*>(out) = (int32_t)cur;
                    }
                }
            } else if (blockIdx.x == 0 && tid == 0) {  // the single-row requests: one thread
                if (body.op == SERVE_OP_BIND) {
                    serve_bind<WIDE>(&s_wcx, &body, &s_bind);
                    for (int j = 0; j < NGS_MAX_CONTAINERS; j++) out[10 + j] = s_bind.gpu_idx[j];
                    out[9] = s_bind.score;
                    out[1] = s_bind.status;
                } else if (body.op == SERVE_OP_APPLY) {  // k_apply's body
                    out[1] = apply_row(s_wcx.T, body.node, dem.C, dem.raw, dem.d, (int)body.mode);
                } else {  // SERVE_OP_LOAD: k_load_update's / k_load_add_node's bodies
                    const Tables& Tt = s_wcx.T;
                    const uint32_t row = body.node - Tt.first;
                    const int metric = (int)body.mode;
                    if (body.gen == 0xFFFFFFFFu) {
                        for (int c = 0; c < (Tt.tenths_hi ? WIDE_MAX : 8); c++) {
                            bool hi;
                            const size_t off = wide_usage_off(Tt, row, metric, c, &hi);
                            (hi ? Tt.tenths_hi : Tt.tenths)[off] = 255;
                            (hi ? Tt.ts_hi : Tt.ts)[off] = 0;
                        }
                    } else {
                        bool hi;
                        const size_t off = wide_usage_off(Tt, row, metric, (int)body.gen, &hi);
                        (hi ? Tt.tenths_hi : Tt.tenths)[off] = (uint8_t)body.n;
                        (hi ? Tt.ts_hi : Tt.ts)[off] = body.idx_lo;
                    }
                    out[1] = NGS_OK;
                }
                __threadfence_system();
                *reinterpret_cast<volatile int32_t*>(out) = (int32_t)cur;
            }
            pref_mode = -1;  // rows may have changed (bind), and the scan's loads are not the pod cycle's
            continue;
        }
        if (!VERBS && body.op != SERVE_OP_POD) {  // never posted to this instantiation: answer rather than hang the caller
            if (blockIdx.x == 0 && tid == 0) {
                out[1] = NGS_ERR_UNSUPPORTED;
                __threadfence_system();
                *reinterpret_cast<volatile int32_t*>(out) = (int32_t)cur;
            }
            continue;
        }

        // ---- K7's body (ngs_k7_one.cuh), one GPU, arg-max policies ----
        const bool load_path = lp.enabled != 0;
        uint32_t row = row_first;
        const int mode = (load_path ? 2 : 0) | (int)(Tq.rl_zero & 1u);
        if (row < Tq.n && (pref_mode != mode || row == body.prev_row)) load_row(Tq, row, load_path, in);
        if (WIDE && tid == 64) {
            s_wcx.T = Tq;
            s_wcx.lp = lp;
        }
        __syncthreads();
        uint64_t best = 0;
        for (; row < Tq.n; row += gridDim.x * blockDim.x) {
            const RowEval ev = eval_row_values<false, false>(Tq, row, in, &dem, lp, true, s_pair, nullptr);
            if (ev.known && ev.feasible) {
                const uint64_t k = make_key(ev.score, Tq.first + row);
                best = k > best ? k : best;
            }
            const uint32_t nxt = row + gridDim.x * blockDim.x;
            if (nxt < Tq.n) load_row(Tq, nxt, load_path, in);
        }
        if (WIDE) {
            for (row = blockIdx.x * blockDim.x + tid; row < Tq.n; row += gridDim.x * blockDim.x) {
                if (Tq.gcount[row] <= 8) continue;
                const uint4 w = eval_wide_row(&s_wcx, row, &dem, WIDE_RATE | (lp.enabled ? WIDE_LOAD : 0u));
                if (w.x & 1u) {
                    const uint64_t k = make_key((int)w.y, Tq.first + row);
                    best = k > best ? k : best;
                }
            }
        }
        K8_STAMP(2);  // rows evaluated
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
        K8_STAMP(3);  // ticket drawn
        if (s_last && warp == 0) {
            // ---- warp 0 of the last CTA to arrive: every CTA's key and RemainLoad rows are visible now ----
            __threadfence();
            unsigned long long key = 0;
            if (lane == 0) {
                key = atomicAdd(&rec->key, 0ull);
                rec->key = 0;  // ready for the next request: nobody touches the record before the host has seen this answer
                rec->arrived = 0;
            }
            key = __shfl_sync(FULL, key, 0);
            int32_t* gout = out + 10;
            if (lane < NGS_MAX_CONTAINERS) s_plan[lane] = INT32_MIN;
            __syncwarp();
            if (key == 0) {
                if (lane == 0) {
                    out[8] = -1;
                    out[9] = NGS_SCORE_MIN;
                }
            } else {
                const uint32_t wrow = key_node(key) - Tq.first;
                const int g = (int)(lane & 7u), gbase = (int)(lane & 24u);
                const int G = Tq.gcount[wrow];
                const int cell = __ldcg(Tq.free + (size_t)wrow * 8 + g);
                const int rlc = (Tq.rl_zero ? 0 : (int)__ldcg(Tq.rl + (size_t)wrow * 8 + g)) + 2;
                uint64_t emit = 0;
                int n_emit = 0, fr = 0;
                if (WIDE && G > 8) {
                    if (lane == 0) {
                        const uint4 r = eval_wide_row(&s_wcx, wrow, &dem, WIDE_APPLY);
                        emit = (uint64_t)r.z | ((uint64_t)r.w << 32);
                    }
                    emit = __shfl_sync(FULL, emit, 0);
                } else {
                    choose_group8<true>(g, gbase, cell, rlc, g < G, dem.policy == NGS_POLICY_SPREAD, &dem, &emit, &n_emit, &fr);
                }
                if (lane < 8) {
                    if (g < G && G <= 8) Tq.free[(size_t)wrow * 8 + lane] = fr;  // GPUs.Allocate (allocate.go:102-118)
                    for (int k = (int)lane; k < dem.C; k += 8) {                // rater.go:102-105
                        const int e = (int)((emit >> (4 * k)) & 0xF);
                        s_plan[dem.order[k]] = e == 0xF ? NGS_NOT_NEED_GPU : e;
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    Tq.tag[wrow] = 0;  // node.go:82
                    out[8] = (int32_t)(Tq.first + wrow);
                    out[9] = key_score(key);
                }
            }
            __syncwarp();
            if (lane < body.max_c) gout[lane] = lane < NGS_MAX_CONTAINERS ? s_plan[lane] : INT32_MIN;
            __syncwarp();
            if (lane == 0) {
#ifdef NGS_K8_TIMING
                K8_STAMP(4);
                const unsigned long long t_seen = *reinterpret_cast<const volatile unsigned long long*>(&dev->t_seen);
                for (int i = 0; i < 5; i++) out[32 + i] = (int32_t)(tg[i] - t_seen);  // ns since CTA 0 saw the request
#endif
                out[1] = NGS_OK;
                __threadfence_system();  // the answer (and the bound row, for whoever reads the table next) before the word that says so
                *reinterpret_cast<volatile int32_t*>(out) = (int32_t)cur;
            }
        }
        // (the other CTAs, and the other warps of that one, go straight back to waiting)
        if (row_first < Tq.n) load_row(Tq, row_first, load_path, in);
        pref_mode = mode;
    }
    if (blockIdx.x == 0 && tid == 0) {
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(exited) = launch_id;
    }
}
