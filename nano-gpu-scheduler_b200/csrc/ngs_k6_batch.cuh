// ngs_k6_batch.cuh — K6 (and K5 inside it): P pods in order in one cooperative launch.
// Included by ngs_dealer.cu inside its anonymous namespace, after ngs_rows.cuh.
#pragma once

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
        // A pod the kernels cannot take is not allowed to stop the batch: it is scheduled as a pod that fits
        // nowhere (one container larger than any cell can be) and reported as NGS_NODE_INVALID_POD in node_out;
        // the call's status word remembers that there was one.  The waits below never look at that word.
        atomicMax(err, NGS_ERR_INVALID);
        for (int k = 0; k < NGS_MAX_CONTAINERS; k++) pc[k] = 0;
        pc[0] = INT32_MAX;
        C = 1;
        pol = NGS_POLICY_BINPACK;
    }
    sort_demand(pc, C, pol, &out[p]);
    if (bad) out[p].flags |= 1;
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
    unsigned long long* rr;    // [P] arg-max exchange records: the first 16 bytes of pod p's record are {atomicMax key, arrival
                               // counter}, zero before launch
    // node-range sharding over the GPUs of one box: every rank runs this same kernel on
    // its own rows; CTA 0 of each rank trades the rank's result with the peers through
    // peer-mapped memory (plain 8-byte stores over NVLink); every CTA reads the ranks' words from there
    uint32_t world, rank;
    unsigned long long* xslots[NGS_MAX_RANKS];  // [rank r] -> r's slot array [P][NGS_MAX_RANKS][2] (this batch's parity half; word 0 is used)
    uint32_t keys_stride, done_stride, rr_stride;  // distance between consecutive pods, in elements of each array
    int32_t* status_out;       // optional host-visible copy of the error word, written when the kernel ends
    int32_t* err;              // device status words: [0] sticky status of the batch (0 = ok), [1] abort flag (set by a wait that timed out)
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

#ifndef NGS_K6_EXCLUSIVE_ROWS
#define NGS_K6_EXCLUSIVE_ROWS 2048
#endif
constexpr uint32_t K6_EXCLUSIVE_ROWS = NGS_K6_EXCLUSIVE_ROWS;
constexpr uint32_t K6_ROW_PER_THREAD_ITERS = 2;  // up to this many rows per worker thread the simple filter loop is the faster one  // up to this many rows per CTA control warp 0 keeps a scheduler to itself
constexpr long long SPIN_TIMEOUT_CYCLES = 2000000000LL;   // ~1 s: waits on CTAs of the same (cooperative, fully resident) launch
constexpr long long PEER_TIMEOUT_CYCLES = 16000000000LL;  // ~8 s: waits on another GPU, whose host may launch its kernel later

// The status word err[0] is sticky (invalid pod, time-out) and is what the host reads; err[1] is the abort flag the
// bounded waits look at: only a wait that has timed out raises it, so a bad pod cannot make other waits give up.
__device__ __forceinline__ bool abort_raised(const int32_t* err) { return *reinterpret_cast<const volatile int32_t*>(err + 1) != 0; }
__device__ __forceinline__ void raise_timeout(int32_t* err) {
    atomicExch(err + 1, 1);
    atomicMax(err, NGS_ERR_TIMEOUT);  // (outranks NGS_ERR_INVALID)
}

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
                                           const SortedDemand* dem_next, int g, int gbase, GroupCommit& cm, long long* stamps = nullptr,
                                           const int16_t* lut = nullptr) {
    (void)bp;  // lut: Binpack.Rate's percent table wherever the caller keeps it (shared memory, global; plain loads), nullptr = fp64
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
    const int pct = (lut != nullptr && used >= 0 && used <= 100 * G) ? (int)lut[G * RATE_LUT_STRIDE + used] : binpack_percent_fp64(used, G);
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
    // less capacity left on the row than the pod asks for in total: no placement exists, so the greedy one fails as well
    if (cls == FEAS_UNDECIDED && sumf < (int)((uint32_t)dem_next->flags >> 1)) cls = FEAS_NO;
    if (cls == FEAS_UNDECIDED) {  // (warp-uniform) the position-free search first; the positional one only on an ambiguous tie
        int frx = fr;
        const int st = greedy_group8<false>(g, gbase, frx, rlc, valid, pol2 == NGS_POLICY_SPREAD, dem_next);
        if (st & 2)
            cls = choose_group8<false>(g, gbase, fr, rlc, valid, pol2 == NGS_POLICY_SPREAD, dem_next, nullptr, nullptr, nullptr) ? FEAS_YES : FEAS_NO;
        else
            cls = (st & 1) ? FEAS_YES : FEAS_NO;
    }
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

__device__ __forceinline__ void publish_none(const BatchParams& bp, uint32_t p, const SortedDemand* dem) {
    bp.node_out[p] = (dem->flags & 1) ? NGS_NODE_INVALID_POD : -1;  // (the pod failed validation, k_prep_demands)
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
            if (abort_raised(err) || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                raise_timeout(err);
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
            if (abort_raised(err) || clock64() - t0 > PEER_TIMEOUT_CYCLES) {
                raise_timeout(err);
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
        st_relaxed_sys_u64(bp.xslots[lane] + ((size_t)p * NGS_MAX_RANKS + bp.rank) * 2, mine | KEY_VALID);
        got = wait_valid(bp.xslots[bp.rank] + ((size_t)p * NGS_MAX_RANKS + lane) * 2, bp.err, dead);
    }
    dead = __any_sync(0xFFFFFFFFu, dead);
    return got;
}

// every other CTA: the ranks' words straight from this rank's slot array (CTA 0 posts our own word there,
// too), so the result needs no relay through a second location.  Same return convention.
__device__ __forceinline__ unsigned long long collect_ranks(const BatchParams& bp, uint32_t p, uint32_t lane, bool& dead) {
    unsigned long long got = 0;
    if (lane < bp.world) got = wait_valid(bp.xslots[bp.rank] + ((size_t)p * NGS_MAX_RANKS + lane) * 2, bp.err, dead);
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
            if (abort_raised(err) || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                raise_timeout(err);
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
// WIDE: the snapshot holds nodes of 9..12 GPUs (ngs_wide.cuh).  Their shared-memory rows are empty (never feasible for
// the code below); widebits marks them, and the lane that owns such a row evaluates it from the global tables instead —
// in its turn, so that a lane still meets its rows in ascending order.  row0: first row of this CTA.
template <bool WIDE>
__device__ __forceinline__ void filter_pass(const SmemRows& S, const SortedDemand* dem, uint32_t wt, uint32_t WT, uint32_t n_iter,
                                            uint32_t my_rows, uint32_t node0, uint32_t lane, int g, int gbase, uint32_t* fb,
                                            uint64_t& k1, uint64_t& k2, uint32_t& cnt, const WideCtx* wcx, const uint32_t* widebits,
                                            uint32_t row0) {
    const int policy = dem->policy;
    const bool spread = policy == NGS_POLICY_SPREAD;
    if (policy == NGS_POLICY_RANDOM) {
        uint32_t c = 0;
        for (uint32_t it = 0; it < n_iter; it++) {
            const uint32_t r = it * WT + wt;
            bool ok = r < my_rows && eval_row_sample(S, r, dem);
            if (WIDE && r < my_rows && ((widebits[r >> 5] >> (r & 31)) & 1u)) ok = (eval_wide_row(wcx, row0 + r, dem, 0u).x & 1u) != 0;
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, ok);
            if (lane == 0 && r < my_rows) {
                fb[r >> 5] = bal;
                c += __popc(bal);
            }
        }
        cnt = c;
        return;
    }
    // Everything the loop needs of the demand sits in registers (the shared-memory stores of the random
    // branch would otherwise make the compiler reload it per row), and a lane tracks its best two rows by
    // score alone: it visits its rows in ascending order, so among equal scores the earlier row is the one
    // the key order (higher score, then lower node index) prefers, and strict comparisons keep it.
    const int nz = dem->nz, d1 = dem->d[0], d2 = dem->d[1];
    const bool generic = nz == 0 || nz > 4 || d1 > 255;  // the top-4 word cannot decide (classify_top4's FEAS_FALLBACK cases)
    const bool may_stay_open = nz >= 3;                   // one or two containers are always settled by the shortcuts
    const uint32_t sh = nz >= 1 ? 8u * (uint32_t)(nz - 1) : 0u;
    const int32_t* __restrict__ rate = spread ? S.scS : S.scB;
    constexpr int NONE = INT32_MIN;
    int s1 = NONE, s2 = NONE;
    uint32_t r1 = 0, r2 = 0;
    const bool may_need_row = generic || nz == 2;  // (warp-uniform) only then can a lane ask for the full row
    const uint32_t* top4p = S.top4 + wt;           // this lane's entries, WT apart: pointers advance, no index arithmetic per row
    const int32_t* ratep = rate + wt;
    uint32_t r = wt;
    for (uint32_t it = 0; it < n_iter; it++, r += WT, top4p += WT, ratep += WT) {
        const bool in = r < my_rows;
        int cls = FEAS_NO;
        bool full = in && generic;
        if (!generic) {
            // classify_top4 with the demand in registers, written without branches: this is the path nearly every
            // row takes, and nested divergent branches cost more than the arithmetic they skip
            const uint32_t t4 = in ? *top4p : 0u;  // rows beyond the range and unknown nodes read as all zeros: infeasible
            const int t0 = (int)(t4 & 0xFFu);
            const bool yes = (int)((t4 >> sh) & 0xFFu) >= d1;  // the nz-th largest cell holds d1
            const bool no = t0 < d1;                            // not even the largest does
            cls = yes ? FEAS_YES : (no ? FEAS_NO : FEAS_UNDECIDED);
            if (nz == 2) {  // (warp-uniform) exactly one cell holds d1: it is the row maximum
                const bool mid = !yes && !no;
                full = mid && t0 == 255;  // saturated: the true maximum is unknown, look at the whole row
                const bool fits = (int)((t4 >> 8) & 0xFFu) >= d2 || t0 - d1 >= d2;
                cls = mid ? (fits ? FEAS_YES : FEAS_NO) : cls;
            }
        }
        if (may_need_row && __any_sync(0xFFFFFFFFu, full)) {  // rare (or the generic case): classify from the full row
            if (full) {
                cls = FEAS_NO;
                const uint32_t meta = S.meta[r];
                if (((meta >> 24) & 15) != 0) {
                    int f[8];
                    unpack_free(S.fA[r], S.fB[r], f);
                    cls = classify_row(f, dem, true);
                }
            }
        }
        if (may_stay_open) {  // (warp-uniform)
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
                // the position-free search first (a third of the latency); the positional one only where it met an ambiguous tie
                int frx = smem_cell(S, rr, g);
                const int st = greedy_group8<false>(g, gbase, frx, (int)((meta >> (3 * g)) & 7), g < G, spread, dem);
                bool ok = (st & 1) != 0;
                if (__any_sync(0xFFFFFFFFu, has && (st & 2))) {  // (warp-uniform)
                    const bool ok_exact = choose_group8<false>(g, gbase, smem_cell(S, rr, g), (int)((meta >> (3 * g)) & 7), g < G, spread, dem,
                                                               nullptr, nullptr, nullptr);
                    if (st & 2) ok = ok_exact;
                }
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
        }
        // best two by score, branch-free: an infeasible row enters as NONE and changes nothing
        int sc = cls == FEAS_YES ? *ratep : NONE;
        if (WIDE && in && ((widebits[r >> 5] >> (r & 31)) & 1u)) {  // Rate (load-aware if the batch is) and Choose, one thread
            const uint4 w = eval_wide_row(wcx, row0 + r, dem, WIDE_RATE | WIDE_LOAD);
            sc = (w.x & 1u) ? (int)w.y : NONE;
        }
        const bool gt1 = sc > s1, gt2 = sc > s2;
        s2 = gt1 ? s1 : (gt2 ? sc : s2);
        r2 = gt1 ? r1 : (gt2 ? r : r2);
        s1 = gt1 ? sc : s1;
        r1 = gt1 ? r : r1;
    }
    k1 = s1 != NONE ? make_key(s1, node0 + r1) : 0;
    k2 = s2 != NONE ? make_key(s2, node0 + r2) : 0;
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

// A pod bound to a node of 9..12 GPUs (one lane): the plan on the row as it stands in the global tables (RemainLoad as
// the filter left it), GPUs.Allocate, the pod's outputs.
__device__ __forceinline__ void publish_wide(const BatchParams& bp, const WideCtx* wcx, uint32_t row, const SortedDemand* dem, uint32_t p,
                                             int score) {
    const uint4 w = eval_wide_row(wcx, row, dem, WIDE_APPLY);
    int32_t* out = bp.gpu_idx_out + (size_t)p * bp.max_c;
    for (int k = dem->C; k < bp.max_c; k++) out[k] = INT32_MIN;
    write_plan(out, dem, dem->policy != NGS_POLICY_RANDOM, true, (uint64_t)w.z | ((uint64_t)w.w << 32), (int)(w.x >> 8));
    bp.T.tag[row] = 0;  // node.go:82
    bp.node_out[p] = (int32_t)(bp.T.first + row);
    bp.score_out[p] = score;
}
// ... and that row's verdict for the next pod, on the row as the bind left it
__device__ __forceinline__ uint64_t fresh_key_of_wide_row(const WideCtx* wcx, uint32_t row, uint32_t node, const SortedDemand* dem_next,
                                                          int* fresh_ok) {
    const uint4 w = eval_wide_row(wcx, row, dem_next, WIDE_RATE | WIDE_LOAD);
    *fresh_ok = (int)(w.x & 1u);
    return ((w.x & 1u) && dem_next->policy != NGS_POLICY_RANDOM) ? make_key((int)w.y, node) : 0ull;
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
// ALL_WORK (chosen by the host from the rows per CTA, schedule_kernel()): every warp below the control warps
// filters; otherwise the warps that share control warp 0's scheduler stay idle.  A compile-time choice: as a
// run-time flag it cost the small-CTA case 1.8 %.
// WIDE: the snapshot holds nodes of 9..12 GPUs; see filter_pass.  The host picks it only then (no cost otherwise).
template <bool ALL_WORK, bool WIDE>
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
    // Binpack.Rate's percent table next to the rows when they leave room (an L2 trip on the plan's path otherwise)
    __shared__ int16_t s_blut[ALL_WORK ? 2 : 9 * RATE_LUT_STRIDE];
    __shared__ WideCtx s_wcx;  // (referenced by the WIDE instantiations only)
    const WideCtx* wcx = &s_wcx;
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
    uint32_t* widebits = S.feasbits + 2 * fwords;  // WIDE: one bit per row, set for nodes of more than 8 GPUs (the allocation has room for four buffers)

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    // the two control warps are the LAST two: the SM's issue arbiter favours high warp ids, and the
    // control path (exchange, plan) is the latency-critical one
    // Warps map to the SM's four schedulers by warp id % 4.  Control warp 0 runs the latency-critical plan
    // search; the worker warps that would share its scheduler stay idle (they own no rows), so that its
    // instructions never queue behind the filter's.
    // When a CTA holds so many rows that the filter, not the exchange, sets the pace (several rows per worker
    // thread), that scheduler is worth more filtering: then every warp below the control warps owns rows.
    const uint32_t NWARP = nwarps - 2;                     // warps below the control warps
    constexpr bool all_work = ALL_WORK;
    const uint32_t csm = NWARP & 3u;                       // scheduler of control warp 0
    const uint32_t idle_below = all_work ? 0u : (warp + 3u - csm) >> 2;    // idle warps among [0, warp)
    const uint32_t NW = all_work ? NWARP : NWARP - ((NWARP + 3u - csm) >> 2); // worker warps that own rows
    const uint32_t WT = NW * 32u;
    const bool is_worker = warp < NWARP && (all_work || (warp & 3u) != csm);
    const bool is_control = warp >= NWARP;
    const uint32_t ww = warp - idle_below;                 // worker index (meaningful for workers)
    const uint32_t wt = ww * 32u + lane;
    const uint32_t cw = warp - NWARP;                      // control warp 0 (plan, bind) / 1 (exchange)
    const uint32_t row0 = blockIdx.x * rpb;
    const uint32_t my_rows = row0 < bp.T.n ? min(rpb, bp.T.n - row0) : 0u;
    const Tables& T = bp.T;
    const int g = lane & 7, gbase = lane & 24;

    // ---- prologue: rows -> shared memory (and the load-aware quantities, frozen for the batch) ----
    if (WIDE) {
        if (tid == 0) {
            s_wcx.T = bp.T;
            s_wcx.lp = bp.lp;
        }
        for (uint32_t i = tid; i < fwords; i += blockDim.x) widebits[i] = 0;
        __syncthreads();
    }
    for (uint32_t r = tid; r < my_rows; r += blockDim.x) {
        const uint32_t row = row0 + r;
        int G = T.gcount[row];
        int f[8];
        const int4 a = ld_row4(T.free + (size_t)row * 8), b = ld_row4(T.free + (size_t)row * 8 + 4);
        f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
        f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
        if (G > 8) {  // 9..12 GPUs (ngs_wide.cuh): the shared-memory row stays empty
            if (WIDE) atomicOr(&widebits[r >> 5], 1u << (r & 31));
            else atomicMax(bp.err, NGS_ERR_UNSUPPORTED);  // a caller-owned view (no second plane): the node is left out, the status says so
            G = 0;
#pragma unroll
            for (int c = 0; c < 8; c++) f[c] = 0;
        }
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
    const int16_t* lut = bp.T.blut;  // where the plan warp looks Binpack.Rate up (nullptr: fp64)
    if (!ALL_WORK && bp.T.blut != nullptr) {
        for (uint32_t i = tid; i < 9 * RATE_LUT_STRIDE; i += blockDim.x) s_blut[i] = __ldg(bp.T.blut + i);
        lut = s_blut;
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
            filter_pass<WIDE>(S, &s_dem[0], wt, WT, n_iter, my_rows, T.first + row0, lane, g, gbase, S.feasbits, k1, k2, cnt, wcx, widebits, row0);
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
        const bool more = p + 1 < bp.P;
        const SortedDemand* dem_next = &s_dem[(p + 1) & 3];
        if (more && rl_pending && dem_next->policy != NGS_POLICY_RANDOM) {
            rl_pending = false;
            write_back_rl();
        }
        if (is_worker) {
            // ---- workers: filter pod p+1 (pod p's bind may still be in flight: its row is dealt with above) ----
#ifdef NGS_PHASE_TIMING
            tq = clock64();  // (profiling) the filter alone, without the wait at the barrier before it
#endif
            if (more) {
                filter_pass<WIDE>(S, dem_next, wt, WT, n_iter, my_rows, T.first + row0, lane, g, gbase, S.feasbits + ((p + 1) & 1) * fwords, k1,
                                  k2, cnt, wcx, widebits, row0);
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
                    const unsigned long long cj = blockIdx.x == 0 ? cross_rank_exchange(bp, p, ltotal, lane, dead) : collect_ranks(bp, p, lane, dead);
                    unsigned long long inc2 = cj;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc2, o);
                        if (lane >= (uint32_t)o) inc2 += t;
                    }
                    total = __shfl_sync(0xFFFFFFFFu, inc2, 31);
                    before_rank = __shfl_sync(0xFFFFFFFFu, inc2 - cj, bp.rank);
                }
                if (total == 0) {
                    if (blockIdx.x == 0 && lane == 0) publish_none(bp, p, dem);
                } else {
                    const uint64_t gpick = splitmix64(bp.seed_seq + p) % total;
                    const bool mine = gpick >= before_rank && gpick < before_rank + ltotal;
                    if (!mine && blockIdx.x == 0 && lane == 0) {
                        publish_none(bp, p, dem);  // another rank owns the pick; it reports node and plan
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
                                    if (WIDE && ((widebits[r >> 5] >> (r & 31)) & 1u)) publish_wide(bp, wcx, row0 + r, dem, p, 100);
                                    else publish_sample(bp, S, r, row0 + r, dem, p);
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
                    uint64_t fk = 0;
                    if (WIDE && ((widebits[br >> 5] >> (br & 31)) & 1u)) {
                        __syncwarp();  // (the lane that bound the pod has stored the row)
                        if (lane == 0) {
                            int fo = 0;
                            fk = fresh_key_of_wide_row(wcx, row0 + (uint32_t)br, T.first + row0 + (uint32_t)br, dem_next, &fo);
                            ok = fo != 0;
                        }
                    } else {
                        fk = fresh_key_of_row(S, (uint32_t)br, T.first + row0 + (uint32_t)br, dem_next, g, gbase, &ok);
                    }
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
                const bool my_wide = WIDE && have && ((widebits[my_r >> 5] >> (my_r & 31)) & 1u);  // planned only if it wins, below
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
                    ulonglong2* xw = reinterpret_cast<ulonglong2*>(bp.rr + (size_t)p * bp.rr_stride);
                    // the demand the workers will need next is fetched into registers now and stored to shared
                    // memory after the exchange, so that its global-load latency hides behind the wait
                    uint32_t stg0 = 0, stg1 = 0;
                    const bool stage = p + 2 < bp.P;
                    if (stage) {
                        const uint32_t* src = reinterpret_cast<const uint32_t*>(&bp.dem[p + 2]);
                        stg0 = __ldcg(src + lane);
                        if (lane + 32 < sizeof(SortedDemand) / 4) stg1 = __ldcg(src + lane + 32);
                    }
                    const bool solo = gridDim.x == 1 && bp.world == 1;  // one CTA holds every row: nothing to trade
                    if (lane == 0 && !solo) {
                        const unsigned long long old = atomicMax(&xw->x, (unsigned long long)v);
                        const unsigned long long one = 1ull + (old >> 63);  // bit 63 is never set: always 1
                        asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(&xw->y), "l"(one) : "memory");
                    }
                    PHASE(2);
                    if (solo) {
                        if (lane == 0) s_win = v;
                    } else if (bp.world == 1) {
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
                        if (lane == 0) s_win = mg;
                    } else {
                        const uint64_t mg = warp_max_u64(collect_ranks(bp, p, lane, dead));
                        if (lane == 0) s_win = mg;
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
                    if (have && !my_wide) plan_group(bp, S, my_r, T.first + row0 + my_r, dem, more ? dem_next : nullptr, g, gbase, cm, st, lut);
                    if (have && cm.fresh != 12345) {
                        ph[0] += st[0] - t_in;   // exact plan (choose_group8<PLAN>)
                        ph[2] += st[1] - st[0];  // post-bind statistics and rates
                        ph[6] -= 0;
                    }
#else
                    if (have && !my_wide) plan_group(bp, S, my_r, T.first + row0 + my_r, dem, more ? dem_next : nullptr, g, gbase, cm, nullptr, lut);
#endif
                }
                PHASE(4);
                asm volatile("bar.sync 1, 64;" ::: "memory");  // the two control warps only
                PHASE_AFTER_BARRIER(5);
                if (cw == 0) {
                    const uint64_t m = s_win;
                    if (!(m & KEY_FEASIBLE)) {
                        if (blockIdx.x == 0 && lane == 0) publish_none(bp, p, dem);
                    } else if (blockIdx.x == 0 && lane == 0 && key_node(m) - T.first >= T.n) {
                        // the winner lives on another rank: node and score are known from the key, the
                        // plan is filled in by the owner (merged on the host side by an allreduce-max)
                        publish_none(bp, p, dem);
                        bp.node_out[p] = (int32_t)key_node(m);
                        bp.score_out[p] = key_score(m);
                    }
                    if (have && m == v) {  // warp-uniform
                        if (my_wide) {
                            if (lane == 0) {
                                publish_wide(bp, wcx, row0 + my_r, dem, p, key_score(m));
                                if (more) cm.fresh = fresh_key_of_wide_row(wcx, row0 + my_r, T.first + row0 + my_r, dem_next, &cm.fresh_ok);
                            }
                        } else {
                            // ---- bind on the owner: stores only, everything was worked out during the exchange ----
                            publish_group(bp, S, my_r, row0 + my_r, dem, p, key_score(m), cm, lane);
                        }
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
    if (is_worker && lane == 0 && bp.dbg) bp.dbg[4096 + blockIdx.x * 32 + ww] = (unsigned long long)ph[0];  // filter time of every worker warp
#endif
}
