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
    unsigned long long* rr;    // [P] arg-max round records {xa, x1, x2, arrivals, announcement} (the first four in one 32-byte
                               // sector; the round that starts at pod p uses pod p's record), zero before launch
    // node-range sharding over the GPUs of one box: every rank runs this same kernel on
    // its own rows; CTA 0 of each rank trades the rank's result with the peers through
    // peer-mapped memory (plain 8-byte stores over NVLink); every CTA reads the ranks' words from there
    uint32_t world, rank;
    unsigned long long* xslots[NGS_MAX_RANKS];  // [rank r] -> r's slot array [P][NGS_MAX_RANKS][2] (this batch's parity half)
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
    uint32_t tok;    // bit j: the row after the bind takes target pod j (PlanTargets)
    uint64_t fresh;  // target 0: key of the changed row for that pod (0 = not feasible, or a random-policy pod)
    int32_t fresh_ok;  // target 0: feasible
};
// the pods that may come to a row right after a bind, in the order the caller wants their verdicts (at most three)
struct PlanTargets {
    const SortedDemand* t0;
    const SortedDemand* t1;
    const SortedDemand* t2;
    int n;
};
// the Binpack.Rate percent table may sit in shared memory (K6 with few rows per CTA): plain loads, not ld.global.nc
__device__ __forceinline__ int binpack_pct_any(const int16_t* lut, int used, int G) {
    return (lut != nullptr && used >= 0 && used <= 100 * G) ? (int)lut[G * RATE_LUT_STRIDE + used] : binpack_percent_fp64(used, G);
}

__device__ __noinline__ void plan_group(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t node, const SortedDemand* dem,
                                           const PlanTargets& pt, const int16_t* lut, int g, int gbase, GroupCommit& cm) {
    (void)bp;
    const uint32_t lane = (uint32_t)(gbase + g);
    const uint32_t meta = S.meta[r];
    const int G = (meta >> 24) & 15;
    const int load_int = (int)(meta >> 28);
    const int rlc = (int)((meta >> (3 * g)) & 7);
    const bool valid = g < G;
    choose_group8<true>(g, gbase, smem_cell(S, r, g), rlc, valid, dem->policy == NGS_POLICY_SPREAD, dem, &cm.emit, &cm.n_emit, &cm.fr);
    const int fr = valid ? cm.fr : 0;
    cm.fr = fr;
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
    cm.top4 = t4;
    cm.scB = binpack_pct_any(lut, used, G) + load_int * 50 - G;  // rater.go:69
    cm.scS = 100 * cnt100 + sumf / 10 - G - load_int;            // rater.go:122
    cm.tok = 0;
    cm.fresh = 0;
    cm.fresh_ok = 0;
    if (pt.n == 0) return;
    // The target pods on the row as it will stand after this bind.  First what the cells tell outright (every lane
    // computes the same, all groups hold the same row); whatever stays open is then searched exactly in ONE call,
    // group j taking target j.
    uint32_t tok = 0, open = 0;
    int f[8];
    bool have_f = false;
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (j >= pt.n) break;  // (warp-uniform)
        const SortedDemand* dn = j == 0 ? pt.t0 : (j == 1 ? pt.t1 : pt.t2);
        const int pol2 = dn->policy;
        int cls = pol2 == NGS_POLICY_RANDOM ? FEAS_FALLBACK : classify_top4(t4, dn, true);
        if (cls == FEAS_FALLBACK || dn->nz == 0) {  // (warp-uniform)
            if (!have_f) {
#pragma unroll
                for (int i = 0; i < 8; i++) f[i] = __shfl_sync(FULL, fr, gbase + i);
                have_f = true;
            }
            cls = classify_row(f, dn, pol2 != NGS_POLICY_RANDOM);
            if (cls == FEAS_UNDECIDED && pol2 == NGS_POLICY_RANDOM) {
                int f2[8];
#pragma unroll
                for (int i = 0; i < 8; i++) f2[i] = f[i];
                cls = choose_sample<false>(f2, G, dn, nullptr, nullptr) ? FEAS_YES : FEAS_NO;
            }
        }
        if (cls == FEAS_UNDECIDED && sumf < (int)((uint32_t)dn->flags >> 1)) cls = FEAS_NO;  // less is left than the pod asks for
        tok |= (cls == FEAS_YES ? 1u : 0u) << j;
        open |= (cls == FEAS_UNDECIDED ? 1u : 0u) << j;
    }
    if (open) {  // (warp-uniform)
        const uint32_t j = lane >> 3;
        const SortedDemand* mine = ((open >> j) & 1u) ? (j == 0 ? pt.t0 : (j == 1 ? pt.t1 : pt.t2)) : nullptr;
        const bool ok = choose_group8_multi(g, gbase, fr, rlc, valid, mine);
        const uint32_t okb = __ballot_sync(FULL, ok && g == 0);  // group j's verdict at bit 8j
#pragma unroll
        for (int jj = 0; jj < 3; jj++) tok |= (((open >> jj) & 1u) & ((okb >> (8 * jj)) & 1u)) << jj;
    }
    cm.tok = tok;
    cm.fresh_ok = (int32_t)(tok & 1u);
    if ((tok & 1u) && pt.t0->policy != NGS_POLICY_RANDOM) cm.fresh = make_key(pt.t0->policy == NGS_POLICY_SPREAD ? cm.scS : cm.scB, node);
}

// publish a winning decision (all 32 lanes of control warp 0): shared row, cached rates, the global
// table and the pod's outputs -- stores only, everything was computed by plan_group
__device__ __noinline__ void publish_group(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t row,
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
__device__ __noinline__ void publish_sample(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t row,
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
// bounded wait for a word that is VALID as soon as it is not zero (pod a's packed word uses all 64 bits)
__device__ __forceinline__ unsigned long long wait_nonzero(const unsigned long long* p, int32_t* err, bool& dead) {
    unsigned long long v = ld_relaxed_sys_u64(p);
    if (v) return v;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!(v = ld_relaxed_sys_u64(p))) {
        if ((++spins & 255u) == 0) {
            if (abort_raised(err) || clock64() - t0 > PEER_TIMEOUT_CYCLES) {
                raise_timeout(err);
                dead = true;
                return 0;
            }
        }
    }
    return v;
}
// bounded wait for a VALID word written by a CTA of the same launch
__device__ __forceinline__ unsigned long long wait_valid_gpu(const unsigned long long* p, int32_t* err, bool& dead) {
    unsigned long long v = ld_relaxed_u64(p);
    if (v & KEY_VALID) return v & ~KEY_VALID;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!((v = ld_relaxed_u64(p)) & KEY_VALID)) {
        if ((++spins & 255u) == 0) {
            if (abort_raised(err) || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                raise_timeout(err);
                dead = true;
                return 0;
            }
        }
    }
    return v & ~KEY_VALID;
}
// Cross-GPU slot arrays: every rank owns an array of [pod][rank][2] 64-bit words (bit 63 = VALID); word w of pod p written
// BY rank j lives at index (p * NGS_MAX_RANKS + j) * 2 + w of the array OF the rank that is to read it.
__device__ __forceinline__ unsigned long long* xslot(unsigned long long* base, uint32_t p, uint32_t writer, uint32_t w) {
    return base + ((size_t)p * NGS_MAX_RANKS + writer) * 2 + w;
}
// CTA 0, one full warp: post this rank's word to every rank's slot array, collect all ranks'
// words from our own array.  Returns lane j's word of rank j (0 for lanes >= world).
__device__ __forceinline__ unsigned long long cross_rank_exchange(const BatchParams& bp, uint32_t p, uint32_t w, unsigned long long mine,
                                                                  uint32_t lane, bool& dead) {
    unsigned long long got = 0;
    if (lane < bp.world) {
        st_relaxed_sys_u64(xslot(bp.xslots[lane], p, bp.rank, w), mine | KEY_VALID);
        got = wait_valid(xslot(bp.xslots[bp.rank], p, lane, w), bp.err, dead);
    }
    dead = __any_sync(0xFFFFFFFFu, dead);
    return got;
}
// CTA 0: post only (the words are collected afterwards, together, by collect_ranks)
__device__ __forceinline__ void post_ranks(const BatchParams& bp, uint32_t p, uint32_t w, unsigned long long mine, uint32_t lane) {
    if (lane < bp.world) st_relaxed_sys_u64(xslot(bp.xslots[lane], p, bp.rank, w), mine | KEY_VALID);
}

// every CTA: the ranks' words straight from this rank's slot array (CTA 0 posts our own word there,
// too), so the result needs no relay through a second location.  Same return convention.
__device__ __forceinline__ unsigned long long collect_ranks(const BatchParams& bp, uint32_t p, uint32_t w, uint32_t lane, bool& dead) {
    unsigned long long got = 0;
    if (lane < bp.world) got = wait_valid(xslot(bp.xslots[bp.rank], p, lane, w), bp.err, dead);
    dead = __any_sync(0xFFFFFFFFu, dead);
    return got;
}

// Arg-max round record {xa, x1, x2, arrivals}: three atomicMax words and the arrival counter share one 32-byte sector, and
// ONE 256-bit load (LDG.E.256.STRONG.GPU) returns all four.  A CTA bumps the counter only after its atomicMax operations
// have returned (data dependence), i.e. after L2 has performed them, hence a load that sees the counter full also sees the
// final keys.  (This leans on the hardware serving the 32-byte sector as one access - the PTX memory model alone would
// not order the relaxed accesses of a vector load; DESIGN.md "exchange word".)  Returns false on time-out.
__device__ __forceinline__ bool wait_round(const unsigned long long* rr, uint32_t target, int32_t* err, unsigned long long (&w)[3]) {
    unsigned long long a, b, c, d;
    const long long t0 = clock64();
    uint32_t spins = 0;
    for (;;) {
        asm volatile("ld.relaxed.gpu.global.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(rr) : "memory");
        if ((uint32_t)d >= target) {
            w[0] = a;
            w[1] = b;
            w[2] = c;
            return true;
        }
        if ((++spins & 255u) == 0) {
            if (abort_raised(err) || clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                raise_timeout(err);
                w[0] = w[1] = w[2] = 0;
                return false;
            }
        }
    }
}

// Pod a's word of a two-pod round: the CTA's best key for pod a and, in the low bits, what that row will offer pod b
// once pod a is bound to it (F).  The maximum over CTAs is decided by the key part alone (two CTAs never hold the same
// node), so the winning word carries the winner's own F:
//   [63] feasible   [62:45] score + 2^17   [44:20] 2^25-1 - node   [19:18] F state   [17:0] F score + 2^17
// F has to be in the word when it is posted, i.e. BEFORE the row's exact plan is known (the plan is worked out while
// the exchange is in flight), so it is what can be said about the row from its cells alone (f_certificate):
//   F_NO     pod b cannot go to the row after the bind, whatever the plan           -> score unused
//   F_YES    it can, and the score the row will then have is known exactly          -> that score
//   F_MAYBE  open; the score is an upper bound of what the row could offer.  If even that bound loses against the best
//            other row, pod b is settled all the same; only otherwise the owner has to say (announce word), once its
//            plan has told it.
// Scores fit 18 bits because cells are bounded by NGS_MAX_FREE (binpack >= -65543, spread <= 53228 + load terms); node
// indices have to stay below 2^25 (checked by the host before the launch).
constexpr int PK_BIAS = 1 << 17;
constexpr uint32_t PK_NODE_MAX = (1u << 25) - 1u;
enum { F_NO = 0, F_YES = 1, F_MAYBE = 2 };
__device__ __forceinline__ uint64_t pack_a(uint64_t key, int f_state, int f_score) {
    if (!(key & KEY_FEASIBLE)) return 0;
    return (1ull << 63) | ((uint64_t)(uint32_t)(key_score(key) + PK_BIAS) << 45) | ((uint64_t)(PK_NODE_MAX - key_node(key)) << 20) |
           ((uint64_t)(uint32_t)f_state << 18) | (uint64_t)(uint32_t)((f_state == F_NO ? 0 : f_score) + PK_BIAS);
}
__device__ __forceinline__ uint64_t unpack_a_key(uint64_t pk) {  // the ordinary 64-bit key (make_key), 0 if none
    if (!(pk >> 63)) return 0;
    return make_key((int)((pk >> 45) & 0x3FFFFu) - PK_BIAS, PK_NODE_MAX - (uint32_t)((pk >> 20) & 0x1FFFFFFu));
}
__device__ __forceinline__ int unpack_a_fstate(uint64_t pk) { return (int)((pk >> 18) & 3u); }
__device__ __forceinline__ int unpack_a_fscore(uint64_t pk) { return (int)(pk & 0x3FFFFu) - PK_BIAS; }

// What row r will offer pod b once pod a (which it is known to take) has been bound to it, from the row's cells alone.
// GPUs.Allocate takes exactly the sum of a's percents off the row, wherever the plan puts them, so
//   * Binpack.Rate of the row after the bind is known exactly (it only depends on the sum, rater.go:59-70);
//     Spread.Rate has the count of untouched (100 %) cells in it, which can only drop: an upper bound (rater.go:113-123);
//   * b cannot fit if no cell holds its largest container now (cells only shrink) or if less capacity is left than b asks for;
//   * b fits if at least nz_a + nz_b cells hold its largest container: pod a touches at most nz_a of them, and nz_b
//     untouched cells that hold d1 are enough (classify_row's second statement).
// Every lane computes the same values.
__device__ __noinline__ void f_certificate(const SmemRows& S, uint32_t r, const SortedDemand* da, const SortedDemand* db,
                                              const int16_t* blut, int& state, int& score) {
    const uint32_t meta = S.meta[r];
    const int G = (meta >> 24) & 15;
    const int load_int = (int)(meta >> 28);
    int f[8];
    unpack_free(S.fA[r], S.fB[r], f);
    int sum = 0, mx = 0, cnt100 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        sum += f[i];
        mx = max(mx, f[i]);
        cnt100 += (i < G && f[i] == 100) ? 1 : 0;
    }
    const int sum2 = sum - (int)((uint32_t)da->flags >> 1);
    const int sb = (int)((uint32_t)db->flags >> 1);
    const bool binpack = db->policy == NGS_POLICY_BINPACK;
    if (binpack) {
        const int used = 100 * G - sum2;
        score = binpack_pct_any(blut, used, G) + load_int * 50 - G;
    } else {
        score = 100 * cnt100 + sum2 / 10 - G - load_int;
    }
    const int nzb = db->nz, d1 = db->d[0], m = da->nz + nzb;
    state = F_MAYBE;
    if (nzb == 0) {
        state = binpack ? F_YES : F_MAYBE;  // fits for sure; only binpack knows the score exactly
    } else if (mx < d1 || sum2 < sb) {
        state = F_NO;
    } else if (binpack && m <= 4 && d1 <= 255 && (int)((S.top4[r] >> (8 * (m - 1))) & 0xFFu) >= d1) {
        state = F_YES;
    }
}

__device__ __forceinline__ unsigned long long read_slot(const unsigned long long* slot, bool dead) {
    unsigned long long v = ld_relaxed_u64(slot);
    while (!(v & KEY_VALID) && !dead) v = ld_relaxed_u64(slot);  // normally valid at the first read
    return v & ~KEY_VALID;
}

// filter (and rate) this worker warp's rows for one pod, on the rows as they stand.  Every policy: one
// feasibility bit per row in the pod's feasbits words.  Argmax policies: every thread keeps the best two
// keys of its rows (k1 >= k2).  Random policy: the warp's count in cnt (lane 0).
// Must be called by all 32 lanes.
__device__ __noinline__ void filter_pass(const SmemRows& S, const SortedDemand* dem, uint32_t wt, uint32_t WT, uint32_t n_iter,
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
        }
        // the verdicts as one bit per row (the 32 rows of this warp's iteration are one word): what the control warps
        // recompute this warp's keys from when one of its rows has changed in between
        {
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, cls == FEAS_YES);
            if (lane == 0 && r < my_rows) fb[r >> 5] = bal;
        }
        // best two by score, branch-free: an infeasible row enters as NONE and changes nothing
        const int sc = cls == FEAS_YES ? *ratep : NONE;
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
__device__ __noinline__ uint64_t fresh_key_of_row(const SmemRows& S, uint32_t r, uint32_t node, const SortedDemand* dem, int g,
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

// The best two keys of worker warp ww's rows for one arg-max pod, recomputed from the pod's feasibility bits and the rows'
// current rates: what the control warps do for a warp one of whose rows has just been bound (bit and rate already patched).
// Same row order as filter_pass (row = it * WT + ww * 32 + lane), same tie rule.  All 32 lanes.
__device__ __noinline__ void warp_keys_from_bits(const uint32_t* fb, const int32_t* __restrict__ rate, uint32_t ww, uint32_t WT,
                                                    uint32_t n_iter, uint32_t my_rows, uint32_t node0, uint32_t lane, uint64_t& m1, uint64_t& m2) {
    constexpr int NONE = INT32_MIN;
    int s1 = NONE, s2 = NONE;
    uint32_t r1 = 0, r2 = 0;
    uint32_t r = ww * 32u + lane;
    for (uint32_t it = 0; it < n_iter; it++, r += WT) {
        const bool ok = r < my_rows && ((fb[r >> 5] >> lane) & 1u);
        const int sc = ok ? rate[r] : NONE;
        const bool gt1 = sc > s1, gt2 = sc > s2;
        s2 = gt1 ? s1 : (gt2 ? sc : s2);
        r2 = gt1 ? r1 : (gt2 ? r : r2);
        s1 = gt1 ? sc : s1;
        r1 = gt1 ? r : r1;
    }
    const uint64_t k1 = s1 != NONE ? make_key(s1, node0 + r1) : 0;
    const uint64_t k2 = s2 != NONE ? make_key(s2, node0 + r2) : 0;
    warp_top2(k1, k2, m1, m2);
}

// K6.  One CTA per SM (cooperative launch, so every CTA is resident), each owning a contiguous
// node range held in shared memory for the whole batch.  The last two warps are control warps, the
// others ("workers") own the rows.  Pods are decided strictly in order, in ROUNDS of up to two pods
// per grid-wide exchange, and the work is pipelined: while the control warps settle the round's pods,
// the workers already filter the next round's pods on the rows as they stand.
//
//   filter (workers): one thread per row settles feasibility with the exact shortcuts of
//      classify_row; the ~1% of rows they leave open are searched exactly right away, 8 lanes per
//      row (choose_group8).  Rate is a function of the row only (allocate.go:96), so both
//      policies' rates sit beside the row and are refreshed when the row changes.  Per pod a worker
//      warp leaves one feasibility bit per row and its best two keys.
//
//   a round of two arg-max pods (a, b = a + 1) - ONE exchange decides both:
//      control warp 0:  this CTA's best key for a and F, what that row will offer pod b after the bind as far as the
//         row's cells tell (f_certificate: no / yes with its exact score / open with a score bound) -> atomicMax of
//         [key | F] into the round's word xa (pack_a) -> arrival counter.  THEN, while the exchange is in flight, the
//         exact plan of that row (choose_group8<PLAN>), the row as GPUs.Allocate would leave it, its rates, and the
//         exact answer for pod b.
//      control warp 1:  this CTA's best TWO keys for b (rows as they stand) -> old = atomicMax(x1, k1);
//         atomicMax(x2, max(min(old, k1), k2)) -> arrival counter.  Whatever order the CTAs arrive in, x1 ends up the
//         best and x2 the second best key of the whole grid for pod b: every key but the final maximum is pushed
//         into x2 either by its own CTA (it lost against `old`) or by the CTA that displaced it (it got `old` back).
//         While the first atomic is in flight the warp works out the exact plan of k1's row for pod b.
//      everyone (one polling lane per CTA, one 256-bit load per poll): xa names pod a's winner r* and carries its F;
//         binding a changes r* and nothing else, so pod b's best key over all OTHER rows is
//         Y = x1 if x1 is not r*'s (stale) key, else x2, and pod b's winner is max(r*'s key after the bind, Y).
//         F says no: Y.  F says yes: compare.  F open: if even its score bound loses against Y, Y; otherwise - the one
//         case that needs a second message - the owner of r* announces the exact answer (one word, written once its
//         plan is done, polled by everyone: one L2 trip instead of a second grid-wide exchange).
//      The owner of each winner binds it with stores only; in the one case nobody prepared (pod b goes to the row
//      pod a was just bound to) the owner computes that plan then.
//   rounds of one pod: an arg-max pod alone (the last pod, or the next one is a random-policy pod) uses xa only;
//      a random-policy pod uses per-CTA feasible counts and a seeded k-th pick over the feasible set.
//
//   after a bind the owner evaluates the changed row for the next round's pods; at the start of that round the control
//   warps put the verdicts into the pods' feasibility bits and recompute, from bits and rates, the best two keys of
//   the one worker warp that owns the row (warp_keys_from_bits).  Every other warp's keys are exact as they are: its
//   rows did not change.
//
// K5 (node-range sharding): after the local exchange CTA 0 posts the rank's (xa, x1, x2) to every peer; every CTA
// reads the ranks' words from its own rank's slot array and folds them (xa: max, which keeps the winner's F; x1 / x2:
// best two of the union).  One NVLink round per TWO pods; an announcement is stored by the owner CTA straight into
// every rank's array.
//
// The workers filter the next round's pods side by side: with two pods, one half of the worker warps takes pod a over
// all rows and the other half pod b, so that the round's filter costs one pod's latency, not two.
//
// ALL_WORK (chosen by the host from the rows per CTA, schedule_kernel()): every warp below the control warps
// filters; otherwise the warps that share control warp 0's scheduler stay idle.  A compile-time choice: as a
// run-time flag it cost the small-CTA case 1.8 %.
constexpr uint32_t DEM_RING = 8;  // demands staged in shared memory: the round's pods and the five after its first
struct RoundResult {
    uint64_t ka;      // pod a's winning key (0: no node takes it)
    uint64_t kb;      // pod b's winning key (final once need_ann is 0)
    uint64_t y;       // pod b's best key over all rows but pod a's winner
    int b_on_a;       // pod b goes to the row pod a was bound to
    int need_ann;     // pod b waits for the word of the owner of pod a's winner
    int owner_rank;   // rank that owns pod a's winner (sharded launches)
};

template <bool ALL_WORK>
__global__ void __launch_bounds__(1024, 1) k_schedule(BatchParams bp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ SortedDemand s_dem[DEM_RING];
    __shared__ uint64_t s_red[4][32][2];  // per pod (ring of 4), per worker warp: best two keys (random policy: {count, -})
    __shared__ RoundResult s_res;         // the round's outcome, from the polling warp to the other control warp
    __shared__ uint64_t s_loc[3];         // single-CTA launches: the words that would otherwise travel through L2
    // what the round's binds leave for the next round (written at the end of a round, read at the start of the next):
    __shared__ int s_dirty_row[2];        // [slot]: row of this CTA a pod was bound to, -1 if none (slot 0: pod a, 1: pod b)
    __shared__ int s_dirty_ok[2][2];      // [slot][j]: that row's verdict for the j-th pod of the next round
    __shared__ int s_abort;
    // Binpack.Rate's percent table next to the rows (the control path reads it once or twice per round, and an L2 trip
    // there is an L2 trip on the round's critical path); CTAs that need all of shared memory for rows keep it in L2
    __shared__ int16_t s_blut[ALL_WORK ? 2 : 9 * RATE_LUT_STRIDE];
    const uint32_t rpb = bp.rows_per_block;
    if (threadIdx.x == 0) {
        s_abort = 0;
        s_dirty_row[0] = s_dirty_row[1] = -1;
    }
    const uint32_t fwords = (rpb + 31) / 32;
    SmemRows S;
    S.fA = reinterpret_cast<uint2*>(smem_raw);
    S.fB = S.fA + rpb;
    S.meta = reinterpret_cast<uint32_t*>(S.fB + rpb);
    S.scB = reinterpret_cast<int32_t*>(S.meta + rpb);
    S.scS = S.scB + rpb;
    S.top4 = reinterpret_cast<uint32_t*>(S.scS + rpb);
    S.feasbits = S.top4 + rpb;  // four buffers of fwords words (pod & 3)

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    // The two control warps are the LAST two: the SM's issue arbiter favours high warp ids, and the control path
    // (exchange, plan) is the latency-critical one.  Warps map to the SM's four schedulers by warp id % 4; the worker
    // warps that would share control warp 0's scheduler stay idle (they own no rows), so that its instructions never
    // queue behind the filter's - unless a CTA holds so many rows that the filter, not the exchange, sets the pace
    // (ALL_WORK): then every warp below the control warps owns rows.
    const uint32_t NWARP = nwarps - 2;                     // warps below the control warps
    constexpr bool all_work = ALL_WORK;
    const uint32_t csm = NWARP & 3u;                       // scheduler of control warp 0
    const uint32_t idle_below = all_work ? 0u : (warp + 3u - csm) >> 2;    // idle warps among [0, warp)
    const uint32_t NW = all_work ? NWARP : NWARP - ((NWARP + 3u - csm) >> 2); // worker warps that own rows
    const bool is_worker = warp < NWARP && (all_work || (warp & 3u) != csm);
    const bool is_control = warp >= NWARP;
    const uint32_t ww = warp - idle_below;                 // worker index (meaningful for workers)
    const uint32_t cw = warp - NWARP;                      // control warp 0 (pod a: plan, bind) / 1 (pod b; polls the exchange)
    const uint32_t row0 = blockIdx.x * rpb;
    const uint32_t my_rows = row0 < bp.T.n ? min(rpb, bp.T.n - row0) : 0u;
    const Tables& T = bp.T;
    const int g = lane & 7, gbase = lane & 24;
    const uint32_t node0 = T.first + row0;
    const uint32_t P = bp.P;

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
    const int16_t* lut = nullptr;  // what the control warps look Binpack.Rate up in (nullptr: fp64)
    if (!ALL_WORK && bp.T.blut != nullptr) {
        for (uint32_t i = tid; i < 9 * RATE_LUT_STRIDE; i += blockDim.x) s_blut[i] = __ldg(bp.T.blut + i);
        lut = s_blut;
    }
    constexpr uint32_t DEM_WORDS = sizeof(SortedDemand) / 4;
    for (uint32_t i = tid; i < 6 * DEM_WORDS; i += blockDim.x)  // the first round's window: pods 0..5
        if (i / DEM_WORDS < P) reinterpret_cast<uint32_t*>(&s_dem[0])[i] = reinterpret_cast<const uint32_t*>(&bp.dem[0])[i];
    __syncthreads();

#ifdef NGS_PHASE_TIMING
    long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tq = clock64();
#define PHASE(i) do { const long long t_ = clock64(); ph[i] += t_ - tq; tq = t_; } while (0)
    long long ev[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // event counts / owner-only cycles (see phase_timing.py)
#define EVENT(i, n) do { ev[i] += (n); } while (0)
    // per-round trace in GPU-wide time (ns): [round < 256][CTA][8] = cw0: post a, plan done, binds done | cw1: round start,
    // posted, resolved, binds done  (profiles/tools/phase_timing.py turns it into who-was-last statistics)
    uint32_t rnd = 0;
#define TRACE(k) do { if (lane == 0 && bp.dbg && rnd < 256) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); \
                      bp.dbg[32768 + ((size_t)rnd * gridDim.x + blockIdx.x) * 8 + (k)] = t_; } } while (0)
#else
#define TRACE(k) do { } while (0)
#define PHASE(i) do { } while (0)
#define EVENT(i, n) do { } while (0)
#endif
    // Worker geometry of a round: a round of one pod is filtered by all NW worker warps, a round of two pods by two
    // groups of NW / 2 warps, one pod each, side by side (every group covers all rows).  The control warps read a
    // pod's per-warp results with the geometry of the round the pod belongs to.
    const bool can_split = NW >= 2;
    auto geom_nw = [&](uint32_t Lr) -> uint32_t { return (Lr == 2 && can_split) ? NW / 2 : NW; };
    bool rl_pending = bp.lp.enabled != 0;  // RemainLoad write-back of the first load-aware Rate

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
    auto dem_of = [&](uint32_t p) -> const SortedDemand* { return &s_dem[p & (DEM_RING - 1)]; };
    // pods of the round that starts at pod p: two when both are arg-max pods, else one
    auto round_len = [&](uint32_t p) -> uint32_t {
        return (p + 1 < P && dem_of(p)->policy != NGS_POLICY_RANDOM && dem_of(p + 1)->policy != NGS_POLICY_RANDOM) ? 2u : 1u;
    };
    // workers: filter one pod on the rows as they stand and hand the warp's result to the control warps
    // (wg: this warp's index among the gnw warps that work on the pod)
    auto filter_pod = [&](uint32_t p, uint32_t wg, uint32_t gnw) {
        const SortedDemand* dem = dem_of(p);
        uint64_t k1 = 0, k2 = 0;
        uint32_t cnt = 0;
        const uint32_t gwt = gnw * 32u;
        filter_pass(S, dem, wg * 32u + lane, gwt, (my_rows + gwt - 1) / gwt, my_rows, node0, lane, g, gbase, S.feasbits + (p & 3) * fwords, k1, k2, cnt);
        if (dem->policy != NGS_POLICY_RANDOM) {
            uint64_t m1, m2;
            warp_top2(k1, k2, m1, m2);
            if (lane == 0) {
                s_red[p & 3][wg][0] = m1;
                s_red[p & 3][wg][1] = m2;
            }
        } else if (lane == 0) {
            s_red[p & 3][wg][0] = cnt;
        }
    };
    auto filter_round = [&](uint32_t q0, uint32_t Lr) {  // all worker warps
        const uint32_t gnw = geom_nw(Lr);
        if (Lr == 2 && can_split) {
            if (ww < 2 * gnw) filter_pod(q0 + ww / gnw, ww % gnw, gnw);
        } else {
            for (uint32_t j = 0; j < Lr; j++) filter_pod(q0 + j, ww, gnw);
        }
    };
    // the owner of a bind: the changed row and its verdicts for the pods of the next round (worked out with the plan)
    auto leave_dirty = [&](int slot, int row_or_none, uint32_t ok0, uint32_t ok1) {
        if (lane == 0) {
            s_dirty_row[slot] = row_or_none;
            s_dirty_ok[slot][0] = (int)ok0;
            s_dirty_ok[slot][1] = (int)ok1;
        }
    };
    // random policy: the bound row is evaluated for the next round's pods after the fact (all 32 lanes)
    auto leave_dirty_eval = [&](int slot, int row_or_none, uint32_t q0, uint32_t Lr) {
        uint32_t ok0 = 0, ok1 = 0;
        if (row_or_none >= 0) {  // (warp-uniform)
            __syncwarp();  // the row as publish_sample left it
            for (uint32_t j = 0; j < Lr; j++) {
                bool ok = false;
                fresh_key_of_row(S, (uint32_t)row_or_none, node0 + (uint32_t)row_or_none, dem_of(q0 + j), g, gbase, &ok);
                if (j == 0) ok0 = ok; else ok1 = ok;
            }
        }
        leave_dirty(slot, row_or_none, ok0, ok1);
    };

    // round 0 is filtered before the loop; inside the loop the workers are always one round ahead
    if (P > 0) {
        const uint32_t L0 = round_len(0);
        bool argmax_pod = false;
        for (uint32_t j = 0; j < L0; j++) argmax_pod = argmax_pod || dem_of(j)->policy != NGS_POLICY_RANDOM;
        if (rl_pending && argmax_pod) {
            rl_pending = false;
            write_back_rl();
        }
        if (is_worker) filter_round(0, L0);
    }
    __syncthreads();

    const bool solo = gridDim.x == 1 && bp.world == 1;  // one CTA holds every row: nothing to trade
    // a run-time zero the compiler cannot see through: ties the arrival counter's increment to the value an atomic returned
    const unsigned long long dep_mask = (unsigned long long)(P >> 24);

    // ONE CTA-wide barrier per round.
    for (uint32_t p = 0; p < P;) {
        const uint32_t L = round_len(p);
        const uint32_t q = p + L;                           // first pod of the next round
        const uint32_t Lq = q < P ? round_len(q) : 0u;
        {
            bool argmax_pod = false;
            for (uint32_t j = 0; j < Lq; j++) argmax_pod = argmax_pod || dem_of(q + j)->policy != NGS_POLICY_RANDOM;
            if (rl_pending && argmax_pod) {
                rl_pending = false;
                write_back_rl();
            }
        }
        if (is_worker) {
            // ---- workers: filter the next round's pods (this round's binds may still be in flight: the rows they
            // change are dealt with by the control warps at the start of the next round) ----
#ifdef NGS_PHASE_TIMING
            tq = clock64();
#endif
            filter_round(q, Lq);
            PHASE(0);
        } else if (is_control) {
            const SortedDemand* dem_a = dem_of(p);
            const bool random_round = dem_a->policy == NGS_POLICY_RANDOM;
            const bool two = L == 2;
            const uint32_t gnw = geom_nw(L), gwt = gnw * 32u, gn_iter = (my_rows + gwt - 1) / gwt;  // how this round's pods were filtered
            PHASE(7);
            if (cw == 1) TRACE(3);
            // ---- what the previous round's binds changed: patch this round's pods (control warp j looks after pod p + j) ----
            if (cw < L) {
                const uint32_t pj = p + cw;
                const SortedDemand* dem = dem_of(pj);
                uint32_t* fb = S.feasbits + (pj & 3) * fwords;
#pragma unroll
                for (int slot = 0; slot < 2; slot++) {
                    const int dr = s_dirty_row[slot];
                    if (dr < 0) continue;  // (warp-uniform)
                    const bool ok = s_dirty_ok[slot][cw] != 0;
                    const uint32_t wi = (uint32_t)dr >> 5, bit = 1u << (dr & 31);
                    const uint32_t old = fb[wi];
                    __syncwarp();
                    if (lane == 0) fb[wi] = ok ? (old | bit) : (old & ~bit);
                    __syncwarp();
                    if (dem->policy == NGS_POLICY_RANDOM) {
                        // counts only matter as a sum: correct any warp's
                        if (lane == 0) s_red[pj & 3][0][0] += (uint64_t)(ok ? 1 : 0) - (uint64_t)((old & bit) ? 1 : 0);
                    } else {
                        const uint32_t dw = ((uint32_t)dr % gwt) >> 5;  // the worker warp (of the pod's group) that owns the row
                        uint64_t m1, m2;
                        warp_keys_from_bits(fb, dem->policy == NGS_POLICY_SPREAD ? S.scS : S.scB, dw, gwt, gn_iter, my_rows, node0, lane, m1, m2);
                        if (lane == 0) {
                            s_red[pj & 3][dw][0] = m1;
                            s_red[pj & 3][dw][1] = m2;
                        }
                    }
                    __syncwarp();
                }
            }
            PHASE(1);
            // the demands the next rounds will need (pods p + 6 .. p + 5 + L) travel to shared memory in the shadow of the
            // exchange: control warp 1's lanes fetch them now and store them at the end of the round
            uint32_t stg[2][2] = {{0, 0}, {0, 0}};
            if (cw == 1) {
#pragma unroll
                for (uint32_t j = 0; j < 2; j++) {
                    const uint32_t ps = p + 6 + j;
                    if (j < L && ps < P) {
                        const uint32_t* src = reinterpret_cast<const uint32_t*>(&bp.dem[ps]);
                        stg[j][0] = __ldcg(src + lane);
                        if (lane + 32 < DEM_WORDS) stg[j][1] = __ldcg(src + lane + 32);
                    }
                }
            }
            if (random_round) {
              if (cw == 0) {
                // ---- control, random policy: counts -> seeded pick -> bind ----
                bool dead = false;
                unsigned long long* krow = bp.keys + (size_t)p * bp.keys_stride;
                uint32_t* done = bp.done + (size_t)p * bp.done_stride;
                uint32_t* fb = S.feasbits + (p & 3) * fwords;
                const uint64_t v = lane < gnw ? s_red[p & 3][lane][0] : 0;
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
                    const unsigned long long cj = blockIdx.x == 0 ? cross_rank_exchange(bp, p, 0, ltotal, lane, dead) : collect_ranks(bp, p, 0, lane, dead);
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
                    if (blockIdx.x == 0 && lane == 0) publish_none(bp, p, dem_a);
                } else {
                    const uint64_t gpick = splitmix64(bp.seed_seq + p) % total;
                    const bool mine = gpick >= before_rank && gpick < before_rank + ltotal;
                    if (!mine && blockIdx.x == 0 && lane == 0) {
                        publish_none(bp, p, dem_a);  // another rank owns the pick; it reports node and plan
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
                                    for (uint32_t qq = 0; qq < need; qq++) bits &= bits - 1;
                                    const uint32_t r = wi * 32 + (__ffs(bits) - 1);
                                    publish_sample(bp, S, r, row0 + r, dem_a, p);
                                    bound_row = (int)r;
                                    break;
                                }
                                need -= pc;
                            }
                        }
                    }
                }
                }
                // if this CTA bound the pod, evaluate that row for the next round's pods right away
                const uint32_t bmask = __ballot_sync(0xFFFFFFFFu, bound_row >= 0);
                const int br = bmask ? __shfl_sync(0xFFFFFFFFu, bound_row, __ffs(bmask) - 1) : -1;
                leave_dirty_eval(0, br, q, Lq);
                if (lane == 0) s_dirty_row[1] = -1;
                if (dead && lane == 0) s_abort = 1;
              }
            } else {
                // ---- control, arg-max policies ----
                const SortedDemand* dem_b = dem_of(p + 1);  // (meaningful when two)
                bool dead = false;
                unsigned long long* rr = bp.rr + (size_t)p * bp.rr_stride;  // {xa, x1, x2, arrivals, announcement}
                GroupCommit cm;
                cm.fr = 0; cm.emit = 0; cm.n_emit = 0; cm.top4 = 0; cm.scB = cm.scS = 0; cm.tok = 0; cm.fresh = 0; cm.fresh_ok = 0;
                // whom a row bound in this round has to be ready for: the next round's pods (and, for pod a's row, pod b first)
                PlanTargets nx;
                nx.t0 = Lq >= 1 ? dem_of(q) : nullptr;
                nx.t1 = Lq >= 2 ? dem_of(q + 1) : nullptr;
                nx.t2 = nullptr;
                nx.n = (int)Lq;
                PlanTargets nxa = nx;  // pod a's row: pod b, then the next round's pods
                if (two) {
                    nxa.t0 = dem_b;
                    nxa.t1 = nx.t0;
                    nxa.t2 = nx.t1;
                    nxa.n = (int)Lq + 1;
                }
                const uint32_t sh_a = two ? 1u : 0u;  // where the next round's verdicts start in pod a's plan
                uint64_t mine_key = 0;   // cw 0: this CTA's key for pod a;  cw 1: its best key for pod b
                uint32_t my_r = 0;
                if (cw == 0) {
                    // pod a: this CTA's best row and what its cells say about pod b afterwards (F); the exact plan follows
                    // once the word is on its way
                    uint64_t v = lane < gnw ? s_red[p & 3][lane][0] : 0;
                    v = warp_max_u64(v);
                    mine_key = v;
                    const bool have = (v & KEY_FEASIBLE) != 0;
                    my_r = have ? key_node(v) - node0 : 0u;
                    int f_state = F_NO, f_score = 0;
                    if (have && two) f_certificate(S, my_r, dem_a, dem_b, lut, f_state, f_score);
                    const uint64_t pk = pack_a(v, f_state, f_score);
                    PHASE(2);
                    if (solo) {
                        if (lane == 0) s_loc[0] = pk;
                    } else if (lane == 0) {
                        unsigned long long old = 0;
                        if (pk) old = atomicMax(rr + 0, (unsigned long long)pk);
                        const unsigned long long one = 1ull + (old & dep_mask);  // (always 1; issued once the atomic has returned)
                        asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(rr + 3), "l"(one) : "memory");
                    }
                    TRACE(0);
                    if (have) plan_group(bp, S, my_r, node0 + my_r, dem_a, nxa, lut, g, gbase, cm);
                    TRACE(1);
                    PHASE(3);
                } else {
                    // pod b: this CTA's best two keys on the rows as they stand; x1 / x2 become the grid's best two
                    uint64_t k1 = 0, k2 = 0;
                    GroupCommit& cmb = cm;
                    if (two) {
                        const uint64_t e0 = lane < gnw ? s_red[(p + 1) & 3][lane][0] : 0, e1 = lane < gnw ? s_red[(p + 1) & 3][lane][1] : 0;
                        warp_top2(e0, e1, k1, k2);
                    }
                    mine_key = k1;
                    const bool have = (k1 & KEY_FEASIBLE) != 0;
                    my_r = have ? key_node(k1) - node0 : 0u;
                    unsigned long long old = 0;
                    if (two && !solo && lane == 0 && k1) old = atomicMax(rr + 1, (unsigned long long)k1);
                    PHASE(2);
                    if (two && solo) {
                        if (lane == 0) {
                            s_loc[1] = k1;
                            s_loc[2] = k2;
                        }
                    } else if (two && lane == 0) {
                        const unsigned long long lost = old < k1 ? old : k1;  // what this CTA's arrival pushed out of first place
                        const unsigned long long second = lost > k2 ? lost : k2;
                        unsigned long long old2 = 0;
                        if (second) old2 = atomicMax(rr + 2, second);
                        const unsigned long long one = 1ull + ((old | old2) & dep_mask);
                        asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(rr + 3), "l"(one) : "memory");
                    }
                    TRACE(4);
                    // while the exchange is in flight: the exact plan of k1's row for pod b (used if the row wins as it stands)
                    if (two && have) plan_group(bp, S, my_r, node0 + my_r, dem_b, nx, lut, g, gbase, cmb);
                    PHASE(3);
                    // ---- the exchange: wait for every CTA's words, fold in the other ranks', resolve the round ----
                    unsigned long long w[3] = {0, 0, 0};
                    if (solo) {
                        asm volatile("bar.sync 2, 64;" ::: "memory");  // control warp 0 has left its word
                        w[0] = s_loc[0];
                        w[1] = s_loc[1];
                        w[2] = s_loc[2];
                    } else if (bp.world == 1 || blockIdx.x == 0) {
                        if (lane == 0) dead = !wait_round(rr, gridDim.x * L, bp.err, w);
                        w[0] = __shfl_sync(0xFFFFFFFFu, w[0], 0);
                        w[1] = __shfl_sync(0xFFFFFFFFu, w[1], 0);
                        w[2] = __shfl_sync(0xFFFFFFFFu, w[2], 0);
                        dead = __any_sync(0xFFFFFFFFu, dead);
                    }
                    uint64_t xa = w[0], x1 = w[1], x2 = w[2];
                    int owner_rank = (int)bp.rank;
                    if (bp.world > 1) {
                        // K5: CTA 0 posts the rank's words to every rank (its own array included); every CTA reads the
                        // ranks' words from its rank's array.  pod a: the best word (its F rides along); pod b: best two of the union.
                        // (xa always has bit 63 set when there is a key, so "no key" travels as 1 and VALID is "non-zero")
                        if (blockIdx.x == 0) {
                            if (lane < bp.world) st_relaxed_sys_u64(xslot(bp.xslots[lane], p, bp.rank, 0), xa ? xa : 1ull);
                            if (two) {
                                post_ranks(bp, p + 1, 0, x1, lane);
                                post_ranks(bp, p + 1, 1, x2, lane);
                            }
                        }
                        unsigned long long rxa = 0;
                        if (lane < bp.world) rxa = wait_nonzero(xslot(bp.xslots[bp.rank], p, lane, 0), bp.err, dead);
                        dead = __any_sync(0xFFFFFFFFu, dead);
                        xa = warp_max_u64(rxa);
                        const uint32_t wl = __ballot_sync(0xFFFFFFFFu, lane < bp.world && rxa == xa);
                        owner_rank = wl ? __ffs(wl) - 1 : 0;
                        if (two) {
                            const unsigned long long r1 = collect_ranks(bp, p + 1, 0, lane, dead);
                            const unsigned long long r2 = collect_ranks(bp, p + 1, 1, lane, dead);
                            warp_top2(r1, r2, x1, x2);
                        }
                        if (dead) xa = x1 = x2 = 0;
                    }
                    // pod b's winner: binding pod a changed its winner's row r* and nothing else
                    const uint64_t ka = unpack_a_key(xa);
                    uint64_t kb = 0, y = 0;
                    int b_on_a = 0, need_ann = 0;
                    if (two) {
                        const bool a_ok = (ka & KEY_FEASIBLE) != 0;
                        y = (a_ok && (x1 & KEY_FEASIBLE) && key_node(x1) == key_node(ka)) ? x2 : x1;  // best over all other rows
                        const int f_state = a_ok ? unpack_a_fstate(xa) : F_NO;
                        const uint64_t fk = f_state != F_NO ? make_key(unpack_a_fscore(xa), key_node(ka)) : 0;  // r* after the bind (F_MAYBE: at best)
                        kb = y;
                        if (f_state == F_YES && fk > y) {
                            kb = fk;
                            b_on_a = 1;
                        } else if (f_state == F_MAYBE && fk > y) {
                            need_ann = 1;  // the owner of r* has to say whether, and with which score, the row takes pod b
                        }
                    }
                    if (two) {
                        EVENT(0, 1);
                        EVENT(1, need_ann);
                        EVENT(8 + ((ka & KEY_FEASIBLE) ? unpack_a_fstate(xa) : 3), 1);
                    }
                    if (lane == 0) {
                        s_res.ka = ka;
                        s_res.kb = kb;
                        s_res.y = y;
                        s_res.b_on_a = b_on_a;
                        s_res.need_ann = need_ann;
                        s_res.owner_rank = owner_rank;
                        if (dead) s_abort = 1;
                    }
                    TRACE(5);
                    PHASE(4);
                }
                if (solo && cw == 0) asm volatile("bar.sync 2, 64;" ::: "memory");
                asm volatile("bar.sync 1, 64;" ::: "memory");  // the two control warps: the exchange is resolved
                if (s_res.need_ann) {  // (uniform over both warps)
                    const uint64_t ka1 = s_res.ka;
                    if (cw == 0) {
                        if ((mine_key & KEY_FEASIBLE) && ka1 == mine_key && lane == 0) {
                            // this CTA owns pod a's winner and its plan is done: say what the row offers pod b afterwards
                            const unsigned long long word = KEY_VALID | ((unsigned long long)(cm.tok & 1u) << 32) |
                                                            (uint32_t)(dem_b->policy == NGS_POLICY_SPREAD ? cm.scS : cm.scB);
                            if (solo) s_loc[0] = word;
                            else if (bp.world == 1) st_relaxed_u64(rr + 4, word);
                            else
                                for (uint32_t j = 0; j < bp.world; j++) st_relaxed_sys_u64(xslot(bp.xslots[j], p, bp.rank, 1), word);
                        }
                        if (solo) asm volatile("bar.sync 2, 64;" ::: "memory");
                    } else {
                        unsigned long long word = 0;
                        if (solo) {
                            asm volatile("bar.sync 2, 64;" ::: "memory");
                            word = s_loc[0];
                        } else if (lane == 0) {
                            word = bp.world == 1 ? wait_valid_gpu(rr + 4, bp.err, dead)
                                                 : wait_valid(xslot(bp.xslots[bp.rank], p, (uint32_t)s_res.owner_rank, 1), bp.err, dead);
                        }
                        word = __shfl_sync(0xFFFFFFFFu, word, 0);
                        dead = __any_sync(0xFFFFFFFFu, dead);
                        const bool f_ok = (word >> 32) & 1u;
                        const uint64_t fk = f_ok ? make_key((int)(uint32_t)word, key_node(ka1)) : 0;
                        const uint64_t y = s_res.y;
                        if (lane == 0) {
                            s_res.kb = fk > y ? fk : y;
                            s_res.b_on_a = fk > y ? 1 : 0;
                            if (dead) s_abort = 1;
                        }
                    }
                    asm volatile("bar.sync 3, 64;" ::: "memory");  // the announcement has settled pod b
                }
                PHASE(5);
                const uint64_t ka = s_res.ka, kb = s_res.kb;
                const bool b_on_a = s_res.b_on_a != 0;
#ifdef NGS_PHASE_TIMING
                const long long t_bind = clock64();
                bool owner_evt = false;
#endif
                // ---- bind: the owner of a winner publishes it (stores only where the plan was prepared) ----
                auto report_elsewhere = [&](uint32_t pod, const SortedDemand* dem, uint64_t k) {  // CTA 0, lane 0
                    if (!(k & KEY_FEASIBLE)) {
                        publish_none(bp, pod, dem);
                    } else if (key_node(k) - T.first >= T.n) {
                        // the winner lives on another rank: node and score are known from the key, the
                        // plan is filled in by the owner (merged on the host side by an allreduce-max)
                        publish_none(bp, pod, dem);
                        bp.node_out[pod] = (int32_t)key_node(k);
                        bp.score_out[pod] = key_score(k);
                    }
                };
                if (cw == 0) {
                    if (blockIdx.x == 0 && lane == 0) {
                        report_elsewhere(p, dem_a, ka);
                        if (two && b_on_a) report_elsewhere(p + 1, dem_b, kb);
                    }
                    int dirty = -1;
                    uint32_t vk = cm.tok >> sh_a;  // the next round's verdicts of the row this warp binds
                    if ((mine_key & KEY_FEASIBLE) && ka == mine_key) {  // (warp-uniform) this CTA owns pod a's winner
                        publish_group(bp, S, my_r, row0 + my_r, dem_a, p, key_score(ka), cm, lane);
                        dirty = (int)my_r;
#ifdef NGS_PHASE_TIMING
                        owner_evt = true;
                        EVENT(3, 1);
                        if (two && b_on_a) EVENT(2, 1);
#endif
                        if (two && b_on_a) {
                            // pod b goes to the very row pod a has just been bound to: nobody could prepare that plan
                            __syncwarp();
                            GroupCommit c2;
                            plan_group(bp, S, my_r, node0 + my_r, dem_b, nx, lut, g, gbase, c2);
                            publish_group(bp, S, my_r, row0 + my_r, dem_b, p + 1, key_score(kb), c2, lane);
                            vk = c2.tok;
                        }
                    }
                    if (two && b_on_a) {  // both binds (if any) hit one row: one entry, with the verdicts of the final row
                        leave_dirty(1, dirty, vk & 1u, (vk >> 1) & 1u);
                        if (lane == 0) s_dirty_row[0] = -1;
                    } else {
                        leave_dirty(0, dirty, vk & 1u, (vk >> 1) & 1u);
                    }
                } else {
                    if (two && !b_on_a) {
                        if (blockIdx.x == 0 && lane == 0) report_elsewhere(p + 1, dem_b, kb);
                        int dirty = -1;
                        const uint32_t rb = key_node(kb) - node0;  // (garbage unless kb is feasible and ours)
                        if ((kb & KEY_FEASIBLE) && key_node(kb) >= node0 && rb < my_rows) {  // (warp-uniform) this CTA owns pod b's winner
#ifdef NGS_PHASE_TIMING
                            owner_evt = true;
                            EVENT(5, 1);
                            if (kb != mine_key) EVENT(7, 1);
#endif
                            if (kb != mine_key)  // not the row this warp prepared (the grid's second best was needed): plan it now
                                plan_group(bp, S, rb, node0 + rb, dem_b, nx, lut, g, gbase, cm);
                            publish_group(bp, S, rb, row0 + rb, dem_b, p + 1, key_score(kb), cm, lane);
                            dirty = (int)rb;
                        }
                        leave_dirty(1, dirty, cm.tok & 1u, (cm.tok >> 1) & 1u);
                    } else if (!two) {
                        if (lane == 0) s_dirty_row[1] = -1;
                    }
                }
#ifdef NGS_PHASE_TIMING
                if (owner_evt) EVENT(cw == 0 ? 4 : 6, clock64() - t_bind);
#endif
                if (cw == 0) TRACE(2); else TRACE(6);
                PHASE(6);
            }
            if (cw == 1) {
#pragma unroll
                for (uint32_t j = 0; j < 2; j++) {
                    const uint32_t ps = p + 6 + j;
                    if (j < L && ps < P) {
                        uint32_t* dst = reinterpret_cast<uint32_t*>(&s_dem[ps & (DEM_RING - 1)]);
                        dst[lane] = stg[j][0];
                        if (lane + 32 < DEM_WORDS) dst[lane + 32] = stg[j][1];
                    }
                }
            }
        }
        __syncthreads();  // the round is settled; the workers' results for the next round are in s_red / feasbits
        if (s_abort) break;  // a peer never posted: give up instead of hanging the GPU
        p = q;
#ifdef NGS_PHASE_TIMING
        rnd++;
#endif
    }
    if (blockIdx.x == 0 && tid == 0 && bp.status_out) *bp.status_out = *reinterpret_cast<volatile int32_t*>(bp.err);
#ifdef NGS_PHASE_TIMING
    if (tid == NWARP * 32 && bp.dbg)
        for (int i = 0; i < 8; i++) bp.dbg[blockIdx.x * 16 + i] = (unsigned long long)ph[i];
    if (tid == NWARP * 32 + 32 && bp.dbg)
        for (int i = 0; i < 8; i++) bp.dbg[blockIdx.x * 16 + 8 + i] = (unsigned long long)ph[i];
    if (is_worker && lane == 0 && bp.dbg) bp.dbg[4096 + blockIdx.x * 32 + ww] = (unsigned long long)ph[0];  // filter time of every worker warp
    if (is_control && lane == 0 && bp.dbg)
        for (int i = 0; i < 16; i++)
            if (ev[i]) atomicAdd(bp.dbg + 20000 + i, (unsigned long long)ev[i]);
#endif
}
