// ngs_device.cuh — device-side building blocks of the dealer kernels (sm_100a).
//
// Everything here is integer / fp64-scalar work on one node row held in registers; the
// kernels in ngs_dealer.cu decide how rows reach the registers (int4 global loads for the
// per-verb scans, shared-memory resident rows for the persistent batch kernel).
//
// Reference semantics (paths relative to /root/reference/pkg/dealer/):
//   GPUs.Choose            allocate.go:92-100   Rate on the pre-placement state, then Choose
//   Binpack/Spread.Choose  rater.go:74-110,127-163
//   SampleRater.Choose     rater.go:28-50 (scratch copy; SURVEY §8 a9')
//   Binpack/Spread.Rate    rater.go:59-70,113-123
//   GPUResource.LoadUsage  allocate.go:173-195 (+ nodeusage.go:82-111, stats.go:30-55)
//   sort.Sort              Go 1.16 standard library (go.mod:3): for <= 12 elements one gap-6
//                          shell pass, then insertion sort (stable).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/ngs_dealer.h"

namespace ngs {

// ---------------------------------------------------------------------------------------------
// Packed GPU cell.  One 32-bit word per cell so that Go's "re-sort, then first fit" becomes
// min/max compare-exchanges on plain unsigned integers:
//   [31:9]  key + 128, key = free + 50*RemainLoad        (SortableGPUs.Less, allocate.go:247)
//   [8:6]   position in the current arrangement           (makes equal keys keep their order)
//   [5:3]   RemainLoad + 2                                (RemainLoad in [-2, 2])
//   [2:0]   original GPU index                            (GPUResourceWithIndex.index)
// Cells beyond the node's GPU count are PAD (all ones): they sort last and stay at
// positions >= G through the gap pass, so a position mask excludes them from every fit test.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t W_KEY_SHIFT = 9;
constexpr uint32_t W_POS_MASK = 7u << 6;
constexpr uint32_t W_PAD = 0xFFFFFFFFu;
constexpr int KEY_BIAS = 128;
constexpr uint32_t RL0_META = 0x492492u;  // eight 3-bit RemainLoad codes, all == 2 (RemainLoad 0)

__device__ __forceinline__ uint32_t pack_cell(int free, int rlc, int g) {
    // 50*(rlc-2) + 128 == 50*rlc + 28
    return ((uint32_t)(free + 50 * rlc + 28) << W_KEY_SHIFT) | ((uint32_t)rlc << 3) | (uint32_t)(g * 65);
}
__device__ __forceinline__ int cell_free(uint32_t w) { return (int)(w >> W_KEY_SHIFT) - 28 - 50 * (int)((w >> 3) & 7u); }

// Sorted demand of one pod, staged in shared memory (BASELINE north_star: "staging each pod's
// demand vector in shared memory").  d[] is the processing order of rater.go:82 (largest
// first) restricted to non-zero percents; order[k] is the container the k-th processed entry
// belongs to (zeros, which emit NotNeedGPU, follow the non-zeros).
struct SortedDemand {
    int32_t C;       // containers
    int32_t nz;      // containers with percent != 0
    int32_t policy;  // NGS_POLICY_*
    int32_t flags;   // bit 0: the pod failed validation (k_prep_demands); bits 1..: sum of the percents
    int32_t raw[NGS_MAX_CONTAINERS];    // container order (SampleRater walks this)
    int32_t d[NGS_MAX_CONTAINERS];      // processing order, non-zero entries first
    int32_t order[NGS_MAX_CONTAINERS];  // container index of each processed entry
};
static_assert(sizeof(SortedDemand) == 160, "SortedDemand layout");

// By-value pod demand for the per-verb kernels.
struct PodDemand {
    int32_t n;
    int32_t policy;
    int32_t percent[NGS_MAX_CONTAINERS];
};

// isLoadSchedule context resolved on the host: a timestamp is inside the update window
// (stats.go:48-53: now.Before(t + activeDuration)) iff ts_sec >= min_ts[e].
struct LoadParams {
    int32_t enabled;  // isLoadSchedule
    int32_t n_periods;
    int32_t metric[NGS_MAX_PERIODS];
    uint32_t min_ts[NGS_MAX_PERIODS];
    uint32_t never[NGS_MAX_PERIODS];  // window lies beyond the representable range: nothing valid
    int32_t canonical;                // entries are exactly (core, memory), both usable: unrolled fast path
};

// sort.Sort(sortableDemand) (rater.go:81,134) with the Go 1.16 small-slice rule, then the
// traversal order of rater.go:82.  Serial, executed by one thread per block (or by the host, for the resident kernel).
__host__ __device__ inline void sort_demand(const int32_t* percent, int C, int policy, SortedDemand* out) {
    int key[NGS_MAX_CONTAINERS], idx[NGS_MAX_CONTAINERS];
    for (int i = 0; i < NGS_MAX_CONTAINERS; i++) {
        key[i] = i < C ? percent[i] : 0;
        idx[i] = i;
    }
    for (int i = 6; i < C; i++)  // gap-6 shell pass
        if (key[i] < key[i - 6]) {
            int t = key[i]; key[i] = key[i - 6]; key[i - 6] = t;
            t = idx[i]; idx[i] = idx[i - 6]; idx[i - 6] = t;
        }
    for (int i = 1; i < C; i++)  // insertionSort
        for (int j = i; j > 0 && key[j] < key[j - 1]; j--) {
            int t = key[j]; key[j] = key[j - 1]; key[j - 1] = t;
            t = idx[j]; idx[j] = idx[j - 1]; idx[j - 1] = t;
        }
    out->C = C;
    out->policy = policy;
    uint32_t sum = 0;
    int nz = 0;
    for (int k = 0; k < NGS_MAX_CONTAINERS; k++) {
        int j = C - 1 - k;  // rater.go:82: j runs from the largest demand down
        out->raw[k] = k < C ? percent[k] : 0;
        sum += (uint32_t)key[k];  // (entries beyond C are 0)
        out->d[k] = j >= 0 ? key[j] : 0;
        out->order[k] = j >= 0 ? idx[j] : 0;
        if (j >= 0 && key[j] != 0) nz++;
    }
    out->nz = nz;  // percents are >= 0 (checked at the boundary), so non-zeros come first
    out->flags = (int32_t)(sum << 1);
}

__device__ __forceinline__ void cex(uint32_t& a, uint32_t& b) {
    uint32_t lo = min(a, b), hi = max(a, b);
    a = lo;
    b = hi;
}
// gap-6 step of Go 1.16 quickSort: if Less(i, i-6) then Swap — compares keys only
__device__ __forceinline__ void gap_swap(uint32_t& hi_pos, uint32_t& lo_pos) {
    bool sw = (hi_pos >> W_KEY_SHIFT) < (lo_pos >> W_KEY_SHIFT);
    uint32_t a = sw ? hi_pos : lo_pos, b = sw ? lo_pos : hi_pos;
    lo_pos = a;
    hi_pos = b;
}

// sort.Sort(sortableGpus) (rater.go:87,140) on the current arrangement: gap-6 pass, stamp
// positions, then a 19-comparator network.  Positions make all words distinct, so the network
// reproduces the (stable) insertion sort exactly.
__device__ __forceinline__ void go_sort8(uint32_t (&w)[8]) {
    gap_swap(w[6], w[0]);
    gap_swap(w[7], w[1]);
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = (w[i] & ~W_POS_MASK) | ((uint32_t)i << 6);
    cex(w[0], w[2]); cex(w[1], w[3]); cex(w[4], w[6]); cex(w[5], w[7]);
    cex(w[0], w[4]); cex(w[1], w[5]); cex(w[2], w[6]); cex(w[3], w[7]);
    cex(w[0], w[1]); cex(w[2], w[3]); cex(w[4], w[5]); cex(w[6], w[7]);
    cex(w[2], w[4]); cex(w[3], w[5]);
    cex(w[1], w[4]); cex(w[3], w[6]);
    cex(w[1], w[2]); cex(w[3], w[4]); cex(w[5], w[6]);
}

// bit i set iff the cell at position i can take `d` (GPUResource.CanAllocate, allocate.go:159)
__device__ __forceinline__ uint32_t fit_mask(const uint32_t (&w)[8], int d, uint32_t gmask, bool rl0) {
    uint32_t m = 0;
    if (rl0) {
        const uint32_t thr = (uint32_t)(d + KEY_BIAS) << W_KEY_SHIFT;
#pragma unroll
        for (int i = 0; i < 8; i++) m |= (uint32_t)(w[i] >= thr) << i;
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) m |= (uint32_t)(cell_free(w[i]) >= d) << i;
    }
    return m & gmask;
}

// Exact shortcuts for the feasibility question (rater.go:98-100), valid for every policy because
// they only use "each container goes to SOME cell that fits it" (first fit in whatever order):
//   * no cell holds the largest demand d1           -> that container can never be placed: NO
//   * at least nz cells hold d1 (nz = non-zero containers) -> when the j-th container is placed at
//     most j-1 cells have been touched, so an untouched cell with free >= d1 >= dj remains: YES
//   * sorted policies, nz == 2, exactly one cell c holds d1: d1 is forced onto c, so the pod fits
//     iff c still holds d2 afterwards or a second cell holds d2.
// Everything else runs the exact greedy search.  Cells beyond the GPU count hold free = 0 and
// demands here are > 0, so they never count.
enum { FEAS_NO = 0, FEAS_YES = 1, FEAS_UNDECIDED = 2 };
__device__ __forceinline__ int classify_row(const int (&f)[8], const SortedDemand* dem, bool sorted_policy) {
    const int nz = dem->nz;
    if (nz == 0) return FEAS_YES;
    const int d1 = dem->d[0];
    int cnt = 0, fmax = 0;
#pragma unroll
    for (int g = 0; g < 8; g++) {
        cnt += f[g] >= d1 ? 1 : 0;
        fmax = max(fmax, f[g]);
    }
    if (cnt == 0) return FEAS_NO;
    if (cnt >= nz) return FEAS_YES;
    if (sorted_policy && nz == 2) {  // cnt == 1: the cell holding d1 is the row maximum
        const int d2 = dem->d[1];
        int cnt2 = 0;
#pragma unroll
        for (int g = 0; g < 8; g++) cnt2 += f[g] >= d2 ? 1 : 0;
        return (cnt2 >= 2 || fmax - d1 >= d2) ? FEAS_YES : FEAS_NO;
    }
    return FEAS_UNDECIDED;
}

// The same three statements read off the row's four LARGEST free values, kept beside the row as
// one word (byte k = k-th largest, saturated at 255; refreshed whenever the row changes):
// "at least nz cells hold d1" is "the nz-th largest holds d1", "no cell holds d1" is "the largest
// does not".  Needs nz <= 4 and d1 <= 255, and a non-saturated maximum for the nz == 2 case;
// FEAS_FALLBACK asks the caller to classify from the full row instead.
constexpr int FEAS_FALLBACK = 3;
__device__ __forceinline__ uint32_t top4_of(const int (&f)[8]) {
    int a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = f[i];
#define NGS_CEX_DESC(i, j) { const int hi_ = max(a[i], a[j]), lo_ = min(a[i], a[j]); a[i] = hi_; a[j] = lo_; }
    NGS_CEX_DESC(0, 2) NGS_CEX_DESC(1, 3) NGS_CEX_DESC(4, 6) NGS_CEX_DESC(5, 7)
    NGS_CEX_DESC(0, 4) NGS_CEX_DESC(1, 5) NGS_CEX_DESC(2, 6) NGS_CEX_DESC(3, 7)
    NGS_CEX_DESC(0, 1) NGS_CEX_DESC(2, 3) NGS_CEX_DESC(4, 5) NGS_CEX_DESC(6, 7)
    NGS_CEX_DESC(2, 4) NGS_CEX_DESC(3, 5)
    NGS_CEX_DESC(1, 4) NGS_CEX_DESC(3, 6)
    NGS_CEX_DESC(1, 2) NGS_CEX_DESC(3, 4) NGS_CEX_DESC(5, 6)
#undef NGS_CEX_DESC
    return (uint32_t)min(a[0], 255) | ((uint32_t)min(a[1], 255) << 8) | ((uint32_t)min(a[2], 255) << 16) | ((uint32_t)min(a[3], 255) << 24);
}
__device__ __forceinline__ int classify_top4(uint32_t top4, const SortedDemand* dem, bool sorted_policy) {
    const int nz = dem->nz;
    if (nz == 0) return FEAS_YES;
    const int d1 = dem->d[0];
    if (nz > 4 || d1 > 255) return FEAS_FALLBACK;
    const int t0 = (int)(top4 & 0xFFu);
    if (t0 < d1) return FEAS_NO;  // (a saturated 255 is >= d1, which is what the true value is, too)
    if ((int)((top4 >> (8 * (nz - 1))) & 0xFFu) >= d1) return FEAS_YES;
    if (sorted_policy && nz == 2) {  // exactly one cell holds d1: it is the row maximum
        if (t0 == 255) return FEAS_FALLBACK;
        const int d2 = dem->d[1];
        return ((int)((top4 >> 8) & 0xFFu) >= d2 || t0 - d1 >= d2) ? FEAS_YES : FEAS_NO;
    }
    return FEAS_UNDECIDED;
}

// Binpack.Choose / Spread.Choose on the packed row.  Returns feasibility (rater.go:98-100).
// PLAN: `emit` receives the `indexes` slice in emission order, 4 bits per entry (0xF = -1),
// and n_emit its length; the greedy loop then runs to the end even after a miss, exactly as
// the reference does.  Without PLAN only feasibility is wanted: a miss returns at once and the
// last non-zero container needs no sort (any cell that fits will do).
template <bool PLAN>
__device__ __forceinline__ bool choose_sorted(uint32_t (&w)[8], int G, bool spread, bool rl0, const SortedDemand* dem,
                                              uint64_t* emit, int* n_emit, int (*placed)[8] = nullptr) {
    const uint32_t gmask = (1u << G) - 1u;
    const int nz = dem->nz;
    uint64_t em = 0;
    int ne = 0;
    bool ok = true;
    for (int k = 0; k < nz; k++) {
        const int d = dem->d[k];
        if (!PLAN && k == nz - 1) return fit_mask(w, d, gmask, rl0) != 0;
        go_sort8(w);
        const uint32_t m = fit_mask(w, d, gmask, rl0);
        if (m == 0) {
            if (!PLAN) return false;
            ok = false;
            continue;  // rater.go:88-95: nothing appended, later containers still placed
        }
        const int sel = spread ? 31 - __clz(m) : __ffs(m) - 1;  // rater.go:141 / :88
        const uint32_t dec = (uint32_t)d << W_KEY_SHIFT;
        uint32_t chosen = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const bool hit = sel == i;
            if (PLAN && hit) chosen = w[i] & 7u;
            w[i] -= hit ? dec : 0u;
        }
        if (PLAN) {
            em |= (uint64_t)chosen << (4 * ne);
            ne++;
            if (placed) {  // percent taken from each GPU, by original index (GPUs.Allocate's effect)
#pragma unroll
                for (int g = 0; g < 8; g++) (*placed)[g] += (chosen == (uint32_t)g) ? d : 0;
            }
        }
    }
    if (PLAN) {
        for (int k = nz; k < dem->C; k++) {  // zero-percent containers: NotNeedGPU (rater.go:83-86)
            em |= (uint64_t)0xF << (4 * ne);
            ne++;
        }
        *emit = em;
        *n_emit = ne;
    }
    return ok;
}

// ---------------------------------------------------------------------------------------------
// Binpack.Choose / Spread.Choose with one GPU cell per lane (8 lanes = one node, 4 nodes per
// warp).  The O(G^2) comparisons of the re-sort are spread over the G lanes: each lane counts the
// cells that sort before its own (7 shuffles), which is exactly the position the stable
// insertion sort gives it.  `pos` persists between containers, as the slice does in the
// reference.  All 32 lanes of the warp must call this together (the four groups may look at four
// different rows); every warp-synchronous op runs with the full mask and only ever exchanges
// data inside the caller's 8-lane group, and no such op sits under divergent control flow.
//   g      cell index (lane & 7)        gbase   first lane of the group (lane & 24)
//   fr     free percent of the cell (cells >= G: anything)   rlc   RemainLoad + 2
// Returns feasibility; with PLAN the emission list (4 bits per entry, GPU index, 0xF =
// NotNeedGPU) and, in *fr_out, the cell's free percent after GPUs.Allocate.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t FULL = 0xFFFFFFFFu;
template <bool PLAN>
__device__ __forceinline__ bool choose_group8(int g, int gbase, int fr, int rlc, bool valid, bool spread, const SortedDemand* dem,
                                              uint64_t* emit, int* n_emit, int* fr_out) {
    uint32_t key = valid ? (uint32_t)(fr + 50 * rlc + 28) : 0x0FFFFFFFu;  // biased key; padding sorts last
    if (!valid) fr = -1;                                                   // and never fits
    int pos = g;
    const int nz = dem->nz;
    bool ok = true;
    uint64_t em = 0;
    int ne = 0;
    for (int k = 0; k < nz; k++) {
        const int d = dem->d[k];
        if (!PLAN && k == nz - 1) {  // last container, feasibility only: any cell that fits will do
            // (the ballot must not sit behind the short-circuit of `ok &&`: every lane has to take part)
            const uint32_t anyfit = (__ballot_sync(FULL, fr >= d) >> gbase) & 0xFFu;
            ok = ok && anyfit != 0u;
            break;
        }
        // gap-6 pass of Go 1.16 sort.Sort on the current arrangement: (6,0) then (7,1)
        const uint32_t b0 = (__ballot_sync(FULL, pos == 0) >> gbase) & 0xFFu;
        const uint32_t b1 = (__ballot_sync(FULL, pos == 1) >> gbase) & 0xFFu;
        const uint32_t b6 = (__ballot_sync(FULL, pos == 6) >> gbase) & 0xFFu;
        const uint32_t b7 = (__ballot_sync(FULL, pos == 7) >> gbase) & 0xFFu;
        const uint32_t k0 = __shfl_sync(FULL, key, gbase + __ffs(b0) - 1);
        const uint32_t k1 = __shfl_sync(FULL, key, gbase + __ffs(b1) - 1);
        const uint32_t k6 = __shfl_sync(FULL, key, gbase + __ffs(b6) - 1);
        const uint32_t k7 = __shfl_sync(FULL, key, gbase + __ffs(b7) - 1);
        const bool s06 = k6 < k0, s17 = k7 < k1;
        int np = pos;
        np = (s06 && pos == 0) ? 6 : np;
        np = (s06 && pos == 6) ? 0 : np;
        np = (s17 && pos == 1) ? 7 : np;
        np = (s17 && pos == 7) ? 1 : np;
        // insertion sort (stable): new position = number of cells ordered before this one
        const uint32_t c = (key << 3) | (uint32_t)np;
        int rank = 0;
#pragma unroll
        for (int sft = 1; sft < 8; sft++) rank += __shfl_xor_sync(FULL, c, sft) < c ? 1 : 0;
        pos = rank;
        // first fit from the low end (binpack, rater.go:88) or the high end (spread, rater.go:141):
        // min (max) position over the cells that fit, by a 3-step butterfly inside the group
        const bool fit = fr >= d;
        int v = fit ? (spread ? pos : 7 - pos) : -1;  // larger is better in both cases
        v = max(v, __shfl_xor_sync(FULL, v, 1));
        v = max(v, __shfl_xor_sync(FULL, v, 2));
        v = max(v, __shfl_xor_sync(FULL, v, 4));
        const int sel = spread ? v : 7 - v;  // 8 / -1 when nothing fits: matches no position
        const bool chosen = fit && pos == sel;
        ok = ok && v >= 0;
        fr -= chosen ? d : 0;
        key -= chosen ? (uint32_t)d : 0u;
        if (PLAN) {
            const uint32_t cb = (__ballot_sync(FULL, chosen) >> gbase) & 0xFFu;
            if (cb) {
                em |= (uint64_t)(__ffs(cb) - 1) << (4 * ne);
                ne++;
            }
        }
    }
    if (PLAN) {
        for (int k = nz; k < dem->C; k++) {  // zero-percent containers: NotNeedGPU (rater.go:83-86)
            em |= (uint64_t)0xF << (4 * ne);
            ne++;
        }
        *emit = em;
        *n_emit = ne;
        *fr_out = fr;
    }
    return ok;
}

// ---------------------------------------------------------------------------------------------
// The greedy search without positions: what it does to the row as a set of cells, and whether it
// succeeds - not which GPU index each container gets.  The order Go's sort leaves among cells of
// DIFFERENT keys is the key order (the sort is correct, only not stable), so "first fit from the low
// end" is "the fitting cell of the smallest key" (spread: of the largest).  Among cells of equal key
// the reference goes by position; if those cells also hold the same free percent (hence the same
// RemainLoad) they are interchangeable and any of them leaves the same row-as-a-set, so the outcome
// is still exact.  Only a tie between equal keys with DIFFERENT free percents (free 60 / RemainLoad 0
// against free 10 / RemainLoad 1) needs the positions: then bit 1 of the result is set and the caller
// has to run the positional search (choose_group8).  A third of the shuffles and votes of
// choose_group8 per container, and that is what the control path of the batch kernel waits for.
//   APPLY: every container is placed and fr holds the cell as GPUs.Allocate would leave it;
//   otherwise feasibility only (the last container only has to fit somewhere).
// Returns bit 0: every non-zero container found a cell; bit 1: ambiguous tie met (result void).
// One row per 8-lane group, all 32 lanes together, as choose_group8.
// ---------------------------------------------------------------------------------------------
template <bool APPLY>
__device__ __forceinline__ int greedy_group8(int g, int gbase, int& fr, int rlc, bool valid, bool spread, const SortedDemand* dem) {
    const int nz = dem->nz;
    bool ok = true, amb = false;
    for (int k = 0; k < nz; k++) {
        const int d = dem->d[k];
        const bool fit = valid && fr >= d;
        if (!APPLY && k == nz - 1) {
            const uint32_t anyfit = (__ballot_sync(FULL, fit) >> gbase) & 0xFFu;
            ok = ok && anyfit != 0u;
            break;
        }
        const uint32_t key = (uint32_t)(fr + 50 * rlc);
        const uint32_t mine = fit ? (spread ? key + 1u : 0x7FFFFFFFu - key) : 0u;  // larger is better; 0 = does not fit
        uint32_t v = mine;
        v = max(v, __shfl_xor_sync(FULL, v, 1));
        v = max(v, __shfl_xor_sync(FULL, v, 2));
        v = max(v, __shfl_xor_sync(FULL, v, 4));
        const bool cand = fit && mine == v;
        const uint32_t cb = (__ballot_sync(FULL, cand) >> gbase) & 0xFFu;
        const int first = cb ? __ffs(cb) - 1 : 0;
        const int f0 = __shfl_sync(FULL, fr, gbase + first);
        const uint32_t differ = (__ballot_sync(FULL, cand && fr != f0) >> gbase) & 0xFFu;
        amb = amb || differ != 0u;
        ok = ok && cb != 0u;  // rater.go:88-95: nothing fits, nothing appended; later containers are still placed
        if (cb != 0u && g == first) fr -= d;
    }
    return (ok ? 1 : 0) | (amb ? 2 : 0);
}

// The same search, feasibility only, for FOUR (row, pod) pairs at once: every 8-lane group brings its own row and its
// own demand (dem == nullptr: the group has nothing to ask and gets false).  Policies may differ between the groups.
// The loop runs to the largest container count among the groups; a group that is through keeps executing the
// warp-synchronous ops with a demand of zero, which places nothing.  One call costs what one search costs, so the
// verdicts a bind owes to several waiting pods come at the price of one.  All 32 lanes must call this together.
__device__ __forceinline__ bool choose_group8_multi(int g, int gbase, int fr, int rlc, bool valid, const SortedDemand* dem) {
    uint32_t key = valid ? (uint32_t)(fr + 50 * rlc + 28) : 0x0FFFFFFFu;  // biased key; padding sorts last
    if (!valid) fr = -1;                                                   // and never fits
    int pos = g;
    const int nz = dem != nullptr ? dem->nz : 0;
    const bool spread = dem != nullptr && dem->policy == NGS_POLICY_SPREAD;
    const int nzmax = (int)__reduce_max_sync(FULL, (unsigned)nz);
    bool ok = dem != nullptr;
    for (int k = 0; k < nzmax; k++) {
        const bool active = k < nz;
        const int d = active ? dem->d[k] : 0;
        // (this group's last container only has to fit somewhere: what the vote below answers)
        const uint32_t anyfit = (__ballot_sync(FULL, fr >= d) >> gbase) & 0xFFu;
        if (active && k == nz - 1) ok = ok && anyfit != 0u;
        const bool place = active && k < nz - 1;  // earlier containers are really placed: sort, first fit, subtract
        const uint32_t b0 = (__ballot_sync(FULL, pos == 0) >> gbase) & 0xFFu;
        const uint32_t b1 = (__ballot_sync(FULL, pos == 1) >> gbase) & 0xFFu;
        const uint32_t b6 = (__ballot_sync(FULL, pos == 6) >> gbase) & 0xFFu;
        const uint32_t b7 = (__ballot_sync(FULL, pos == 7) >> gbase) & 0xFFu;
        const uint32_t k0 = __shfl_sync(FULL, key, gbase + __ffs(b0) - 1);
        const uint32_t k1 = __shfl_sync(FULL, key, gbase + __ffs(b1) - 1);
        const uint32_t k6 = __shfl_sync(FULL, key, gbase + __ffs(b6) - 1);
        const uint32_t k7 = __shfl_sync(FULL, key, gbase + __ffs(b7) - 1);
        const bool s06 = k6 < k0, s17 = k7 < k1;
        int np = pos;
        np = (s06 && pos == 0) ? 6 : np;
        np = (s06 && pos == 6) ? 0 : np;
        np = (s17 && pos == 1) ? 7 : np;
        np = (s17 && pos == 7) ? 1 : np;
        const uint32_t c = (key << 3) | (uint32_t)np;
        int rank = 0;
#pragma unroll
        for (int sft = 1; sft < 8; sft++) rank += __shfl_xor_sync(FULL, c, sft) < c ? 1 : 0;
        const bool fit = fr >= d;
        int v = fit ? (spread ? rank : 7 - rank) : -1;  // larger is better in both cases
        v = max(v, __shfl_xor_sync(FULL, v, 1));
        v = max(v, __shfl_xor_sync(FULL, v, 2));
        v = max(v, __shfl_xor_sync(FULL, v, 4));
        const int sel = spread ? v : 7 - v;  // 8 / -1 when nothing fits: matches no position
        const bool chosen = place && fit && rank == sel;
        if (place) {
            pos = rank;
            ok = ok && v >= 0;
        }
        fr -= chosen ? d : 0;
        key -= chosen ? (uint32_t)d : 0u;
    }
    return ok;
}

// SampleRater.Choose: containers in pod order, GPUs in index order, no sorting.
template <bool PLAN>
__device__ __forceinline__ bool choose_sample(int (&f)[8], int G, const SortedDemand* dem, uint64_t* emit, int* n_emit) {
    uint64_t em = 0;
    int ne = 0;
    bool ok = true;
    for (int k = 0; k < dem->C; k++) {
        const int d = dem->raw[k];
        if (d == 0) {
            if (PLAN) { em |= (uint64_t)0xF << (4 * ne); ne++; }
            continue;
        }
        uint32_t m = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) m |= (uint32_t)(i < G && f[i] >= d) << i;
        if (m == 0) {
            if (!PLAN) return false;
            ok = false;
            continue;
        }
        const int sel = __ffs(m) - 1;
#pragma unroll
        for (int i = 0; i < 8; i++) f[i] -= (sel == i) ? d : 0;
        if (PLAN) { em |= (uint64_t)sel << (4 * ne); ne++; }
    }
    if (PLAN) { *emit = em; *n_emit = ne; }
    return ok;
}

// Un-permute the emission list into container order (rater.go:102-105) or expose the partial
// list as the error text of rater.go:99 does.  out has dem->C entries.
__device__ inline void write_plan(int32_t* out, const SortedDemand* dem, bool sorted_policy, bool ok, uint64_t emit, int n_emit) {
    const int C = dem->C;
    if (ok) {
        for (int k = 0; k < C; k++) {
            int e = (int)((emit >> (4 * k)) & 0xF);
            int v = e == 0xF ? NGS_NOT_NEED_GPU : e;
            out[sorted_policy ? dem->order[k] : k] = v;
        }
    } else {
        for (int k = 0; k < C; k++) {
            int e = (int)((emit >> (4 * k)) & 0xF);
            out[k] = k < n_emit ? (e == 0xF ? NGS_NOT_NEED_GPU : e) : INT32_MIN;
        }
    }
}

// Binpack.Rate (rater.go:59-70) with GPUs.Usage (allocate.go:164-171): the fp64 division and
// multiplication are done in IEEE round-to-nearest with explicit intrinsics (no contraction),
// then truncated toward zero like Go's int(float64).
__device__ __forceinline__ int binpack_percent_fp64(int used, int G) {
    const double usage = __ddiv_rn((double)used, (double)(100 * G));  // allocate.go:170
    return __double2int_rz(__dmul_rn(usage, 100.0));                  // rater.go:69
}
// The table is filled on the device by the very same fp64 sequence (k_build_rate_lut), one entry
// per (G, used) with 0 <= used <= 100*G; anything outside (a cell above 100 after a stray
// Release) takes the fp64 path.  lut == nullptr: always fp64.
constexpr int RATE_LUT_STRIDE = 801;
__device__ __forceinline__ int rate_binpack(const int (&f)[8], int G, int load_int, const int16_t* __restrict__ lut = nullptr) {
    int used = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) used += i < G ? 100 - f[i] : 0;
    const int pct = (lut != nullptr && used >= 0 && used <= 100 * G) ? (int)__ldg(lut + G * RATE_LUT_STRIDE + used)
                                                                      : binpack_percent_fp64(used, G);
    return pct + load_int * 50 - G;
}
// Spread.Rate (rater.go:113-123) with PercentAvailableAndFreeGpuCount (allocate.go:205-213)
__device__ __forceinline__ int rate_spread(const int (&f)[8], int G, int load_int) {
    int avail = 0, cnt = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        avail += i < G ? f[i] : 0;
        cnt += (i < G && f[i] == 100) ? 1 : 0;
    }
    return 100 * cnt + avail / 10 - G - load_int;
}
__device__ __forceinline__ int rate_policy(int policy, const int (&f)[8], int G, int load_int, const int16_t* __restrict__ lut = nullptr) {
    if (policy == NGS_POLICY_BINPACK) return rate_binpack(f, G, load_int, lut);
    if (policy == NGS_POLICY_SPREAD) return rate_spread(f, G, load_int);
    return 100;  // SampleRater.Rate -> ScoreMax (rater.go:24-26)
}

// ceil(10*usage)/10 for ceil(10*usage) = 0..10: the literal k/10 and the fp64 quotient k/10.0 are the
// same double (both correctly rounded), so the table replaces a division per usage sample
__device__ __constant__ double c_tenth[11] = {0.0, 0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0};

// GPUResource.LoadUsage for the 8 cells of one row.  tenths[m] holds ceil(10*usage) per card
// (255 = the reference would skip this entry: node or card missing, unparsable or stale
// timestamp, unparsable or out-of-range value), ts[m] the update time in unix seconds.
// Per card the SyncPeriod entries are summed in order, the node sum runs in GPU-index order
// (rater.go:63-65), both in fp64 exactly as the reference does.  rlc[] = RemainLoad + 2.
// pair[a*11+b] = fl(a/10 + b/10) for a, b in 0..10 — a 121-entry table the caller keeps in shared
// memory (fill_usage_pair_table); nullptr = add the two table values with __dadd_rn instead.
__device__ __forceinline__ void fill_usage_pair_table(double* pair, int tid, int nthreads) {
    for (int i = tid; i < 121; i += nthreads) pair[i] = __dadd_rn(c_tenth[i / 11], c_tenth[i % 11]);
}
__device__ __forceinline__ void load_usage_row(const uint8_t (&tenths)[2][8], const uint32_t (&ts)[2][8], const LoadParams& lp,
                                               int G, int (&rlc)[8], int* load_int, const double* pair = nullptr) {
    double lu = 0.0;
    if (lp.canonical && pair != nullptr) {
        // as below, with the per-card sum looked up: one shared-memory load instead of two divergent
        // constant loads and an add.  A skipped entry contributes index 0 (= 0.0).
        const uint32_t min0 = lp.min_ts[0], min1 = lp.min_ts[1];
#pragma unroll
        for (int g = 0; g < 8; g++) {
            const uint32_t a = tenths[0][g], b = tenths[1][g];
            const uint32_t ia = (a <= 10u && ts[0][g] >= min0) ? a : 0u;
            const uint32_t ib = (b <= 10u && ts[1][g] >= min1) ? b : 0u;
            const double u = pair[ia * 11u + ib];
            rlc[g] = g < G ? 4 - __double2int_rz(u) : 2;
            if (g < G) lu = __dadd_rn(lu, u);
        }
        *load_int = __double2int_rz(lu) / G;
        return;
    }
    if (lp.canonical) {
        // the deployed policy (deploy/dynamic-scheduler-node-annotator-cm.yaml:11-16): core then memory.
        // A skipped entry adds nothing, which for these non-negative values is the same as adding 0.0.
        const uint32_t min0 = lp.min_ts[0], min1 = lp.min_ts[1];
#pragma unroll
        for (int g = 0; g < 8; g++) {
            const uint32_t a = tenths[0][g], b = tenths[1][g];
            const double xa = (a != 255u && ts[0][g] >= min0) ? c_tenth[a <= 10u ? a : 10u] : 0.0;
            const double xb = (b != 255u && ts[1][g] >= min1) ? c_tenth[b <= 10u ? b : 10u] : 0.0;
            const double u = __dadd_rn(xa, xb);
            rlc[g] = g < G ? 4 - __double2int_rz(u) : 2;  // RemainLoad + 2 = (2 - int(usage)) + 2, allocate.go:193
            if (g < G) lu = __dadd_rn(lu, u);
        }
        *load_int = __double2int_rz(lu) / G;  // rater.go:68,121
        return;
    }
#pragma unroll
    for (int g = 0; g < 8; g++) {
        double u = 0.0;
        for (int e = 0; e < lp.n_periods; e++) {
            const int m = lp.metric[e];
            const uint32_t t10 = m ? tenths[1][g] : tenths[0][g];
            const uint32_t t = m ? ts[1][g] : ts[0][g];
            if (t10 == 255u || lp.never[e] || t < lp.min_ts[e]) continue;
            u = __dadd_rn(u, c_tenth[t10 <= 10u ? t10 : 10u]);  // ceil(10*u)/10, allocate.go:190-191
        }
        const int rl = 2 - __double2int_rz(u);  // allocate.go:193 (LoadTotal = 2)
        rlc[g] = g < G ? rl + 2 : 2;
        if (g < G) lu = __dadd_rn(lu, u);
    }
    *load_int = __double2int_rz(lu) / G;  // rater.go:68,121
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

// arg-max key: highest score wins, ties go to the lowest global node index.
constexpr uint64_t KEY_VALID = 1ULL << 63;
constexpr uint64_t KEY_FEASIBLE = 1ULL << 62;
constexpr int SCORE_BIAS = 1 << 29;
__device__ __forceinline__ uint64_t make_key(int score, uint32_t global_node) {
    return KEY_FEASIBLE | ((uint64_t)(uint32_t)(score + SCORE_BIAS) << 32) | (uint64_t)(0xFFFFFFFFu - global_node);
}
__device__ __forceinline__ uint32_t key_node(uint64_t k) { return 0xFFFFFFFFu - (uint32_t)k; }
__device__ __forceinline__ int key_score(uint64_t k) { return (int)((k >> 32) & 0x3FFFFFFFu) - SCORE_BIAS; }

}  // namespace ngs
