// ngs_dealer.cu — CUDA kernels (sm_100a) and the C ABI of the B200-native dealer.
//
// Snapshot layout in HBM (struct-of-arrays, one row per node, rows padded to 8 GPUs):
//   free    int32 [N][8]   GPUResource.Percent per cell (32 B row = 2 x int4)   allocate.go:141-145
//   gcount  uint8 [N]      len(NodeInfo.GPUs); 0 = node unknown                 node.go:25-42
//   rl      int8  [N][8]   GPUResource.RemainLoad (written by load-aware Rate)  allocate.go:193
//   tenths  uint8 [2][N][8] ceil(10*usage) per (metric, node, card); 255 = skip  nodeusage.go:82-111
//   ts      uint32[2][N][8] update time, unix seconds                            stats.go:30-55
//   tag     uint32[N], cscore int32[N]   per-node plan cache (NodeInfo.PlanCache, node.go:22)
//   free_hi / rl_hi / tenths_hi / ts_hi   [..][N][4]: cells 8..11 of nodes with 9..12 GPUs, allocated with the first such node
//
// Kernels (one translation unit; the device code lives in headers included below):
//   ngs_device.cuh    building blocks on one row held in registers (sort emulation, rates, usage, keys)
//   ngs_tables.cuh    the tables as the kernels see them;  ngs_wide.cuh  one-thread evaluation of rows with 9..12 GPUs
//   ngs_rows.cuh      evaluation of one row (GPUs.Choose)
//   ngs_k1_scan.cuh   K1  k_assume (one thread per candidate: feasibility / score / plan out) and
//                         k_scan_tma (all rows, tiles staged by TMA bulk copies)
//   ngs_k3_rows.cuh   K3  k_bind, k_apply, k_release_batch;  K4  k_load_update, k_load_add_node, k_upsert
//   ngs_k6_batch.cuh  K6  k_schedule: persistent cooperative kernel, rows resident in shared memory, per pod
//                         filter + score every row, grid-wide arg-max (K2), bind on the owner; K5 inside it
//   ngs_k7_one.cuh    K7  k_schedule_one[_random]: one pod's whole cycle in one ordinary launch
//   ngs_k8_serve.cuh  K8  k_serve: the resident form - launched once, serves pods (and the verbs) from requests in mapped host memory
// This file: the handle, launch code and the C ABI.
// All paths relative to /root/reference/pkg/dealer/.
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <pthread.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <chrono>
#include <cmath>
#include <atomic>
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

#include "ngs_rows.cuh"
#include "ngs_k1_scan.cuh"
#include "ngs_k3_rows.cuh"
#include "ngs_k6_batch.cuh"
#include "ngs_k7_one.cuh"
#include "ngs_k8_serve.cuh"

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
    int8_t* h_vplan = nullptr;   // mapped pinned: packed plans [cap][NGS_MAX_CONTAINERS] bytes (+ a word of padding), entries per node of a call: its C
    uint8_t* h_vnem = nullptr;   // mapped pinned: emitted counts
    int8_t* d_vplan = nullptr;   // device aliases
    uint8_t* d_vnem = nullptr;
    size_t vplan_cap = 0;        // candidates the two hold
    size_t verb_cap = 0;   // candidates the verb scratch holds
    BindResult* d_bind = nullptr;  // device alias of h_bind
    BindResult* h_bind = nullptr;  // mapped pinned: k_bind writes its answer here, the host polls the status word
    // verb staging in mapped pinned memory: the scan kernels write feasible / reason / score straight into host memory
    // and raise a flag; a call needs no copy command and no stream synchronisation
    int8_t* h_vreason = nullptr;
    int32_t* h_vscore = nullptr;
    int32_t* h_vflag = nullptr;
    int8_t* d_vreason = nullptr;  // device aliases
    int32_t* d_vscore = nullptr;
    int32_t* d_vflag = nullptr;
    uint32_t* d_vticket = nullptr;
    size_t vstage_cap = 0;
    int32_t vseq = 0;
    // NodeInfo.Score after NodeInfo.Assume (node.go:59-68: the cached Plan.Score of the filter call): the scores of the last
    // ngs_assume stay in the staging area; an ngs_score for the same demand and candidates, with no row changed in between, is a copy
    bool sc_valid = false;
    uint32_t sc_gen = 0, sc_n = 0;
    uint64_t sc_sig = 0, sc_epoch = 0;
    uint64_t epoch = 1;       // bumped by every call that changes a row
    // the candidate list resident in d_idx, kept on the host as well: the same list (kube-scheduler sends one list for
    // filter and prioritize, and mostly the same from pod to pod) is recognised by one memcmp and not uploaded again
    std::vector<uint32_t> idx_host;
    uint64_t idx_sig = 0;     // serial number of that list (0: none); part of the score cache's key
    int32_t* d_small = nullptr;    // gpu_idx in / status out for k_apply
    int32_t* h_small = nullptr;    // pinned
    // batch
    uint32_t batch_cap = 0;  // pods per launch chunk
    SortedDemand* d_dem = nullptr;
    OneRec* d_one = nullptr;               // K7's arg-max word and ticket counter (zero between calls)
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
    uint64_t xbatch = 0;                                  // launches so far: selects the parity half
    // K8, the resident single-pod kernel (ngs_k8_serve.cuh)
    bool serve_running = false;       // a k_serve launch has been queued and not yet been waited for
    uint32_t serve_seq = 0;           // sequence number of the last request posted
    uint32_t serve_launch = 0;        // id of the current launch (the kernel writes it to *h_serve_exit when it leaves)
    uint32_t* h_serve_req = nullptr;  // mapped pinned: request block [SERVE_REQ_WORDS], exit word at [SERVE_REQ_WORDS]
    uint32_t* d_serve_req = nullptr;  // device alias
    int32_t* h_serve_out = nullptr;   // mapped pinned: answer block
    int32_t* d_serve_out = nullptr;
    ServeDev* d_serve_dev = nullptr;
    int serve_grid_cap[4] = {0, 0, 0, 0};  // co-resident CTAs of k_serve<WIDE, VERBS>, index WIDE + 2 * VERBS
    bool serve_kind_verbs = false;    // which instantiation is running
    double serve_idle_ms = 20.0;      // the kernel leaves after this long without a request
    bool serve_verbs = false;         // ngs_assume / ngs_score / ngs_bind go through the resident kernel, too (ngs_serve_verbs)
    uint32_t serve_prev_row = 0xFFFFFFFFu;  // the row the last served pod was bound to (told to the kernel with the next pod)
    cudaStream_t copy_stream = nullptr;  // uploads while the resident kernel holds the handle's stream
    // nodes with 9..12 GPUs (ngs_wide.cuh): which rows, so that the batch kernel can refuse them before it binds anything
    std::vector<uint8_t> wide_row;
    uint32_t n_wide = 0;
    ngs_stats_t stats{};
    unsigned long long* d_dbg = nullptr;  // profiling builds only
    uint32_t dbg_grid = 0;
    char err[512] = {0};
};

namespace {

constexpr uint32_t BATCH_CHUNK = 4096;
constexpr size_t XSLOT_HALF = (size_t)BATCH_CHUNK * NGS_MAX_RANKS * 2 + XMAIL_WORDS;  // u64 words per parity half: pod slots [pod][rank][2], then K7's plan mailbox
static_assert(XMAIL_OFFSET == (size_t)BATCH_CHUNK * NGS_MAX_RANKS * 2, "K7's mailbox sits right behind the pod slots");

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

int ensure_verb_stage(ngs_dealer* h, size_t n) {
    if (n <= h->vstage_cap) return NGS_OK;
    const size_t cap = n < 4096 ? 4096 : n;
    cudaFreeHost(h->h_vreason); cudaFreeHost(h->h_vscore);
    h->h_vreason = nullptr; h->h_vscore = nullptr;
    h->vstage_cap = 0;
    h->sc_valid = false;
    CU(h, cudaHostAlloc(&h->h_vreason, cap, cudaHostAllocMapped));
    CU(h, cudaHostAlloc(&h->h_vscore, cap * 4, cudaHostAllocMapped));
    CU(h, cudaHostGetDevicePointer(&h->d_vreason, h->h_vreason, 0));
    CU(h, cudaHostGetDevicePointer(&h->d_vscore, h->h_vscore, 0));
    h->vstage_cap = cap;
    return NGS_OK;
}

// feasible[i] = (status[i] == NGS_REASON_OK), eight nodes per step (NGS_REASON_OK is 0: a byte-is-zero test in a 64-bit word)
static_assert(NGS_REASON_OK == 0, "expand_feasible tests for zero bytes");
void expand_feasible(uint8_t* feasible, const int8_t* status, size_t n) {
    const uint64_t lo7 = 0x7F7F7F7F7F7F7F7FULL, hi = 0x8080808080808080ULL;
    size_t i = 0;
#if defined(__SSE2__)
    const __m128i zero = _mm_setzero_si128(), one = _mm_set1_epi8(1);
    for (; i + 16 <= n; i += 16) {
        const __m128i x = _mm_loadu_si128(reinterpret_cast<const __m128i*>(status + i));
        _mm_storeu_si128(reinterpret_cast<__m128i*>(feasible + i), _mm_and_si128(_mm_cmpeq_epi8(x, zero), one));
    }
#endif
    for (; i + 8 <= n; i += 8) {
        uint64_t x;
        memcpy(&x, status + i, 8);
        const uint64_t nonzero = (((x & lo7) + lo7) | x) & hi;  // bit 7 of every byte that is not 0
        const uint64_t ok = (nonzero ^ hi) >> 7;                 // 1 in every byte that is 0
        memcpy(feasible + i, &ok, 8);
    }
    for (; i < n; i++) feasible[i] = status[i] == NGS_REASON_OK ? 1 : 0;
}

// plan entries travel as one byte each (PLAN_NONE = INT32_MIN), emitted counts as one byte: widen into the caller's int32 arrays
void widen_plan(int32_t* dst, const int8_t* src, size_t n) {
    size_t i = 0;
#if defined(__SSE2__)
    const __m128i zero = _mm_setzero_si128(), none8 = _mm_set1_epi32(PLAN_NONE), none32 = _mm_set1_epi32(INT32_MIN);
    for (; i + 16 <= n; i += 16) {
        const __m128i x = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
        const __m128i lo16 = _mm_unpacklo_epi8(zero, x), hi16 = _mm_unpackhi_epi8(zero, x);  // byte in the high half of each 16-bit lane
        __m128i q[4] = {_mm_unpacklo_epi16(zero, lo16), _mm_unpackhi_epi16(zero, lo16), _mm_unpacklo_epi16(zero, hi16), _mm_unpackhi_epi16(zero, hi16)};
        for (int k = 0; k < 4; k++) {
            const __m128i v = _mm_srai_epi32(q[k], 24);  // sign-extended
            const __m128i m = _mm_cmpeq_epi32(v, none8);
            _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + i + 4 * k), _mm_or_si128(_mm_andnot_si128(m, v), _mm_and_si128(m, none32)));
        }
    }
#endif
    for (; i < n; i++) dst[i] = src[i] == PLAN_NONE ? INT32_MIN : (int32_t)src[i];
}
void widen_u8(int32_t* dst, const uint8_t* src, size_t n) {
    size_t i = 0;
#if defined(__SSE2__)
    const __m128i zero = _mm_setzero_si128();
    for (; i + 16 <= n; i += 16) {
        const __m128i x = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
        const __m128i lo16 = _mm_unpacklo_epi8(x, zero), hi16 = _mm_unpackhi_epi8(x, zero);
        _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + i), _mm_unpacklo_epi16(lo16, zero));
        _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + i + 4), _mm_unpackhi_epi16(lo16, zero));
        _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + i + 8), _mm_unpacklo_epi16(hi16, zero));
        _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + i + 12), _mm_unpackhi_epi16(hi16, zero));
    }
#endif
    for (; i < n; i++) dst[i] = src[i];
}

// wait for a verb kernel's flag in mapped host memory (no stream synchronisation); keeps an eye on the stream
int wait_flag(ngs_dealer* h, volatile int32_t* flag, int32_t want, const char* what) {
    for (uint32_t spins = 1; *flag != want; spins++) {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
        if ((spins & 0xFFFu) == 0) {
            const cudaError_t q = cudaStreamQuery(h->stream);
            if (q == cudaSuccess) {
                if (*flag == want) break;
                return fail(h, NGS_ERR_CUDA, "%s ended without raising its flag", what);
            }
            if (q != cudaErrorNotReady) return fail(h, NGS_ERR_CUDA, "%s: %s", what, cudaGetErrorString(q));
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return NGS_OK;
}

int ensure_verb_scratch(ngs_dealer* h, size_t n, size_t plan_entries) {
    if (n > h->verb_cap) {
        size_t cap = n < 1024 ? 1024 : n;
        cudaFree(h->d_idx); cudaFree(h->d_feas); cudaFree(h->d_reason); cudaFree(h->d_score);
        h->d_idx = nullptr; h->d_feas = nullptr; h->d_reason = nullptr; h->d_score = nullptr;
        h->verb_cap = 0;
        h->idx_sig = 0;
        CU(h, cudaMalloc(&h->d_idx, cap * 4));
        CU(h, cudaMalloc(&h->d_feas, cap));
        CU(h, cudaMalloc(&h->d_reason, cap));
        CU(h, cudaMalloc(&h->d_score, cap * 4));
        h->verb_cap = cap;
    }
    if (plan_entries && n > h->vplan_cap) {  // packed plans and emitted counts, written by the kernel straight into host memory
        const size_t cap = n < 4096 ? 4096 : n;
        cudaFreeHost(h->h_vplan); cudaFreeHost(h->h_vnem);
        h->h_vplan = nullptr; h->h_vnem = nullptr;
        h->vplan_cap = 0;
        CU(h, cudaHostAlloc(&h->h_vplan, cap * NGS_MAX_CONTAINERS + 16, cudaHostAllocMapped));
        CU(h, cudaHostAlloc(&h->h_vnem, cap + 16, cudaHostAllocMapped));
        CU(h, cudaHostGetDevicePointer(&h->d_vplan, h->h_vplan, 0));
        CU(h, cudaHostGetDevicePointer(&h->d_vnem, h->h_vnem, 0));
        h->vplan_cap = cap;
    }
    return NGS_OK;
}

int ensure_batch(ngs_dealer* h) {
    if (h->batch_cap) return NGS_OK;
    const uint32_t cap = BATCH_CHUNK;
    const size_t in_words = (size_t)cap * (NGS_MAX_CONTAINERS + 2), out_words = (size_t)cap * (NGS_MAX_CONTAINERS + 2);
    CU(h, cudaMalloc(&h->d_dem, sizeof(SortedDemand) * cap));
    // one exchange record per pod, pod-major, so that a batch of P pods clears ONE contiguous range:
    //   [0,32) arg-max round record {xa, x1, x2, arrivals}   [32,36) random-policy arrival counter   [64, 64+8*SMs) per-CTA slots
    h->pod_rec_bytes = (64 + sizeof(unsigned long long) * (size_t)h->num_sms + 31) / 32 * 32;
    CU(h, cudaMalloc(&h->d_keys, h->pod_rec_bytes * cap));
    CU(h, cudaMalloc(&h->d_one, sizeof(OneRec) + sizeof(uint32_t) * ONE_MAX_CTAS));  // record, then K7's per-CTA counts
    CU(h, cudaMemsetAsync(h->d_one, 0, sizeof(OneRec) + sizeof(uint32_t) * ONE_MAX_CTAS, h->stream));
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

// dynamic shared memory of K6: 32 bytes per row plus four feasibility-bit buffers (one per pod in flight); its static part
// (staged demands, per-warp keys, round state) stays below K6_STATIC_SMEM
size_t schedule_smem_bytes(uint32_t rpb) { return (size_t)rpb * 32 + (size_t)((rpb + 31) / 32) * 16 + 16; }
constexpr size_t K6_STATIC_SMEM = 4096;
// which instantiation of K6 serves CTAs of `rpb` rows, and its block size
// (wide: the snapshot holds nodes of more than NGS_MAX_GPUS GPUs)
const void* schedule_kernel(uint32_t rpb, bool wide) {
    if (wide) return rpb > K6_EXCLUSIVE_ROWS ? (const void*)k_schedule<true, true> : (const void*)k_schedule<false, true>;
    return rpb > K6_EXCLUSIVE_ROWS ? (const void*)k_schedule<true, false> : (const void*)k_schedule<false, false>;
}
uint32_t schedule_block(uint32_t rpb) { return rpb > K6_EXCLUSIVE_ROWS ? 1024u : schedule_block_size(rpb); }
// the dynamic shared memory limit is a per-device upper bound shared by every launch: allow the device maximum for every instantiation
cudaError_t schedule_allow_max_smem(int smem_optin) {
    const void* fns[4] = {(const void*)k_schedule<false, false>, (const void*)k_schedule<true, false>, (const void*)k_schedule<false, true>,
                          (const void*)k_schedule<true, true>};
    for (const void* fn : fns) {
        cudaFuncAttributes fa;
        cudaError_t e = cudaFuncGetAttributes(&fa, fn);  // the opt-in limit covers static + dynamic shared memory
        if (e == cudaSuccess) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin - (int)fa.sharedSizeBytes);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// launch K6 for one chunk of <= BATCH_CHUNK pods whose inputs are in device memory
int launch_schedule(ngs_dealer* h, uint32_t P, const int32_t* d_percent, const int32_t* d_ncont, int max_c, const int32_t* d_policy,
                    const LoadParams& lp, uint64_t seed_seq, int32_t* d_node, int32_t* d_score, int32_t* d_gidx,
                    int32_t* status_out = nullptr) {
    if (!h->coop) return fail(h, NGS_ERR_CUDA, "device lacks cooperative launch");
    h->epoch++;  // the batch binds pods
    const uint32_t N = h->T.n;
    uint32_t grid = (N + 63) / 64;
    if (grid < 1) grid = 1;
    if (grid > (uint32_t)h->num_sms) grid = (uint32_t)h->num_sms;
#ifdef NGS_K6_GRID_ENV
    if (const char* e = getenv("NGS_K6_GRID")) grid = (uint32_t)atoi(e);  // (experiments only)
#endif
    const uint32_t rpb = (N + grid - 1) / grid;
    const uint32_t block = schedule_block(rpb);
    const size_t smem = schedule_smem_bytes(rpb);
    if (smem + K6_STATIC_SMEM > (size_t)h->smem_optin)
        return fail(h, NGS_ERR_UNSUPPORTED, "%u rows per SM need %zu B shared memory (> %d): shard the cluster over more GPUs", rpb, smem, h->smem_optin);
    {
        // the attribute is a per-device upper bound shared by every handle: allow the device maximum, once
        static bool s_done[64] = {false};
        static pthread_mutex_t s_mu = PTHREAD_MUTEX_INITIALIZER;
        const int dev = h->device & 63;
        pthread_mutex_lock(&s_mu);
        cudaError_t e = cudaSuccess;
        if (!s_done[dev]) {
            e = schedule_allow_max_smem(h->smem_optin);
            s_done[dev] = e == cudaSuccess;
        }
        pthread_mutex_unlock(&s_mu);
        CU(h, e);
    }
    // the abort flag belongs to one launch (a batch that timed out must not make the next one give up); the status word
    // beside it stays until ngs_sync() has reported it
    CU(h, cudaMemsetAsync(h->d_err + 1, 0, 4, h->stream));
    k_prep_demands<<<(P + 127) / 128, 128, 0, h->stream>>>(d_percent, d_ncont, max_c, d_policy, P, h->d_dem, h->d_err);
    CU(h, cudaGetLastError());
    unsigned char* rec = reinterpret_cast<unsigned char*>(h->d_keys);
    uint32_t* d_done = reinterpret_cast<uint32_t*>(rec + 32);
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
    bp.keys = reinterpret_cast<unsigned long long*>(rec + 64);
    bp.done = d_done;
    bp.rr = reinterpret_cast<unsigned long long*>(rec);
    bp.keys_stride = (uint32_t)(h->pod_rec_bytes / 8);
    bp.done_stride = (uint32_t)(h->pod_rec_bytes / 4);
    bp.rr_stride = (uint32_t)(h->pod_rec_bytes / 8);
    bp.status_out = status_out;
    bp.world = (uint32_t)h->world;
    bp.rank = (uint32_t)h->rank;
    for (int r = 0; r < NGS_MAX_RANKS; r++) bp.xslots[r] = nullptr;
    if (h->world > 1) {
        // Parity halves instead of a cross-rank barrier per batch: batch i uses half i&1 on every
        // rank; the other half is cleared now, before this batch starts.  A peer can only write into
        // that half once it has finished batch i, which needs this rank's batch-i posts first.
        const size_t cur = (size_t)(h->xbatch & 1), nxt = cur ^ 1;
        CU(h, cudaMemsetAsync(h->d_xslots + nxt * XSLOT_HALF, 0, sizeof(unsigned long long) * XSLOT_HALF, h->stream));
        for (int r = 0; r < h->world; r++) bp.xslots[r] = h->peer_xslots[r] + cur * XSLOT_HALF;
        h->xbatch++;
    }
    bp.err = h->d_err;
    bp.rows_per_block = rpb;
#ifdef NGS_PHASE_TIMING
    if (!h->d_dbg) CU(h, cudaMalloc(&h->d_dbg, sizeof(unsigned long long) * (4096 + 256 * 256 * 8)));
    CU(h, cudaMemsetAsync(h->d_dbg, 0, sizeof(unsigned long long) * (4096 + 256 * 256 * 8), h->stream));
    bp.dbg = h->d_dbg;
    h->dbg_grid = grid;
#endif
    void* args[] = {&bp};
    CU(h, cudaLaunchCooperativeKernel(schedule_kernel(rpb, h->n_wide != 0), dim3(grid), dim3(block), args, smem, h->stream));
    h->stats.kernel_launches += 2;
    h->stats.node_evals += (uint64_t)P * N;
    h->stats.pods_scheduled += P;
    return NGS_OK;
}

// ---- K8: the resident kernel's life cycle ----
// first thing in every other entry point of a handle
#define SERVE_OFF(h)                   \
    do {                               \
        if ((h)->serve_running) {      \
            const int rc_ = serve_stop(h); \
            if (rc_) return rc_;       \
        }                              \
    } while (0)

uint32_t serve_next_seq(uint32_t s) {
    s++;
    return (s == 0 || s == SERVE_QUIT) ? 1u : s;
}

// stop the resident kernel (if any) before anything else touches the tables or the stream: it owns both while it runs
int serve_stop(ngs_dealer* h) {
    if (!h->serve_running) return NGS_OK;
    cudaSetDevice(h->device);
    volatile uint32_t* rq = h->h_serve_req;
    if (rq[SERVE_REQ_WORDS] != h->serve_launch) {  // still there: a request with quit set
        ServeBody b;
        memset(&b, 0, sizeof(b));
        b.quit = 1;
        const uint32_t* w = reinterpret_cast<const uint32_t*>(&b);
        const uint32_t seq = h->serve_seq = serve_next_seq(h->serve_seq);
        for (int i = 0; i < SERVE_BODY_WORDS; i++) rq[serve_block_word(i)] = i < SERVE_BODY_USED ? w[i] : 0u;
        std::atomic_thread_fence(std::memory_order_seq_cst);
        for (int q = 0; q < SERVE_REQ_WORDS / 4; q++) rq[4 * q + 3] = seq;
        std::atomic_thread_fence(std::memory_order_seq_cst);
    }
    h->serve_running = false;
    CU(h, cudaStreamSynchronize(h->stream));
    return NGS_OK;
}

// last_seq: the sequence number the kernel regards as already answered
int serve_start(ngs_dealer* h, uint32_t last_seq) {
    if (!h->coop) return fail(h, NGS_ERR_CUDA, "device lacks cooperative launch");
    if (!h->h_serve_req) {
        CU(h, cudaHostAlloc(&h->h_serve_req, (SERVE_REQ_WORDS + 16) * 4, cudaHostAllocMapped));
        CU(h, cudaHostGetDevicePointer(&h->d_serve_req, h->h_serve_req, 0));
        memset(h->h_serve_req, 0, (SERVE_REQ_WORDS + 16) * 4);
        CU(h, cudaHostAlloc(&h->h_serve_out, 64 * 4, cudaHostAllocMapped));
        CU(h, cudaHostGetDevicePointer(&h->d_serve_out, h->h_serve_out, 0));
        memset(h->h_serve_out, 0, 64 * 4);
        CU(h, cudaMalloc(&h->d_serve_dev, sizeof(ServeDev)));
        int per_sm = 0;
        CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_serve<false, false>, 256, 0));
        h->serve_grid_cap[0] = per_sm * h->num_sms;
        CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_serve<true, false>, 256, 0));
        h->serve_grid_cap[1] = per_sm * h->num_sms;
        CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_serve<false, true>, 256, 0));
        h->serve_grid_cap[2] = per_sm * h->num_sms;
        CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_serve<true, true>, 256, 0));
        h->serve_grid_cap[3] = per_sm * h->num_sms;
    }
    const int wide = (h->n_wide ? 1 : 0) + (h->serve_verbs ? 2 : 0);
    h->serve_kind_verbs = h->serve_verbs;
    uint32_t grid = (h->T.n + 255) / 256;
    if (grid > (uint32_t)h->num_sms * 4) grid = (uint32_t)h->num_sms * 4;
    if (grid > (uint32_t)h->serve_grid_cap[wide]) grid = (uint32_t)h->serve_grid_cap[wide];
    if (grid < 1) return fail(h, NGS_ERR_CUDA, "the resident kernel does not fit the device");
    // the published word starts at the last sequence number: the kernel waits for the next one
    ServeDev init;
    memset(&init, 0, sizeof(init));
    for (int q = 0; q < SERVE_REQ_WORDS / 4; q++) init.blk[4 * q + 3] = last_seq;
    CU(h, cudaMemcpyAsync(h->d_serve_dev, &init, sizeof(init), cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));  // (pageable source; and nothing else of this handle is in flight from here on)
    h->serve_launch = h->serve_launch + 1 ? h->serve_launch + 1 : 1;
    Tables T = h->T;
    OneRec* rec = h->d_one;
    ServeDev* dev = h->d_serve_dev;
    const uint32_t* req = h->d_serve_req;
    int32_t* out = h->d_serve_out;
    uint32_t* exited = h->d_serve_req + SERVE_REQ_WORDS;
    uint32_t launch_id = h->serve_launch;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, h->device);
    long long idle_cycles = (long long)(h->serve_idle_ms * (khz > 0 ? khz : 1900000));
    void* args[] = {&T, &rec, &dev, &req, &out, &exited, &last_seq, &launch_id, &idle_cycles};
    const void* fns[4] = {(const void*)k_serve<false, false>, (const void*)k_serve<true, false>, (const void*)k_serve<false, true>,
                          (const void*)k_serve<true, true>};
    const void* fn = fns[wide];
    CU(h, cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(256), args, 0, h->stream));
    h->serve_running = true;
    h->stats.kernel_launches++;
    return NGS_OK;
}

// K8: make sure the resident kernel runs, post one request and wait for its answer (out[0] == sequence number)
int serve_request(ngs_dealer* h, ServeBody& b) {
    volatile uint32_t* rq = h->h_serve_req;
    if (h->serve_running && rq[SERVE_REQ_WORDS] == h->serve_launch) {  // it left (idle): collect the launch
        h->serve_running = false;
        CU(h, cudaStreamSynchronize(h->stream));
    }
    if (h->serve_running && h->serve_kind_verbs != h->serve_verbs) {  // the other instantiation is wanted now
        const int rc = serve_stop(h);
        if (rc) return rc;
    }
    if (!h->serve_running) {
        int rc = ensure_batch(h);
        if (rc) return rc;
        rc = serve_start(h, h->serve_seq);
        if (rc) return rc;
        rq = h->h_serve_req;
    }
    b.rl_zero = h->T.rl_zero;
    const uint32_t seq = h->serve_seq = serve_next_seq(h->serve_seq);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&b);
    for (int i = 0; i < SERVE_BODY_USED; i++) rq[serve_block_word(i)] = w[i];
    std::atomic_thread_fence(std::memory_order_seq_cst);  // payload before the sequence words
    for (int q = 0; q < SERVE_REQ_WORDS / 4; q++) rq[4 * q + 3] = seq;
    std::atomic_thread_fence(std::memory_order_seq_cst);
    volatile int32_t* out = h->h_serve_out;
    int restarts = 0;
    const auto t_start = std::chrono::steady_clock::now();
    for (uint64_t spins = 1; (uint32_t)out[0] != seq; spins++) {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
        if ((spins & 0x3FFu) == 0) {
            if (rq[SERVE_REQ_WORDS] == h->serve_launch && (uint32_t)out[0] != seq) {
                // the kernel left (idle) just before the request arrived: start it again, the request is still in the block
                h->serve_running = false;
                CU(h, cudaStreamSynchronize(h->stream));
                if (++restarts > 3) return fail(h, NGS_ERR_TIMEOUT, "the resident kernel keeps leaving before it answers");
                const int rc = serve_start(h, seq == 1 ? SERVE_QUIT - 1 : seq - 1);
                if (rc) return rc;
            } else if ((spins & 0xFFFFFu) == 0) {
                const cudaError_t q = cudaStreamQuery(h->stream);
                if (q != cudaErrorNotReady && (uint32_t)out[0] != seq && rq[SERVE_REQ_WORDS] != h->serve_launch) {
                    h->serve_running = false;
                    return fail(h, NGS_ERR_CUDA, "the resident kernel ended without an answer: %s", cudaGetErrorString(q));
                }
                if (std::chrono::steady_clock::now() - t_start > std::chrono::seconds(10)) {
                    h->serve_running = false;  // (whatever is on the stream is left to ngs_destroy / the next synchronising call)
                    return fail(h, NGS_ERR_TIMEOUT, "the resident kernel did not answer within 10 s");
                }
            }
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return NGS_OK;
}

// the second plane of the tables (cells 8..11), created when the first node with more than 8 GPUs arrives
int ensure_wide(ngs_dealer* h) {
    if (h->T.free_hi) return NGS_OK;
    const size_t NP = h->T.nstride;
    int32_t* fh = nullptr;
    int8_t* rh = nullptr;
    uint8_t* th = nullptr;
    uint32_t* sh = nullptr;
    if (cudaMalloc(&fh, NP * 4 * 4) != cudaSuccess || cudaMalloc(&rh, NP * 4) != cudaSuccess || cudaMalloc(&th, 2 * NP * 4) != cudaSuccess ||
        cudaMalloc(&sh, 2 * NP * 4 * 4) != cudaSuccess) {
        const cudaError_t e = cudaGetLastError();
        cudaFree(fh); cudaFree(rh); cudaFree(th); cudaFree(sh);
        return fail(h, NGS_ERR_CUDA, "tables for nodes of more than %d GPUs: %s", NGS_MAX_GPUS, cudaGetErrorString(e));
    }
    CU(h, cudaMemsetAsync(fh, 0, NP * 4 * 4, h->stream));
    CU(h, cudaMemsetAsync(rh, 0, NP * 4, h->stream));
    CU(h, cudaMemsetAsync(th, 0xFF, 2 * NP * 4, h->stream));
    CU(h, cudaMemsetAsync(sh, 0, 2 * NP * 4 * 4, h->stream));
    h->T.free_hi = fh;
    h->T.rl_hi = rh;
    h->T.tenths_hi = th;
    h->T.ts_hi = sh;
    h->wide_row.assign(h->T.n, 0);
    return NGS_OK;
}
void mark_width(ngs_dealer* h, uint32_t row, int G) {
    if (h->wide_row.empty()) return;
    const uint8_t w = G > NGS_MAX_GPUS ? 1 : 0;
    h->n_wide += (uint32_t)w - (uint32_t)h->wide_row[row];
    h->wide_row[row] = w;
}

int check_dev_err(ngs_dealer* h) {
    if (*h->h_err != NGS_OK) {
        int e = *h->h_err;
        *h->h_err = NGS_OK;
        cudaMemsetAsync(h->d_err, 0, 8, h->stream);
        if (e == NGS_ERR_TIMEOUT) return fail(h, e, "batch kernel gave up waiting for a peer CTA (results of this batch are incomplete)");
        if (e == NGS_ERR_UNSUPPORTED) return fail(h, e, "the batch kernel met a node with more than NGS_MAX_GPUS GPUs and left it out (schedule such snapshots through the verbs)");
        return fail(h, e, "a batch held a pod with a bad container count, percent or policy: that pod was left unscheduled "
                          "(node_out = NGS_NODE_INVALID_POD), every other pod was scheduled as usual");
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
    CU(h, cudaHostAlloc(&h->h_bind, sizeof(BindResult), cudaHostAllocMapped));
    CU(h, cudaHostGetDevicePointer(&h->d_bind, h->h_bind, 0));
    CU(h, cudaHostAlloc(&h->h_vflag, 64, cudaHostAllocMapped));
    CU(h, cudaHostGetDevicePointer(&h->d_vflag, h->h_vflag, 0));
    *h->h_vflag = 0;
    CU(h, cudaMalloc(&h->d_vticket, 64));
    CU(h, cudaMemsetAsync(h->d_vticket, 0, 64, h->stream));
    CU(h, cudaMalloc(&h->d_small, 64 * 4));
    CU(h, cudaMallocHost(&h->h_small, 64 * 4));
    CU(h, cudaMalloc(&h->d_err, 8));  // [0] status, [1] abort flag
    CU(h, cudaMallocHost(&h->h_err, 4));
    *h->h_err = 0;
    CU(h, cudaMemsetAsync(h->d_err, 0, 8, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return NGS_OK;
}

int ngs_destroy(ngs_dealer* h) {
    if (!h) return NGS_OK;
    cudaSetDevice(h->device);
    if (h->serve_running) serve_stop(h);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFreeHost(h->h_serve_req); cudaFreeHost(h->h_serve_out); cudaFree(h->d_serve_dev);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    cudaFree(h->T.free); cudaFree(h->T.gcount); cudaFree(h->T.rl); cudaFree(h->T.tenths); cudaFree(h->T.ts);
    cudaFree(h->T.tag); cudaFree(h->T.cscore); cudaFree(const_cast<int16_t*>(h->T.blut));
    cudaFree(h->T.free_hi); cudaFree(h->T.rl_hi); cudaFree(h->T.tenths_hi); cudaFree(h->T.ts_hi);
    cudaFree(h->d_idx); cudaFree(h->d_feas); cudaFree(h->d_reason); cudaFree(h->d_score); cudaFreeHost(h->h_vplan); cudaFreeHost(h->h_vnem);
    cudaFreeHost(h->h_bind); cudaFree(h->d_small); cudaFreeHost(h->h_small);
    cudaFreeHost(h->h_vreason); cudaFreeHost(h->h_vscore); cudaFreeHost(h->h_vflag); cudaFree(h->d_vticket);
    cudaFree(h->d_dem); cudaFree(h->d_keys); cudaFree(h->d_one);
    cudaFreeHost(h->h_pod_in); cudaFreeHost(h->h_pod_out);
    cudaFree(h->d_err); cudaFreeHost(h->h_err);
    for (int r = 0; r < NGS_MAX_RANKS; r++)
        if (h->peer_xslots[r] && r != h->rank) cudaIpcCloseMemHandle(h->peer_xslots[r]);
    cudaFree(h->d_xslots); cudaFree(h->d_rel);
    cudaFree(h->d_dbg);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return NGS_OK;
}

void* ngs_stream(ngs_dealer* h) { return h ? (void*)h->stream : nullptr; }

int ngs_sync(ngs_dealer* h) {
    if (!h) return NGS_ERR_INVALID;
    SERVE_OFF(h);
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaMemcpyAsync(h->h_err, h->d_err, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return check_dev_err(h);
}

#ifdef NGS_PHASE_TIMING
// profiling builds only: per-CTA cycles per phase of the last k_schedule launch ([grid][16])
int ngs_debug_phases(ngs_dealer* h, unsigned long long* out, uint32_t* grid) {
    if (!h || !h->d_dbg) return NGS_ERR_INVALID;
    CU(h, cudaMemcpyAsync(out, h->d_dbg, sizeof(unsigned long long) * (4096 + 256 * 256 * 8), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    *grid = h->dbg_grid;
    return NGS_OK;
}
#endif

#ifdef NGS_K8_TIMING
// profiling builds only: nanoseconds from CTA 0 seeing the last served request to each stamp of the CTA that answered it
int ngs_debug_k8(ngs_dealer* h, int32_t* out5) {
    if (!h || !h->h_serve_out) return NGS_ERR_INVALID;
    for (int i = 0; i < 5; i++) out5[i] = h->h_serve_out[32 + i];
    return NGS_OK;
}
#endif
#ifdef NGS_K7_TIMING
// profiling builds only: cycles from kernel entry to each stamp of the last k_schedule_one launch (its last CTA)
int ngs_debug_k7(ngs_dealer* h, int32_t* out5) {
    if (!h || !h->h_pod_out) return NGS_ERR_INVALID;
    for (int i = 0; i < 5; i++) out5[i] = h->h_pod_out[32 + i];
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
    SERVE_OFF(h);
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    if (gpu_count < 0) return fail(h, NGS_ERR_INVALID, "gpu_count %d", gpu_count);
    if (gpu_count > NGS_MAX_GPUS_WIDE) return fail(h, NGS_ERR_UNSUPPORTED, "gpu_count %d > NGS_MAX_GPUS_WIDE", gpu_count);
    CU(h, cudaSetDevice(h->device));
    if (gpu_count > NGS_MAX_GPUS) {
        const int rc = ensure_wide(h);
        if (rc) return rc;
    }
    mark_width(h, idx - h->T.first, gpu_count);
    h->epoch++;
    k_upsert<<<1, 32, 0, h->stream>>>(h->T, idx - h->T.first, gpu_count);
    CU(h, cudaGetLastError());
    h->stats.kernel_launches++;
    return NGS_OK;
}

int ngs_node_set_wide(ngs_dealer* h, uint32_t idx, int gpu_count, const int32_t* free, const int8_t* remain_load) {
    if (!h) return NGS_ERR_INVALID;
    SERVE_OFF(h);
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    if (gpu_count < 0) return fail(h, NGS_ERR_INVALID, "gpu_count %d", gpu_count);
    if (gpu_count > NGS_MAX_GPUS_WIDE) return fail(h, NGS_ERR_UNSUPPORTED, "gpu_count %d > NGS_MAX_GPUS_WIDE", gpu_count);
    int32_t fr[NGS_MAX_GPUS_WIDE];
    int8_t rl[NGS_MAX_GPUS_WIDE];
    for (int g = 0; g < NGS_MAX_GPUS_WIDE; g++) {
        fr[g] = g < gpu_count ? (free ? free[g] : 100) : 0;
        if (fr[g] < 0 || fr[g] > NGS_MAX_FREE) return fail(h, NGS_ERR_RANGE, "node %u cell %d: free %d out of range", idx, g, fr[g]);
        const int r = (g < gpu_count && remain_load) ? remain_load[g] : 0;
        if (r < -2 || r > 2) return fail(h, NGS_ERR_RANGE, "node %u cell %d: remain_load %d out of range", idx, g, r);
        rl[g] = (int8_t)r;
    }
    CU(h, cudaSetDevice(h->device));
    if (gpu_count > NGS_MAX_GPUS) {
        const int rc = ensure_wide(h);
        if (rc) return rc;
    }
    const uint32_t row = idx - h->T.first;
    mark_width(h, row, gpu_count);
    h->epoch++;
    if (remain_load) h->T.rl_zero = 0;
    const uint8_t G8 = (uint8_t)gpu_count;
    // (pageable sources: the copies have left the host buffers when the calls return)
    CU(h, cudaMemcpyAsync(h->T.free + (size_t)row * 8, fr, 32, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(h->T.rl + (size_t)row * 8, rl, 8, cudaMemcpyHostToDevice, h->stream));
    if (h->T.free_hi) {
        CU(h, cudaMemcpyAsync(h->T.free_hi + (size_t)row * 4, fr + 8, 16, cudaMemcpyHostToDevice, h->stream));
        CU(h, cudaMemcpyAsync(h->T.rl_hi + (size_t)row * 4, rl + 8, 4, cudaMemcpyHostToDevice, h->stream));
    }
    CU(h, cudaMemcpyAsync(h->T.gcount + row, &G8, 1, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemsetAsync(h->T.tag + row, 0, 4, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.h2d_bytes += 61;
    return NGS_OK;
}

int ngs_node_get_wide(ngs_dealer* h, uint32_t idx, int* gpu_count, int32_t* free, int8_t* remain_load) {
    if (!h) return NGS_ERR_INVALID;
    SERVE_OFF(h);
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    const uint32_t row = idx - h->T.first;
    CU(h, cudaSetDevice(h->device));
    uint8_t G8 = 0;
    int32_t fr[NGS_MAX_GPUS_WIDE] = {0};
    int8_t rl[NGS_MAX_GPUS_WIDE] = {0};
    CU(h, cudaMemcpyAsync(&G8, h->T.gcount + row, 1, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(fr, h->T.free + (size_t)row * 8, 32, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(rl, h->T.rl + (size_t)row * 8, 8, cudaMemcpyDeviceToHost, h->stream));
    if (h->T.free_hi) {
        CU(h, cudaMemcpyAsync(fr + 8, h->T.free_hi + (size_t)row * 4, 16, cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaMemcpyAsync(rl + 8, h->T.rl_hi + (size_t)row * 4, 4, cudaMemcpyDeviceToHost, h->stream));
    }
    CU(h, cudaStreamSynchronize(h->stream));
    h->stats.d2h_bytes += 61;
    if (gpu_count) *gpu_count = G8;
    for (int g = 0; g < NGS_MAX_GPUS_WIDE; g++) {
        if (free) free[g] = g < (int)G8 ? fr[g] : 0;
        if (remain_load) remain_load[g] = g < (int)G8 ? rl[g] : 0;
    }
    return NGS_OK;
}

int ngs_snapshot_load(ngs_dealer* h, uint32_t first, uint32_t n, const uint8_t* gpu_count, const int32_t* free, const int8_t* remain_load) {
    if (!h || !gpu_count) return NGS_ERR_INVALID;
    SERVE_OFF(h);
    if (first < h->T.first || (uint64_t)(first - h->T.first) + n > h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "rows [%u,%u) outside this shard", first, first + n);
    if (n == 0) return NGS_OK;
    const uint32_t r0 = first - h->T.first;
    std::vector<int32_t> fr((size_t)n * 8);
    std::vector<int8_t> rl((size_t)n * 8, 0);
    for (uint32_t i = 0; i < n; i++) {
        const int G = gpu_count[i];
        if (G > NGS_MAX_GPUS) return fail(h, NGS_ERR_UNSUPPORTED, "row %u: gpu_count %d > NGS_MAX_GPUS (a node that wide goes through ngs_node_set_wide)", first + i, G);
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
    h->epoch++;
    for (uint32_t i = 0; i < n && h->n_wide; i++) mark_width(h, r0 + i, gpu_count[i]);
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
    SERVE_OFF(h);
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
    const bool resident = h->serve_verbs && h->world == 1;  // K8: a request to the resident kernel, which keeps running
    if (!resident) SERVE_OFF(h);
    PodDemand pd;
    int rc = make_demand(h, percent, n_containers, NGS_POLICY_BINPACK, &pd);
    if (rc) return rc;
    if (n_containers > 0 && !gpu_idx) return fail(h, NGS_ERR_INVALID, "apply: gpu_idx is NULL");
    CU(h, cudaSetDevice(h->device));
    if (resident) {
        ServeBody b;
        memset(&b, 0, sizeof(b));
        b.dem.C = n_containers;
        for (int i = 0; i < n_containers; i++) {
            b.dem.raw[i] = percent[i];
            b.dem.d[i] = gpu_idx[i];
        }
        b.op = SERVE_OP_APPLY;
        b.node = idx;
        b.mode = (uint32_t)(int32_t)(sign > 0 ? 1 : -1);
        h->epoch++;
        rc = serve_request(h, b);
        if (rc) return rc;
        rc = h->h_serve_out[1];
        if (rc) return fail(h, rc, "apply on node %u: %s", idx, ngs_strerror(rc));
        return NGS_OK;
    }
    for (int i = 0; i < n_containers; i++) h->h_small[i] = gpu_idx[i];
    CU(h, cudaMemcpyAsync(h->d_small, h->h_small, NGS_MAX_CONTAINERS * 4, cudaMemcpyHostToDevice, h->stream));
    h->epoch++;
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
    SERVE_OFF(h);
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
    h->epoch++;
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
    if (h->serve_running && h->serve_verbs && h->serve_kind_verbs) {  // K8: as a request, the resident kernel keeps running
        ServeBody b;
        memset(&b, 0, sizeof(b));
        b.op = SERVE_OP_LOAD;
        b.node = idx;
        b.mode = (uint32_t)metric;
        b.gen = 0xFFFFFFFFu;
        return serve_request(h, b);
    }
    SERVE_OFF(h);
    k_load_add_node<<<1, 32, 0, h->stream>>>(h->T, idx - h->T.first, metric);
    CU(h, cudaGetLastError());
    h->stats.kernel_launches++;
    return NGS_OK;
}

int ngs_load_update(ngs_dealer* h, uint32_t idx, int metric, int card, double value, int value_ok, int64_t ts_sec, int ts_ok) {
    if (!h || metric < 0 || metric > 1) return NGS_ERR_INVALID;
    if (idx < h->T.first || idx - h->T.first >= h->T.n) return fail(h, NGS_ERR_UNKNOWN_NODE, "node %u outside this shard", idx);
    if (card < 0) return fail(h, NGS_ERR_INVALID, "card %d", card);
    if (card >= NGS_MAX_GPUS_WIDE) return NGS_OK;  // a card the kernels never look at (LoadUsage walks gpus only)
    CU(h, cudaSetDevice(h->device));
    const uint8_t t10 = encode_tenths(value, value_ok, ts_sec, ts_ok);
    // K8: as a request while the resident kernel (verbs instantiation) runs - it keeps running; a card beyond 7 before the
    // second table plane exists takes the launched path, which allocates it
    if (h->serve_running && h->serve_verbs && h->serve_kind_verbs && (card < NGS_MAX_GPUS || h->T.free_hi)) {
        ServeBody b;
        memset(&b, 0, sizeof(b));
        b.op = SERVE_OP_LOAD;
        b.node = idx;
        b.mode = (uint32_t)metric;
        b.gen = (uint32_t)card;
        b.n = t10;
        b.idx_lo = t10 == 255 ? 0u : (uint32_t)ts_sec;
        return serve_request(h, b);
    }
    SERVE_OFF(h);
    if (card >= NGS_MAX_GPUS) {  // (an entry may arrive before its node does: nodeusage.go keeps them apart, too)
        const int rc = ensure_wide(h);
        if (rc) return rc;
    }
    k_load_update<<<1, 32, 0, h->stream>>>(h->T, idx - h->T.first, metric, card, t10, t10 == 255 ? 0u : (uint32_t)ts_sec);
    CU(h, cudaGetLastError());
    h->stats.kernel_launches++;
    return NGS_OK;
}

int ngs_load_update_bulk(ngs_dealer* h, int metric, uint32_t first, uint32_t n, const double* value, const uint8_t* value_ok,
                         const int64_t* ts_sec, const uint8_t* ts_ok) {
    if (!h || metric < 0 || metric > 1 || !value || !ts_sec) return NGS_ERR_INVALID;
    SERVE_OFF(h);
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
                       int32_t* d_score, const VerbDone& vd) {
    const uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (lp.enabled) {
        // load-aware pass: instruction-heavier per row; measured best with direct 256-bit loads at full
        // occupancy (k_assume, 73 % of HBM peak) rather than TMA staging at 2-3 CTAs per SM (66-69 %)
        int grid = (int)((n + 255) / 256);
        const int max_grid = h->num_sms * NGS_K1_BLOCKS;
        if (grid > max_grid) grid = max_grid;
        (h->n_wide ? k_assume<false, true> : k_assume<false, false>)<<<grid, 256, 0, h->stream>>>(h->T, nullptr, n, pd, lp, MODE_ASSUME, gen, d_feas, d_reason, d_score, nullptr, nullptr, vd, AssumePad{});
    } else {
        const size_t smem = (size_t)NGS_SCAN_STAGES * SCAN_STAGE_BYTES;
        uint32_t grid = (uint32_t)h->num_sms * NGS_SCAN_BLOCKS;
        if (grid > ntiles) grid = ntiles;
        (h->n_wide ? k_scan_tma<true> : k_scan_tma<false>)<<<grid, SCAN_TILE, smem, h->stream>>>(h->T, n, pd, lp, d_feas, d_reason, d_score, vd, ScanPad{});
    }
    CU(h, cudaGetLastError());
    return NGS_OK;
}

static int run_assume(ngs_dealer* h, int mode, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
                      const ngs_load_ctx* load, uint8_t* feasible, int8_t* reason, int32_t* score, int32_t* plan, int32_t* n_emitted,
                      const int32_t** score_view = nullptr, bool failed_plans_only = false) {
    if (!h) return NGS_ERR_INVALID;
    // K8: with ngs_serve_verbs on, the scan goes to the resident kernel as a request (no plans there: those calls stop it)
    const bool resident = h->serve_verbs && h->world == 1 && (plan != nullptr) == (n_emitted != nullptr);
    if (!resident || n > h->verb_cap || n > h->vstage_cap || (plan && n > h->vplan_cap)) SERVE_OFF(h);  // (growing the scratch frees memory: not under a running kernel)
    if (score_view) *score_view = nullptr;
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
    rc = ensure_verb_stage(h, n);
    if (rc) return rc;
    const uint32_t gen = demand_generation(h, pd, lp.enabled != 0);
    uint64_t sig = 0;
    if (idx) {
        if (h->idx_sig == 0 || h->idx_host.size() != n || memcmp(h->idx_host.data(), idx, (size_t)n * 4) != 0) {
            // a list the device does not hold yet
            h->idx_host.assign(idx, idx + n);
            if (resident) {  // the handle's stream belongs to the resident kernel: the list goes up on a stream of its own
                if (!h->copy_stream) CU(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
                CU(h, cudaMemcpyAsync(h->d_idx, h->idx_host.data(), (size_t)n * 4, cudaMemcpyHostToDevice, h->copy_stream));
                CU(h, cudaStreamSynchronize(h->copy_stream));
            } else {
                CU(h, cudaMemcpyAsync(h->d_idx, h->idx_host.data(), (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
            }
            h->stats.h2d_bytes += (size_t)n * 4;
            h->idx_sig = h->idx_sig + 1 ? h->idx_sig + 1 : 1;
        }
        sig = h->idx_sig;
    }
    // NodeInfo.Score right after NodeInfo.Assume: the filter call's scores are still in the staging area
    if (mode == MODE_SCORE && h->sc_valid && h->sc_gen == gen && h->sc_n == n && h->sc_sig == sig && h->sc_epoch == h->epoch) {
        if (score) memcpy(score, h->h_vscore, (size_t)n * 4);
        if (score_view) *score_view = h->h_vscore;
        h->stats.node_evals += n;
        return NGS_OK;
    }
    h->sc_valid = false;
    if (resident) {
        ServeBody b;
        memset(&b, 0, sizeof(b));
        sort_demand(pd.percent, pd.n, pd.policy, &b.dem);
        b.lp = lp;
        b.op = SERVE_OP_SCAN;
        b.mode = (uint32_t)mode;
        b.gen = gen;
        b.n = n;
        const unsigned long long p_idx = (unsigned long long)(idx ? h->d_idx : nullptr),
                                 p_reason = (unsigned long long)((reason || feasible || failed_plans_only) ? h->d_vreason : nullptr),
                                 p_score = (unsigned long long)h->d_vscore;
        b.idx_lo = (uint32_t)p_idx; b.idx_hi = (uint32_t)(p_idx >> 32);
        b.reason_lo = (uint32_t)p_reason; b.reason_hi = (uint32_t)(p_reason >> 32);
        b.score_lo = (uint32_t)p_score; b.score_hi = (uint32_t)(p_score >> 32);
        if (want_plan) {
            const unsigned long long p_plan = (unsigned long long)h->d_vplan, p_nem = (unsigned long long)h->d_vnem;
            b.plan_lo = (uint32_t)p_plan; b.plan_hi = (uint32_t)(p_plan >> 32);
            b.nem_lo = (uint32_t)p_nem; b.nem_hi = (uint32_t)(p_nem >> 32);
            if (failed_plans_only) b.mode |= MODE_FAILED_PLANS_ONLY;
        }
        rc = serve_request(h, b);
        if (rc) return rc;
        if (h->h_serve_out[1] != NGS_OK) return fail(h, h->h_serve_out[1], "resident scan: %s", ngs_strerror(h->h_serve_out[1]));
    } else {
        const int block = 256;
        int grid = (int)((n + block - 1) / block);
        const int max_grid = h->num_sms * (want_plan ? 2 : NGS_K1_BLOCKS);  // exactly the resident blocks: measured best
        if (grid > max_grid) grid = max_grid;
        // per-node outputs go straight into mapped host memory; the scores are always produced (an ngs_score usually follows)
        VerbDone vd;
        vd.ticket = h->d_vticket;
        vd.flag = h->d_vflag;
        vd.seq = ++h->vseq;
        // one status byte per node travels (NGS_REASON_*); `feasible` is derived from it on the host
        uint8_t* o_feas = nullptr;
        int8_t* o_reason = (reason || feasible || failed_plans_only) ? h->d_vreason : nullptr;
        int32_t* o_score = h->d_vscore;
        if (!idx && mode == MODE_ASSUME && !want_plan) {
            rc = launch_scan(h, n, pd, lp, gen, o_feas, o_reason, o_score, vd);
            if (rc) return rc;
        } else if (want_plan)
            (h->n_wide ? k_assume<true, true> : k_assume<true, false>)<<<grid, block, 0, h->stream>>>(h->T, idx ? h->d_idx : nullptr, n, pd, lp,
                                                         mode | (failed_plans_only ? MODE_FAILED_PLANS_ONLY : 0), gen, o_feas, o_reason, o_score,
                                                         plan ? h->d_vplan : nullptr, n_emitted ? h->d_vnem : nullptr, vd, AssumePad{});
        else
            (h->n_wide ? k_assume<false, true> : k_assume<false, false>)<<<grid, block, 0, h->stream>>>(h->T, idx ? h->d_idx : nullptr, n, pd, lp, mode, gen, o_feas, o_reason, o_score, nullptr,
                                                          nullptr, vd, AssumePad{});
        CU(h, cudaGetLastError());
        rc = wait_flag(h, h->h_vflag, vd.seq, "scan kernel");
        if (rc) return rc;
    }
    if (failed_plans_only) {
        // only the infeasible candidates' rows are written: filling 4 bytes per entry for all of them is what a call with plans costs
        const int8_t* st = h->h_vreason;
        auto take = [&](uint32_t i) {
            n_emitted[i] = h->h_vnem[i];
            for (int k = 0; k < n_containers; k++) {
                const int8_t v = h->h_vplan[(size_t)i * n_containers + k];
                plan[(size_t)i * n_containers + k] = v == PLAN_NONE ? INT32_MIN : (int32_t)v;
            }
        };
        uint32_t i = 0;
#if defined(__SSE2__)
        const __m128i bad = _mm_set1_epi8((char)NGS_REASON_INFEASIBLE);
        for (; i + 16 <= n; i += 16) {  // sixteen status bytes per step: mostly nothing to do
            unsigned m = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i*>(st + i)), bad));
            while (m) {
                take(i + (uint32_t)__builtin_ctz(m));
                m &= m - 1;
            }
        }
#endif
        for (; i < n; i++)
            if (st[i] == NGS_REASON_INFEASIBLE) take(i);
    } else {
        if (plan && n_containers > 0) widen_plan(plan, h->h_vplan, (size_t)n * n_containers);  // one byte per entry in the staging area
        if (n_emitted) widen_u8(n_emitted, h->h_vnem, n);
    }
    if (feasible) expand_feasible(feasible, h->h_vreason, n);
    if (reason) memcpy(reason, h->h_vreason, n);
    if (score) memcpy(score, h->h_vscore, (size_t)n * 4);
    if (score_view) *score_view = h->h_vscore;
    if (mode == MODE_ASSUME) {  // (a Score pass serves cached values per node: not a filter call's result)
        h->sc_valid = true;
        h->sc_gen = gen;
        h->sc_n = n;
        h->sc_sig = sig;
        h->sc_epoch = h->epoch;
    }
    if (!resident) h->stats.kernel_launches++;
    h->stats.node_evals += n;
    h->stats.d2h_bytes += ((feasible || reason) ? n : 0) + (size_t)n * 4 + (plan ? (size_t)n * n_containers : 0) + (n_emitted ? n : 0);
    return NGS_OK;
}

int ngs_assume(ngs_dealer* h, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
               const ngs_load_ctx* load, uint8_t* feasible, int8_t* reason, int32_t* score, int32_t* plan, int32_t* n_emitted) {
    return run_assume(h, MODE_ASSUME, idx, n, percent, n_containers, policy, load, feasible, reason, score, plan, n_emitted);
}

int ngs_assume_failed_plans(ngs_dealer* h, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
                            const ngs_load_ctx* load, uint8_t* feasible, int8_t* reason, int32_t* score, int32_t* plan, int32_t* n_emitted) {
    if (!plan || !n_emitted) return NGS_ERR_INVALID;
    return run_assume(h, MODE_ASSUME, idx, n, percent, n_containers, policy, load, feasible, reason, score, plan, n_emitted, nullptr, true);
}

// DealerImpl.Assume + Score over rows [0, n) with the per-node outputs left in device memory
// (no copies, no synchronisation): what a caller that keeps the arg-max on the GPU uses, and what
// bench.py times for the scan kernel's own roofline.
int ngs_assume_dev(ngs_dealer* h, uint32_t n, const int32_t* percent, int n_containers, int policy, const ngs_load_ctx* load,
                   uint8_t* d_feasible, int32_t* d_score) {
    if (!h) return NGS_ERR_INVALID;
    SERVE_OFF(h);
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
    VerbDone none;
    none.ticket = nullptr;
    none.flag = nullptr;
    none.seq = 0;
    h->sc_valid = false;  // (a load-aware scan rewrites the per-node plan cache)
    rc = launch_scan(h, n, pd, lp, gen, d_feasible, nullptr, d_score, none);
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

int ngs_score_view(ngs_dealer* h, const uint32_t* idx, uint32_t n, const int32_t* percent, int n_containers, int policy,
                   const ngs_load_ctx* load, const int32_t** view) {
    if (!view) return NGS_ERR_INVALID;
    return run_assume(h, MODE_SCORE, idx, n, percent, n_containers, policy, load, nullptr, nullptr, nullptr, nullptr, nullptr, view);
}

int ngs_bind(ngs_dealer* h, uint32_t idx, const int32_t* percent, int n_containers, int policy, const ngs_load_ctx* load,
             int32_t* gpu_idx_out, int32_t* score_out) {
    if (!h) return NGS_ERR_INVALID;
    const bool resident = h->serve_verbs && h->world == 1;  // K8: the bind travels to the resident kernel as a request
    if (!resident) SERVE_OFF(h);
    PodDemand pd;
    LoadParams lp;
    int rc = make_demand(h, percent, n_containers, policy, &pd);
    if (rc) return rc;
    rc = make_load(h, load, &lp);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    const uint32_t gen = demand_generation(h, pd, lp.enabled != 0);
    if (resident) {
        ServeBody b;
        memset(&b, 0, sizeof(b));
        sort_demand(pd.percent, pd.n, pd.policy, &b.dem);
        b.lp = lp;
        b.op = SERVE_OP_BIND;
        b.gen = gen;
        b.node = idx;
        h->epoch++;
        rc = serve_request(h, b);
        if (rc) return rc;
        volatile int32_t* out = h->h_serve_out;
        h->stats.node_evals++;
        h->stats.d2h_bytes += sizeof(BindResult);
        rc = out[1];
        if (rc) return fail(h, rc, "bind on node %u: %s", idx, ngs_strerror(rc));
        for (int i = 0; i < n_containers; i++)
            if (gpu_idx_out) gpu_idx_out[i] = out[10 + i];
        if (score_out) *score_out = out[9];
        h->stats.pods_scheduled++;
        return NGS_OK;
    }
    constexpr int32_t BIND_PENDING = 0x7FFFFFFE;
    *reinterpret_cast<volatile int32_t*>(&h->h_bind->status) = BIND_PENDING;
    h->epoch++;  // a row changes: scores kept from the last filter call are no longer the cache's
    (h->n_wide ? k_bind<true> : k_bind<false>)<<<1, 32, 0, h->stream>>>(h->T, idx, pd, lp, gen, h->d_bind);
    CU(h, cudaGetLastError());
    {   // the answer arrives in mapped host memory, status word last
        volatile int32_t* st = &h->h_bind->status;
        for (uint32_t spins = 1; *st == BIND_PENDING; spins++) {
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
            if ((spins & 0xFFFu) == 0) {
                const cudaError_t q = cudaStreamQuery(h->stream);
                if (q == cudaSuccess) break;
                if (q != cudaErrorNotReady) return fail(h, NGS_ERR_CUDA, "k_bind: %s", cudaGetErrorString(q));
            }
        }
        std::atomic_thread_fence(std::memory_order_acquire);
        if (*st == BIND_PENDING) return fail(h, NGS_ERR_CUDA, "k_bind ended without an answer");
    }
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
    SERVE_OFF(h);
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
    SERVE_OFF(h);
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
    if (n_pods == 1) {
        // one pod: K7, a single ordinary launch; the answer arrives in mapped host memory and the
        // status word behind it is polled, so the call returns without a stream synchronisation
        PodDemand pd;
        rc = make_demand(h, percent, n_containers[0], policy[0], &pd);
        if (rc) return rc;
        if (lp.enabled && policy[0] != NGS_POLICY_RANDOM) h->T.rl_zero = 0;  // the load-aware Rate writes RemainLoad
        volatile int32_t* hout = h->h_pod_out;
        hout[0] = ONE_PENDING;
        h->epoch++;  // the call binds the pod
        uint32_t grid = (h->T.n + 255) / 256;
        if (grid > (uint32_t)h->num_sms * 4) grid = (uint32_t)h->num_sms * 4;
        if (grid < 1) grid = 1;
        OneShard sh;
        memset(&sh, 0, sizeof(sh));
        sh.world = (uint32_t)h->world;
        sh.rank = (uint32_t)h->rank;
        sh.err = h->d_err;
        if (h->world > 1) {  // same parity discipline as K6 (launch_schedule): this call uses one half, the other is cleared now
            const size_t cur = (size_t)(h->xbatch & 1), nxt = cur ^ 1;
            CU(h, cudaMemsetAsync(h->d_xslots + nxt * XSLOT_HALF, 0, sizeof(unsigned long long) * XSLOT_HALF, h->stream));
            for (int r = 0; r < h->world; r++) sh.xs[r] = h->peer_xslots[r] + cur * XSLOT_HALF;
            h->xbatch++;
        }
        if (policy[0] == NGS_POLICY_RANDOM)
            (h->n_wide ? k_schedule_one_random<true> : k_schedule_one_random<false>)<<<grid, 256, 0, h->stream>>>(h->T, pd, seed + seq0, h->d_one, reinterpret_cast<uint32_t*>(h->d_one + 1),
                                                               h->d_pod_out_dev, max_c, sh);
        else
            (h->n_wide ? k_schedule_one<true> : k_schedule_one<false>)<<<grid, 256, 0, h->stream>>>(h->T, pd, lp, h->d_one, h->d_pod_out_dev, max_c, sh, OnePad{});
        CU(h, cudaGetLastError());
        for (uint32_t spins = 1; hout[0] == ONE_PENDING; spins++) {
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
            if ((spins & 0xFFFu) == 0) {  // now and then make sure the kernel is still alive
                const cudaError_t q = cudaStreamQuery(h->stream);
                if (q == cudaSuccess) break;
                if (q != cudaErrorNotReady) return fail(h, NGS_ERR_CUDA, "k_schedule_one: %s", cudaGetErrorString(q));
            }
        }
        std::atomic_thread_fence(std::memory_order_acquire);
        if (hout[0] == NGS_ERR_TIMEOUT) {
            cudaMemsetAsync(h->d_err, 0, 8, h->stream);
            return fail(h, NGS_ERR_TIMEOUT, "single-pod kernel gave up waiting for a peer GPU");
        }
        if (hout[0] != NGS_OK) return fail(h, NGS_ERR_CUDA, "k_schedule_one ended without an answer");
        node_out[0] = hout[8];
        score_out[0] = hout[9];
        for (int k = 0; k < max_c; k++) gpu_idx_out[k] = hout[10 + k];
        h->stats.kernel_launches += 1;
        h->stats.node_evals += h->T.n;
        h->stats.pods_scheduled += 1;
        h->stats.h2d_bytes += sizeof(PodDemand);
        h->stats.d2h_bytes += (size_t)(3 + max_c) * 4;
        return NGS_OK;
    }
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

// K8: ngs_schedule_batch for one pod through the resident kernel (ngs_k8_serve.cuh)
int ngs_serve_pod(ngs_dealer* h, const int32_t* percent, int n_containers, int max_c, int policy, const ngs_load_ctx* load, uint64_t seed,
                  uint64_t seq0, int32_t* node_out, int32_t* score_out, int32_t* gpu_idx_out) {
    if (!h || !percent || !node_out || !score_out || !gpu_idx_out) return NGS_ERR_INVALID;
    if (max_c < 1 || max_c > NGS_MAX_CONTAINERS) return fail(h, NGS_ERR_UNSUPPORTED, "max_c %d outside [1, NGS_MAX_CONTAINERS]", max_c);
    if (n_containers < 0 || n_containers > max_c) return fail(h, NGS_ERR_INVALID, "%d containers", n_containers);
    if (policy == NGS_POLICY_RANDOM || h->world > 1) {  // not what the resident kernel does: the launched form (K7)
        const int32_t nc = n_containers, pol = policy;
        return ngs_schedule_batch(h, 1, percent, &nc, max_c, &pol, load, seed, seq0, node_out, score_out, gpu_idx_out);
    }
    ServeBody b;
    memset(&b, 0, sizeof(b));
    PodDemand pd;
    int rc = make_demand(h, percent, n_containers, policy, &pd);
    if (rc) return rc;
    sort_demand(pd.percent, pd.n, pd.policy, &b.dem);
    rc = make_load(h, load, &b.lp);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    if (b.lp.enabled) h->T.rl_zero = 0;  // the load-aware Rate writes RemainLoad
    b.max_c = (uint32_t)max_c;
    b.op = SERVE_OP_POD;
    b.prev_row = h->serve_prev_row;
    h->serve_prev_row = 0xFFFFFFFFu;
    h->epoch++;  // the call binds the pod
    rc = serve_request(h, b);
    if (rc) return rc;
    volatile int32_t* out = h->h_serve_out;
    rc = out[1];
    *node_out = out[8];
    if (rc == NGS_OK && out[8] >= 0) h->serve_prev_row = (uint32_t)out[8] - h->T.first;
    *score_out = out[9];
    for (int k = 0; k < max_c; k++) gpu_idx_out[k] = out[10 + k];
    h->stats.node_evals += h->T.n;
    h->stats.pods_scheduled += 1;
    h->stats.d2h_bytes += (size_t)(2 + max_c) * 4;
    h->stats.h2d_bytes += sizeof(ServeBody);
    if (rc != NGS_OK) return fail(h, rc, "resident kernel: %s", ngs_strerror(rc));
    return NGS_OK;
}

int ngs_serve_stop(ngs_dealer* h) {
    if (!h) return NGS_ERR_INVALID;
    h->serve_verbs = false;
    return serve_stop(h);
}

int ngs_serve_verbs(ngs_dealer* h, int on) {
    if (!h) return NGS_ERR_INVALID;
    if (on && h->world > 1) return fail(h, NGS_ERR_UNSUPPORTED, "the resident kernel serves unsharded handles");
    h->serve_verbs = on != 0;
    return on ? NGS_OK : serve_stop(h);
}

int ngs_serve_idle_ms(ngs_dealer* h, double ms) {
    if (!h || !(ms > 0) || ms > 2000) return NGS_ERR_INVALID;
    const int rc = serve_stop(h);
    h->serve_idle_ms = ms;
    return rc;
}

int ngs_xchg_export(ngs_dealer* h, void* handle_out) {
    if (!h || !handle_out) return NGS_ERR_INVALID;
    SERVE_OFF(h);
    static_assert(sizeof(cudaIpcMemHandle_t) == NGS_IPC_HANDLE_BYTES, "IPC handle size");
    CU(h, cudaSetDevice(h->device));
    if (!h->d_xslots) {
        CU(h, cudaMalloc(&h->d_xslots, sizeof(unsigned long long) * 2 * XSLOT_HALF));
        CU(h, cudaMemset(h->d_xslots, 0, sizeof(unsigned long long) * 2 * XSLOT_HALF));
        CU(h, cudaDeviceSynchronize());  // cleared before any peer can learn the handle
    }
    cudaIpcMemHandle_t mh;
    CU(h, cudaIpcGetMemHandle(&mh, h->d_xslots));
    memcpy(handle_out, &mh, sizeof(mh));
    return NGS_OK;
}

int ngs_xchg_attach(ngs_dealer* h, int rank, int world, const void* handles) {
    if (!h || !handles || world < 1 || world > NGS_MAX_RANKS || rank < 0 || rank >= world) return NGS_ERR_INVALID;
    SERVE_OFF(h);
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
constexpr size_t NA_MX_BYTES = 64 * (size_t)NGS_NOALLOC_MAX_PODS;                                                   // arg-max round records
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

// once per device, with that device current: resolve the scratch array's address there and opt the batch kernels in to
// the device's full shared memory (driver calls that do not belong on the per-call path)
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
            cudaDeviceGetAttribute(&d->coop, cudaDevAttrCooperativeLaunch, device) != cudaSuccess ||
            // (an upper bound per device: asking for the device maximum once keeps it valid for every later launch, also of handles)
            schedule_allow_max_smem(d->smem_optin) != cudaSuccess) {
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
    {
        // whole rows are read with vector loads: a table at an odd offset would fault inside the kernel
        auto misaligned = [](const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) != 0; };
        if (misaligned(snap->free, 32) || misaligned(snap->ts, 32) || misaligned(snap->tenths, 8) || misaligned(snap->remain_load, 8) ||
            misaligned(snap->plan_tag, 4) || misaligned(snap->plan_score, 4) || misaligned(d_percent, 4) || misaligned(d_n_containers, 4) ||
            misaligned(d_policy, 4) || misaligned(d_node_out, 4) || misaligned(d_score_out, 4) || misaligned(d_gpu_idx_out, 4))
            return NGS_ERR_INVALID;
    }
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
    const uint32_t block = schedule_block(rpb);
    const size_t smem = schedule_smem_bytes(rpb);
    if (smem + K6_STATIC_SMEM > (size_t)dv->smem_optin) return NGS_ERR_UNSUPPORTED;

    unsigned char* sc = dv->scratch;
    SortedDemand* dem = reinterpret_cast<SortedDemand*>(sc + NA_DEM_OFF);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(sc + NA_KEYS_OFF);
    unsigned long long* mx = reinterpret_cast<unsigned long long*>(sc + NA_MX_OFF);
    uint32_t* done = reinterpret_cast<uint32_t*>(sc + NA_DONE_OFF);
    int32_t* err = reinterpret_cast<int32_t*>(sc + NA_ERR_OFF);
    // the array still holds the previous call's words: everything the kernel expects to find at
    // zero is cleared here, in stream order, before the launch
    if (cudaMemsetAsync(keys, 0, sizeof(unsigned long long) * (size_t)n_pods * grid, st) != cudaSuccess ||
        cudaMemsetAsync(mx, 0, 64 * (size_t)n_pods, st) != cudaSuccess ||
        cudaMemsetAsync(done, 0, sizeof(uint32_t) * (size_t)n_pods, st) != cudaSuccess ||
        cudaMemsetAsync(err, 0, 2 * sizeof(int32_t), st) != cudaSuccess)
        return NGS_ERR_CUDA;
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
    bp.rr = mx;
    bp.keys_stride = grid;
    bp.done_stride = 1;
    bp.rr_stride = 8;
    bp.status_out = nullptr;
    bp.world = 1;
    bp.rank = 0;
    bp.err = err;
    bp.rows_per_block = rpb;
    void* args[] = {&bp};
    if (cudaLaunchCooperativeKernel(schedule_kernel(rpb, false), dim3(grid), dim3(block), args, smem, st) != cudaSuccess) return NGS_ERR_CUDA;
    int32_t dev_status = NGS_OK;  // device-side validation / timeout word of this batch
    if (cudaMemcpyAsync(&dev_status, err, sizeof(dev_status), cudaMemcpyDeviceToHost, st) != cudaSuccess) return NGS_ERR_CUDA;
    if (cudaStreamSynchronize(st) != cudaSuccess) return NGS_ERR_CUDA;
    return dev_status;
}
