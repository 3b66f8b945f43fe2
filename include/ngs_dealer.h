/*
 * ngs_dealer.h — C ABI of the B200-native dealer (filter / prioritize / bind evaluation).
 *
 * This is the drop-in boundary for the one data-parallel hot path of nano-gpu-scheduler:
 * the numeric part of pkg/dealer (Dealer interface, /root/reference/pkg/dealer/dealer.go:23-43).
 * The Go host keeps node names, pod maps, apiserver calls and error strings; everything that
 * evaluates or mutates per-GPU capacity goes through the calls below.  Nodes are dense
 * uint32 indices (the shim owns the name<->index map).  All outputs are caller-allocated;
 * the library never retains a caller pointer after a call returns (cgo pointer rules).
 *
 * Every entry point returns an NGS_* status (0 = ok).  One call at a time per handle (the
 * reference serialises every Dealer method behind one mutex, dealer.go:81,90,139,...); calls
 * may come from any OS thread.  There is no CPU fallback: without a CUDA device every call
 * that touches the snapshot fails with NGS_ERR_CUDA.
 *
 * Citations are relative to /root/reference/pkg/dealer/ unless stated.
 */
#ifndef NGS_DEALER_H
#define NGS_DEALER_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NGS_ABI_VERSION 3u /* 2: ngs_load_ctx holds 8 periods; NGS_NODE_INVALID_POD.  3: nodes of 9..12 GPUs (ngs_node_*_wide), ngs_score_view, ngs_serve_* */

#define NGS_MAX_GPUS 8         /* cells per row of the bulk arrays (row = 8 x int32 = 2 x int4): what a node usually has */
#define NGS_MAX_GPUS_WIDE 12   /* most GPUs a node may have (Go sort.Sort's <=12-element regime, as for containers); cells
                                * 8..11 travel through ngs_node_set_wide / ngs_node_get_wide */
#define NGS_MAX_CONTAINERS 12  /* containers per pod (Go sort.Sort's <=12-element regime) */
#define NGS_MAX_PERIODS 8      /* usable PolicySpec.SyncPeriod entries (the deployed policy has 2; more than 8 -> NGS_ERR_UNSUPPORTED) */
#define NGS_MAX_FREE 65535     /* upper bound of a cell's free percent kept in the snapshot */
#define NGS_MAX_BATCH 65536    /* pods per ngs_schedule_batch call */
#define NGS_MAX_RANKS 8        /* GPUs of one box a cluster can be sharded over */
#define NGS_IPC_HANDLE_BYTES 64

/* status codes */
enum {
    NGS_OK = 0,
    NGS_ERR_INVALID = 1,      /* bad argument (null pointer, count out of range, negative percent) */
    NGS_ERR_CUDA = 2,         /* CUDA runtime error; ngs_last_error() has the text */
    NGS_ERR_UNKNOWN_NODE = 3, /* node index out of range or never upserted */
    NGS_ERR_INFEASIBLE = 4,   /* bind: "assume ... failed" (node.go:74-76) */
    NGS_ERR_APPLY = 5,        /* apply: "can't apply plan" (allocate.go:113) / index >= G (allocate.go:126) */
    NGS_ERR_RANGE = 6,        /* a cell would leave [0, NGS_MAX_FREE] */
    NGS_ERR_UNSUPPORTED = 7,  /* gpu_count > NGS_MAX_GPUS_WIDE, too many containers or periods */
    NGS_ERR_TIMEOUT = 8,      /* the batch kernel gave up waiting for a peer (never expected) */
    NGS_ERR_NOALLOC_LIMIT = 9 /* ngs_schedule_batch_noalloc: batch or grid larger than its fixed scratch */
};

/* policies: Binpack (rater.go:52), Spread (rater.go:55), random = SampleRater plan search
 * (rater.go:21-50) on a scratch copy + seeded uniform pick over the feasible set */
enum { NGS_POLICY_BINPACK = 0, NGS_POLICY_SPREAD = 1, NGS_POLICY_RANDOM = 2 };

/* per-node reason codes of ngs_assume (the shim renders the FailedNodes strings) */
enum {
    NGS_REASON_OK = 0,
    NGS_REASON_UNKNOWN_NODE = 1, /* "nano gpu scheduler get node failed" (dealer.go:102) */
    NGS_REASON_INFEASIBLE = 2    /* "can't allocate ... indexes=" (rater.go:99,152) */
};

/* load metric tables: DealerImpl.CoreUsage / MemoryUsage (dealer.go:84-85) */
enum { NGS_METRIC_CORE = 0, NGS_METRIC_MEMORY = 1 };

#define NGS_NOT_NEED_GPU (-1) /* allocate.go:15 */
#define NGS_NODE_NONE (-1)         /* node_out: no node takes the pod */
#define NGS_NODE_INVALID_POD (-2)  /* node_out: the pod itself was malformed (device-resident batches only, see below) */
#define NGS_SCORE_MIN 0       /* rater.go:12 */

/* isLoadSchedule=true context: PolicySpec.SyncPeriod (type.go:16-27) after the host resolved
 * getActiveDuration (stats.go:57-66) for each entry, plus the evaluation's time.Now()
 * (stats.go:48).  Passing NULL where a ngs_load_ctx* is expected means isLoadSchedule=false. */
typedef struct {
    int32_t n_periods;                   /* entries whose active duration is non-zero, in order */
    int32_t metric[NGS_MAX_PERIODS];     /* NGS_METRIC_CORE iff name == "gpu_core_usage_avg" (nodeusage.go:84) */
    int64_t window_ns[NGS_MAX_PERIODS];  /* period + ExtenderAtivePeriod (type.go:6) */
    int64_t now_ns;                      /* unix nanoseconds */
} ngs_load_ctx;

typedef struct {
    uint64_t node_evals;    /* (pod, node) evaluations done so far */
    uint64_t kernel_launches;
    uint64_t pods_scheduled;
    uint64_t h2d_bytes;
    uint64_t d2h_bytes;
} ngs_stats_t;

typedef struct ngs_dealer ngs_dealer;

uint32_t ngs_abi_version(void);
const char *ngs_strerror(int status);
const char *ngs_last_error(const ngs_dealer *h);

/* replaces NewDealer's cache construction (dealer.go:45-75): an empty snapshot of max_nodes rows
 * resident on CUDA device `device`.  stream: a cudaStream_t the handle should use, or NULL to
 * create its own.  first_node: global index of row 0 (node-range sharding). */
int ngs_create(ngs_dealer **out, int device, uint32_t max_nodes, uint32_t first_node, void *stream);
int ngs_destroy(ngs_dealer *h);
void *ngs_stream(ngs_dealer *h); /* the cudaStream_t every kernel of this handle is launched on */
int ngs_sync(ngs_dealer *h);
int ngs_stats(ngs_dealer *h, ngs_stats_t *out);

/* ---- snapshot ownership ---- */
/* NewNodeInfo (node.go:25-42): gpu_count cells of (100,100,0); gpu_count from
 * capacity["nano-gpu/gpu-percent"]/100 (pkg/utils/node.go:8-14). */
int ngs_node_upsert(ngs_dealer *h, uint32_t idx, int gpu_count);
/* Nodes with 9..12 GPUs (the reference takes any count, pkg/utils/node.go:8-14; above 12 Go 1.16's sort.Sort leaves
 * its insertion-sort regime and NGS_ERR_UNSUPPORTED comes back).  ngs_node_upsert registers them like any node;
 * these two set / read one node of ANY width with all its cells: free[gpu_count] (NULL = all 100),
 * remain_load[gpu_count] (NULL = 0) — the per-node form of ngs_snapshot_load / ngs_snapshot_read, whose
 * [n][NGS_MAX_GPUS] arrays hold cells 0..7 only (ngs_snapshot_load refuses gpu_count > NGS_MAX_GPUS,
 * ngs_snapshot_read returns such a node's gpu_count and its first eight cells).  Every verb (assume, score, bind,
 * apply, release, load updates with card < 12) and ngs_schedule_batch / _dev handle such nodes: their rows are
 * evaluated by one thread each, from the global tables, so a snapshot full of them runs several times slower than
 * one of 8-GPU nodes.  ngs_schedule_batch_noalloc takes caller-owned tables of 8 cells per row: a node with more
 * GPUs is left out there and NGS_ERR_UNSUPPORTED comes back. */
int ngs_node_set_wide(ngs_dealer *h, uint32_t idx, int gpu_count, const int32_t *free, const int8_t *remain_load);
int ngs_node_get_wide(ngs_dealer *h, uint32_t idx, int *gpu_count, int32_t *free, int8_t *remain_load); /* free / remain_load: [NGS_MAX_GPUS_WIDE] */
/* bulk rebuild on start (dealer.go:58-72,279-299 replayed host-side, then one H2D copy).
 * free is [n][NGS_MAX_GPUS] (NULL = all 100), remain_load [n][NGS_MAX_GPUS] (NULL = 0). */
int ngs_snapshot_load(ngs_dealer *h, uint32_t first, uint32_t n, const uint8_t *gpu_count, const int32_t *free,
                      const int8_t *remain_load);
/* Status() / PrintStatus (dealer.go:303-309,321-323): read rows back. Any output may be NULL. */
int ngs_snapshot_read(ngs_dealer *h, uint32_t first, uint32_t n, uint8_t *gpu_count, int32_t *free, int8_t *remain_load);
/* GPUs.Allocate (sign=+1, allocate.go:102-118; all-or-nothing) / GPUs.Release (sign=-1,
 * allocate.go:120-131) via NodeInfo.Allocate/Release (node.go:86-94).  Drops the node's cached plan. */
int ngs_node_apply(ngs_dealer *h, uint32_t idx, const int32_t *gpu_idx, const int32_t *percent, int n_containers, int sign);

/* DealerImpl.Release for n finished pods in one call (dealer.go:230-255 -> node.go:91-94 ->
 * allocate.go:120-131; the informer path pkg/controller/controller.go:210-243 handles them one by
 * one).  node[n], gpu_idx[n][max_c], percent[n][max_c], n_containers[n]; status_out[n] (nullable)
 * receives each pod's status.  Releases only add and additions commute, so the resulting snapshot
 * equals releasing the pods one after the other.  Returns the first non-OK pod status, if any. */
int ngs_release_batch(ngs_dealer *h, uint32_t n, const uint32_t *node, const int32_t *gpu_idx, const int32_t *percent,
                      const int32_t *n_containers, int max_c, int32_t *status_out);

/* ---- load-usage tables ---- */
/* AddCoreUsage / AddMemoryUsage (nodeusage.go:58-68): fresh empty per-card map for the node */
int ngs_load_add_node(ngs_dealer *h, uint32_t idx, int metric);
/* UpdateCoreUsage / UpdateMemoryUsage (nodeusage.go:70-80) after string parsing: value_ok =
 * strconv.ParseFloat succeeded (nodeusage.go:103), ts_ok = the time string parsed (stats.go:32-44). */
int ngs_load_update(ngs_dealer *h, uint32_t idx, int metric, int card, double value, int value_ok, int64_t ts_sec, int ts_ok);
/* bulk form of the above for [n][NGS_MAX_GPUS] cells (NULL flags = all ok) */
int ngs_load_update_bulk(ngs_dealer *h, int metric, uint32_t first, uint32_t n, const double *value,
                         const uint8_t *value_ok, const int64_t *ts_sec, const uint8_t *ts_ok);
/* host-side parsing helpers for the two strings the metric sync hands over
 * (pkg/controller/node.go:111-135, pkg/prometheus/prometheus.go:57-62) */
int ngs_parse_usage(const char *s, double *value_out);      /* strconv.ParseFloat(s, 64); 1 = ok */
int ngs_parse_update_time(const char *s, int64_t *ts_out);  /* "2006-01-02T15:04:05Z" in Asia/Shanghai; 1 = ok */

/* ---- verbs ---- */
/* DealerImpl.Assume (dealer.go:89-136): per candidate NodeInfo.Assume -> GPUs.Choose = Rate then
 * Choose (node.go:44-57, allocate.go:92-100).  idx == NULL means rows 0..n-1.  Outputs (each may
 * be NULL): feasible[n], reason[n], score[n] (Plan.Score; NGS_SCORE_MIN when not feasible,
 * node.go:59-68), plan[n][n_containers] (GPUIndexes when feasible, else the partial `indexes` of
 * rater.go:99 in emission order padded with INT32_MIN), n_emitted[n]. */
int ngs_assume(ngs_dealer *h, const uint32_t *idx, uint32_t n, const int32_t *percent, int n_containers, int policy,
               const ngs_load_ctx *load, uint8_t *feasible, int8_t *reason, int32_t *score, int32_t *plan,
               int32_t *n_emitted);
/* ngs_assume for a caller that wants plans only to render the FailedNodes texts (predicate.go:30, rater.go:99,152):
 * plan / n_emitted (both required) are written for the INFEASIBLE candidates only - their partial `indexes` - and the
 * rows of every other candidate are LEFT AS THEY WERE (look at reason[] / feasible[] first).  The others' exact plan search
 * is skipped and their 4 bytes per entry are never written, so the call costs about what one without plans costs, where
 * ngs_assume with plans costs about twice as much at 100k candidates. */
int ngs_assume_failed_plans(ngs_dealer *h, const uint32_t *idx, uint32_t n, const int32_t *percent, int n_containers,
                            int policy, const ngs_load_ctx *load, uint8_t *feasible, int8_t *reason, int32_t *score,
                            int32_t *plan, int32_t *n_emitted);
/* ngs_assume over rows 0..n-1 with feasible[n] / score[n] (either may be NULL) left in DEVICE memory:
 * no copies and no synchronisation; results are valid after ngs_sync() (dealer.go:89-153). */
int ngs_assume_dev(ngs_dealer *h, uint32_t n, const int32_t *percent, int n_containers, int policy,
                   const ngs_load_ctx *load, uint8_t *d_feasible, int32_t *d_score);
/* DealerImpl.Score (dealer.go:138-153): cached Plan.Score of the last ngs_assume with the same
 * demand, else a fresh evaluation; unknown / infeasible -> NGS_SCORE_MIN. */
int ngs_score(ngs_dealer *h, const uint32_t *idx, uint32_t n, const int32_t *percent, int n_containers, int policy,
              const ngs_load_ctx *load, int32_t *score);
/* ngs_score without the copy into a caller's buffer: *view receives a pointer to the n scores where the kernels left them
 * (pinned host memory owned by the handle), readable until the next call on this handle.  What a binding that converts the
 * scores element by element anyway (HostPriorityList, priority.go:29-37) reads from directly. */
int ngs_score_view(ngs_dealer *h, const uint32_t *idx, uint32_t n, const int32_t *percent, int n_containers, int policy,
                   const ngs_load_ctx *load, const int32_t **view);
/* NodeInfo.Bind (node.go:70-84): cached or fresh plan -> GPUs.Allocate -> drop the node's cache. */
int ngs_bind(ngs_dealer *h, uint32_t idx, const int32_t *percent, int n_containers, int policy, const ngs_load_ctx *load,
             int32_t *gpu_idx_out, int32_t *score_out);
/* One launch for P pods in order: filter + prioritize over every row, arg-max (highest score,
 * ties -> lowest node index; NGS_POLICY_RANDOM: splitmix64(seed+seq0+p) mod |feasible|), bind.
 * Equivalent to P x (Assume, Score, pick, Bind) on the reference dealer.  percent is
 * [P][max_c]; node_out[p] = global node index or -1; gpu_idx_out [P][max_c] padded with INT32_MIN.
 * A call with ONE pod is the low-latency form: a single ordinary launch over the tables in place,
 * the answer polled from mapped host memory (sharded handles: every rank makes the same call and
 * gets the whole decision, every policy).
 * Two things a caller has to know about a batch:
 *  - load-aware mode: `load->now_ns` and the usage tables are read ONCE, when the launch starts
 *    (LoadUsage of every row is evaluated in the kernel's prologue).  The whole batch - up to
 *    NGS_MAX_BATCH pods, in chunks of 4096 per launch - is one "pod cycle" in the sense of
 *    SURVEY hard part 6: an ngs_load_update between two pods of one batch is not seen, and no pod
 *    of the batch crosses an update-window edge.  Callers that need a fresher clock cut the batch.
 *  - size: the batch kernel keeps every row of the shard in shared memory (32 B per row, one CTA
 *    per SM), which holds about 1.04 M rows on a B200 (148 SMs x 227 KB).  A larger shard gets
 *    NGS_ERR_UNSUPPORTED from this call - there is no streaming fallback; split the cluster over
 *    more handles / GPUs (node-range sharding below).  The verb calls have no such limit.
 * The host-buffer form validates every pod before anything is launched (NGS_ERR_INVALID, nothing
 * scheduled). */
int ngs_schedule_batch(ngs_dealer *h, uint32_t n_pods, const int32_t *percent, const int32_t *n_containers, int max_c,
                       const int32_t *policy, const ngs_load_ctx *load, uint64_t seed, uint64_t seq0, int32_t *node_out,
                       int32_t *score_out, int32_t *gpu_idx_out);
/* same, with every array already resident in device memory and no synchronisation:
 * results are valid after ngs_sync() or once work queued later on ngs_stream() runs.
 * The pods cannot be checked on the host here, so the device does it: a pod with a bad container
 * count (< 0, > max_c), percent (< 0, > NGS_MAX_FREE) or policy is left unscheduled with
 * node_out = NGS_NODE_INVALID_POD, score NGS_SCORE_MIN and gpu_idx INT32_MIN; every other pod of the
 * batch is scheduled exactly as if the bad one were not there, and the next ngs_sync() returns
 * NGS_ERR_INVALID once to say that some batch since the last sync held such a pod. */
int ngs_schedule_batch_dev(ngs_dealer *h, uint32_t n_pods, const int32_t *d_percent, const int32_t *d_n_containers,
                           int max_c, const int32_t *d_policy, const ngs_load_ctx *load, uint64_t seed, uint64_t seq0,
                           int32_t *d_node_out, int32_t *d_score_out, int32_t *d_gpu_idx_out);

/* ---- the resident form of the one-pod call (no reference counterpart) ----
 * ngs_schedule_batch with one pod costs one kernel launch per pod, and on this hardware a launch whose result the host
 * can see takes longer than the kernel's own work.  ngs_serve_pod gives the same answer (same arguments as one pod of
 * ngs_schedule_batch, bit for bit the same decision and table update) through a kernel that is launched once and then
 * polls a request block in pinned host memory: the call writes the request there and spins on the answer.
 *  - The kernel stays resident (it occupies the GPU's SMs with spinning CTAs) until ngs_serve_stop, until any other
 *    entry point of this handle is called (each of them stops it first, which costs that call a stream synchronisation),
 *    or until no request came for ngs_serve_idle_ms (default 20 ms); the next ngs_serve_pod starts it again (one
 *    cooperative launch).  Work other handles or the caller queue on this GPU meanwhile runs only in whatever the
 *    resident CTAs leave free, and work queued on ngs_stream() waits for the kernel to leave.
 *  - Arg-max policies on unsharded handles are served by the resident kernel; NGS_POLICY_RANDOM and sharded handles
 *    take the launched path (ngs_schedule_batch) from inside the call.
 *  - One caller at a time, like every call on a handle. */
int ngs_serve_pod(ngs_dealer *h, const int32_t *percent, int n_containers, int max_c, int policy, const ngs_load_ctx *load,
                  uint64_t seed, uint64_t seq0, int32_t *node_out, int32_t *score_out, int32_t *gpu_idx_out);
int ngs_serve_stop(ngs_dealer *h);               /* also switches ngs_serve_verbs off */
int ngs_serve_idle_ms(ngs_dealer *h, double ms); /* 0 < ms <= 2000; stops the kernel if it runs */
/* on != 0: from now on ngs_assume[_failed_plans] / ngs_score / ngs_score_view, ngs_bind, ngs_node_apply, ngs_load_update and
 * ngs_load_add_node of this handle are served by the resident kernel as requests, too - same results, no launch per call -
 * and do not stop it; it is started by the first verb and again after every call that had to stop it (bulk snapshot and
 * load-table updates, node upserts, batches, ngs_sync, growing scratch).  Unsharded handles.  The occupancy note above
 * applies for as long as this is on. */
int ngs_serve_verbs(ngs_dealer *h, int on);

/* ---- node-range sharding over the GPUs of one box (no reference counterpart: the reference is a
 * single process) ----
 * One process and one handle per GPU, each created with its own first_node / max_nodes range.
 * ngs_schedule_batch[_dev] then has to be called by every rank with the same pods: each rank scans
 * its own rows, the per-pod best (score, node) keys are traded between the ranks' kernels through
 * peer-mapped memory over NVLink, and the rank owning the winner binds it.  On return node_out and
 * score_out are complete on every rank; gpu_idx_out holds the plan only on the owning rank and
 * INT32_MIN elsewhere (merge with an element-wise max, e.g. one small NCCL allreduce-max).
 *   ngs_xchg_export: writes this rank's NGS_IPC_HANDLE_BYTES-byte handle of its exchange buffer;
 *   ngs_xchg_attach: takes the `world` handles gathered from all ranks (rank-major) and maps them. */
int ngs_xchg_export(ngs_dealer *h, void *handle_out);
int ngs_xchg_attach(ngs_dealer *h, int rank, int world, const void *handles);

/* ---- scheduling without any allocation (no counterpart in the reference) ----
 * Everything above works on a handle that owns its device memory (ngs_create calls cudaMalloc, and
 * the batch path grows scratch buffers on first use).  Some hosts of this library are not allowed
 * to do that: an extender embedded in a process whose whole device arena was handed to another
 * allocator (a co-located serving runtime, a fixed pool owned by the embedding application), where
 * a cudaMalloc of ours can fail or fragment that pool; and callers on the scheduling fast path that
 * must never hit the driver's allocator, because cudaMalloc / cudaFree synchronise the whole
 * device and take a process-wide lock, stalling every other stream while a pod waits for its
 * decision.  For them there is this one call.  It is ngs_schedule_batch_dev without a handle:
 * the caller owns the snapshot tables (described by ngs_snapshot_view, laid out exactly as the
 * handle keeps them) and passes the pods and the outputs as device pointers; the scratch the batch
 * kernel needs (sorted demands, per-pod exchange words, error word) lives in a fixed array inside
 * the library's device image, 386 304 bytes (about 377 KiB) that every GPU loading the library keeps
 * until exit.
 *
 * Limits: at most NGS_NOALLOC_MAX_PODS pods per call and a launch of at most
 * NGS_NOALLOC_MAX_CTAS CTAs (one per SM); anything larger is refused with NGS_ERR_NOALLOC_LIMIT
 * and nothing is launched.  One GPU only (no node-range sharding).
 * The call clears the part of the scratch that has to start at zero, queues the work on `stream`
 * (NULL = the default stream), waits for it with cudaStreamSynchronize - every call does, not only
 * the first - and returns the batch's status (NGS_ERR_INVALID if a pod was malformed: see
 * NGS_NODE_INVALID_POD above).  It calls no cudaMalloc, cudaFree, stream / event creation or any
 * other allocating routine.
 * Alignment: the kernels read whole rows with vector loads, so `free` and `ts` must be 32-byte
 * aligned, `tenths` and `remain_load` 8-byte aligned, `plan_tag` / `plan_score` 4-byte aligned
 * (anything cudaMalloc returns is; a sub-allocation at an odd offset may not be).  A view that is
 * not gets NGS_ERR_INVALID and nothing is launched.
 * The first call on each device looks up the scratch array's device address
 * (cudaGetSymbolAddress); make that first call outside of any stream capture.
 * All calls on a device share the one scratch array, which keeps the previous call's contents:
 * two calls that overlap on the GPU (different streams, different threads) race on it - issue
 * them one after the other.
 *
 * Table encodings (what ngs_snapshot_load / ngs_load_update_bulk write for a handle):
 *   free        int32  [n][8]    free percent per GPU cell, 0 for cells >= gpu_count
 *   gpu_count   uint8  [n]       0 = node unknown
 *   remain_load int8   [n][8]    GPUResource.RemainLoad (allocate.go:144), normally 0
 *   tenths      uint8  [2][n][8] ceil(10 * usage) per (metric, node, card); 255 = no usable sample
 *   ts          uint32 [2][n][8] sample time, unix seconds
 *   plan_tag    uint32 [n]       plan-cache tag, cleared for a node when it is bound; start at 0
 *   plan_score  int32  [n]       plan-cache score; start at 0 */
#define NGS_NOALLOC_MAX_PODS 256
#define NGS_NOALLOC_MAX_CTAS 160
typedef struct {
    int32_t *free;
    uint8_t *gpu_count;
    int8_t *remain_load;
    uint8_t *tenths;
    uint32_t *ts;
    uint32_t *plan_tag;
    int32_t *plan_score;
    uint32_t n_nodes;
    uint32_t first_node;
} ngs_snapshot_view;
int ngs_schedule_batch_noalloc(int device, void *stream, const ngs_snapshot_view *snap, uint32_t n_pods,
                               const int32_t *d_percent, const int32_t *d_n_containers, int max_c, const int32_t *d_policy,
                               const ngs_load_ctx *load, uint64_t seed, uint64_t seq0, int32_t *d_node_out,
                               int32_t *d_score_out, int32_t *d_gpu_idx_out);

#ifdef __cplusplus
}
#endif
#endif
