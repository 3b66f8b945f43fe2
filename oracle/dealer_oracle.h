/*
 * dealer_oracle.h — CPU restatement of nano-gpu-scheduler's pkg/dealer hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library, and only as the checker / CPU baseline.  The CUDA library never links it.
 *
 * Parity status: the reference is Go and cannot be compiled in this image (no Go toolchain,
 * k8s.io deps not vendored).  This restatement is pinned against every golden vector the
 * reference's own tests hold for the path (pkg/dealer/rater_test.go, allocate_test.go; see
 * tests/golden/reference_kats.json).  What those tests do NOT pin — sort.Sort tie order for
 * 7..12 elements (Go 1.16 gap-6 shell pass), the fp64 truncation points of Binpack.Rate and
 * the whole load-aware mode — is "parity unpinned": restated from the published Go 1.16
 * standard library algorithm (src/sort/sort.go) and IEEE-754 fp64 semantics.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/pkg/dealer/ unless stated).
 */
#ifndef DEALER_ORACLE_H
#define DEALER_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OR_MAX_GPUS 64       /* oracle supports G up to 64 (Go quicksort path for G > 12) */
#define OR_MAX_CONTAINERS 64
#define OR_MAX_PERIODS 8

/* policies (rater.go:21,52,55; "random" = SURVEY §8 a9') */
enum { OR_POLICY_BINPACK = 0, OR_POLICY_SPREAD = 1, OR_POLICY_RANDOM = 2 };
/* which Go toolchain's sort.Sort small-slice rule to emulate */
enum { OR_SORT_GO116 = 0, OR_SORT_GO119 = 1 };
/* per-node reason codes */
enum { OR_REASON_OK = 0, OR_REASON_UNKNOWN_NODE = 1, OR_REASON_INFEASIBLE = 2 };
/* load metric tables (type.go:7-8, dealer.go:84-85) */
enum { OR_METRIC_CORE = 0, OR_METRIC_MEMORY = 1 };

#define OR_NOT_NEED_GPU (-1) /* allocate.go:15 */
#define OR_SCORE_MIN 0       /* rater.go:12 */
#define OR_SCORE_MAX 100     /* rater.go:13 */

/* PolicySpec.SyncPeriod after host-side resolution of getActiveDuration (stats.go:57-66):
 * one entry per SyncPeriod element whose active duration resolved to non-zero, in order. */
typedef struct {
    int32_t n_periods;
    int32_t metric[OR_MAX_PERIODS];     /* OR_METRIC_CORE iff name == "gpu_core_usage_avg" (nodeusage.go:84) */
    int64_t window_ns[OR_MAX_PERIODS];  /* period + 5 min (type.go:6, stats.go:61) */
    int64_t now_ns;                     /* time.Now() of this evaluation (stats.go:48) */
} or_load_ctx;

typedef struct or_cluster or_cluster;

or_cluster *or_create(uint32_t n_nodes, int sort_rule);
void or_destroy(or_cluster *c);
void or_set_buggy_restore(or_cluster *c, int on); /* reproduce allocate.go:110-112 arithmetic */

/* NewNodeInfo (node.go:25-42): G cells of (100,100,0).  G == 0 marks the node unknown. */
int or_node_upsert(or_cluster *c, uint32_t idx, int gpu_count);
/* direct state poke for tests / synthetic snapshots */
int or_node_set(or_cluster *c, uint32_t idx, int gpu_count, const int64_t *free, const int64_t *rl);
int or_node_get(const or_cluster *c, uint32_t idx, int *gpu_count, int64_t *free, int64_t *rl);
/* bulk forms over the [n][8] int32 layout the GPU snapshot uses */
int or_snapshot_load(or_cluster *c, uint32_t first, uint32_t n, const uint8_t *gpu_count, const int32_t *free8);
int or_snapshot_read(const or_cluster *c, uint32_t first, uint32_t n, uint8_t *gpu_count, int32_t *free8, int8_t *rl8);
int or_load_update_bulk(or_cluster *c, int metric, uint32_t first, uint32_t n, const double *value8, const uint8_t *ok8,
                        const int64_t *ts8);

/* GPUs.Allocate (sign=+1, allocate.go:102-118) / GPUs.Release (sign=-1, allocate.go:120-131).
 * returns 0, or -1 on the reference's error paths. */
int or_node_apply(or_cluster *c, uint32_t idx, const int32_t *gpu_idx, const int32_t *percent, int n_containers, int sign);

/* Update{Core,Memory}Usage (nodeusage.go:70-80) after host-side string parsing:
 * value_ok = strconv.ParseFloat succeeded; ts_ok = len>=5 && time.ParseInLocation succeeded. */
int or_load_update(or_cluster *c, uint32_t idx, int metric, int card, double value, int value_ok, int64_t ts_sec, int ts_ok);
/* Add{Core,Memory}Usage (nodeusage.go:58-68): node gets an (empty) per-card map */
int or_load_add_node(or_cluster *c, uint32_t idx, int metric);

/* DealerImpl.Assume + DealerImpl.Score over a candidate list (dealer.go:89-153):
 * per node NodeInfo.Assume -> GPUs.Choose -> Rate then Choose (node.go:44-57, allocate.go:92-100).
 * idx == NULL means nodes 0..n-1.  plan (nullable) is [n][n_containers]: full plan when feasible,
 * otherwise the partial `indexes` list of rater.go:99 in emission order, padded with INT32_MIN;
 * n_emitted (nullable) its length.  score follows NodeInfo.Score: ScoreMin when infeasible. */
int or_assume(or_cluster *c, const uint32_t *idx, uint32_t n, const int32_t *percent, int n_containers, int policy,
              const or_load_ctx *load, uint8_t *feasible, int8_t *reason, int32_t *score, int32_t *plan, int32_t *n_emitted);

/* NodeInfo.Bind (node.go:70-84) on a cache miss: Assume then GPUs.Allocate. */
int or_bind(or_cluster *c, uint32_t idx, const int32_t *percent, int n_containers, int policy, const or_load_ctx *load,
            int32_t *gpu_idx_out, int32_t *score_out);

/* In-order scheduling of P pods against all nodes: filter -> score -> arg-max (highest score,
 * ties -> lowest node index; random policy: seeded uniform pick over the feasible set) -> bind.
 * percent is [P][max_c]; node_out[p] = -1 when no node is feasible.  n_threads > 1 fans the
 * per-node loop out over host threads (the reference uses 4 goroutines, dealer.go:113). */
int or_schedule_batch(or_cluster *c, uint32_t n_pods, const int32_t *percent, const int32_t *n_containers, int max_c,
                      const int32_t *policy, const or_load_ctx *load, uint64_t seed, uint64_t seq0, int32_t *node_out,
                      int32_t *score_out, int32_t *gpu_idx_out, int n_threads);

/* seeded pick used by the random policy: splitmix64(seed + seq) */
uint64_t or_splitmix64(uint64_t x);

/* --- unit-level entry points (golden-vector tests call these directly) --- */
/* sort.Sort on SortableGPUs (allocate.go:239-247): keys free+50*rl, payload idx; in place */
void or_go_sort(int64_t *free, int64_t *rl, int32_t *idx, int n, int sort_rule);
/* Rater.Choose on a bare GPU list; returns 1 feasible / 0 not; out has n_containers entries */
int or_choose(int policy, const int64_t *free, const int64_t *rl, int gpu_count, const int32_t *percent, int n_containers,
              int sort_rule, int32_t *out, int32_t *n_emitted);
/* Rater.Rate on a bare GPU list with a given fp64 loadUsage sum */
int64_t or_rate(int policy, const int64_t *free, int gpu_count, double load_usage);

#ifdef __cplusplus
}
#endif
#endif
