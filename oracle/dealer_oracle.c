/*
 * dealer_oracle.c — CPU restatement of nano-gpu-scheduler's pkg/dealer hot path.
 * TEST INFRASTRUCTURE ONLY (see dealer_oracle.h).  Plain C11 + pthreads (the thread pool exists
 * only for the multi-threaded CPU baseline).  Citations are relative to /root/reference/pkg/dealer/.
 */
#include "dealer_oracle.h"

#include <limits.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * Go 1.16 sort.Sort (third-party to the reference: Go standard library, toolchain pinned by
 * /root/reference/go.mod:3 and Dockerfile:1 to go 1.16; call sites rater.go:81,87,134,140).
 * Restated from the published algorithm in go1.16 src/sort/sort.go: quickSort with a
 * median-of-three / ninther pivot, heapsort when the depth limit hits, and for slices of
 * <= 12 elements one gap-6 shell pass followed by insertion sort.  Go >= 1.19 replaced this
 * with pdqsort whose <= 12 element case is plain insertion sort (OR_SORT_GO119; only that
 * case is modelled).  The sort is driven purely through Less/Swap, as sort.Interface is.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int64_t free; /* GPUResource.Percent  (allocate.go:142) */
    int64_t rl;   /* GPUResource.RemainLoad (allocate.go:144) */
    int32_t idx;  /* GPUResourceWithIndex.index (allocate.go:238-241) */
} cell_t;

/* SortableGPUs.Less (allocate.go:247) */
static inline int cell_less(const cell_t *a, int i, int j) { return a[i].free + a[i].rl * 50 < a[j].free + a[j].rl * 50; }
/* SortableGPUs.Swap (allocate.go:246) */
static inline void cell_swap(cell_t *a, int i, int j) {
    cell_t t = a[i];
    a[i] = a[j];
    a[j] = t;
}

static void go_insertion_sort(cell_t *d, int a, int b) {
    for (int i = a + 1; i < b; i++)
        for (int j = i; j > a && cell_less(d, j, j - 1); j--) cell_swap(d, j, j - 1);
}

static void go_sift_down(cell_t *d, int lo, int hi, int first) {
    int root = lo;
    for (;;) {
        int child = 2 * root + 1;
        if (child >= hi) return;
        if (child + 1 < hi && cell_less(d, first + child, first + child + 1)) child++;
        if (!cell_less(d, first + root, first + child)) return;
        cell_swap(d, first + root, first + child);
        root = child;
    }
}

static void go_heap_sort(cell_t *d, int a, int b) {
    int first = a, lo = 0, hi = b - a;
    for (int i = (hi - 1) / 2; i >= 0; i--) go_sift_down(d, i, hi, first);
    for (int i = hi - 1; i >= 0; i--) {
        cell_swap(d, first, first + i);
        go_sift_down(d, lo, i, first);
    }
}

static void go_median_of_three(cell_t *d, int m1, int m0, int m2) {
    if (cell_less(d, m1, m0)) cell_swap(d, m1, m0);
    if (cell_less(d, m2, m1)) {
        cell_swap(d, m2, m1);
        if (cell_less(d, m1, m0)) cell_swap(d, m1, m0);
    }
}

static void go_do_pivot(cell_t *d, int lo, int hi, int *midlo, int *midhi) {
    int m = (int)((unsigned)(lo + hi) >> 1);
    if (hi - lo > 40) {
        int s = (hi - lo) / 8;
        go_median_of_three(d, lo, lo + s, lo + 2 * s);
        go_median_of_three(d, m, m - s, m + s);
        go_median_of_three(d, hi - 1, hi - 1 - s, hi - 1 - 2 * s);
    }
    go_median_of_three(d, lo, m, hi - 1);

    int pivot = lo;
    int a = lo + 1, c = hi - 1;
    for (; a < c && cell_less(d, a, pivot); a++) {
    }
    int b = a;
    for (;;) {
        for (; b < c && !cell_less(d, pivot, b); b++) {
        }
        for (; b < c && cell_less(d, pivot, c - 1); c--) {
        }
        if (b >= c) break;
        cell_swap(d, b, c - 1);
        b++;
        c--;
    }
    int protect = hi - c < 5;
    if (!protect && hi - c < (hi - lo) / 4) {
        int dups = 0;
        if (!cell_less(d, pivot, hi - 1)) {
            cell_swap(d, c, hi - 1);
            c++;
            dups++;
        }
        if (!cell_less(d, b - 1, pivot)) {
            b--;
            dups++;
        }
        if (!cell_less(d, m, pivot)) {
            cell_swap(d, m, b - 1);
            b--;
            dups++;
        }
        protect = dups > 1;
    }
    if (protect) {
        for (;;) {
            for (; a < b && !cell_less(d, b - 1, pivot); b--) {
            }
            for (; a < b && cell_less(d, a, pivot); a++) {
            }
            if (a >= b) break;
            cell_swap(d, a, b - 1);
            a++;
            b--;
        }
    }
    cell_swap(d, pivot, b - 1);
    *midlo = b - 1;
    *midhi = c;
}

static void go_quick_sort(cell_t *d, int a, int b, int max_depth) {
    while (b - a > 12) {
        if (max_depth == 0) {
            go_heap_sort(d, a, b);
            return;
        }
        max_depth--;
        int mlo, mhi;
        go_do_pivot(d, a, b, &mlo, &mhi);
        if (mlo - a < b - mhi) {
            go_quick_sort(d, a, mlo, max_depth);
            a = mhi;
        } else {
            go_quick_sort(d, mhi, b, max_depth);
            b = mlo;
        }
    }
    if (b - a > 1) {
        for (int i = a + 6; i < b; i++)
            if (cell_less(d, i, i - 6)) cell_swap(d, i, i - 6);
        go_insertion_sort(d, a, b);
    }
}

static int go_max_depth(int n) {
    int depth = 0;
    for (int i = n; i > 0; i >>= 1) depth++;
    return depth * 2;
}

static void go_sort(cell_t *d, int n, int rule) {
    if (rule == OR_SORT_GO119) {
        /* pdqsort_func: n <= 12 -> insertionSort only.  Larger n is not modelled. */
        go_insertion_sort(d, 0, n);
        return;
    }
    go_quick_sort(d, 0, n, go_max_depth(n));
}

void or_go_sort(int64_t *free, int64_t *rl, int32_t *idx, int n, int sort_rule) {
    cell_t tmp[OR_MAX_GPUS];
    if (n > OR_MAX_GPUS) n = OR_MAX_GPUS;
    for (int i = 0; i < n; i++) tmp[i] = (cell_t){free[i], rl ? rl[i] : 0, idx[i]};
    go_sort(tmp, n, sort_rule);
    for (int i = 0; i < n; i++) {
        free[i] = tmp[i].free;
        if (rl) rl[i] = tmp[i].rl;
        idx[i] = tmp[i].idx;
    }
}

/* ------------------------------------------------------------------------------------------
 * Rater.Choose
 * ------------------------------------------------------------------------------------------ */

/* Binpack.Choose (rater.go:74-110) / Spread.Choose (rater.go:127-163).
 * emitted[] receives the `indexes` slice in emission order (largest demand first); out[] the
 * un-permuted resultIndexs (rater.go:102-105) when feasible. */
static int choose_sorted(int spread, const int64_t *free, const int64_t *rl, int G, const int32_t *percent, int C, int rule,
                         int32_t *out, int32_t *emitted, int *n_emitted) {
    cell_t gp[OR_MAX_GPUS], dm[OR_MAX_CONTAINERS];
    /* gpus.ToSortableGPUs (allocate.go:225-236): scratch copies, node state is not mutated */
    for (int i = 0; i < G; i++) gp[i] = (cell_t){free[i], rl[i], i};
    /* d.ToSortableGPUs (allocate.go:77-88): RemainLoad forced to 0 */
    for (int i = 0; i < C; i++) dm[i] = (cell_t){percent[i], 0, i};
    go_sort(dm, C, rule); /* rater.go:81 */
    int ne = 0;
    for (int j = C - 1; j >= 0; j--) { /* rater.go:82 */
        if (dm[j].free == 0) {
            emitted[ne++] = OR_NOT_NEED_GPU; /* rater.go:83-86 */
            continue;
        }
        go_sort(gp, G, rule); /* rater.go:87 */
        if (!spread) {
            for (int i = 0; i < G; i++) { /* rater.go:88-95: first fit from the low end */
                if (!(gp[i].free >= dm[j].free)) continue;
                emitted[ne++] = gp[i].idx;
                gp[i].free -= dm[j].free;
                break;
            }
        } else {
            for (int i = G - 1; i >= 0; i--) { /* rater.go:141-148: first fit from the high end */
                if (!(gp[i].free >= dm[j].free)) continue;
                emitted[ne++] = gp[i].idx;
                gp[i].free -= dm[j].free;
                break;
            }
        }
    }
    *n_emitted = ne;
    if (ne != C) return 0; /* rater.go:98-100 */
    for (int j = C - 1; j >= 0; j--) out[dm[j].idx] = emitted[C - 1 - j]; /* rater.go:102-105 */
    return 1;
}

/* SampleRater.Choose (rater.go:28-50) on a scratch copy (SURVEY §8 a9': the write-through of
 * rater.go:41 into live node state is a reference defect and is deliberately not reproduced). */
static int choose_sample(const int64_t *free, int G, const int32_t *percent, int C, int32_t *out, int32_t *emitted,
                         int *n_emitted) {
    int64_t f[OR_MAX_GPUS];
    for (int i = 0; i < G; i++) f[i] = free[i];
    int ne = 0;
    for (int k = 0; k < C; k++) {
        if (percent[k] == 0) {
            emitted[ne++] = OR_NOT_NEED_GPU;
            continue;
        }
        for (int j = 0; j < G; j++) {
            if (!(f[j] >= percent[k])) continue;
            emitted[ne++] = j;
            f[j] -= percent[k];
            break;
        }
    }
    *n_emitted = ne;
    if (ne != C) return 0;
    for (int k = 0; k < C; k++) out[k] = emitted[k];
    return 1;
}

static int choose_any(int policy, const int64_t *free, const int64_t *rl, int G, const int32_t *percent, int C, int rule,
                      int32_t *out, int32_t *emitted, int *n_emitted) {
    if (policy == OR_POLICY_RANDOM) return choose_sample(free, G, percent, C, out, emitted, n_emitted);
    return choose_sorted(policy == OR_POLICY_SPREAD, free, rl, G, percent, C, rule, out, emitted, n_emitted);
}

int or_choose(int policy, const int64_t *free, const int64_t *rl, int gpu_count, const int32_t *percent, int n_containers,
              int sort_rule, int32_t *out, int32_t *n_emitted) {
    int64_t zero[OR_MAX_GPUS] = {0};
    int32_t emitted[OR_MAX_CONTAINERS];
    int ne = 0;
    int ok = choose_any(policy, free, rl ? rl : zero, gpu_count, percent, n_containers, sort_rule, out, emitted, &ne);
    if (!ok) /* expose the partial list, as the error text of rater.go:99 does */
        for (int i = 0; i < n_containers; i++) out[i] = i < ne ? emitted[i] : INT32_MIN;
    if (n_emitted) *n_emitted = ne;
    return ok;
}

/* ------------------------------------------------------------------------------------------
 * Rater.Rate
 * ------------------------------------------------------------------------------------------ */

/* Binpack.Rate (rater.go:59-70) with GPUs.Usage (allocate.go:164-171);
 * Spread.Rate (rater.go:113-123) with PercentAvailableAndFreeGpuCount (allocate.go:205-213);
 * SampleRater.Rate (rater.go:24-26).  PercentTotal is 100 for every cell (types.go:10, node.go:31-34). */
int64_t or_rate(int policy, const int64_t *free, int G, double load_usage) {
    if (policy == OR_POLICY_RANDOM) return OR_SCORE_MAX;
    int64_t load_int = (int64_t)load_usage / G; /* int(loadUsage) / len(gpus), Go truncating division */
    if (policy == OR_POLICY_BINPACK) {
        int64_t sum = 0, used = 0;
        for (int i = 0; i < G; i++) {
            sum += 100;
            used += 100 - free[i];
        }
        double usage = (double)used / (double)sum; /* allocate.go:170 */
        double scaled = usage * 100.0;              /* rater.go:69 (no FMA: nothing to fuse with) */
        return (int64_t)scaled + load_int * 50 - G;
    }
    int64_t avail = 0, free_cnt = 0;
    for (int i = 0; i < G; i++) {
        avail += free[i];
        if (free[i] == 100) free_cnt++;
    }
    return 100 * free_cnt + avail / 10 - G - load_int; /* rater.go:122 */
}

/* ------------------------------------------------------------------------------------------
 * cluster state
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    double value;
    int64_t ts_sec;
    uint8_t value_ok, ts_ok, has_card;
} usage_cell;

struct or_cluster {
    uint32_t n;
    int sort_rule;
    int buggy_restore;
    uint8_t *gpu_count;   /* [n]; 0 = node unknown */
    int64_t *free;        /* [n][OR_MAX_GPUS] */
    int64_t *rl;          /* [n][OR_MAX_GPUS] */
    usage_cell *usage[2]; /* [n][OR_MAX_GPUS] per metric; CoreUsage / MemoryUsage maps (dealer.go:84-85) */
    uint8_t *present[2];  /* node has an entry in the metric's outer map */
};

or_cluster *or_create(uint32_t n_nodes, int sort_rule) {
    or_cluster *c = (or_cluster *)calloc(1, sizeof(*c));
    if (!c) return NULL;
    c->n = n_nodes;
    c->sort_rule = sort_rule;
    size_t cells = (size_t)n_nodes * OR_MAX_GPUS;
    c->gpu_count = (uint8_t *)calloc(n_nodes ? n_nodes : 1, 1);
    c->free = (int64_t *)calloc(cells ? cells : 1, sizeof(int64_t));
    c->rl = (int64_t *)calloc(cells ? cells : 1, sizeof(int64_t));
    for (int m = 0; m < 2; m++) {
        c->usage[m] = (usage_cell *)calloc(cells ? cells : 1, sizeof(usage_cell));
        c->present[m] = (uint8_t *)calloc(n_nodes ? n_nodes : 1, 1);
    }
    return c;
}

void or_destroy(or_cluster *c) {
    if (!c) return;
    free(c->gpu_count);
    free(c->free);
    free(c->rl);
    for (int m = 0; m < 2; m++) {
        free(c->usage[m]);
        free(c->present[m]);
    }
    free(c);
}

void or_set_buggy_restore(or_cluster *c, int on) { c->buggy_restore = on; }

int or_node_upsert(or_cluster *c, uint32_t idx, int G) {
    if (idx >= c->n || G < 0 || G > OR_MAX_GPUS) return -1;
    c->gpu_count[idx] = (uint8_t)G;
    for (int i = 0; i < OR_MAX_GPUS; i++) {
        c->free[(size_t)idx * OR_MAX_GPUS + i] = i < G ? 100 : 0; /* node.go:30-35 */
        c->rl[(size_t)idx * OR_MAX_GPUS + i] = 0;
    }
    return 0;
}

int or_node_set(or_cluster *c, uint32_t idx, int G, const int64_t *free, const int64_t *rl) {
    if (idx >= c->n || G < 0 || G > OR_MAX_GPUS) return -1;
    c->gpu_count[idx] = (uint8_t)G;
    for (int i = 0; i < G; i++) {
        c->free[(size_t)idx * OR_MAX_GPUS + i] = free[i];
        c->rl[(size_t)idx * OR_MAX_GPUS + i] = rl ? rl[i] : 0;
    }
    return 0;
}

int or_snapshot_load(or_cluster *c, uint32_t first, uint32_t n, const uint8_t *gpu_count, const int32_t *free8) {
    if ((uint64_t)first + n > c->n) return -1;
    for (uint32_t k = 0; k < n; k++) {
        uint32_t i = first + k;
        int G = gpu_count[k];
        if (G > 8) return -1;
        c->gpu_count[i] = (uint8_t)G;
        for (int g = 0; g < OR_MAX_GPUS; g++) {
            c->free[(size_t)i * OR_MAX_GPUS + g] = g < G ? (free8 ? free8[(size_t)k * 8 + g] : 100) : 0;
            c->rl[(size_t)i * OR_MAX_GPUS + g] = 0;
        }
    }
    return 0;
}

int or_snapshot_read(const or_cluster *c, uint32_t first, uint32_t n, uint8_t *gpu_count, int32_t *free8, int8_t *rl8) {
    if ((uint64_t)first + n > c->n) return -1;
    for (uint32_t k = 0; k < n; k++) {
        uint32_t i = first + k;
        int G = c->gpu_count[i];
        if (gpu_count) gpu_count[k] = (uint8_t)G;
        for (int g = 0; g < 8; g++) {
            if (free8) free8[(size_t)k * 8 + g] = g < G ? (int32_t)c->free[(size_t)i * OR_MAX_GPUS + g] : 0;
            if (rl8) rl8[(size_t)k * 8 + g] = g < G ? (int8_t)c->rl[(size_t)i * OR_MAX_GPUS + g] : 0;
        }
    }
    return 0;
}

int or_node_get(const or_cluster *c, uint32_t idx, int *G, int64_t *free, int64_t *rl) {
    if (idx >= c->n) return -1;
    int g = c->gpu_count[idx];
    if (G) *G = g;
    for (int i = 0; i < g; i++) {
        if (free) free[i] = c->free[(size_t)idx * OR_MAX_GPUS + i];
        if (rl) rl[i] = c->rl[(size_t)idx * OR_MAX_GPUS + i];
    }
    return 0;
}

int or_node_apply(or_cluster *c, uint32_t idx, const int32_t *gpu_idx, const int32_t *percent, int C, int sign) {
    if (idx >= c->n || c->gpu_count[idx] == 0) return -1;
    int G = c->gpu_count[idx];
    int64_t *f = &c->free[(size_t)idx * OR_MAX_GPUS];
    if (sign > 0) {
        /* GPUs.Allocate (allocate.go:102-118) */
        for (int i = 0; i < C; i++) {
            if (gpu_idx[i] < 0) continue;
            if (gpu_idx[i] >= G || !(f[gpu_idx[i]] >= percent[i])) { /* index >= G panics in Go */
                if (c->buggy_restore) {
                    /* allocate.go:110-112 as written: adds Demand[i] (not [j]) back to the cells of
                     * containers j<i; a -1 among them would panic in Go, here it is skipped */
                    for (int j = 0; j < i; j++)
                        if (gpu_idx[j] >= 0 && gpu_idx[j] < G) f[gpu_idx[j]] += percent[i];
                } else {
                    /* what the restore intends: undo containers j<i exactly (all-or-nothing) */
                    for (int j = 0; j < i; j++)
                        if (gpu_idx[j] >= 0) f[gpu_idx[j]] += percent[j];
                }
                return -1;
            }
            f[gpu_idx[i]] -= percent[i];
        }
        return 0;
    }
    /* GPUs.Release (allocate.go:120-131): partial application before the error is kept */
    for (int i = 0; i < C; i++) {
        if (gpu_idx[i] < 0) continue;
        if (gpu_idx[i] >= G) return -1;
        f[gpu_idx[i]] += percent[i];
    }
    return 0;
}

int or_load_add_node(or_cluster *c, uint32_t idx, int metric) {
    if (idx >= c->n || metric < 0 || metric > 1) return -1;
    /* nodeusage.go:61,67: a fresh empty per-card map replaces whatever was there */
    c->present[metric][idx] = 1;
    memset(&c->usage[metric][(size_t)idx * OR_MAX_GPUS], 0, sizeof(usage_cell) * OR_MAX_GPUS);
    return 0;
}

int or_load_update(or_cluster *c, uint32_t idx, int metric, int card, double value, int value_ok, int64_t ts_sec,
                   int ts_ok) {
    if (idx >= c->n || metric < 0 || metric > 1 || card < 0 || card >= OR_MAX_GPUS) return -1;
    /* nodeusage.go:73,79 writes into d.CoreUsage[node][card]; a nil inner map would panic, so
     * the controller always calls Add*Usage first (controller/node.go:121-133) */
    c->present[metric][idx] = 1;
    usage_cell *u = &c->usage[metric][(size_t)idx * OR_MAX_GPUS + card];
    u->value = value;
    u->value_ok = (uint8_t)(value_ok != 0);
    u->ts_sec = ts_sec;
    u->ts_ok = (uint8_t)(ts_ok != 0);
    u->has_card = 1;
    return 0;
}

int or_load_update_bulk(or_cluster *c, int metric, uint32_t first, uint32_t n, const double *value8, const uint8_t *ok8,
                        const int64_t *ts8) {
    if ((uint64_t)first + n > c->n || metric < 0 || metric > 1) return -1;
    for (uint32_t k = 0; k < n; k++)
        for (int g = 0; g < 8; g++) {
            int ok = ok8 ? ok8[(size_t)k * 8 + g] : 1;
            or_load_update(c, first + k, metric, g, value8[(size_t)k * 8 + g], ok, ts8[(size_t)k * 8 + g], ok);
        }
    return 0;
}

/* GPUResource.LoadUsage (allocate.go:173-195) with DealerImpl.GetUsage (nodeusage.go:82-111)
 * and inUpdateTimePeriod (stats.go:30-55).  Writes RemainLoad into live node state. */
static double load_usage_of(or_cluster *c, uint32_t idx, int card, const or_load_ctx *L) {
    double usage = 0.0;
    for (int e = 0; e < L->n_periods; e++) { /* allocate.go:175 */
        int m = L->metric[e];
        if (!c->present[m][idx]) continue; /* nodeusage.go:85-88,93-96: !exist */
        const usage_cell *u = &c->usage[m][(size_t)idx * OR_MAX_GPUS + card];
        /* a missing card yields the zero struct: time "" -> len < 5 -> false (stats.go:32-35) */
        if (!u->has_card || !u->ts_ok) continue;
        /* stats.go:48-53: now.Before(updateTime + activeDuration), nanosecond granular */
        if (!(L->now_ns < u->ts_sec * 1000000000LL + L->window_ns[e])) continue;
        if (!u->value_ok) continue;                       /* nodeusage.go:103-106 */
        if (u->value < 0 || u->value > 1) continue;       /* nodeusage.go:107-110 */
        if (isnan(u->value)) continue; /* out of domain: Go's int(NaN) is unspecified; the metric sync
                                          never stores NaN (pkg/prometheus/prometheus.go:57-60) */
        double t = 10 * u->value;                /* allocate.go:190 */
        double pu = ceil(t) / 10;
        usage += pu; /* allocate.go:191 */
    }
    c->rl[(size_t)idx * OR_MAX_GPUS + card] = 2 - (int64_t)usage; /* allocate.go:193, LoadTotal = 2 */
    return usage;
}

/* GPUs.Choose (allocate.go:92-100) for one node: Rate first, then Choose. */
static int node_choose(or_cluster *c, uint32_t idx, const int32_t *percent, int C, int policy, const or_load_ctx *load,
                       int64_t *score, int32_t *out, int32_t *emitted, int *n_emitted) {
    int G = c->gpu_count[idx];
    int64_t *f = &c->free[(size_t)idx * OR_MAX_GPUS];
    int64_t *rl = &c->rl[(size_t)idx * OR_MAX_GPUS];
    double lu = 0.0;
    if (load && policy != OR_POLICY_RANDOM) /* rater.go:62-66,116-120; SampleRater.Rate ignores load */
        for (int i = 0; i < G; i++) lu += load_usage_of(c, idx, i, load);
    *score = or_rate(policy, f, G, lu);
    return choose_any(policy, f, rl, G, percent, C, c->sort_rule, out, emitted, n_emitted);
}

int or_assume(or_cluster *c, const uint32_t *idx, uint32_t n, const int32_t *percent, int C, int policy,
              const or_load_ctx *load, uint8_t *feasible, int8_t *reason, int32_t *score, int32_t *plan,
              int32_t *n_emitted) {
    if (C < 0 || C > OR_MAX_CONTAINERS) return -1;
    for (uint32_t k = 0; k < n; k++) {
        uint32_t i = idx ? idx[k] : k;
        if (i >= c->n || c->gpu_count[i] == 0) { /* dealer.go:97-103 / :143-148 */
            feasible[k] = 0;
            if (reason) reason[k] = OR_REASON_UNKNOWN_NODE;
            if (score) score[k] = OR_SCORE_MIN;
            if (plan)
                for (int j = 0; j < C; j++) plan[(size_t)k * C + j] = INT32_MIN;
            if (n_emitted) n_emitted[k] = 0;
            continue;
        }
        int64_t s;
        int32_t out[OR_MAX_CONTAINERS], em[OR_MAX_CONTAINERS];
        int ne = 0;
        int ok = node_choose(c, i, percent, C, policy, load, &s, out, em, &ne);
        feasible[k] = (uint8_t)ok;
        if (reason) reason[k] = ok ? OR_REASON_OK : OR_REASON_INFEASIBLE;
        if (score) score[k] = ok ? (int32_t)s : OR_SCORE_MIN; /* node.go:59-68 */
        if (plan)
            for (int j = 0; j < C; j++) plan[(size_t)k * C + j] = ok ? out[j] : (j < ne ? em[j] : INT32_MIN);
        if (n_emitted) n_emitted[k] = ne;
    }
    return 0;
}

int or_bind(or_cluster *c, uint32_t idx, const int32_t *percent, int C, int policy, const or_load_ctx *load,
            int32_t *gpu_idx_out, int32_t *score_out) {
    if (idx >= c->n || c->gpu_count[idx] == 0 || C < 0 || C > OR_MAX_CONTAINERS) return -1;
    int64_t s;
    int32_t out[OR_MAX_CONTAINERS], em[OR_MAX_CONTAINERS];
    int ne = 0;
    if (!node_choose(c, idx, percent, C, policy, load, &s, out, em, &ne)) return -1; /* node.go:74-76 */
    if (or_node_apply(c, idx, out, percent, C, +1) != 0) return -1;                  /* node.go:79-81 */
    for (int j = 0; j < C; j++) gpu_idx_out[j] = out[j];
    if (score_out) *score_out = (int32_t)s;
    return 0;
}

uint64_t or_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

/* per-node filter+score of one pod over a node slice (dealer.go:109-134, 138-153) */
typedef struct {
    or_cluster *c;
    const int32_t *d;
    int C, pol;
    const or_load_ctx *load;
    uint8_t *feas;
    int32_t *sc;
} scan_job;

static void scan_slice(const scan_job *j, uint32_t lo, uint32_t hi) {
    or_cluster *c = j->c;
    for (uint32_t i = lo; i < hi; i++) {
        if (c->gpu_count[i] == 0) {
            j->feas[i] = 0;
            j->sc[i] = OR_SCORE_MIN;
            continue;
        }
        int64_t s;
        int32_t out[OR_MAX_CONTAINERS], em[OR_MAX_CONTAINERS];
        int ne = 0;
        int ok = node_choose(c, i, j->d, j->C, j->pol, j->load, &s, out, em, &ne);
        j->feas[i] = (uint8_t)ok;
        j->sc[i] = ok ? (int32_t)s : OR_SCORE_MIN;
    }
}

/* tiny persistent pool: the reference fans the filter out over 4 goroutines (dealer.go:113) */
typedef struct {
    pthread_barrier_t start, done;
    volatile int quit;
    int n_threads;
    scan_job job;
    uint32_t n_nodes;
} scan_pool;

typedef struct {
    scan_pool *pool;
    int tid;
} scan_worker_arg;

static void pool_run_slice(scan_pool *p, int tid) {
    uint64_t N = p->n_nodes;
    uint32_t lo = (uint32_t)(N * (uint64_t)tid / (uint64_t)p->n_threads);
    uint32_t hi = (uint32_t)(N * (uint64_t)(tid + 1) / (uint64_t)p->n_threads);
    scan_slice(&p->job, lo, hi);
}

static void *scan_worker(void *argp) {
    scan_worker_arg *a = (scan_worker_arg *)argp;
    for (;;) {
        pthread_barrier_wait(&a->pool->start);
        if (a->pool->quit) return NULL;
        pool_run_slice(a->pool, a->tid);
        pthread_barrier_wait(&a->pool->done);
    }
}

int or_schedule_batch(or_cluster *c, uint32_t P, const int32_t *percent, const int32_t *n_containers, int max_c,
                      const int32_t *policy, const or_load_ctx *load, uint64_t seed, uint64_t seq0, int32_t *node_out,
                      int32_t *score_out, int32_t *gpu_idx_out, int n_threads) {
    uint32_t N = c->n;
    uint8_t *feas = (uint8_t *)malloc(N ? N : 1);
    int32_t *sc = (int32_t *)malloc(sizeof(int32_t) * (N ? N : 1));
    if (!feas || !sc) return -1;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    scan_pool pool;
    pthread_t th[256];
    scan_worker_arg wa[256];
    memset(&pool, 0, sizeof(pool));
    pool.n_threads = n_threads;
    pool.n_nodes = N;
    if (n_threads > 1) {
        pthread_barrier_init(&pool.start, NULL, (unsigned)n_threads);
        pthread_barrier_init(&pool.done, NULL, (unsigned)n_threads);
        for (int t = 1; t < n_threads; t++) {
            wa[t].pool = &pool;
            wa[t].tid = t;
            pthread_create(&th[t], NULL, scan_worker, &wa[t]);
        }
    }
    int rc = 0;
    for (uint32_t p = 0; p < P; p++) {
        const int32_t *d = percent + (size_t)p * max_c;
        int C = n_containers[p];
        int pol = policy[p];
        pool.job = (scan_job){c, d, C, pol, load, feas, sc};
        if (n_threads > 1) {
            pthread_barrier_wait(&pool.start);
            pool_run_slice(&pool, 0);
            pthread_barrier_wait(&pool.done);
        } else {
            scan_slice(&pool.job, 0, N);
        }
        int64_t best = -1;
        if (pol == OR_POLICY_RANDOM) {
            uint64_t F = 0;
            for (uint32_t i = 0; i < N; i++) F += feas[i];
            if (F) {
                uint64_t r = or_splitmix64(seed + seq0 + p) % F;
                for (uint32_t i = 0; i < N; i++)
                    if (feas[i] && r-- == 0) {
                        best = i;
                        break;
                    }
            }
        } else {
            for (uint32_t i = 0; i < N; i++)
                if (feas[i] && (best < 0 || sc[i] > sc[best])) best = i;
        }
        node_out[p] = (int32_t)best;
        score_out[p] = best >= 0 ? sc[best] : OR_SCORE_MIN;
        for (int j = 0; j < max_c; j++) gpu_idx_out[(size_t)p * max_c + j] = INT32_MIN;
        if (best >= 0) {
            /* bind = cached plan of the filter pass (node.go:70-84): re-derive it without touching
             * the load tables again (RemainLoad already holds the filter-time value) */
            int32_t out[OR_MAX_CONTAINERS], em[OR_MAX_CONTAINERS];
            int ne = 0;
            int G = c->gpu_count[best];
            int ok = choose_any(pol, &c->free[(size_t)best * OR_MAX_GPUS], &c->rl[(size_t)best * OR_MAX_GPUS], G, d, C,
                                c->sort_rule, out, em, &ne);
            if (!ok || or_node_apply(c, (uint32_t)best, out, d, C, +1) != 0) {
                rc = -2; /* cannot happen: the plan was feasible on this very state */
                break;
            }
            for (int j = 0; j < C; j++) gpu_idx_out[(size_t)p * max_c + j] = out[j];
        }
    }
    if (n_threads > 1) {
        pool.quit = 1;
        pthread_barrier_wait(&pool.start);
        for (int t = 1; t < n_threads; t++) pthread_join(th[t], NULL);
        pthread_barrier_destroy(&pool.start);
        pthread_barrier_destroy(&pool.done);
    }
    free(feas);
    free(sc);
    return rc;
}
