// ngs_wide.cuh — nodes with 9..12 GPUs.
//
// The reference takes any GPU count (/root/reference/pkg/utils/node.go:8-14, pkg/dealer/node.go:25-42); up to 12 cells
// sort.Sort stays in its small-slice regime (Go 1.16 sort.go quickSort: one gap-6 pass over the pairs (6,0)..(n-1,n-7),
// then insertion sort), which is what every kernel here reproduces.  Such nodes are rare (no stock GPU server
// carries more than 8 or 16 cards), so the 32-byte row of the main tables stays 8 cells wide and cells 8..11 live in a
// second plane ("hi": free_hi / rl_hi / tenths_hi / ts_hi, 4 cells per row).  A row with gcount > 8 leaves the
// register-resident fast paths at their first look at gcount and is evaluated here, by one thread, straight from
// the global tables, in plain loops over local arrays: the same Rate / LoadUsage / Choose statements, written for
// clarity, not speed.
// Included by ngs_rows.cuh inside ngs_dealer.cu's anonymous namespace, after Tables and RowEval.
#pragma once

constexpr int WIDE_MAX = NGS_MAX_GPUS_WIDE;  // 12

__device__ __forceinline__ int32_t* wide_free_cell(const Tables& T, uint32_t row, int g) {
    return g < 8 ? T.free + (size_t)row * 8 + g : T.free_hi + (size_t)row * 4 + (g - 8);
}
__device__ __forceinline__ int8_t* wide_rl_cell(const Tables& T, uint32_t row, int g) {
    return g < 8 ? T.rl + (size_t)row * 8 + g : T.rl_hi + (size_t)row * 4 + (g - 8);
}
__device__ __forceinline__ size_t wide_usage_off(const Tables& T, uint32_t row, int metric, int g, bool* hi) {
    *hi = g >= 8;
    return g < 8 ? ((size_t)metric * T.nstride + row) * 8 + g : ((size_t)metric * T.nstride + row) * 4 + (g - 8);
}

// sort.Sort(sortableGpus) on key = Percent + 50*RemainLoad (allocate.go:247), Go 1.16, n <= 12
__device__ inline void wide_go_sort(int* key, int* fr, int* ix, int n) {
    for (int i = 6; i < n; i++)
        if (key[i] < key[i - 6]) {
            int t = key[i]; key[i] = key[i - 6]; key[i - 6] = t;
            t = fr[i]; fr[i] = fr[i - 6]; fr[i - 6] = t;
            t = ix[i]; ix[i] = ix[i - 6]; ix[i - 6] = t;
        }
    for (int i = 1; i < n; i++)
        for (int j = i; j > 0 && key[j] < key[j - 1]; j--) {
            int t = key[j]; key[j] = key[j - 1]; key[j - 1] = t;
            t = fr[j]; fr[j] = fr[j - 1]; fr[j - 1] = t;
            t = ix[j]; ix[j] = ix[j - 1]; ix[j - 1] = t;
        }
}

// What the evaluation reads, staged once per CTA in shared memory (wide_ctx_fill) by the kernels that may meet such a
// row: the out-of-line call then takes one pointer instead of two by-value structs, which would otherwise be copied
// to every thread's stack whether or not a wide row ever shows up.
struct WideCtx {
    Tables T;
    LoadParams lp;
};
// call before the kernel's first __syncthreads(); snapshots without wide nodes (free_hi == nullptr) skip the copy
__device__ __forceinline__ void wide_ctx_fill(WideCtx* cx, const Tables& T, const LoadParams& lp, bool one_thread) {
    if (T.free_hi != nullptr && one_thread) {
        cx->T = T;
        cx->lp = lp;
    }
}
enum { WIDE_RATE = 1, WIDE_LOAD = 2, WIDE_APPLY = 4 };

// GPUs.Choose (allocate.go:92-100) on one wide row.  flags: WIDE_RATE = run Rate (do_rate of eval_row_values),
// WIDE_LOAD = the call is load-aware if the context says so (cx->lp.enabled), WIDE_APPLY = when the plan is feasible
// store the cells as GPUs.Allocate leaves them (allocate.go:102-118).
// Returns x: bit 0 feasible, bits 8.. n_emit; y: score; (z, w): the emission list, 4 bits per entry as elsewhere.
__device__ __noinline__ uint4 eval_wide_row(const WideCtx* cx, uint32_t row, const SortedDemand* dem, uint32_t flags) {
    const Tables& T = cx->T;
    const LoadParams& lp = cx->lp;
    const bool do_rate = (flags & WIDE_RATE) != 0;
    int score = NGS_SCORE_MIN;
    const int policy = dem->policy;
    const int G = min((int)T.gcount[row], WIDE_MAX);
    const bool load_path = (flags & WIDE_LOAD) && lp.enabled && do_rate && policy != NGS_POLICY_RANDOM;
    int f[WIDE_MAX], rl[WIDE_MAX];
    for (int g = 0; g < G; g++) f[g] = __ldcg(wide_free_cell(T, row, g));
    int load_int = 0;
    if (load_path) {
        // GPUResource.LoadUsage (allocate.go:173-195) per card, summed in GPU-index order (rater.go:63-65)
        double lu = 0.0;
        for (int g = 0; g < G; g++) {
            double u = 0.0;
            for (int e = 0; e < lp.n_periods; e++) {
                bool hi;
                const size_t off = wide_usage_off(T, row, lp.metric[e], g, &hi);
                const uint32_t t10 = hi ? T.tenths_hi[off] : T.tenths[off];
                const uint32_t t = hi ? T.ts_hi[off] : T.ts[off];
                if (t10 == 255u || lp.never[e] || t < lp.min_ts[e]) continue;
                u = __dadd_rn(u, c_tenth[t10 <= 10u ? t10 : 10u]);
            }
            rl[g] = 2 - __double2int_rz(u);  // allocate.go:193
            *wide_rl_cell(T, row, g) = (int8_t)rl[g];
            lu = __dadd_rn(lu, u);
        }
        load_int = __double2int_rz(lu) / G;  // rater.go:68,121
    } else {
        for (int g = 0; g < G; g++) rl[g] = T.rl_zero ? 0 : (int)*wide_rl_cell(T, row, g);
    }
    if (do_rate) {
        if (policy == NGS_POLICY_BINPACK) {  // rater.go:59-70
            int used = 0;
            for (int g = 0; g < G; g++) used += 100 - f[g];
            score = binpack_percent_fp64(used, G) + load_int * 50 - G;
        } else if (policy == NGS_POLICY_SPREAD) {  // rater.go:113-123
            int avail = 0, cnt = 0;
            for (int g = 0; g < G; g++) {
                avail += f[g];
                cnt += f[g] == 100 ? 1 : 0;
            }
            score = 100 * cnt + avail / 10 - G - load_int;
        } else {
            score = 100;  // rater.go:24-26
        }
    }
    uint64_t em = 0;
    int ne = 0;
    bool ok = true;
    if (policy == NGS_POLICY_RANDOM) {  // SampleRater.Choose (rater.go:28-50): pod order, GPU-index order
        for (int k = 0; k < dem->C; k++) {
            const int d = dem->raw[k];
            if (d == 0) {
                em |= (uint64_t)0xF << (4 * ne++);
                continue;
            }
            int sel = -1;
            for (int g = 0; g < G && sel < 0; g++)
                if (f[g] >= d) sel = g;
            if (sel < 0) {
                ok = false;
                continue;
            }
            f[sel] -= d;
            em |= (uint64_t)sel << (4 * ne++);
        }
    } else {  // Binpack.Choose / Spread.Choose (rater.go:74-110, 127-163): re-sort before every container
        int key[WIDE_MAX], fr[WIDE_MAX], ix[WIDE_MAX];
        for (int g = 0; g < G; g++) {
            fr[g] = f[g];
            key[g] = f[g] + 50 * rl[g];
            ix[g] = g;
        }
        const bool spread = policy == NGS_POLICY_SPREAD;
        for (int k = 0; k < dem->nz; k++) {
            const int d = dem->d[k];
            wide_go_sort(key, fr, ix, G);
            int sel = -1;
            if (spread) {
                for (int i = G - 1; i >= 0 && sel < 0; i--)
                    if (fr[i] >= d) sel = i;
            } else {
                for (int i = 0; i < G && sel < 0; i++)
                    if (fr[i] >= d) sel = i;
            }
            if (sel < 0) {
                ok = false;
                continue;
            }
            fr[sel] -= d;
            key[sel] -= d;
            em |= (uint64_t)ix[sel] << (4 * ne++);
        }
        for (int k = dem->nz; k < dem->C; k++) em |= (uint64_t)0xF << (4 * ne++);
        for (int i = 0; i < G; i++) f[ix[i]] = fr[i];
    }
    if (ok && (flags & WIDE_APPLY))
        for (int g = 0; g < G; g++) *wide_free_cell(T, row, g) = f[g];
    return make_uint4((ok ? 1u : 0u) | ((uint32_t)ne << 8), (uint32_t)score, (uint32_t)em, (uint32_t)(em >> 32));
}
// the same, unpacked into what the 8-cell paths return
__device__ __forceinline__ RowEval wide_row_eval(const WideCtx* cx, uint32_t row, const SortedDemand* dem, uint32_t flags) {
    const uint4 r = eval_wide_row(cx, row, dem, flags);
    RowEval ev;
    ev.known = true;
    ev.feasible = (r.x & 1u) != 0;
    ev.score = (int)r.y;
    ev.emit = (uint64_t)r.z | ((uint64_t)r.w << 32);
    ev.n_emit = (int)(r.x >> 8);
    return ev;
}
