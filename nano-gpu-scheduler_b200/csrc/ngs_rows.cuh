// ngs_rows.cuh — the snapshot tables as the kernels see them and the evaluation of one row
// (GPUs.Choose = Rate then Choose, /root/reference/pkg/dealer/allocate.go:92-100).
// Included by ngs_dealer.cu inside its anonymous namespace.
#pragma once

#include "ngs_tables.cuh"
#include "ngs_wide.cuh"

__device__ __forceinline__ int4 ld_row4(const int32_t* p) { return __ldcg(reinterpret_cast<const int4*>(p)); }
// one 256-bit load per 32-byte row (LDG.E.256 on sm_100a): every sector a warp touches is fully used
__device__ __forceinline__ void ld_row8(const void* p, uint32_t (&v)[8]) {
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "l"(p));
}
// exact plan search kept out of line: the scan's common path (shortcuts decide) stays small
// (arguments by value — four words of u16 pairs and the 3-bit RemainLoad codes — so that the caller's
// row stays in registers)
__device__ __noinline__ bool choose_sorted_slow(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3, uint32_t rlmeta, int G, bool spread,
                                                const SortedDemand* dem) {
    const uint32_t pk[4] = {p0, p1, p2, p3};
    uint32_t w[8];
#pragma unroll
    for (int g = 0; g < 8; g++) {
        const int fr = (int)((pk[g >> 1] >> (16 * (g & 1))) & 0xFFFFu);
        w[g] = g < G ? pack_cell(fr, (int)((rlmeta >> (3 * g)) & 7u), g) : W_PAD;
    }
    return choose_sorted<false>(w, G, spread, (rlmeta & 0xFFFFFFu) == RL0_META, dem, nullptr, nullptr);
}
__global__ void k_build_rate_lut(int16_t* lut) {
    const int G = blockIdx.x;  // 0..8
    for (int used = threadIdx.x; used < RATE_LUT_STRIDE; used += blockDim.x)
        lut[G * RATE_LUT_STRIDE + used] = (G > 0 && used <= 100 * G) ? (int16_t)binpack_percent_fp64(used, G) : (int16_t)0;
}

// One full GPUs.Choose (allocate.go:92-100) on a row read from the global tables.
// do_rate=false is the plan-cache hit of NodeInfo.Bind (node.go:72-78): the plan is re-derived
// from the RemainLoad the filter pass left behind, without running Rate / LoadUsage again.
// everything one evaluation reads of its row, as it arrives from memory
struct RowIn {
    int G;
    uint32_t v[8];      // free
    uint2 t8[2];        // tenths per metric (load-aware pass only)
    uint32_t ts[2][8];  // update times per metric (load-aware pass only)
    uint2 r8;           // RemainLoad (other passes; zeros when the table is known to hold only zeros)
};

// wcx: the CTA's context for rows of more than 8 GPUs (ngs_wide.cuh).  WIDE_INLINE = false: such rows come back as
// "unknown": the instantiations for snapshots without such nodes (the host picks per launch, from what the snapshot
// holds) carry none of that code, and K7 evaluates such rows in a pass of its own.
template <bool PLAN, bool WIDE_INLINE = true>
__device__ __forceinline__ RowEval eval_row_values(const Tables& T, uint32_t row, const RowIn& in, const SortedDemand* dem,
                                                   const LoadParams& lp, bool do_rate, const double* pair, const WideCtx* wcx) {
    RowEval ev;
    ev.known = false;
    ev.feasible = false;
    ev.score = NGS_SCORE_MIN;
    ev.emit = 0;
    ev.n_emit = 0;
    const int policy = dem->policy;
    const bool load_path = lp.enabled && do_rate && policy != NGS_POLICY_RANDOM;
    const int G = in.G;
    const uint32_t (&v)[8] = in.v;
    const uint2 (&t8)[2] = in.t8;
    const uint32_t (&ts)[2][8] = in.ts;
    const uint2 r8 = in.r8;
    if (G == 0) return ev;
    if (G > 8) {  // 9..12 GPUs: the row's cells are not all in `in`
        if (!WIDE_INLINE) return ev;
        return wide_row_eval(wcx, row, dem, (do_rate ? WIDE_RATE : 0u) | (lp.enabled ? WIDE_LOAD : 0u));
    }
    ev.known = true;
    int f[8];
#pragma unroll
    for (int g = 0; g < 8; g++) f[g] = (int)v[g];
    int rlc[8];
    int load_int = 0;
    if (load_path) {
        // rater.go:62-66 / 116-120: LoadUsage per GPU, which also rewrites RemainLoad
        uint8_t tenths[2][8];
#pragma unroll
        for (int m = 0; m < 2; m++) {
#pragma unroll
            for (int g = 0; g < 4; g++) {
                tenths[m][g] = (uint8_t)(t8[m].x >> (8 * g));
                tenths[m][g + 4] = (uint8_t)(t8[m].y >> (8 * g));
            }
        }
        load_usage_row(tenths, ts, lp, G, rlc, &load_int, pair);
        uint2 out;
        out.x = out.y = 0;
#pragma unroll
        for (int g = 0; g < 4; g++) {
            out.x |= (uint32_t)(uint8_t)(int8_t)(g < G ? rlc[g] - 2 : 0) << (8 * g);
            out.y |= (uint32_t)(uint8_t)(int8_t)(g + 4 < G ? rlc[g + 4] - 2 : 0) << (8 * g);
        }
        *reinterpret_cast<uint2*>(T.rl + (size_t)row * 8) = out;
    } else {  // (r8 is all zeros when the table is known to be)
#pragma unroll
        for (int g = 0; g < 4; g++) {
            rlc[g] = (int)(int8_t)(r8.x >> (8 * g)) + 2;
            rlc[g + 4] = (int)(int8_t)(r8.y >> (8 * g)) + 2;
        }
    }
    if (do_rate) ev.score = rate_policy(policy, f, G, load_int, T.blut);
    if (!PLAN) {
        const int cls = classify_row(f, dem, policy != NGS_POLICY_RANDOM);
        if (cls != FEAS_UNDECIDED) {
            ev.feasible = cls == FEAS_YES;
            return ev;
        }
        if (policy != NGS_POLICY_RANDOM) {
            uint32_t rlmeta = 0;
#pragma unroll
            for (int g = 0; g < 8; g++) rlmeta |= (uint32_t)(g < G ? rlc[g] : 2) << (3 * g);
            ev.feasible = choose_sorted_slow((uint32_t)f[0] | ((uint32_t)f[1] << 16), (uint32_t)f[2] | ((uint32_t)f[3] << 16),
                                             (uint32_t)f[4] | ((uint32_t)f[5] << 16), (uint32_t)f[6] | ((uint32_t)f[7] << 16), rlmeta, G,
                                             policy == NGS_POLICY_SPREAD, dem);
            return ev;
        }
    }
    if (policy == NGS_POLICY_RANDOM) {
        ev.feasible = choose_sample<PLAN>(f, G, dem, &ev.emit, &ev.n_emit);
    } else {
        uint32_t w[8];
        bool rl0 = true;
#pragma unroll
        for (int g = 0; g < 8; g++) {
            w[g] = g < G ? pack_cell(f[g], rlc[g], g) : W_PAD;
            rl0 = rl0 && (g >= G || rlc[g] == 2);
        }
        ev.feasible = choose_sorted<PLAN>(w, G, policy == NGS_POLICY_SPREAD, rl0, dem, &ev.emit, &ev.n_emit);
    }
    return ev;
}

// One full GPUs.Choose (allocate.go:92-100) on a row read from the global tables.
// do_rate=false is the plan-cache hit of NodeInfo.Bind (node.go:72-78): the plan is re-derived
// from the RemainLoad the filter pass left behind, without running Rate / LoadUsage again.
// every load of a row issued before the first use, so that they are all in flight together
__device__ __forceinline__ void load_row(const Tables& T, uint32_t row, bool load_path, RowIn& in) {
    in.G = T.gcount[row];
    ld_row8(T.free + (size_t)row * 8, in.v);
    in.r8 = make_uint2(0u, 0u);
    if (load_path) {
#pragma unroll
        for (int m = 0; m < 2; m++) {
            const size_t off = ((size_t)m * T.nstride + row) * 8;
            in.t8[m] = __ldcg(reinterpret_cast<const uint2*>(T.tenths + off));
            ld_row8(T.ts + off, in.ts[m]);
        }
    } else if (!T.rl_zero) {
        in.r8 = __ldcg(reinterpret_cast<const uint2*>(T.rl + (size_t)row * 8));
    }
}

template <bool PLAN, bool WIDE_INLINE = true>
__device__ __forceinline__ RowEval eval_row_global(const Tables& T, uint32_t row, const SortedDemand* dem, const LoadParams& lp,
                                                   bool do_rate, const WideCtx* wcx, const double* pair = nullptr) {
    RowIn in;
    load_row(T, row, lp.enabled && do_rate && dem->policy != NGS_POLICY_RANDOM, in);
    return eval_row_values<PLAN, WIDE_INLINE>(T, row, in, dem, lp, do_rate, pair, wcx);
}
