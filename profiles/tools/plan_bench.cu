// plan_bench.cu — micro-benchmark: latency of one exact plan (GPUs.Choose + the row after the bind + the next pod's
// verdict) for a single warp, one-cell-per-lane formulation (choose_group8, shuffles and votes) against the
// thread-serial one (choose_sorted, every lane redundantly), warm instruction cache, rows and demands in shared memory.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nano-gpu-scheduler_b200/csrc -o /tmp/plan_bench profiles/tools/plan_bench.cu && /tmp/plan_bench
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include "ngs_device.cuh"
namespace cg = cooperative_groups;
using namespace ngs;
namespace {
#include "ngs_rows.cuh"
#include "ngs_k6_batch.cuh"
struct GroupCommitL {
    int32_t fr; uint64_t emit; int32_t n_emit; uint32_t top4; int32_t scB, scS; uint64_t fresh; int32_t fresh_ok;
};
__device__ __forceinline__ void plan_group_lanes(const BatchParams& bp, const SmemRows& S, uint32_t r, uint32_t node, const SortedDemand* dem,
                                           const SortedDemand* dem_next, int g, int gbase, GroupCommitL& cm, long long* stamps = nullptr) {
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
    const int pct = (bp.T.blut != nullptr && used >= 0 && used <= 100 * G) ? (int)__ldg(bp.T.blut + G * RATE_LUT_STRIDE + used)
                                                                            : binpack_percent_fp64(used, G);
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
    if (cls == FEAS_UNDECIDED)
        cls = choose_group8<false>(g, gbase, fr, rlc, valid, pol2 == NGS_POLICY_SPREAD, dem_next, nullptr, nullptr, nullptr) ? FEAS_YES
                                                                                                                               : FEAS_NO;
    cm.fresh_ok = cls == FEAS_YES;
    if (cls == FEAS_YES && pol2 != NGS_POLICY_RANDOM) cm.fresh = make_key(pol2 == NGS_POLICY_SPREAD ? cm.scS : cm.scB, node);
}

constexpr int NROWS = 256, NDEM = 64;
template <int MODE>
__global__ void k(const uint2* fA, const uint2* fB, const uint32_t* meta, const SortedDemand* dems, int iters, long long* out, unsigned long long* sink) {
    __shared__ uint2 sA[NROWS], sB[NROWS];
    __shared__ uint32_t sM[NROWS], sT[NROWS];
    __shared__ int32_t sSB[NROWS], sSS[NROWS];
    __shared__ SortedDemand sD[NDEM];
    for (int i = threadIdx.x; i < NROWS; i += blockDim.x) { sA[i] = fA[i]; sB[i] = fB[i]; sM[i] = meta[i]; sT[i] = 0; sSB[i] = 0; sSS[i] = 0; }
    for (int i = threadIdx.x; i < NDEM * (int)sizeof(SortedDemand) / 4; i += blockDim.x) ((uint32_t*)sD)[i] = ((const uint32_t*)dems)[i];
    __syncthreads();
    SmemRows S; S.fA = sA; S.fB = sB; S.meta = sM; S.scB = sSB; S.scS = sSS; S.top4 = sT; S.feasbits = nullptr;
    BatchParams bp; memset(&bp, 0, sizeof bp);
    const int lane = threadIdx.x & 31, g = lane & 7, gbase = lane & 24;
    unsigned long long acc = 0;
    long long t0 = 0;
    for (int it = -8; it < iters; it++) {
        if (it == 0) t0 = clock64();
        const int r = (it * 37 + 5) & (NROWS - 1), d = (it * 11 + 3) & (NDEM - 1), d2 = (d + 1) & (NDEM - 1);
        if (MODE == 0) { GroupCommitL cm; plan_group_lanes(bp, S, r, r, &sD[d], &sD[d2], g, gbase, cm); acc += cm.emit + cm.top4 + cm.scB + cm.fresh + cm.fr; }
        else { GroupCommit cm; plan_group(bp, S, r, r, &sD[d], &sD[d2], g, gbase, cm); acc += cm.emit + cm.top4 + cm.scB + cm.fresh + cm.fr; }
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) { out[0] = (t1 - t0) / iters; }
    sink[threadIdx.x] = acc;
}
}  // namespace
int main() {
    std::vector<uint2> fA(NROWS), fB(NROWS); std::vector<uint32_t> meta(NROWS); std::vector<SortedDemand> dem(NDEM);
    srand(7);
    for (int nzmax = 1; nzmax <= 4; nzmax++) {
        for (int i = 0; i < NROWS; i++) {
            int f[8]; uint32_t m = (8u << 24);
            for (int c = 0; c < 8; c++) { f[c] = (rand() & 1) ? 100 : 5 * (rand() % 20); m |= (uint32_t)(2 + (rand() % 3) - (rand() % 2)) << (3 * c); }
            fA[i] = make_uint2(f[0] | (f[1] << 16), f[2] | (f[3] << 16)); fB[i] = make_uint2(f[4] | (f[5] << 16), f[6] | (f[7] << 16)); meta[i] = m;
        }
        for (int i = 0; i < NDEM; i++) {
            SortedDemand& D = dem[i]; memset(&D, 0, sizeof D);
            int C = nzmax, v[12];
            for (int c = 0; c < C; c++) v[c] = 5 * (1 + rand() % 12);
            for (int a = 0; a < C; a++) for (int b = a + 1; b < C; b++) if (v[b] > v[a]) { int t = v[a]; v[a] = v[b]; v[b] = t; }
            D.C = C; D.nz = C; D.policy = 0;
            for (int c = 0; c < C; c++) { D.raw[c] = v[c]; D.d[c] = v[c]; D.order[c] = c; }
        }
        uint2 *dA, *dB; uint32_t* dM; SortedDemand* dD; long long* dO; unsigned long long* dS;
        cudaMalloc(&dA, NROWS * 8); cudaMalloc(&dB, NROWS * 8); cudaMalloc(&dM, NROWS * 4); cudaMalloc(&dD, NDEM * sizeof(SortedDemand)); cudaMalloc(&dO, 64); cudaMalloc(&dS, 8 * 64);
        cudaMemcpy(dA, fA.data(), NROWS * 8, cudaMemcpyHostToDevice); cudaMemcpy(dB, fB.data(), NROWS * 8, cudaMemcpyHostToDevice);
        cudaMemcpy(dM, meta.data(), NROWS * 4, cudaMemcpyHostToDevice); cudaMemcpy(dD, dem.data(), NDEM * sizeof(SortedDemand), cudaMemcpyHostToDevice);
        long long o0 = 0, o1 = 0;
        k<0><<<1, 32>>>(dA, dB, dM, dD, 512, dO, dS); cudaMemcpy(&o0, dO, 8, cudaMemcpyDeviceToHost);
        k<1><<<1, 32>>>(dA, dB, dM, dD, 512, dO, dS); cudaMemcpy(&o1, dO, 8, cudaMemcpyDeviceToHost);
        printf("containers=%d: one cell per lane (shuffles) %lld cycles per plan, thread-serial %lld cycles per plan (%s)\n", nzmax, o0, o1, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
