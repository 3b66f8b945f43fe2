// xchg_bench.cu — micro-benchmark of grid-wide "every CTA posts a key, every CTA learns the max"
// exchange variants on one GPU (cooperative launch, 1 CTA per SM).  Build & run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/xchg profiles/tools/xchg_bench.cu && /tmp/xchg
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define VALID (1ULL << 63)
__device__ __forceinline__ void st_rlx(unsigned long long* p, unsigned long long v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned long long ld_rlx(const unsigned long long* p) { unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ uint32_t ld_rlx32(const uint32_t* p) { uint32_t v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ uint64_t wmax(uint64_t v) {
    uint32_t hi = v >> 32, mhi = __reduce_max_sync(~0u, hi), lo = hi == mhi ? (uint32_t)v : 0u, mlo = __reduce_max_sync(~0u, lo);
    return ((uint64_t)mhi << 32) | mlo;
}
template <int MODE>
__global__ void k(unsigned long long* slots, uint32_t* cnt, unsigned long long* mx, int rounds, int work, unsigned long long* out) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long acc = 0;
    long long t0 = clock64();
    for (int p = 0; p < rounds; p++) {
        // fake per-round work so that CTAs do not arrive in lock step
        int x = threadIdx.x + p;
        for (int i = 0; i < work + ((blockIdx.x * 7 + p) % 5) * 20; i++) x = x * 1664525 + 1013904223;
        if (x == 42) acc++;
        __syncthreads();
        if (warp == 0) {
            unsigned long long* row = slots + (size_t)p * gridDim.x;
            const unsigned long long key = ((unsigned long long)((blockIdx.x * 2654435761u + p * 40503u) & 0xFFFFFF) << 8) | 1;
            unsigned long long m = 0;
            if (MODE == 0) {  // slots with VALID bit, every lane polls its slots
                if (lane == 0) st_rlx(row + blockIdx.x, key | VALID);
                for (int j = lane; j < gridDim.x; j += 32) { unsigned long long v; do v = ld_rlx(row + j); while (!(v & VALID)); m = max(m, v); }
                m = wmax(m);
            } else if (MODE == 1) {  // slot + relaxed counter, one lane polls the counter, then slots (VALID-checked)
                if (lane == 0) { st_rlx(row + blockIdx.x, key | VALID); asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(cnt + p) : "memory"); while (ld_rlx32(cnt + p) < gridDim.x) {} }
                __syncwarp();
                for (int j = lane; j < gridDim.x; j += 32) { unsigned long long v; do v = ld_rlx(row + j); while (!(v & VALID)); m = max(m, v); }
                m = wmax(m);
            } else if (MODE == 2) {  // atomicMax into one word, then (data-dependent) counter bump; poll counter, read the word
                if (lane == 0) {
                    unsigned long long old = atomicMax(mx + p, key);
                    uint32_t one = 1 + (uint32_t)(old >> 63);  // old never has bit 63: forces the red after the atom returned
                    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(cnt + p), "r"(one) : "memory");
                    while (ld_rlx32(cnt + p) < gridDim.x) {}
                    m = ld_rlx(mx + p);
                }
                m = __shfl_sync(~0u, m, 0);
            } else if (MODE == 3) {  // like 1 but with a short sleep between polls
                if (lane == 0) { st_rlx(row + blockIdx.x, key | VALID); asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(cnt + p) : "memory"); while (ld_rlx32(cnt + p) < gridDim.x) __nanosleep(40); }
                __syncwarp();
                for (int j = lane; j < gridDim.x; j += 32) { unsigned long long v; do v = ld_rlx(row + j); while (!(v & VALID)); m = max(m, v); }
                m = wmax(m);
            } else if (MODE == 4) {  // like 0 but only lanes whose slot is still missing keep polling, with backoff
                if (lane == 0) st_rlx(row + blockIdx.x, key | VALID);
                for (int j = lane; j < gridDim.x; j += 32) { unsigned long long v = ld_rlx(row + j); while (!(v & VALID)) { __nanosleep(20); v = ld_rlx(row + j); } m = max(m, v); }
                m = wmax(m);
            }
            acc += m;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out[blockIdx.x * 2] = clock64() - t0; out[blockIdx.x * 2 + 1] = acc; }
}
template <int MODE> void run(const char* name, int work, int grid_override = 0) {
    int dev = 0, sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (grid_override) sms = grid_override;
    const int rounds = 2000;
    unsigned long long *slots, *mx, *out; uint32_t* cnt;
    cudaMalloc(&slots, 8ull * rounds * sms); cudaMalloc(&mx, 8ull * rounds); cudaMalloc(&cnt, 4ull * rounds); cudaMalloc(&out, 16ull * sms);
    for (int rep = 0; rep < 2; rep++) {
        cudaMemset(slots, 0, 8ull * rounds * sms); cudaMemset(mx, 0, 8ull * rounds); cudaMemset(cnt, 0, 4ull * rounds);
        int r = rounds; void* args[] = {&slots, &cnt, &mx, &r, &work, &out};
        cudaLaunchCooperativeKernel((void*)k<MODE>, dim3(sms), dim3(704), args, 0, 0);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    }
    unsigned long long h[512]; cudaMemcpy(h, out, 16ull * sms, cudaMemcpyDeviceToHost);
    double s = 0; for (int i = 0; i < sms; i++) s += h[2 * i];
    printf("%-44s work=%4d grid=%3d : %8.0f cycles per round\n", name, work, sms, s / sms / rounds);
    cudaFree(slots); cudaFree(mx); cudaFree(cnt); cudaFree(out);
}
int main() {
    for (int work : {0, 400}) {
        run<0>("0 slots+VALID, all lanes poll", work);
        run<1>("1 slots + counter, lane 0 polls counter", work);
        run<2>("2 atomicMax word + counter", work);
        run<3>("3 = 1 with nanosleep(40)", work);
        run<4>("4 = 0 with nanosleep(20) backoff", work);
    }
    for (int g : {16, 37, 74, 148}) run<2>("2 atomicMax word + counter", 0, g);
    return 0;
}
