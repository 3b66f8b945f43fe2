// mapped_write_bench.cu — how fast do kernel stores into mapped pinned host memory drain over PCIe, by store width?
// (decides how the verb kernels should write feasible / reason / score for the host: profiles/r02/mapped_write_bench.txt)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mapped_write_bench mapped_write_bench.cu
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

__device__ unsigned int g_ticket;

template <typename T, bool CTA_FENCE = false>
__global__ void k_write(T* dst, size_t n_elems, volatile int* flag, int seq) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_elems; i += (size_t)gridDim.x * blockDim.x) {
        T v;
        memset(&v, (int)(i & 0x7f), sizeof(T));
        dst[i] = v;
    }
    if (!CTA_FENCE) __threadfence_system();  // every thread fences its own stores ...
    __syncthreads();
    if (threadIdx.x == 0) {
        if (CTA_FENCE) __threadfence_system();  // ... or one thread per CTA does, after the barrier (fences are cumulative)
        if (atomicAdd(&g_ticket, 1u) == gridDim.x - 1) {
            g_ticket = 0;
            __threadfence_system();
            *flag = seq;
        }
    }
}

template <typename T, bool CTA_FENCE = false>
double run(void* d_dst, size_t bytes, int* h_flag, int* d_flag, int& seq, cudaStream_t st) {
    const size_t n = bytes / sizeof(T);
    int grid = (int)std::min<size_t>((n + 255) / 256, 148 * 4);
    if (grid < 1) grid = 1;
    std::vector<double> t;
    for (int it = 0; it < 60; it++) {
        ++seq;
        auto t0 = std::chrono::steady_clock::now();
        k_write<T, CTA_FENCE><<<grid, 256, 0, st>>>((T*)d_dst, n, d_flag, seq);
        while (*(volatile int*)h_flag != seq) {}
        auto t1 = std::chrono::steady_clock::now();
        if (it >= 10) t.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
    }
    std::sort(t.begin(), t.end());
    return t[t.size() / 2];
}

int main() {
    cudaSetDevice(0);
    cudaStream_t st;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    const size_t MAXB = 1 << 20;
    void *h_buf, *d_buf;
    int *h_flag, *d_flag;
    cudaHostAlloc(&h_buf, MAXB, cudaHostAllocMapped);
    cudaHostGetDevicePointer(&d_buf, h_buf, 0);
    cudaHostAlloc((void**)&h_flag, 64, cudaHostAllocMapped);
    cudaHostGetDevicePointer((void**)&d_flag, h_flag, 0);
    *h_flag = 0;
    void* d_dev;
    cudaMalloc(&d_dev, MAXB);
    int seq = 0;
    printf("bytes      target  u8/thread  u32/thread  uint4/thread   (us from launch to flag on the host, p50)\n");
    for (size_t bytes : {(size_t)256, (size_t)100000, (size_t)200000, (size_t)400000, (size_t)600000, (size_t)1000000}) {
        for (int dev = 0; dev < 2; dev++) {
            void* dst = dev ? d_dev : d_buf;
            const double a = run<uint8_t>(dst, bytes, h_flag, d_flag, seq, st);
            const double b = run<uint32_t>(dst, bytes, h_flag, d_flag, seq, st);
            const double c = run<uint4>(dst, bytes / 16 * 16, h_flag, d_flag, seq, st);
            const double e = run<uint32_t, true>(dst, bytes, h_flag, d_flag, seq, st);
            printf("%8zu   %s   %8.1f   %8.1f   %8.1f   | one fence per CTA, u32: %8.1f\n", bytes, dev ? "device" : "host  ", a, b, c, e);
        }
    }
    return 0;
}
