// param_size_bench.cu — does the size of the kernel parameter block change launch-to-completion latency?
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o param_size_bench param_size_bench.cu
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>

template <int N>
struct Pad {
    int v[N / 4];
};
template <int N>
__global__ void k(Pad<N> p, volatile int* flag, int seq) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (p.v[N / 4 - 1] == 12345) seq++;
        __threadfence_system();
        *flag = seq;
    }
}
template <int N>
double run(int* h_flag, int* d_flag, int& seq, cudaStream_t st) {
    Pad<N> p = {};
    std::vector<double> t;
    for (int it = 0; it < 400; it++) {
        ++seq;
        auto t0 = std::chrono::steady_clock::now();
        k<N><<<391, 256, 0, st>>>(p, d_flag, seq);
        while (*(volatile int*)h_flag != seq) {}
        auto t1 = std::chrono::steady_clock::now();
        if (it >= 50) t.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
    }
    std::sort(t.begin(), t.end());
    return t[t.size() / 2];
}
int main() {
    cudaStream_t st;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    int *h_flag, *d_flag;
    cudaHostAlloc((void**)&h_flag, 64, cudaHostAllocMapped);
    cudaHostGetDevicePointer((void**)&d_flag, h_flag, 0);
    *h_flag = 0;
    int seq = 0;
    for (int rep = 0; rep < 3; rep++) {
        printf("%d:%.2f ", 280, run<280>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 288, run<288>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 296, run<296>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 304, run<304>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 312, run<312>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 320, run<320>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 328, run<328>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 336, run<336>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 344, run<344>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 352, run<352>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 360, run<360>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 368, run<368>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 376, run<376>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 384, run<384>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 392, run<392>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 400, run<400>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 408, run<408>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 416, run<416>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 424, run<424>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 432, run<432>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 440, run<440>(h_flag, d_flag, seq, st));
        printf("%d:%.2f ", 448, run<448>(h_flag, d_flag, seq, st));
        printf("\n");
    }
    return 0;
}
