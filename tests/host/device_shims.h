// device_shims.h — TEST HARNESS ONLY: what a host compiler needs in order to compile the device-side headers of
// nano-gpu-scheduler_b200/csrc (the CUDA intrinsics they use, in plain C++) and a single-threaded lockstep warp
// simulator for the warp-synchronous ones.  Included by tests/host/device_on_host.cpp and k6_on_host.cpp; nothing
// in the package includes, links or loads it.  Build with -ffp-contract=off.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

// ---- the intrinsics ngs_device.cuh uses, for a host compiler (built with -ffp-contract=off) ----
static inline int __ffs(unsigned x) { return x ? __builtin_ctz(x) + 1 : 0; }
static inline int __clz(unsigned x) { return x ? __builtin_clz(x) : 32; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __double2int_rz(double x) { return (int)x; }  // truncation toward zero, like Go's int(float64)
static inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
static inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
static inline double __ddiv_rn(double a, double b) { volatile double r = a / b; return r; }
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline int min(int a, int b) { return a < b ? a : b; }  // CUDA's global-namespace overloads
static inline int max(int a, int b) { return a > b ? a : b; }
static inline uint32_t min(uint32_t a, uint32_t b) { return a < b ? a : b; }
static inline uint32_t max(uint32_t a, uint32_t b) { return a > b ? a : b; }

// ---- a lockstep warp, single-threaded: 32 coroutines (ucontext), one per lane.  A lane runs until it reaches a
// warp-synchronous op, posts its operand and yields; once every lane has posted, the results are handed out and
// the lanes run on.  Lanes that reach DIFFERENT ops, or a lane that returns while others wait at an op, would
// hang a real warp (full-mask ops): the simulator reports that instead (warp_sim_error).
#include <ucontext.h>
namespace warpsim {
enum Op { NONE = 0, BALLOT, SHFL, SHFL_XOR, SHFL_UP, REDUCE_MAX, SYNC, DONE };
struct Lane {
    ucontext_t ctx;
    Op op = NONE;
    uint64_t val = 0;   // operand
    int arg = 0;        // source lane / xor mask
    uint64_t res = 0;   // result
    char* stack = nullptr;
};
static Lane g_lane[32];
static ucontext_t g_sched;
static int g_cur = -1;          // lane being run (-1: not inside a simulated warp)
static const char* g_error = nullptr;
static void (*g_body)(int lane, void* arg);
static void* g_body_arg;

static void trampoline() {
    g_body(g_cur, g_body_arg);
    g_lane[g_cur].op = DONE;
    swapcontext(&g_lane[g_cur].ctx, &g_sched);
}
static uint64_t post(Op op, uint64_t v, int arg) {
    if (g_cur < 0) {
        fprintf(stderr, "device_on_host: warp-synchronous op outside a simulated warp\n");
        abort();
    }
    Lane& L = g_lane[g_cur];
    L.op = op;
    L.val = v;
    L.arg = arg;
    swapcontext(&L.ctx, &g_sched);  // back when the op has been resolved
    return L.res;
}
// run body(lane, arg) on 32 lanes in lockstep; nullptr on success, else what went wrong
static const char* run_warp(void (*body)(int, void*), void* arg) {
    constexpr size_t STACK = 256 * 1024;
    g_body = body;
    g_body_arg = arg;
    g_error = nullptr;
    for (int l = 0; l < 32; l++) {
        Lane& L = g_lane[l];
        if (!L.stack) L.stack = (char*)malloc(STACK);
        getcontext(&L.ctx);
        L.ctx.uc_stack.ss_sp = L.stack;
        L.ctx.uc_stack.ss_size = STACK;
        L.ctx.uc_link = &g_sched;
        L.op = NONE;
        makecontext(&L.ctx, trampoline, 0);
    }
    for (;;) {
        for (int l = 0; l < 32; l++) {  // let every live lane run to its next op
            if (g_lane[l].op == DONE) continue;
            g_cur = l;
            swapcontext(&g_sched, &g_lane[l].ctx);
        }
        g_cur = -1;
        int done = 0;
        for (int l = 0; l < 32; l++) done += g_lane[l].op == DONE;
        if (done == 32) return nullptr;
        const Op op = g_lane[0].op;
        for (int l = 0; l < 32; l++)
            if (g_lane[l].op != op) {
                g_error = done ? "a lane returned while others wait at a full-mask warp op (a real warp would hang)"
                               : "lanes reached different warp-synchronous ops (divergent control flow around a sync op)";
                return g_error;  // (the abandoned coroutines are simply dropped: their stacks are reused by the next run)
            }
        if (op == BALLOT) {
            uint32_t m = 0;
            for (int l = 0; l < 32; l++) m |= (g_lane[l].val ? 1u : 0u) << l;
            for (int l = 0; l < 32; l++) g_lane[l].res = m;
        } else if (op == SHFL) {
            for (int l = 0; l < 32; l++) g_lane[l].res = g_lane[g_lane[l].arg & 31].val;
        } else if (op == SHFL_XOR) {
            for (int l = 0; l < 32; l++) g_lane[l].res = g_lane[(l ^ g_lane[l].arg) & 31].val;
        } else if (op == SHFL_UP) {
            for (int l = 0; l < 32; l++) g_lane[l].res = l >= g_lane[l].arg ? g_lane[l - g_lane[l].arg].val : g_lane[l].val;
        } else if (op == REDUCE_MAX) {
            uint64_t m = 0;
            for (int l = 0; l < 32; l++) m = g_lane[l].val > m ? g_lane[l].val : m;
            for (int l = 0; l < 32; l++) g_lane[l].res = m;
        }  // SYNC: nothing to hand out, every lane has arrived
    }
}
template <class T> static inline uint64_t bits(T v) {
    static_assert(sizeof(T) <= 8, "shuffle operand");
    uint64_t b = 0;
    memcpy(&b, &v, sizeof(T));
    return b;
}
template <class T> static inline T from(uint64_t b) {
    T v;
    memcpy(&v, &b, sizeof(T));
    return v;
}
}  // namespace warpsim
static inline unsigned __ballot_sync(unsigned, bool pred) { return (unsigned)warpsim::post(warpsim::BALLOT, pred ? 1 : 0, 0); }
template <class T> static inline T __shfl_sync(unsigned, T v, int src) { return warpsim::from<T>(warpsim::post(warpsim::SHFL, warpsim::bits(v), src)); }
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m) { return warpsim::from<T>(warpsim::post(warpsim::SHFL_XOR, warpsim::bits(v), m)); }


template <class T> static inline T __shfl_up_sync(unsigned, T v, unsigned d) { return warpsim::from<T>(warpsim::post(warpsim::SHFL_UP, warpsim::bits(v), (int)d)); }
static inline unsigned __reduce_max_sync(unsigned, unsigned v) { return (unsigned)warpsim::post(warpsim::REDUCE_MAX, v, 0); }
static inline bool __any_sync(unsigned m, bool pred) { return __ballot_sync(m, pred) != 0u; }
static inline bool __all_sync(unsigned m, bool pred) { return __ballot_sync(m, pred) == 0xFFFFFFFFu; }
static inline void __syncwarp(unsigned = 0xFFFFFFFFu) { warpsim::post(warpsim::SYNC, 0, 0); }
// one simulated warp is the whole "block" here: a block barrier is a warp barrier
static inline void __syncthreads() { warpsim::post(warpsim::SYNC, 0, 0); }
static inline long long clock64() { return 0; }
template <class T> static inline T __ldcg(const T* p) { return *p; }
static inline void __threadfence() {}
static inline void __threadfence_system() {}
static inline void __threadfence_block() {}
// atomics of a single-threaded simulation
static inline int atomicExch(int* p, int v) { int o = *p; *p = v; return o; }
static inline int atomicCAS(int* p, int c, int v) { int o = *p; if (o == c) *p = v; return o; }
static inline unsigned atomicAdd(unsigned* p, unsigned v) { unsigned o = *p; *p = o + v; return o; }
static inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; *p = o + v; return o; }
static inline unsigned long long atomicMax(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; if (v > o) *p = v; return o; }
static inline int atomicMax(int* p, int v) { int o = *p; if (v > o) *p = v; return o; }
static inline unsigned atomicOr(unsigned* p, unsigned v) { unsigned o = *p; *p = o | v; return o; }
#ifndef __noinline__
#define __noinline__ __attribute__((noinline))
#endif
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif
