// ngs_tables.cuh — the snapshot tables as the kernels see them, and what one row evaluation returns.
// Included by ngs_rows.cuh (inside ngs_dealer.cu's anonymous namespace) and by the host-compiled test harness.
#pragma once

struct Tables {
    int32_t* free;
    uint8_t* gcount;
    int8_t* rl;
    uint8_t* tenths;
    uint32_t* ts;
    uint32_t* tag;
    int32_t* cscore;
    // cells 8..11 of nodes with 9..12 GPUs (ngs_wide.cuh): [N][4] each, tenths / ts per metric plane like the
    // main tables; nullptr until the first such node is registered
    int32_t* free_hi;
    int8_t* rl_hi;
    uint8_t* tenths_hi;
    uint32_t* ts_hi;
    const int16_t* blut;  // Binpack.Rate percent table [9][801] (nullptr: compute in fp64)
    uint32_t rl_zero;     // the rl table is known to hold only zeros (no load-aware pass ran yet): skip reading it
    uint32_t nstride;  // rows per metric plane of tenths / ts (>= n; padded so that every plane is 256-row aligned)
    uint32_t n;      // rows in this shard
    uint32_t first;  // global node index of row 0
};

struct RowEval {
    bool known;
    bool feasible;
    int score;
    uint64_t emit;
    int n_emit;
};

// Launch-to-start latency on this driver depends on the size of a kernel's parameter block modulo 64: blocks of 64k+20 ..
// 64k+44 bytes start 0.3-0.45 us sooner than the others (profiles/tools/param_size_bench.cu, profiles/r02/param_size_bench.txt;
// found when four more table pointers made the single-pod kernel 0.5 us slower with identical SASS).  The latency-critical
// kernels take a trailing LaunchPad that puts their block at 64k+32; the static_asserts beside them keep that true when a
// parameter struct changes.
template <int N>
struct LaunchPad {
    unsigned char b[N];
};
template <class... Ts>
constexpr size_t param_block_bytes() {
    size_t off = 0;
    ((off = (off + alignof(Ts) - 1) / alignof(Ts) * alignof(Ts) + sizeof(Ts)), ...);
    return off;
}
