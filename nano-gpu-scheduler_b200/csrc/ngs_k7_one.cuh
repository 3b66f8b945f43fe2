// ngs_k7_one.cuh — K7, one pod's whole cycle (filter + prioritize + arg-max + bind) in ONE ordinary
// launch, for the decision-latency half of the metric.  K6 is built for batches: a demand-prep
// launch, a memset and a cooperative launch that first moves every row into shared memory.  For a
// single pod none of that pays, so this kernel scans the tables where they lie (K1's row
// evaluation), folds the keys with one atomicMax per CTA, and the CTA that arrives last — it can
// tell from a ticket counter, nobody waits for anybody — binds the winner and leaves the exchange
// record zeroed for the next call.  Binpack / spread only; the random policy and node-range
// sharding stay with K6.  Same results as K6 with P = 1 (tests/test_gpu_parity.py).
// Included by ngs_dealer.cu inside its anonymous namespace, after ngs_k6_batch.cuh.
#pragma once

struct OneRec {
    unsigned long long key;  // arg-max over the CTAs' best keys (0 = nothing feasible)
    unsigned int arrived;    // CTAs that have folded their key in
    unsigned int pad;
};

constexpr int32_t ONE_PENDING = 0x7FFFFFFF;  // status word while the kernel has not answered yet

// out (mapped host memory): [0] status, [8] node, [9] score, [10 .. 10+max_c) gpu indexes
__global__ void __launch_bounds__(256, 4) k_schedule_one(Tables T, PodDemand pd, LoadParams lp, OneRec* rec, int32_t* out, int max_c) {
    __shared__ SortedDemand dem;
    __shared__ double s_pair[121];
    __shared__ uint64_t s_wmax[8];
    __shared__ int s_last;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    // the first row of every thread is requested before the demand is sorted: the loads do not depend on it
    const bool load_path = lp.enabled != 0;  // (binpack / spread only)
    uint32_t row = blockIdx.x * blockDim.x + tid;
    RowIn in;
    if (row < T.n) load_row(T, row, load_path, in);
    if (tid == 32) sort_demand(pd.percent, pd.n, pd.policy, &dem);
    if (lp.enabled) fill_usage_pair_table(s_pair, tid, blockDim.x);
    __syncthreads();
    uint64_t best = 0;
    for (; row < T.n; row += gridDim.x * blockDim.x) {
        // Rate + Choose on the row as it stands; in load-aware mode this also leaves RemainLoad behind (allocate.go:193)
        const RowEval ev = eval_row_values<false>(T, row, in, &dem, lp, true, s_pair);
        if (ev.known && ev.feasible) {
            const uint64_t k = make_key(ev.score, T.first + row);
            best = k > best ? k : best;
        }
        const uint32_t nxt = row + gridDim.x * blockDim.x;
        if (nxt < T.n) load_row(T, nxt, load_path, in);
    }
    best = warp_max_u64(best);
    if (lane == 0) s_wmax[warp] = best;
    __syncthreads();
    if (tid == 0) {
        uint64_t m = 0;
        for (uint32_t w = 0; w < blockDim.x / 32; w++) m = s_wmax[w] > m ? s_wmax[w] : m;
        if (m) atomicMax(&rec->key, (unsigned long long)m);
        __threadfence();  // this CTA's RemainLoad stores and its key, before the ticket
        s_last = atomicAdd(&rec->arrived, 1u) == gridDim.x - 1 ? 1 : 0;
    }
    __syncthreads();
    if (!s_last || warp != 0) return;

    // ---- warp 0 of the last CTA to arrive: every other CTA's key and RemainLoad rows are visible now ----
    __threadfence();
    unsigned long long key = 0;
    if (lane == 0) {
        key = atomicAdd(&rec->key, 0ull);
        rec->key = 0;  // ready for the next call (ordered by the kernel boundary)
        rec->arrived = 0;
    }
    key = __shfl_sync(FULL, key, 0);
    int32_t* gout = out + 10;
    for (int k = (int)lane; k < max_c; k += 32) gout[k] = INT32_MIN;
    __syncwarp();
    if (key == 0) {  // no node takes the pod
        if (lane == 0) {
            out[8] = -1;
            out[9] = NGS_SCORE_MIN;
        }
    } else {
        // NodeInfo.Bind (node.go:70-84) on the winner, eight lanes on the row's eight cells (the four
        // groups of the warp hold identical copies): the plan from the RemainLoad the scan left behind
        const uint32_t wrow = key_node(key) - T.first;
        const int g = (int)(lane & 7u), gbase = (int)(lane & 24u);
        const int G = T.gcount[wrow];
        const int cell = __ldcg(T.free + (size_t)wrow * 8 + g);
        const int rlc = (int)__ldcg(T.rl + (size_t)wrow * 8 + g) + 2;
        uint64_t emit = 0;
        int n_emit = 0, fr = 0;
        choose_group8<true>(g, gbase, cell, rlc, g < G, dem.policy == NGS_POLICY_SPREAD, &dem, &emit, &n_emit, &fr);
        if (lane < 8) {
            if (g < G) T.free[(size_t)wrow * 8 + lane] = fr;  // GPUs.Allocate (allocate.go:102-118)
            for (int k = (int)lane; k < dem.C; k += 8) {      // rater.go:102-105: un-permute the emission order
                const int e = (int)((emit >> (4 * k)) & 0xF);
                gout[dem.order[k]] = e == 0xF ? NGS_NOT_NEED_GPU : e;
            }
        }
        if (lane == 0) {
            T.tag[wrow] = 0;  // node.go:82
            out[8] = (int32_t)(T.first + wrow);
            out[9] = key_score(key);
        }
    }
    __syncwarp();
    if (lane == 0) {
        __threadfence_system();  // the answer is in host memory before the status word says so
        *reinterpret_cast<volatile int32_t*>(out) = NGS_OK;
    }
}
