// warp_pipeline.cuh — the whole fp32 forward warp of a batch of units as ONE persistent kernel.
//
// Work items are handed out by a global ticket counter in a fixed order
//     S(0) .. S(lag-1) | S(lag) N(0) | S(lag+1) N(1) | ... | S(n-1) N(n-1-lag) | N(n-lag) .. N(n-1)
// S(u) = the splat items of unit u (8 strips = one CTA pass each), N(u) = its normalise items.
// (lag is chosen so that lag * items-per-unit exceeds the number of resident CTAs: by the time a
// normalise ticket comes up, the splat items it depends on were handed out a full wave earlier).
// A unit's accumulators live in slot u % ring of a small ring (ring = lag + 2 units), so
// they are written by L2 reductions and read back by the normalise items while still L2-resident;
// the normalise items re-zero what they read, which hands the slot to unit u + ring.
// Dependencies are per-unit completion counters (release: __threadfence + atomicAdd, acquire:
// ld.acquire spin by one thread + __syncthreads).  A waiting item only ever waits for items with
// SMALLER tickets, which are finished or held by running CTAs, so the queue cannot deadlock.
// The memory-bound normalise items overlap the issue-bound splat items of the next unit on the
// same SMs, and there is one launch per call instead of two per chunk.
#pragma once

constexpr int PIPE_RING = 8;                   // accumulator slots the workspace provides (upper bound of ring)
constexpr int NORM_ITEM = 1024;                 // quads (VEC) or pixels (scalar) per normalise item

struct PipeArgs {
    SplatArgs A;
    UnitStats* stats_rw;
    int* ticket;
    int n_units, ns, nn, vec;
    int lag, ring;                              // ring >= lag + 2, ring <= PIPE_RING
    float* out_frame; float* out_mask; float* out_depth; int* range_flag;
};

__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void wait_at_least(const int* p, int target) {
    while (ld_acquire(p) < target) __nanosleep(100);
}

__global__ void __launch_bounds__(TILE_THREADS)
k_warp_pipeline(PipeArgs P) {
    __shared__ int s_ticket;
    const int tid = threadIdx.x;
    const int per_unit = P.ns + P.nn;
    const int total = P.n_units * per_unit;
    const size_t hw = (size_t)P.A.h * P.A.w;
    bool bad = false;
    for (;;) {
        if (tid == 0) s_ticket = atomicAdd(P.ticket, 1);
        __syncthreads();
        const int t = s_ticket;
        __syncthreads();
        if (t >= total) break;
        // group k holds S(k) (if k < n) followed by N(k - lag) (if k >= lag)
        int u = 0, item = 0;
        bool splat = true;
        {
            const int lagS = P.lag < P.n_units ? P.lag : P.n_units;       // leading groups: splat only
            const int head = lagS * P.ns;
            if (t < head) { u = t / P.ns; item = t - u * P.ns; }
            else {
                const int mid = (P.n_units - lagS) * per_unit;               // groups lagS .. n-1: S(k) N(k-lag)
                const int tp = t - head;
                if (tp < mid) {
                    const int k = lagS + tp / per_unit, r = tp - (k - lagS) * per_unit;
                    if (r < P.ns) { u = k; item = r; }
                    else { u = k - lagS; item = r - P.ns; splat = false; }
                } else {                                                     // trailing groups: normalise only
                    const int tq = tp - mid;
                    u = P.n_units - lagS + tq / P.nn; item = tq % P.nn; splat = false;
                }
            }
        }
        const int slot = u % P.ring;
        if (splat) {
            if (u >= P.ring) {                                      // the slot's previous tenant must be normalised + re-zeroed
                if (tid == 0) wait_at_least(&P.stats_rw[u - P.ring].norm_done, P.nn);
                __syncthreads();
            }
            const int strip = item * (TILE_THREADS / 32) + (tid >> 5);
            if (strip < P.A.strips_per_unit) splat_strip<MODE_F32>(P.A, u, slot, strip);
            __threadfence();
            __syncthreads();
            if (tid == 0) atomicAdd(&P.stats_rw[u].splat_done, 1);
        } else {
            if (tid == 0) wait_at_least(&P.stats_rw[u].splat_done, P.ns);
            __syncthreads();
            float4* a4 = P.A.acc4 + (size_t)slot * hw;
            float* aw = P.A.accw + (size_t)slot * hw;
            float* of = P.out_frame + (size_t)u * 3 * hw;
            float* om = P.out_mask + (size_t)u * hw;
            float* od = P.out_depth ? P.out_depth + (size_t)u * hw : nullptr;
            const size_t n_i = P.vec ? hw / 4 : hw;
            const size_t i0 = (size_t)item * NORM_ITEM;
            #pragma unroll
            for (int k = 0; k < NORM_ITEM / TILE_THREADS; ++k) {
                const size_t i = i0 + k * TILE_THREADS + tid;
                if (i < n_i) {
                    if (P.vec) normalise_item<true>(a4, aw, hw, i, of, om, od, bad);
                    else normalise_item<false>(a4, aw, hw, i, of, om, od, bad);
                }
            }
            __threadfence();
            __syncthreads();
            if (tid == 0) atomicAdd(&P.stats_rw[u].norm_done, 1);
        }
    }
    if (bad && P.range_flag) atomicOr(P.range_flag, 1);
}
