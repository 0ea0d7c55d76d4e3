// warp_pipeline2.cuh — the whole fp32 forward warp of a batch of units as ONE persistent kernel with
// WARP-granular work items (v4; included by tcwarp.cu after warp_pair.cuh).
//
// Work items: S(u) = the splat items of unit u (one 64 x PAIR_H strip per warp, warp_pair.cuh), N(u) = its
// normalise items (NORM2_QUADS quads per warp).  A unit's accumulators live in slot u % ring of a small
// ring, so they are written by L2 reductions and read back (and re-zeroed) by the normalise items while
// still L2-resident.  Completion: release = __threadfence + __syncwarp + one atomicAdd by lane 0; acquire =
// ld.acquire by lane 0 + shuffle.  Measured dead ends (profiles/r1_v4_notes.md): a fixed ticket order
// S(k) N(k-lag) (warps that drew a not-yet-ready N ticket spun: half of all issued instructions were polls),
// and fetch-add claim counters (three serial round trips to two hot words per 20 us strip, and blocking
// over-claims that could deadlock).  The schedule used now is static, see the kernel.
#pragma once

constexpr int PIPE2_RING = 12;                  // accumulator slots the workspace provides (upper bound of ring)
constexpr int NORM2_QUADS = 256;                // quads (VEC) or pixels (scalar) per normalise item
constexpr int PIPE2_THREADS = 128;
#ifndef TCW_SPLAT_MINB
#define TCW_SPLAT_MINB 4          // resident CTAs per SM the splat kernels are register-capped for
#endif

struct Pipe2Args {
    PairArgs A;
    UnitStats* stats_rw;
    int* ticket;
    int n_units, ns, nn, vec;
    int lag, ring;                              // ring >= lag + 2, ring <= PIPE2_RING
    int diag;                                   // TCW_DIAG bits (measurement only): 1 skip normalise body, 2 skip fences, 4 skip splat body
    float* out_frame; float* out_mask; float* out_depth; int* range_flag;
};

__device__ float log1x_fwd(float a) { return log1x(a); }

__device__ __forceinline__ int ld_relaxed(const int* p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Static schedule (cooperative launch: every warp of the grid is resident).  With W warps in the grid,
//   * warp w splats strips t = w, w + W, w + 2W, ... of the unit-major strip order (unit t / ns, strip t % ns),
//   * warp w owns normalise items q = w, w + W, ... of the unit-major item order (unit q / nn, item q % nn) and
//     runs each one as soon as it finds all splat items of that unit completed: it looks after every strip,
//     while it waits for a ring slot, and in a final drain loop.
// No claim counters, no hot words: per strip one acquire-load round (ring slot free? next duty ready?) and
// one completion increment.
__global__ void __launch_bounds__(PIPE2_THREADS, 4)
k_warp_pipeline2(Pipe2Args P) {
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const size_t hw = (size_t)P.A.h * P.A.w;
    const int W = gridDim.x * (PIPE2_THREADS / 32);
    const int wid = blockIdx.x * (PIPE2_THREADS / 32) + (threadIdx.x >> 5);
    const int total_s = P.n_units * P.ns, total_n = P.n_units * P.nn;
    bool bad = false;
    int q_d = wid;                                                   // my next normalise duty

    // runs my pending duties whose unit is fully splatted; returns true if the next duty (if any) is NOT ready
    auto duties = [&]() {
        while (q_d < total_n) {
            const int u = q_d / P.nn, item = q_d - u * P.nn;
            int rdy = 0;
            if (lane == 0) rdy = ld_acquire(&P.stats_rw[u].splat_done) >= P.ns;
            rdy = __shfl_sync(FULL, rdy, 0);
            if (!rdy) return;
            const int slot = u % P.ring;
            float4* a4 = P.A.acc4 + (size_t)slot * hw;
            float* aw = P.A.accw + (size_t)slot * hw;
            float* of = P.out_frame + (size_t)u * 3 * hw;
            float* om = P.out_mask + (size_t)u * hw;
            float* od = P.out_depth ? P.out_depth + (size_t)u * hw : nullptr;
            const size_t n_i = P.vec ? hw / 4 : hw;
            const size_t i0 = (size_t)item * NORM2_QUADS;
            #pragma unroll 2
            for (int k = 0; k < NORM2_QUADS / 32; ++k) {
                const size_t i = i0 + k * 32 + lane;
                if (i < n_i) {
                    if (P.vec) normalise_quad_fast(a4, aw, hw, i, of, om, od, bad);
                    else normalise_item<false>(a4, aw, hw, i, of, om, od, bad);
                }
            }
            __threadfence();
            __syncwarp();
            if (lane == 0) atomicAdd(&P.stats_rw[u].norm_done, 1);
            q_d += W;
        }
    };

    for (int t = wid; t < total_s; t += W) {
        const int u = t / P.ns, item = t - u * P.ns;
        if (u >= P.ring) {                                           // the slot's previous tenant must be normalised + re-zeroed
            for (;;) {
                int ok = 0;
                if (lane == 0) ok = ld_acquire(&P.stats_rw[u - P.ring].norm_done) >= P.nn;
                ok = __shfl_sync(FULL, ok, 0);
                if (ok) break;
                duties();                                            // never wait without serving my own duties
                __nanosleep(500);
            }
        }
        splat_pair_strip<PAIR_H>(P.A, u, u % P.ring, item);
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicAdd(&P.stats_rw[u].splat_done, 1);
        duties();
    }
    while (q_d < total_n) {                                          // drain: my remaining duties
        duties();
        if (q_d < total_n) __nanosleep(500);
    }
    if (bad && P.range_flag) atomicOr(P.range_flag, 1);
}

// per-chunk launch of the same strip code (used when the ring does not fit the workspace, and by ncu studies)
__global__ void __launch_bounds__(PIPE2_THREADS, TCW_SPLAT_MINB)
k_splat_pair(PairArgs A, int unit0) {
    const int strip = blockIdx.x * (PIPE2_THREADS / 32) + (threadIdx.x >> 5);
    if (strip >= A.strips_per_unit) return;                          // whole warps leave together
    const int u = unit0 + blockIdx.y;
#if TCW_ROWS2
    splat_pair_strip2<PAIR_H_CHUNK>(A, u, blockIdx.y, strip);
#else
    splat_pair_strip<PAIR_H_CHUNK>(A, u, blockIdx.y, strip);
#endif
}

#if TCW_CONST_M
// same kernel reading the matrix blocks from __constant__ memory (tcw_forward_warp copies them there per call)
__global__ void __launch_bounds__(PIPE2_THREADS, TCW_SPLAT_MINB)
k_splat_pair_cm(PairArgs A, int unit0) {
    const int strip = blockIdx.x * (PIPE2_THREADS / 32) + (threadIdx.x >> 5);
    if (strip >= A.strips_per_unit) return;
    const int u = unit0 + blockIdx.y;
    splat_pair_strip<PAIR_H_CHUNK, MODE_F32, true>(A, u, blockIdx.y, strip);
}
#endif

// normalise / mask / clip of a chunk of units as its own launch: each warp takes 128-pixel groups
__global__ void __launch_bounds__(256)
k_normalise_pair(float4* __restrict__ acc4, float* __restrict__ accw, int unit0, size_t hw,
                 float* __restrict__ out_frame, float* __restrict__ out_mask, float* __restrict__ out_depth,
                 int* __restrict__ range_flag) {
    const int uc = blockIdx.y, u = unit0 + uc;
    const size_t p0 = ((size_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 128;
    if (p0 >= hw) return;
    bool bad = false;
    normalise_group128(acc4 + (size_t)uc * hw, accw + (size_t)uc * hw, hw, p0, threadIdx.x & 31, out_frame + (size_t)u * 3 * hw,
                       out_mask + (size_t)u * hw, out_depth ? out_depth + (size_t)u * hw : nullptr, bad);
    if (bad && range_flag) atomicOr(range_flag, 1);
}

// deterministic modes of the same strip code (per-chunk launches): MODE_WMAX = max emitted weight per cell,
// MODE_I64 = order-free fixed-point sums relative to that maximum (see warp_strip.cuh emit<MODE_I64>)
template <int MODE>
__global__ void __launch_bounds__(PIPE2_THREADS, TCW_SPLAT_MINB)
k_splat_pair_det(PairArgs A, int unit0) {
    const int strip = blockIdx.x * (PIPE2_THREADS / 32) + (threadIdx.x >> 5);
    if (strip >= A.strips_per_unit) return;
    const int u = unit0 + blockIdx.y;
    splat_pair_strip<PAIR_H_CHUNK, MODE>(A, u, blockIdx.y, strip);
}
