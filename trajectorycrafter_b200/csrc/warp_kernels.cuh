// warp_kernels.cuh — the forward-warp kernels (included by tcwarp.cu after the device helpers).
//
//   (K0, the soft-z normaliser pre-pass, lives in warp_fast.cuh)
//   (K2, the fused project + splat kernel, lives in warp_strip.cuh)
//   K3 k_normalise<CLEAN, DET>      normalise / mask / clip (+ fused clean_points)   (:566-583, 585-626)
#pragma once

enum { MODE_F32 = 0, MODE_WMAX = 1, MODE_I64 = 2 };

// Per-unit header in the workspace: soft-z normaliser and depth scale, as float bit patterns.
struct UnitStats {
    unsigned lbits;    // max log(1+clamp(z',0,1000))   (>= 0, so uint order == float order)
    unsigned zabs;     // max |z'| (scale of the fixed-point depth channel)
    int splat_done;    // pipeline kernel: splat items of this unit that have finished
    int norm_done;     // pipeline kernel: normalise items of this unit that have finished
    int splat_claim;   // pipeline kernel v4: splat items of this unit handed out so far
    int norm_claim;    // pipeline kernel v4: normalise items of this unit handed out so far
};

// block-wide sum of NV values held by every thread; result valid in thread 0
template <typename T, int NV>
__device__ __forceinline__ void block_sum(T (&v)[NV], T* smem /* [8][NV] */) {
    #pragma unroll
    for (int i = 0; i < NV; ++i) {
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    }
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        #pragma unroll
        for (int i = 0; i < NV; ++i) smem[warp * NV + i] = v[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        #pragma unroll
        for (int i = 0; i < NV; ++i) {
            T t = smem[i];
            for (int wv = 1; wv < TILE_THREADS / 32; ++wv) t += smem[wv * NV + i];
            v[i] = t;
        }
    }
}

// Fixed-point scale of the deterministic path: every |q| <= 64 * 2^S and at most 4*h*w contributions
// can meet in one cell, so S = 56 - ceil(log2(4hw)) keeps every sum below 2^62.
__host__ __device__ inline int det_shift(int h, int w) {
    unsigned long long n = 4ull * (unsigned long long)h * (unsigned long long)w;
    int lg = 0;
    while ((1ull << lg) < n) ++lg;
    int s = 56 - lg;
    return s < 8 ? 8 : s;
}

// ----------------------------------------------------------------------------------------------
// K3: normalise / mask / clip (+ optional fused clean_points)   (models/utils.py:566-583, 585-626)
//   DET = false: fp32 sums;  DET = true: int64 sums (exact), ratio taken in double.
// ----------------------------------------------------------------------------------------------
template <bool CLEAN, bool DET>
__global__ void __launch_bounds__(TILE_THREADS)
k_normalise(const float4* __restrict__ acc4, const float* __restrict__ accw, const long long* __restrict__ acc_i64,
            const UnitStats* __restrict__ stats, int stat_index_global,
            int unit0, int h, int w, int tiles_x,
            void* __restrict__ out_frame, void* __restrict__ out_mask, float* __restrict__ out_depth,
            int* __restrict__ range_flag) {
    const int uc = blockIdx.y, u = unit0 + uc;
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    if (x >= w || y >= h) return;
    const size_t hw = (size_t)h * w, p = (size_t)y * w + x;
    const float4* a4 = DET ? nullptr : acc4 + (size_t)uc * hw;
    const float* aw = DET ? nullptr : accw + (size_t)uc * hw;
    const long long* ai = DET ? acc_i64 + (size_t)uc * NCH * hw : nullptr;
    bool valid;
    float col[3], dep;
    if (!DET) {
        const float W = aw[p];
        const float4 S = a4[p];
        valid = W > 0.0f;
        col[0] = valid ? __fdiv_rn(S.x, W) : -1.0f;
        col[1] = valid ? __fdiv_rn(S.y, W) : -1.0f;
        col[2] = valid ? __fdiv_rn(S.z, W) : -1.0f;
        dep = valid ? __fdiv_rn(S.w, W) : 0.0f;
    } else {
        const long long Q = ai[4 * hw + p];
        valid = Q > 0;
        const double inv = valid ? 1.0 / (double)Q : 0.0;
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) col[ch] = valid ? (float)((double)ai[ch * hw + p] * inv) : -1.0f;
        float zabs = __uint_as_float(stats[stat_index_global >= 0 ? stat_index_global : u].zabs);
        if (!(zabs > 0.0f)) zabs = 1.0f;
        dep = valid ? (float)((double)ai[3 * hw + p] * inv * (double)zabs) : 0.0f;
    }
    bool bad = false;
    #pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        bad |= !(col[ch] >= -1.1f) || !(col[ch] <= 1.1f);      // the reference asserts here (:580-581)
        col[ch] = clampf(col[ch], -1.0f, 1.0f);
    }
    if (bad && range_flag) atomicOr(range_flag, 1);
    if (out_depth) out_depth[(size_t)u * hw + p] = dep;
    if (!CLEAN) {
        float* of = (float*)out_frame;
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) of[((size_t)u * 3 + ch) * hw + p] = col[ch];
        ((float*)out_mask)[(size_t)u * hw + p] = valid ? 1.0f : 0.0f;
    } else {
        // clean_points: hole mask dilated 5x5 (outside pixels never win), float64 outputs
        bool hole = false;
        #pragma unroll
        for (int dy = -2; dy <= 2; ++dy) {
            const int yy = y + dy;
            if (yy < 0 || yy >= h) continue;
            #pragma unroll
            for (int dx = -2; dx <= 2; ++dx) {
                const int xx = x + dx;
                if (xx < 0 || xx >= w) continue;
                const size_t q = (size_t)yy * w + xx;
                hole |= DET ? !(ai[4 * hw + q] > 0) : !(aw[q] > 0.0f);
            }
        }
        const double keep = hole ? 0.0 : 1.0;
        double* of = (double*)out_frame;
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float half = __fdiv_rn(__fadd_rn(col[ch], 1.0f), 2.0f);
            of[((size_t)u * 3 + ch) * hw + p] = __dsub_rn(__dmul_rn(__dmul_rn((double)half, keep), 2.0), 1.0);
        }
        ((double*)out_mask)[(size_t)u * hw + p] = keep;
    }
}
