// warp_kernels.cuh — the forward-warp kernels (included by tcwarp.cu after the device helpers).
//
//   K0 k_zmax          per-unit max of log(1+sat(z')) and max |z'|            (models/utils.py:490-492)
//   K1 k_project_splat<MODE_WMAX>   deterministic path only: per-cell max weight (order-free)
//   K2 k_project_splat<MODE_F32>    fused unproject/transform/project + soft-z splat, fp32 L2 reductions
//      k_project_splat<MODE_I64>    same, order-independent int64 fixed point relative to the cell's max weight
//   K3 k_normalise<CLEAN, DET>      normalise / mask / clip (+ fused clean_points)   (:566-583, 585-626)
#pragma once

enum { MODE_F32 = 0, MODE_WMAX = 1, MODE_I64 = 2 };

// Per-unit header in the workspace: soft-z normaliser and depth scale, as float bit patterns.
struct UnitStats {
    unsigned lbits;    // max log(1+clamp(z',0,1000))   (>= 0, so uint order == float order)
    unsigned zabs;     // max |z'| (scale of the fixed-point depth channel)
};

__global__ void __launch_bounds__(TILE_THREADS)
k_zmax(const float* __restrict__ depth, const int* __restrict__ src_index, const float* __restrict__ Mall,
       int unit0, int h, int w, int tiles_x, UnitStats* __restrict__ stats, int n_units_total) {
    __shared__ float M[M_FLOATS];
    const int u = unit0 + blockIdx.y;
    if (threadIdx.x < M_FLOATS) M[threadIdx.x] = Mall[(size_t)u * M_FLOATS + threadIdx.x];
    __syncthreads();
    const int s = src_index ? src_index[u] : u;
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    float ld = 0.0f, za = 0.0f;
    if (x < w && y < h) {
        float d = depth[(size_t)s * h * w + (size_t)y * w + x];
        float P0, P1, P2;
        project_px(M, (float)x, (float)y, d, P0, P1, P2);
        ld = log_depth(P2);
        za = fabsf(P2);
    }
    const float m = block_max_256(ld);
    __syncthreads();
    const float mz = block_max_256(za);
    if (threadIdx.x == 0) {
        atomicMax(&stats[u].lbits, __float_as_uint(m));
        atomicMax(&stats[n_units_total].lbits, __float_as_uint(m));
        atomicMax(&stats[u].zabs, __float_as_uint(mz));
        atomicMax(&stats[n_units_total].zabs, __float_as_uint(mz));
    }
}

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

template <int MODE>
__global__ void __launch_bounds__(TILE_THREADS)
k_project_splat(const float* __restrict__ frame, const float* __restrict__ mask1,
                const float* __restrict__ depth, const int* __restrict__ src_index,
                const float* __restrict__ Mall, int unit0, int h, int w, int tiles_x,
                const UnitStats* __restrict__ stats, int stat_index_global /* -1: per unit */,
                float* __restrict__ acc_f32, long long* __restrict__ acc_i64, unsigned* __restrict__ wmax,
                float det_scale, float* __restrict__ out_flow) {
    __shared__ float M[M_FLOATS];
    __shared__ long long red[8 * 5];
    const int uc = blockIdx.y, u = unit0 + uc;
    if (threadIdx.x < M_FLOATS) M[threadIdx.x] = Mall[(size_t)u * M_FLOATS + threadIdx.x];
    __syncthreads();
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    const bool inb = x < w && y < h;
    const int s = src_index ? src_index[u] : u;
    const size_t hw = (size_t)h * w, p = inb ? (size_t)y * w + x : 0;
    const UnitStats st = stats[stat_index_global >= 0 ? stat_index_global : u];
    const float L = __uint_as_float(st.lbits);
    float zabs = __uint_as_float(st.zabs);
    if (!(zabs > 0.0f)) zabs = 1.0f;

    float P0 = 0.f, P1 = 0.f, P2 = 1.f, r = 0.f, gch = 0.f, b = 0.f, m = 0.f;
    bool behind = false;
    if (inb) {
        const float d = depth[(size_t)s * hw + p];
        project_px(M, (float)x, (float)y, d, P0, P1, P2);
        // the reference overwrites behind-camera points with exactly (1000,1000,1000): all of them
        // project to (1,1) with identical soft-z weight — a single-address hotspot, aggregated below
        behind = (P0 == 1000.0f) & (P1 == 1000.0f) & (P2 == 1000.0f);
        m = mask1 ? mask1[(size_t)s * hw + p] : 1.0f;
        if (MODE != MODE_WMAX) {
            r = frame[((size_t)s * 3 + 0) * hw + p];
            gch = frame[((size_t)s * 3 + 1) * hw + p];
            b = frame[((size_t)s * 3 + 2) * hw + p];
        }
    }
    // coords = P.xy / P.z ; flow12 = coords - grid   (models/utils.py:273-278)
    const float flx = __fsub_rn(__fdiv_rn(P0, P2), (float)x);
    const float fly = __fsub_rn(__fdiv_rn(P1, P2), (float)y);
    if (MODE != MODE_WMAX && out_flow && inb) {
        out_flow[(size_t)u * 2 * hw + p] = flx;
        out_flow[(size_t)u * 2 * hw + hw + p] = fly;
    }
    const float g = expf(__fmul_rn(__fdiv_rn(log_depth(P2), L), 50.0f));   // :492
    const float zc = MODE == MODE_I64 ? __fdiv_rn(P2, zabs) : P2;
    float* af = MODE == MODE_F32 ? acc_f32 + (size_t)uc * NCH * hw : nullptr;
    long long* ai = MODE == MODE_I64 ? acc_i64 + (size_t)uc * NCH * hw : nullptr;
    unsigned* wm = MODE != MODE_F32 ? wmax + (size_t)uc * hw : nullptr;

    if (inb && !behind) {
        const Corners c = corners_from_flow(flx, fly, x, y, h, w);
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int cx = (k & 2) ? c.ix1 : c.ix0;
            const int cy = (k & 1) ? c.iy1 : c.iy0;
            if (cx < 1 || cx > w || cy < 1 || cy > h) continue;          // pad ring: cropped away (:568-569)
            const float prox = k == 0 ? c.nw : (k == 1 ? c.sw : (k == 2 ? c.ne : c.se));
            const float wk = __fdiv_rn(__fmul_rn(prox, m), g);            // :494-513
            const size_t cell = (size_t)(cy - 1) * w + (cx - 1);
            if (MODE == MODE_F32) {
                atomicAdd(af + 0 * hw + cell, __fmul_rn(r, wk));
                atomicAdd(af + 1 * hw + cell, __fmul_rn(gch, wk));
                atomicAdd(af + 2 * hw + cell, __fmul_rn(b, wk));
                atomicAdd(af + 3 * hw + cell, __fmul_rn(P2, wk));
                atomicAdd(af + 4 * hw + cell, wk);
            } else if (MODE == MODE_WMAX) {
                if (wk > 0.0f) atomicMax(wm + cell, __float_as_uint(wk));
            } else {
                const unsigned top = wm[cell];
                if (top == 0u) continue;                                   // every weight into this cell is 0
                // q = round(w / wmax(cell) * 2^S [* value]) : a pure function of the pixel and the cell
                const float rs = __fmul_rn(__fdiv_rn(wk, __uint_as_float(top)), det_scale);
                atomicAdd((unsigned long long*)(ai + 0 * hw + cell), (unsigned long long)__float2ll_rn(__fmul_rn(rs, clampf(r, -64.f, 64.f))));
                atomicAdd((unsigned long long*)(ai + 1 * hw + cell), (unsigned long long)__float2ll_rn(__fmul_rn(rs, clampf(gch, -64.f, 64.f))));
                atomicAdd((unsigned long long*)(ai + 2 * hw + cell), (unsigned long long)__float2ll_rn(__fmul_rn(rs, clampf(b, -64.f, 64.f))));
                atomicAdd((unsigned long long*)(ai + 3 * hw + cell), (unsigned long long)__float2ll_rn(__fmul_rn(rs, zc)));
                atomicAdd((unsigned long long*)(ai + 4 * hw + cell), (unsigned long long)__float2ll_rn(rs));
            }
        }
    }

    // ---- behind-camera pixels: coords are exactly (1,1) => offset (2,2), floor == ceil, all four
    //      corners hit padded cell (2,2) = output pixel (1,1) with proximity 1.
    if (!__syncthreads_or(behind)) return;
    if (w < 2 || h < 2) return;                                           // that cell is in the pad ring
    const size_t cell = (size_t)w + 1;
    const float wk = behind ? __fdiv_rn(m, g) : 0.0f;                     // prox = 1
    if (MODE == MODE_F32) {
        float v[5] = {4.f * __fmul_rn(r, wk), 4.f * __fmul_rn(gch, wk), 4.f * __fmul_rn(b, wk),
                      4.f * __fmul_rn(P2, wk), 4.f * wk};
        block_sum<float, 5>(v, (float*)red);
        if (threadIdx.x == 0) {
            #pragma unroll
            for (int i = 0; i < 5; ++i) atomicAdd(af + i * hw + cell, v[i]);
        }
    } else if (MODE == MODE_WMAX) {
        const float mx = block_max_256(wk);
        if (threadIdx.x == 0 && mx > 0.0f) atomicMax(wm + cell, __float_as_uint(mx));
    } else {
        const unsigned top = wm[cell];
        long long v[5] = {0, 0, 0, 0, 0};
        if (behind && top != 0u) {
            const float rs = __fmul_rn(__fdiv_rn(wk, __uint_as_float(top)), det_scale);
            v[0] = 4 * __float2ll_rn(__fmul_rn(rs, clampf(r, -64.f, 64.f)));
            v[1] = 4 * __float2ll_rn(__fmul_rn(rs, clampf(gch, -64.f, 64.f)));
            v[2] = 4 * __float2ll_rn(__fmul_rn(rs, clampf(b, -64.f, 64.f)));
            v[3] = 4 * __float2ll_rn(__fmul_rn(rs, zc));
            v[4] = 4 * __float2ll_rn(rs);
        }
        block_sum<long long, 5>(v, red);
        if (threadIdx.x == 0) {
            #pragma unroll
            for (int i = 0; i < 5; ++i) atomicAdd((unsigned long long*)(ai + i * hw + cell), (unsigned long long)v[i]);
        }
    }
}

// ----------------------------------------------------------------------------------------------
// K3: normalise / mask / clip (+ optional fused clean_points)   (models/utils.py:566-583, 585-626)
//   DET = false: fp32 sums;  DET = true: int64 sums (exact), ratio taken in double.
// ----------------------------------------------------------------------------------------------
template <bool CLEAN, bool DET>
__global__ void __launch_bounds__(TILE_THREADS)
k_normalise(const float* __restrict__ acc_f32, const long long* __restrict__ acc_i64,
            const UnitStats* __restrict__ stats, int stat_index_global,
            int unit0, int h, int w, int tiles_x,
            void* __restrict__ out_frame, void* __restrict__ out_mask, float* __restrict__ out_depth,
            int* __restrict__ range_flag) {
    const int uc = blockIdx.y, u = unit0 + uc;
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    if (x >= w || y >= h) return;
    const size_t hw = (size_t)h * w, p = (size_t)y * w + x;
    const float* af = DET ? nullptr : acc_f32 + (size_t)uc * NCH * hw;
    const long long* ai = DET ? acc_i64 + (size_t)uc * NCH * hw : nullptr;
    bool valid;
    float col[3], dep;
    if (!DET) {
        const float W = af[4 * hw + p];
        valid = W > 0.0f;
        #pragma unroll
        for (int ch = 0; ch < 3; ++ch) col[ch] = valid ? __fdiv_rn(af[ch * hw + p], W) : -1.0f;
        dep = valid ? __fdiv_rn(af[3 * hw + p], W) : 0.0f;
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
                const size_t q = 4 * hw + (size_t)yy * w + xx;
                hole |= DET ? !(ai[q] > 0) : !(af[q] > 0.0f);
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
