// warp_strip.cuh — K2 of the forward warp: fused unproject / transform / project + soft-z splat with
// warp-aggregated reductions (included by tcwarp.cu after warp_kernels.cuh).
//
// One warp owns a strip of 32 source columns x STRIP_H source rows and walks it row by row.  The
// four bilinear contributions of a pixel are merged IN REGISTERS with those of its neighbours
// before anything leaves the SM:
//   * horizontally, lane p takes the NE/SE contributions of lane p-1 when they address the same
//     target cells as its own NW/SW contributions (one shuffle round), and
//   * vertically, the merged bottom cell of row j is carried in registers and folded into the top
//     cell of row j+1 when the cell ids agree.
// For a locally smooth warp this turns 4 cells x 5 channels = 20 L2 reductions per source pixel
// into ~1.2 cells x (one F32x4 + one F32 reduction).  Every merge is guarded by an exact cell-id
// comparison, so arbitrary (discontinuous, folding, degenerate floor==ceil) warps stay correct:
// whatever is not merged is emitted on its own.  The merge order is fixed by the strip geometry.
//
// Accumulators per unit: acc4[cell] = (sum w*r, sum w*g, sum w*b, sum w*z), accw[cell] = sum w.
// Reference: models/utils.py:350-420 (projection), :422-564 (weights + scatter).
#pragma once

constexpr int STRIP_H = 16;      // source rows per strip
constexpr int STRIP_W = 31;      // source columns OWNED by a strip (lane 31 is a halo lane on the next column)

struct SplatOut {
    float4* acc4;            // MODE_F32
    float* accw;             // MODE_F32
    long long* acc_i64;      // MODE_I64: [5][hw]
    unsigned* wmax;          // MODE_WMAX (written) / MODE_I64 (read): max emitted weight per cell, float bits
    float det_scale;         // 2^S
    float inv_zabs;          // 1 / max|z'| of the unit (fixed-point depth channel)
    size_t hw;
};

// one merged contribution (5 sums) leaves the SM
template <int MODE>
__device__ __forceinline__ void emit(const SplatOut& o, int id, float sr, float sg, float sb, float sz, float sw) {
    if (id < 0) return;
    if (MODE == MODE_F32) {
        atomicAdd(o.acc4 + id, make_float4(sr, sg, sb, sz));      // REDG.E.ADD.F32x4
        atomicAdd(o.accw + id, sw);
    } else if (MODE == MODE_WMAX) {
        if (sw > 0.0f) atomicMax(o.wmax + id, __float_as_uint(sw));
    } else {
        const unsigned top = o.wmax[id];
        if (top == 0u) return;                                     // every weight into this cell is 0
        // q = round(sum / wmax(cell) * 2^S): a pure function of the strip and the cell -> order-free adds
        const float rs = __fdiv_rn(o.det_scale, __uint_as_float(top));
        const float lim = 64.0f * o.det_scale;
        long long* a = o.acc_i64 + id;
        atomicAdd((unsigned long long*)(a + 0 * o.hw), (unsigned long long)__float2ll_rn(clampf(__fmul_rn(sr, rs), -lim, lim)));
        atomicAdd((unsigned long long*)(a + 1 * o.hw), (unsigned long long)__float2ll_rn(clampf(__fmul_rn(sg, rs), -lim, lim)));
        atomicAdd((unsigned long long*)(a + 2 * o.hw), (unsigned long long)__float2ll_rn(clampf(__fmul_rn(sb, rs), -lim, lim)));
        atomicAdd((unsigned long long*)(a + 3 * o.hw), (unsigned long long)__float2ll_rn(clampf(__fmul_rn(__fmul_rn(sz, o.inv_zabs), rs), -lim, lim)));
        atomicAdd((unsigned long long*)(a + 4 * o.hw), (unsigned long long)__float2ll_rn(__fmul_rn(sw, rs)));
    }
}

// everything one strip needs besides its own coordinates (shared by the per-chunk kernel and the
// persistent pipeline kernel)
struct SplatArgs {
    const float* frame; const float* mask1; const float* depth; const int* src_index; const float* Mall;
    int h, w, strips_x, strips_per_unit;
    const UnitStats* stats; int stat_index_global;      // -1: per-unit soft-z normaliser
    float4* acc4; float* accw; long long* acc_i64; unsigned* wmax; float det_scale;
    float* out_flow;
};

// Bilinear corner cells and proximity weights of one pixel (models/utils.py:450-488), as flat cell
// ids into the cropped h x w target (-1: the cell lies in the 1-px pad ring that :568-569 crop away).
struct CornerIds {
    int nw, sw, ne, se;
    float pnw, psw, pne, pse;
};

// Interior case: 1 <= off <= w (h) in both axes, so no clamp can change anything and all four cells
// are valid.  Same arithmetic as corners_from_flow without the clamps.
__device__ __forceinline__ CornerIds corners_inside(float ox, float oy, int w) {
    const float fx0 = floorf(ox), fx1 = ceilf(ox), fy0 = floorf(oy), fy1 = ceilf(oy);
    const float ax = __fsub_rn(1.0f, __fsub_rn(ox, fx0)), bx = __fsub_rn(1.0f, __fsub_rn(fx1, ox));
    const float ay = __fsub_rn(1.0f, __fsub_rn(oy, fy0)), by = __fsub_rn(1.0f, __fsub_rn(fy1, oy));
    const int ix0 = (int)fx0, ix1 = (int)fx1, iy0 = (int)fy0, iy1 = (int)fy1;
    const int rowT = (iy0 - 1) * w - 1, rowB = (iy1 - 1) * w - 1;
    CornerIds c;
    c.nw = rowT + ix0; c.sw = rowB + ix0; c.ne = rowT + ix1; c.se = rowB + ix1;
    c.pnw = __fmul_rn(ay, ax); c.psw = __fmul_rn(by, ax); c.pne = __fmul_rn(ay, bx); c.pse = __fmul_rn(by, bx);
    return c;
}

// General case: clamps to the padded grid exactly like the reference, pad-ring cells become -1.
__device__ __forceinline__ CornerIds corners_general(float flx, float fly, int x, int y, int h, int w) {
    const Corners k = corners_from_flow(flx, fly, x, y, h, w);
    const bool x0ok = k.ix0 >= 1 && k.ix0 <= w, x1ok = k.ix1 >= 1 && k.ix1 <= w;
    const bool y0ok = k.iy0 >= 1 && k.iy0 <= h, y1ok = k.iy1 >= 1 && k.iy1 <= h;
    const int rowT = (k.iy0 - 1) * w - 1, rowB = (k.iy1 - 1) * w - 1;
    CornerIds c;
    c.nw = x0ok && y0ok ? rowT + k.ix0 : -1;
    c.sw = x0ok && y1ok ? rowB + k.ix0 : -1;
    c.ne = x1ok && y0ok ? rowT + k.ix1 : -1;
    c.se = x1ok && y1ok ? rowB + k.ix1 : -1;
    c.pnw = k.nw; c.psw = k.sw; c.pne = k.ne; c.pse = k.se;
    return c;
}

// one warp processes strip `strip` of unit `u`; its accumulators live in slot `uc` of the chunk buffers.
// A strip owns STRIP_W = 31 columns; lane 31 is a HALO lane on the first column of the next strip: it
// projects that pixel only to learn its cell ids, contributes nothing of its own (weights 0, no NE/SE),
// and thereby gives lane 30 a regular right-hand neighbour to merge into.
template <int MODE>
__device__ __forceinline__ void splat_strip(const SplatArgs& A, int u, int uc, int strip) {
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int h = A.h, w = A.w;
    const int sy = strip / A.strips_x, sx = strip - sy * A.strips_x;
    const int x = sx * STRIP_W + lane;
    const int y0 = sy * STRIP_H, y1 = y0 + STRIP_H < h ? y0 + STRIP_H : h;
    const bool colok = x < w;
    const bool own = colok && lane < STRIP_W;                    // this lane's pixel belongs to this strip
    const size_t hw = (size_t)h * w;
    const float* __restrict__ Mu = A.Mall + (size_t)u * M_FLOATS;
    const int s = A.src_index ? A.src_index[u] : u;
    const UnitStats* st = A.stats + (A.stat_index_global >= 0 ? A.stat_index_global : u);
    const float L = __uint_as_float(st->lbits);
    SplatOut o;
    o.acc4 = MODE == MODE_F32 ? A.acc4 + (size_t)uc * hw : nullptr;
    o.accw = MODE == MODE_F32 ? A.accw + (size_t)uc * hw : nullptr;
    o.acc_i64 = MODE == MODE_I64 ? A.acc_i64 + (size_t)uc * NCH * hw : nullptr;
    o.wmax = MODE != MODE_F32 ? A.wmax + (size_t)uc * hw : nullptr;
    o.det_scale = A.det_scale;
    o.inv_zabs = 1.0f;
    if (MODE == MODE_I64) {
        float zabs = __uint_as_float(st->zabs);
        if (!(zabs > 0.0f)) zabs = 1.0f;
        o.inv_zabs = __fdiv_rn(1.0f, zabs);
    }
    o.hw = hw;

    // projection constants (models/utils.py:388-399).  The reference's dot products start from
    // "0 + m0*v0"; that first add only turns a -0 product into +0, which can change nothing but the
    // sign of an exactly-zero flow component, so it is dropped.  K1^-1 column 0 times x is hoisted.
    const float fx = (float)x;
    const float a0 = __fmul_rn(__ldg(Mu + 0), fx), a1 = __fmul_rn(__ldg(Mu + 3), fx), a2 = __fmul_rn(__ldg(Mu + 6), fx);
    const float m01 = __ldg(Mu + 1), m11 = __ldg(Mu + 4), m21 = __ldg(Mu + 7);
    const float m02 = __ldg(Mu + 2), m12 = __ldg(Mu + 5), m22 = __ldg(Mu + 8);
    float T[12], K[9];
    #pragma unroll
    for (int i = 0; i < 12; ++i) T[i] = __ldg(Mu + 9 + i);
    #pragma unroll
    for (int i = 0; i < 9; ++i) K[i] = __ldg(Mu + 21 + i);
    const float wf = (float)w, hf = (float)h, wf1 = (float)(w + 1), hf1 = (float)(h + 1);

    const int col = colok ? x : 0;
    const float* __restrict__ dp = A.depth + (size_t)s * hw + col;
    const float* __restrict__ fp = A.frame + (size_t)s * 3 * hw + col;
    const float* __restrict__ mp = A.mask1 ? A.mask1 + (size_t)s * hw + col : nullptr;
    float* __restrict__ flp = (MODE != MODE_WMAX && A.out_flow && own) ? A.out_flow + (size_t)u * 2 * hw + col : nullptr;

    int idC = -1;                                               // carried bottom cell of the previous row
    float cr = 0.f, cg = 0.f, cb = 0.f, cz = 0.f, cw = 0.f;
    float hr = 0.f, hg = 0.f, hb = 0.f, hz = 0.f, hwt = 0.f;    // behind-camera sums of this lane
    bool any_behind = false;

    // software pipeline: the loads of row y+1 are issued before row y is processed
    float d_n, r_n = 0.f, g_n = 0.f, b_n = 0.f, m_n = 1.f;
    {
        const int q = y0 * w;
        d_n = __ldg(dp + q);
        if (MODE != MODE_WMAX) { r_n = __ldg(fp + q); g_n = __ldg(fp + hw + q); b_n = __ldg(fp + 2 * hw + q); }
        if (mp) m_n = __ldg(mp + q);
    }
    for (int y = y0; y < y1; ++y) {
        const float d = d_n, r = r_n, gch = g_n, b = b_n, m = m_n;
        const int q = y * w;
        if (y + 1 < y1) {
            const int qn = q + w;
            d_n = __ldg(dp + qn);
            if (MODE != MODE_WMAX) { r_n = __ldg(fp + qn); g_n = __ldg(fp + hw + qn); b_n = __ldg(fp + 2 * hw + qn); }
            if (mp) m_n = __ldg(mp + qn);
        }
        // ---- unproject, transform, project (never fused: -fmad=false and explicit _rn intrinsics)
        const float fy = (float)y;
        const float u0 = __fadd_rn(__fadd_rn(a0, __fmul_rn(m01, fy)), m02);
        const float u1 = __fadd_rn(__fadd_rn(a1, __fmul_rn(m11, fy)), m12);
        const float u2 = __fadd_rn(__fadd_rn(a2, __fmul_rn(m21, fy)), m22);
        const float X0 = __fmul_rn(d, u0), X1 = __fmul_rn(d, u1), X2 = __fmul_rn(d, u2);
        const float Y0 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0], X0), __fmul_rn(T[1], X1)), __fmul_rn(T[2], X2)), T[3]);
        const float Y1 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[4], X0), __fmul_rn(T[5], X1)), __fmul_rn(T[6], X2)), T[7]);
        const float Y2 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[8], X0), __fmul_rn(T[9], X1)), __fmul_rn(T[10], X2)), T[11]);
        float P0 = __fadd_rn(__fadd_rn(__fmul_rn(K[0], Y0), __fmul_rn(K[1], Y1)), __fmul_rn(K[2], Y2));
        float P1 = __fadd_rn(__fadd_rn(__fmul_rn(K[3], Y0), __fmul_rn(K[4], Y1)), __fmul_rn(K[5], Y2));
        float P2 = __fadd_rn(__fadd_rn(__fmul_rn(K[6], Y0), __fmul_rn(K[7], Y1)), __fmul_rn(K[8], Y2));
        const bool bh = Y2 <= 0.01f;                                         // behind the camera (:404-417)
        if (bh) { P0 = 1000.0f; P1 = 1000.0f; P2 = 1000.0f; }
        const float flx = __fsub_rn(__fdiv_rn(P0, P2), fx);                 // flow12 (:273-278)
        const float fly = __fsub_rn(__fdiv_rn(P1, P2), fy);
        if (flp) {
            __stcs(flp + q, flx);
            __stcs(flp + hw + q, fly);
        }
        // ---- soft-z weight: w_k = prox_k * m / exp(50 * log(1+sat(z)) / L)   (:490-513)
        const float tl = __fmul_rn(__fdiv_rn(log_depth(P2), L), 50.0f);
        const float mg = __fmul_rn(m, expf(-tl));

        // ---- corner cells and proximity weights (:450-488)
        const float ox = __fadd_rn(__fadd_rn(flx, fx), 1.0f), oy = __fadd_rn(__fadd_rn(fly, fy), 1.0f);
        const bool live = colok && !bh;
        const bool dead = !live || ox <= 0.0f || ox >= wf1 || oy <= 0.0f || oy >= hf1;   // every corner in the pad ring
        const bool inside = ox >= 1.0f && ox <= wf && oy >= 1.0f && oy <= hf;
        CornerIds c;
        if (__all_sync(FULL, dead || inside)) c = corners_inside(ox, oy, w);
        else c = corners_general(flx, fly, x, y, h, w);
        const bool halo = lane == STRIP_W;
        const int idNW = dead ? -1 : c.nw, idSW = dead ? -1 : c.sw;
        const int idNE = dead || halo ? -1 : c.ne, idSE = dead || halo ? -1 : c.se;
        const float mgo = halo ? 0.0f : mg;                                  // the halo lane adds nothing of its own
        const float wNW = __fmul_rn(c.pnw, mgo), wSW = __fmul_rn(c.psw, mgo), wNE = __fmul_rn(c.pne, mgo), wSE = __fmul_rn(c.pse, mgo);
        if (own && bh) {     // all four corners hit padded cell (2,2) with proximity 1
            any_behind = true;
            hr += __fmul_rn(r, mg); hg += __fmul_rn(gch, mg); hb += __fmul_rn(b, mg); hz += __fmul_rn(P2, mg); hwt += mg;
        }

        // ---- horizontal merge: take the left lane's NE / SE when they address my NW / SW cells
        int idNE_l = __shfl_up_sync(FULL, idNE, 1), idSE_l = __shfl_up_sync(FULL, idSE, 1);
        const float wNE_l = __shfl_up_sync(FULL, wNE, 1), wSE_l = __shfl_up_sync(FULL, wSE, 1);
        const float z_l = __shfl_up_sync(FULL, P2, 1);
        float r_l = 0.f, g_l = 0.f, b_l = 0.f;
        if (MODE != MODE_WMAX) { r_l = __shfl_up_sync(FULL, r, 1); g_l = __shfl_up_sync(FULL, gch, 1); b_l = __shfl_up_sync(FULL, b, 1); }
        if (lane == 0) { idNE_l = -1; idSE_l = -1; }
        const int idNW_r = __shfl_down_sync(FULL, idNW, 1), idSW_r = __shfl_down_sync(FULL, idSW, 1);
        const bool takeNE = idNE_l >= 0 && idNE_l == idNW;
        const bool takeSE = idSE_l >= 0 && idSE_l == idSW;
        const float aNE = takeNE ? wNE_l : 0.0f, aSE = takeSE ? wSE_l : 0.0f;
        float tr = 0.f, tg = 0.f, tb = 0.f, br = 0.f, bg = 0.f, bb = 0.f;
        if (MODE != MODE_WMAX) {
            tr = __fmul_rn(r, wNW) + __fmul_rn(r_l, aNE); tg = __fmul_rn(gch, wNW) + __fmul_rn(g_l, aNE); tb = __fmul_rn(b, wNW) + __fmul_rn(b_l, aNE);
            br = __fmul_rn(r, wSW) + __fmul_rn(r_l, aSE); bg = __fmul_rn(gch, wSW) + __fmul_rn(g_l, aSE); bb = __fmul_rn(b, wSW) + __fmul_rn(b_l, aSE);
        }
        float tz = __fmul_rn(P2, wNW) + __fmul_rn(z_l, aNE), tw = wNW + aNE;
        const float bz = __fmul_rn(P2, wSW) + __fmul_rn(z_l, aSE), bw = wSW + aSE;

        // my NE / SE when the right lane did not take them (lane 31 is the halo: its ids are -1)
        const bool emNE = idNE >= 0 && idNE != idNW_r;
        const bool emSE = idSE >= 0 && idSE != idSW_r;
        if (__any_sync(FULL, emNE || emSE)) {
            if (emNE) emit<MODE>(o, idNE, __fmul_rn(r, wNE), __fmul_rn(gch, wNE), __fmul_rn(b, wNE), __fmul_rn(P2, wNE), wNE);
            if (emSE) emit<MODE>(o, idSE, __fmul_rn(r, wSE), __fmul_rn(gch, wSE), __fmul_rn(b, wSE), __fmul_rn(P2, wSE), wSE);
        }
        // ---- vertical merge: fold the carried bottom cell of the previous row into my top cell
        const bool fold = idC >= 0 && idC == idNW;
        if (__any_sync(FULL, idC >= 0 && !fold)) {
            if (!fold) emit<MODE>(o, idC, cr, cg, cb, cz, cw);
        }
        if (fold) { tr += cr; tg += cg; tb += cb; tz += cz; tw += cw; }
        emit<MODE>(o, idNW, tr, tg, tb, tz, tw);
        idC = idSW; cr = br; cg = bg; cb = bb; cz = bz; cw = bw;
    }
    emit<MODE>(o, idC, cr, cg, cb, cz, cw);

    // behind-camera pixels of the strip: one reduction per warp onto output pixel (1,1)
    if (__any_sync(FULL, any_behind)) {
        #pragma unroll
        for (int ofs = 16; ofs > 0; ofs >>= 1) {
            hr += __shfl_xor_sync(FULL, hr, ofs); hg += __shfl_xor_sync(FULL, hg, ofs); hb += __shfl_xor_sync(FULL, hb, ofs);
            hz += __shfl_xor_sync(FULL, hz, ofs); hwt += __shfl_xor_sync(FULL, hwt, ofs);
        }
        if (lane == 0 && w >= 2 && h >= 2) emit<MODE>(o, w + 1, 4.f * hr, 4.f * hg, 4.f * hb, 4.f * hz, 4.f * hwt);
    }
}

template <int MODE>
__global__ void __launch_bounds__(TILE_THREADS)
k_splat_strip(SplatArgs A, int unit0) {
    const int strip = blockIdx.x * (TILE_THREADS / 32) + (threadIdx.x >> 5);
    if (strip >= A.strips_per_unit) return;                     // whole warps leave together
    splat_strip<MODE>(A, unit0 + blockIdx.y, blockIdx.y, strip);
}
