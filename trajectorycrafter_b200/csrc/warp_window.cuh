// warp_window.cuh — the DEFAULT forward-warp path: deterministic splat through warp-private shared-memory windows.
// Included by tcwarp.cu after warp_pair.cuh (packed-fp32x2 helpers, log2x / exp2x, division helpers).
//
//   K1 k_win_prepass<PROJECT>  unproject / transform / project of every source pixel (models/utils.py:350-420):
//                              writes flow12 (an output of the path anyway) and z', the per-unit soft-z normaliser,
//                              and the EXACT target bounding box of every 32 x 8 source half-strip
//   K2 k_splat_window          bilinear soft-z splat (models/utils.py:422-564) of one 64 x 8 source strip per warp into
//                              a window of shared memory that only this warp touches: plain LDS / FADD / STS in
//                              program order, no atomics; the window is then stored as one dense slab
//   K3 k_gather_normalise      per 32 x 8 target tile: adds the slabs that overlap it in ascending half-strip
//                              order, normalises, writes colour / mask / depth (models/utils.py:566-583)
//
// Determinism: every target cell is the sum of (a) contributions one warp added in its own program order and (b) the
// slabs K3 adds in a fixed order, so the result does not depend on how warps are scheduled.  There is no float atomic on
// this path when every half-strip's footprint fits the window (WIN_CAP cells); a larger footprint uses a slab in global
// memory that still belongs to one warp only (vector reductions issued by that warp in program order); only a
// footprint that does not fit the unit's slab pool at all is reduced into a shared plane (flagged in the status word).
// Lanes that hit the same cell in ONE step are serialised through a tag array (see emit_window).
#pragma once

// -DTCW_BOUNDS_CHECK=1 (scripts/r2_checked.sh): every window / slab / pool access of the default path is range-checked, an
// access outside its allocation is SKIPPED and status bit 8 is raised.  compute-sanitizer is closed on the measurement pool, so
// this build is the memory-safety evidence; the product build compiles the checks away.
#ifndef TCW_BOUNDS_CHECK
#define TCW_BOUNDS_CHECK 0
#endif
#if TCW_BOUNDS_CHECK
#define TCW_INRANGE(st, cond) ((cond) ? true : (((st) ? (void)atomicOr((st), 8) : (void)0), false))
#else
#define TCW_INRANGE(st, cond) true
#endif
#ifndef TCW_WIN_CAP
#define TCW_WIN_CAP 768            // cells of shared-memory window per warp (both half-strips together)
#endif
constexpr int WIN_CAP = TCW_WIN_CAP;
constexpr int WIN_SCRATCH = 32;                               // scratch cells behind the window: one per lane, for pixels that add nothing
constexpr int WIN_P4 = (WIN_CAP + WIN_SCRATCH) + (WIN_CAP + WIN_SCRATCH) / 8 + 1;      // float4 slots: one pad slot per 8 cells (stride-2 lanes stay conflict-free)
constexpr int WIN_PW = WIN_P4;                                // float slots: the weight plane uses the float4 plane's padded index
constexpr int WIN_BYTES = ((WIN_P4 * 16 + WIN_PW * 4 + WIN_CAP) + 15) / 16 * 16;
#ifndef TCW_WIN_WARPS
#define TCW_WIN_WARPS 2
#endif
constexpr int WIN_WARPS = TCW_WIN_WARPS;                     // warps (strips) per CTA of K2
constexpr int WIN_GROUP = 8;                                 // half-strips per K1 CTA = per box group
constexpr int WIN_PRE_WARPS = WIN_GROUP / 2;                 // K1: one warp per 64 x 8 strip = two half-strips
constexpr int WIN_POOL_MULT = 6;                             // shared part of a unit's slab pool: WIN_POOL_MULT * h * w cells (rough depth maps need ~3: DESIGN.md section 5)
constexpr int WIN_MAXC = 64;                                 // candidate slabs one K3 warp lists at a time

struct WinBox { short x0, y0, bw, bh; int flags, pad; };     // cropped target coords; bw == 0: no cell.  flags bit 0: has pile pixels
struct WinGrp { int x0, y0, x1, y1; };                       // union of WIN_GROUP boxes, [x0, x1) x [y0, y1); x1 <= x0: empty

// What K1/K2/K3 share.  "prep slot" = where K1's per-unit data lives, "splat slot" = slab pool / spill plane of a unit in flight.
struct WinArgs {
    // sources
    const float* frame; const float* mask1; const float* depth; const int* src_index; const float* Mall;
    const float* gflow; const float* gflow_mask; int nv;     // generic splat (tcw_splat): flow given, nv value channels, depth = weight depth
    int fused;                                               // 1: forward warp (values = RGB of frame[s] + z'), 0: generic splat
    int n_src, h, w, strips_x, ns, nh, ng;
    const UnitStats* stats; UnitStats* stats_rw; int stat_index_global; int n_units_total;
    // K1 products
    float* flow_out;                                         // (n,2,h,w) caller's flow12, or nullptr -> flowbuf
    float* flowbuf; float* zbuf; WinBox* box; WinGrp* grp;   // per prep slot
    // K2 products
    int* counter; int* spill_flag; float* pile;                  // per splat slot
    WinBox* sbox;                                            // per splat slot: the slabs K2 actually wrote (box + offset in `pad`; bw == 0: none)
    float4* pool4; float* poolw; long long pool_cells; int n_slots;
    float4* spill4; float* spillw;
    int* status;                                             // device word: bit 1 = a footprint was reduced into the shared plane, bit 2 = src_index out of range
    // tcw_forward_warp_scatter: unit u's results go to unit slot dst_index[u] of the (n_dst, ...) output tensors (nullptr: slot u)
    const int* dst_index; int n_dst;
};

// output slot of unit u (an index outside [0, n_dst) writes slot 0 and is flagged: never write out of bounds)
__device__ __forceinline__ int win_dst(const WinArgs& A, int u) {
    if (!A.dst_index) return u;
    int d = A.dst_index[u];
    if ((unsigned)d >= (unsigned)A.n_dst) { if (A.status) atomicOr(A.status, 4); d = 0; }
    return d;
}

__device__ __forceinline__ int win_src(const WinArgs& A, int u) {
    int s = A.src_index ? A.src_index[u] : u;
    if ((unsigned)s >= (unsigned)A.n_src) { if (A.status) atomicOr(A.status, 4); s = 0; }     // never read out of bounds
    return s;
}

// ------------------------------------------------------------------------------------------------
// K1: per source pixel flow12 = P.xy / P.z - grid and z' = P.z, bit-identical to the reference (scalar un-fused adds,
// packed products, the correctly rounded packed division of warp_pair.cuh); per warp the target box of its half-strip.
// One warp = 32 columns x 8 rows, lane = column, two rows (y, y + 1) per iteration as the packed pair.
// PROJECT = false (generic splat): flow and depth are inputs; only the boxes and the normaliser are computed.
// ------------------------------------------------------------------------------------------------
template <bool PROJECT>
__global__ void __launch_bounds__(WIN_PRE_WARPS * 32)
k_win_prepass(WinArgs A, int unit0, int prep0, int slot0) {
    constexpr unsigned FULL = 0xffffffffu;
    __shared__ int gb[4];
    const int lane = threadIdx.x & 31, wv = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);      // the shuffle tells the compiler the warp index is uniform
    const int uc = blockIdx.y, u = unit0 + uc, ps = prep0 + uc;
    const int h = A.h, w = A.w;
    const size_t hw = (size_t)h * w;
    if (threadIdx.x == 0) { gb[0] = INT_MAX; gb[1] = INT_MAX; gb[2] = INT_MIN; gb[3] = INT_MIN; }
    if (blockIdx.x == 0 && threadIdx.x == 0 && slot0 >= 0) { A.counter[slot0 + uc] = 0; A.spill_flag[slot0 + uc] = 0; }
    __syncthreads();
    const int strip = blockIdx.x * WIN_PRE_WARPS + wv;
    // box of this lane's pixels in float (floor / ceil are monotone: the box of the floats gives the box of the cells)
    const float INF = __int_as_float(0x7f800000);
    float bx0 = INF, by0 = INF, bx1 = -INF, by1 = -INF;
    bool pile = false;
    float zs = 0.0f;
    if (strip < A.ns) {
        const int sy = strip / A.strips_x, sx = strip - sy * A.strips_x;
        const int xA = sx * 64 + 2 * lane, xB = xA + 1;
        const bool okA = xA < w, okB = xB < w;
        const int cA = okA ? xA : 0, cB = okB ? xB : 0;
        const int y0 = sy * 8, y1 = y0 + 8 < h ? y0 + 8 : h;
        const int s = PROJECT ? win_src(A, u) : u;
        const float fxA = (float)xA, fxB = (float)xB;
        const u64 fx2 = pk(fxA, fxB);
        const float wf1 = (float)(w + 1), hf1 = (float)(h + 1);
        const float* __restrict__ dp = A.depth + (size_t)s * hw;
        float* __restrict__ flp = nullptr;
        float* __restrict__ zp = nullptr;
        const float* __restrict__ gfl = nullptr;
        float m[9], T[12], K[9];
        bool PIN = false;
        u64 u0p = 0;
        float a0A = 0.f, a1A = 0.f, a2A = 0.f, a0B = 0.f, a1B = 0.f, a2B = 0.f;
        if (PROJECT) {
            flp = A.flow_out ? A.flow_out + (size_t)win_dst(A, u) * 2 * hw : A.flowbuf + (size_t)ps * 2 * hw;
            zp = A.zbuf + (size_t)ps * hw;
            const float* __restrict__ Mu = A.Mall + (size_t)u * M_FLOATS;
            #pragma unroll
            for (int i = 0; i < 9; ++i) m[i] = __ldg(Mu + i);
            #pragma unroll
            for (int i = 0; i < 12; ++i) T[i] = __ldg(Mu + 9 + i);
            #pragma unroll
            for (int i = 0; i < 9; ++i) K[i] = __ldg(Mu + 21 + i);
            // pinhole zero pattern of K1^-1 and K2: the dropped terms are products with an exact 0 (see warp_pair.cuh)
            PIN = m[1] == 0.0f && m[3] == 0.0f && m[6] == 0.0f && m[7] == 0.0f && m[8] == 1.0f &&
                  K[1] == 0.0f && K[3] == 0.0f && K[6] == 0.0f && K[7] == 0.0f && K[8] == 1.0f;
            a0A = __fmul_rn(m[0], fxA); a1A = __fmul_rn(m[3], fxA); a2A = __fmul_rn(m[6], fxA);
            a0B = __fmul_rn(m[0], fxB); a1B = __fmul_rn(m[3], fxB); a2B = __fmul_rn(m[6], fxB);
            u0p = pk(__fadd_rn(a0A, m[2]), __fadd_rn(a0B, m[2]));
        } else {
            gfl = A.gflow + (size_t)u * 2 * hw;
        }
        // float2 accesses when both pixels exist in every lane that has any, rows stay 8-byte aligned
        const bool vec2 = __all_sync(FULL, okB || !okA) && !(w & 1) &&
            !((((size_t)dp) | ((size_t)flp) | ((size_t)zp) | ((size_t)gfl)) & 7);
        // the strip's depth values up front: eight independent loads in flight instead of one per row
        // (32-bit byte offsets from the unit's plane bases: 2 * h * w * 4 < 2^32, h * w < 2^28 is checked at submission)
        const unsigned w4 = (unsigned)w * 4u, hw4 = (unsigned)hw * 4u;
        const unsigned oA0 = ((unsigned)y0 * (unsigned)w + (unsigned)cA) * 4u, oB0 = ((unsigned)y0 * (unsigned)w + (unsigned)cB) * 4u;
        const int nr = y1 - y0;
        float2 dpre[8];
        if (PROJECT) {
            #pragma unroll
            for (int r = 0; r < 8; ++r) {
                dpre[r] = make_float2(0.f, 0.f);
                if (r < nr) {
                    if (vec2) dpre[r] = __ldg((const float2*)((const char*)dp + (oA0 + r * w4)));
                    else dpre[r] = make_float2(__ldg((const float*)((const char*)dp + (oA0 + r * w4))), __ldg((const float*)((const char*)dp + (oB0 + r * w4))));
                }
            }
        }
        #pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int y = y0 + r;
            if (r >= nr) break;
            const float fy = (float)y;
            const unsigned oA = oA0 + r * w4, oB = oB0 + r * w4;
            float flxA, flxB, flyA, flyB, zA, zB;
            if (PROJECT) {
                const float dA = dpre[r].x, dB = dpre[r].y;
                float P0A, P1A, P2A, P0B, P1B, P2B, Y2A, Y2B;
                if (PIN) {
                    const float u1 = __fadd_rn(__fmul_rn(m[4], fy), m[5]);
                    const u64 d2 = pk(dA, dB);
                    const u64 X0 = mul2(d2, u0p), X1 = mul2(d2, pk1(u1));
                    float a, b, c, e, f, g;
                    upk(mul2(X0, pk1(T[0])), a, b); upk(mul2(X1, pk1(T[1])), c, e); upk(mul2(d2, pk1(T[2])), f, g);
                    const float Y0A = __fadd_rn(__fadd_rn(__fadd_rn(a, c), f), T[3]), Y0B = __fadd_rn(__fadd_rn(__fadd_rn(b, e), g), T[3]);
                    upk(mul2(X0, pk1(T[4])), a, b); upk(mul2(X1, pk1(T[5])), c, e); upk(mul2(d2, pk1(T[6])), f, g);
                    const float Y1A = __fadd_rn(__fadd_rn(__fadd_rn(a, c), f), T[7]), Y1B = __fadd_rn(__fadd_rn(__fadd_rn(b, e), g), T[7]);
                    upk(mul2(X0, pk1(T[8])), a, b); upk(mul2(X1, pk1(T[9])), c, e); upk(mul2(d2, pk1(T[10])), f, g);
                    Y2A = __fadd_rn(__fadd_rn(__fadd_rn(a, c), f), T[11]); Y2B = __fadd_rn(__fadd_rn(__fadd_rn(b, e), g), T[11]);
                    const u64 Y22 = pk(Y2A, Y2B);
                    upk(mul2(pk(Y0A, Y0B), pk1(K[0])), a, b); upk(mul2(Y22, pk1(K[2])), c, e);
                    P0A = __fadd_rn(a, c); P0B = __fadd_rn(b, e);
                    upk(mul2(pk(Y1A, Y1B), pk1(K[4])), a, b); upk(mul2(Y22, pk1(K[5])), c, e);
                    P1A = __fadd_rn(a, c); P1B = __fadd_rn(b, e);
                    P2A = Y2A; P2B = Y2B;
                } else {
                    #pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        const float d = k ? dB : dA;
                        const float u0 = __fadd_rn(__fadd_rn(k ? a0B : a0A, __fmul_rn(m[1], fy)), m[2]);
                        const float u1 = __fadd_rn(__fadd_rn(k ? a1B : a1A, __fmul_rn(m[4], fy)), m[5]);
                        const float u2 = __fadd_rn(__fadd_rn(k ? a2B : a2A, __fmul_rn(m[7], fy)), m[8]);
                        const float X0 = __fmul_rn(d, u0), X1 = __fmul_rn(d, u1), X2 = __fmul_rn(d, u2);
                        const float Y0 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0], X0), __fmul_rn(T[1], X1)), __fmul_rn(T[2], X2)), T[3]);
                        const float Y1 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[4], X0), __fmul_rn(T[5], X1)), __fmul_rn(T[6], X2)), T[7]);
                        const float Y2 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[8], X0), __fmul_rn(T[9], X1)), __fmul_rn(T[10], X2)), T[11]);
                        const float P0 = __fadd_rn(__fadd_rn(__fmul_rn(K[0], Y0), __fmul_rn(K[1], Y1)), __fmul_rn(K[2], Y2));
                        const float P1 = __fadd_rn(__fadd_rn(__fmul_rn(K[3], Y0), __fmul_rn(K[4], Y1)), __fmul_rn(K[5], Y2));
                        const float P2 = __fadd_rn(__fadd_rn(__fmul_rn(K[6], Y0), __fmul_rn(K[7], Y1)), __fmul_rn(K[8], Y2));
                        if (k) { P0B = P0; P1B = P1; P2B = P2; Y2B = Y2; } else { P0A = P0; P1A = P1; P2A = P2; Y2A = Y2; }
                    }
                }
                if (Y2A <= 0.01f) { P0A = 1000.0f; P1A = 1000.0f; P2A = 1000.0f; }      // behind the camera (:404-417)
                if (Y2B <= 0.01f) { P0B = 1000.0f; P1B = 1000.0f; P2B = 1000.0f; }
                u64 qx, qy;
                {   // two IEEE divisions per pixel off one refined reciprocal (preconditions: warp_pair.cuh)
                    const float zlo = fminf(P2A, P2B), zhi = fmaxf(P2A, P2B);
                    const float nlo = fminf(fminf(fabsf(P0A), fabsf(P0B)), fminf(fabsf(P1A), fabsf(P1B)));
                    const float nhi = fmaxf(fmaxf(fabsf(P0A), fabsf(P0B)), fmaxf(fabsf(P1A), fabsf(P1B)));
                    const bool fast = zlo >= 9.5367431640625e-07f && zhi <= 1048576.0f && nlo >= 7.888609052210118e-31f && nhi <= 1.2676506002282294e30f;
                    if (__all_sync(FULL, fast)) {
                        const u64 b2 = pk(P2A, P2B), nb2 = pk(-P2A, -P2B);
                        const u64 r = refine_rcp2(b2, nb2);
                        qx = div2_with_rcp(pk(P0A, P0B), nb2, r);
                        qy = div2_with_rcp(pk(P1A, P1B), nb2, r);
                    } else {
                        qx = pk(__fdiv_rn(P0A, P2A), __fdiv_rn(P0B, P2B));
                        qy = pk(__fdiv_rn(P1A, P2A), __fdiv_rn(P1B, P2B));
                    }
                }
                const u64 flx2 = sub2(qx, fx2), fly2 = sub2(qy, pk1(fy));
                upk(flx2, flxA, flxB);
                upk(fly2, flyA, flyB);
                zA = P2A; zB = P2B;
                char* __restrict__ fb = (char*)flp;
                char* __restrict__ zb = (char*)zp;
                if (vec2) {
                    if (okA) {
                        __stcg((float2*)(fb + oA), make_float2(flxA, flxB)); __stcg((float2*)(fb + (oA + hw4)), make_float2(flyA, flyB));
                        __stcg((float2*)(zb + oA), make_float2(zA, zB));
                    }
                } else {
                    if (okA) { __stcg((float*)(fb + oA), flxA); __stcg((float*)(fb + (oA + hw4)), flyA); __stcg((float*)(zb + oA), zA); }
                    if (okB) { __stcg((float*)(fb + oB), flxB); __stcg((float*)(fb + (oB + hw4)), flyB); __stcg((float*)(zb + oB), zB); }
                }
            } else {
                const char* __restrict__ gb = (const char*)gfl;
                const char* __restrict__ db = (const char*)dp;
                if (vec2) {
                    const float2 f0 = __ldg((const float2*)(gb + oA)), f1 = __ldg((const float2*)(gb + (oA + hw4))), z2 = __ldg((const float2*)(db + oA));
                    flxA = f0.x; flxB = f0.y; flyA = f1.x; flyB = f1.y; zA = z2.x; zB = z2.y;
                } else {
                    flxA = __ldg((const float*)(gb + oA)); flyA = __ldg((const float*)(gb + (oA + hw4))); zA = __ldg((const float*)(db + oA));
                    flxB = __ldg((const float*)(gb + oB)); flyB = __ldg((const float*)(gb + (oB + hw4))); zB = __ldg((const float*)(db + oB));
                }
            }
            // same arithmetic as K2 (models/utils.py:450-452): off = (flow + grid) + 1
            float oxA, oxB, oyA, oyB;
            upk(add2(add2(pk(flxA, flxB), fx2), pk1(1.0f)), oxA, oxB);
            upk(add2(add2(pk(flyA, flyB), pk1(fy)), pk1(1.0f)), oyA, oyB);
            if (okA) zs = fmaxf(zs, zA);
            if (okB) zs = fmaxf(zs, zB);
            const bool inA = okA && oxA > 0.0f && oxA < wf1 && oyA > 0.0f && oyA < hf1;      // else: every corner in the pad ring
            const bool inB = okB && oxB > 0.0f && oxB < wf1 && oyB > 0.0f && oyB < hf1;
            const bool pA = inA && oxA == 2.0f && oyA == 2.0f && w >= 2 && h >= 2;
            const bool pB = inB && oxB == 2.0f && oyB == 2.0f && w >= 2 && h >= 2;
            pile |= pA | pB;
            if (inA && !pA) { bx0 = fminf(bx0, oxA); bx1 = fmaxf(bx1, oxA); by0 = fminf(by0, oyA); by1 = fmaxf(by1, oyA); }
            if (inB && !pB) { bx0 = fminf(bx0, oxB); bx1 = fmaxf(bx1, oxB); by0 = fminf(by0, oyB); by1 = fmaxf(by1, oyB); }
        }
    }
    // cells a pixel at (ox, oy) touches: floor .. ceil, clipped to the real cells 1 .. w / 1 .. h; in cropped coordinates
    // [max(floor(ox), 1) - 1, min(ceil(ox), w) - 1]
    int xmin = INT_MAX, ymin = INT_MAX, xmax = INT_MIN, ymax = INT_MIN;
    if (bx1 >= bx0) {
        xmin = max((int)floorf(bx0), 1) - 1; xmax = min((int)ceilf(bx1), w) - 1;
        ymin = max((int)floorf(by0), 1) - 1; ymax = min((int)ceilf(by1), h) - 1;
        if (xmin > xmax || ymin > ymax) { xmin = INT_MAX; ymin = INT_MAX; xmax = INT_MIN; ymax = INT_MIN; }
    }
    // lanes 0-15 hold the left half-strip (columns 0-31 of the strip), lanes 16-31 the right one
    const unsigned hm = lane < 16 ? 0x0000ffffu : 0xffff0000u;
    xmin = __reduce_min_sync(hm, xmin); ymin = __reduce_min_sync(hm, ymin);
    xmax = __reduce_max_sync(hm, xmax); ymax = __reduce_max_sync(hm, ymax);
    const bool anyp = __ballot_sync(FULL, pile) & hm;
    zs = clampf(zs, 0.0f, 1000.0f);                                                 // clamp is monotone: once, after the max
    const unsigned zbits = __reduce_max_sync(FULL, __float_as_uint(zs));          // zs >= 0: uint order == float order
    if ((lane & 15) == 0 && strip < A.ns) {
        WinBox b;
        const bool some = xmax >= xmin;
        b.x0 = some ? (short)xmin : 0; b.y0 = some ? (short)ymin : 0;
        b.bw = some ? (short)(xmax - xmin + 1) : 0; b.bh = some ? (short)(ymax - ymin + 1) : 0;
        b.flags = anyp ? 1 : 0; b.pad = 0;
        A.box[(size_t)ps * A.nh + 2 * strip + (lane >> 4)] = b;
        if (some) { atomicMin(&gb[0], xmin); atomicMin(&gb[1], ymin); atomicMax(&gb[2], xmax + 1); atomicMax(&gb[3], ymax + 1); }
    }
    if (lane == 0) {
        // log is monotone: one log per warp.  The polynomial K2 evaluates per pixel, so ld / L == 1 at the arg-max pixel
        const float ld = log1x(__fadd_rn(1.0f, __uint_as_float(zbits)));
        atomicMax(&A.stats_rw[u].lbits, __float_as_uint(ld));
        atomicMax(&A.stats_rw[A.n_units_total].lbits, __float_as_uint(ld));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        WinGrp g; g.x0 = gb[0]; g.y0 = gb[1]; g.x1 = gb[2]; g.y1 = gb[3];
        if (g.x1 == INT_MIN) { g.x0 = 0; g.y0 = 0; g.x1 = 0; g.y1 = 0; }
        A.grp[(size_t)ps * A.ng + blockIdx.x] = g;
    }
}

// ------------------------------------------------------------------------------------------------
// K2: one warp = one 64 x 8 source strip, two horizontally adjacent pixels per lane (A = x0 + 2 lane, B = A + 1).
// Lanes 0-15 cover the left half-strip, lanes 16-31 the right one; each half has its own window (its box from K1).
// ------------------------------------------------------------------------------------------------
enum { WM_SMEM = 0, WM_GLOBAL = 1 };

struct WinOut {
    // WM_SMEM
    float4* s4; float* sw; unsigned char* tag;
    int* status; int glimit;      // checked builds: the call's status word; cells of the global slab / plane o.g4 points at
    int scratch;          // first scratch cell (= cells of the window): lane l may use cell scratch + l
    // WM_GLOBAL: this warp's slab (or the unit's shared plane)
    float4* g4; float* gw;
};

// One merged contribution.  Must be reached by all 32 lanes (id < 0: nothing to add).
// WM_SMEM: read-modify-write of the warp's own window.  Two lanes with the same cell in this step are serialised:
// every pending lane writes its number to the cell's tag, the lane that reads its own number back adds and leaves,
// the others go round again.  Which of several same-address stores of one instruction survives is a fixed
// function of that instruction (the hardware keeps the highest lane), so the order of the additions is fixed.
template <int MODE>
__device__ __forceinline__ void emit_window(const WinOut& o, int lane, int id, u64 rg, u64 bz, float sw) {
    constexpr unsigned FULL = 0xffffffffu;
    if (MODE == WM_GLOBAL) {
        if (id >= 0 && TCW_INRANGE(o.status, id < o.glimit)) {
            float r, g, b, z;
            upk(rg, r, g); upk(bz, b, z);
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n\t"
                         "red.global.add.f32 [%5], %6;"
                         :: "l"(o.g4 + id), "f"(r), "f"(g), "f"(b), "f"(z), "l"(o.gw + id), "f"(sw) : "memory");
        }
        return;
    }
    bool pend = id >= 0 && TCW_INRANGE(o.status, id < o.scratch);
    if (!__any_sync(FULL, pend)) return;
    const int p4 = id + (id >> 3), pw = p4;
    do {
        if (pend) o.tag[id] = (unsigned char)lane;
        __syncwarp();
        const bool win = pend && o.tag[id] == (unsigned char)lane;
        if (win) {
            float4 a = o.s4[p4];
            float r, g, b, z;
            upk(add2(pk(a.x, a.y), rg), r, g); upk(add2(pk(a.z, a.w), bz), b, z);
            o.s4[p4] = make_float4(r, g, b, z);
            o.sw[pw] += sw;
        }
        pend = pend && !win;
        __syncwarp();
    } while (__any_sync(FULL, pend));
}
template <int MODE>
__device__ __forceinline__ void emit_window_single(const WinOut& o, int lane, bool take, int id, u64 rg, u64 bz, float w) {
    const u64 w2 = pk1(w);
    emit_window<MODE>(o, lane, take ? id : -1, mul2(rg, w2), mul2(bz, w2), w);
}


// One corner of one pixel column set, WM_SMEM fast path: the caller has established that no two lanes address the
// same cell in this step (win_steps_distinct), so the update is a plain read-modify-write of the warp's own window.
__device__ __forceinline__ void win_step(const WinOut& o, int lane, bool on, int id, u64 rg, u64 bz, float w) {
    // a pixel that adds nothing (its weight and values are 0) goes through the same instructions into its lane's scratch cell:
    // no divergent branch in the step
    const int c = (on && TCW_INRANGE(o.status, (unsigned)id < (unsigned)o.scratch)) ? id : o.scratch + lane;
    const int p4 = c + (c >> 3);
    const float4 a = o.s4[p4];
    const float s = o.sw[p4];
    const u64 w2 = pk1(w);
    float r, g, b, z;
    upk(fma2(rg, w2, pk(a.x, a.y)), r, g); upk(fma2(bz, w2, pk(a.z, a.w)), b, z);
    o.s4[p4] = make_float4(r, g, b, z);
    o.sw[p4] = s + w;
    __syncwarp();
}
// The top and the bottom cell of one column of one pixel in ONE step: both reads first, then both writes, so the two
// shared-memory round trips overlap.  The caller guarantees that the two cells differ (a pixel that sits exactly on a cell row
// has folded its bottom weight into the top one and sends the bottom to the scratch cell) and that no other lane addresses
// either of them in this step.
__device__ __forceinline__ void win_step2(const WinOut& o, int lane, bool onT, int idT, bool onB, int idB, u64 rg, u64 bz, float wT, float wB) {
    const int cT = (onT && TCW_INRANGE(o.status, (unsigned)idT < (unsigned)o.scratch)) ? idT : o.scratch + lane;
    const int cB = (onB && TCW_INRANGE(o.status, (unsigned)idB < (unsigned)o.scratch)) ? idB : o.scratch + lane;
    const int pT = cT + (cT >> 3), pB = cB + (cB >> 3);
    const float4 a = o.s4[pT], c = o.s4[pB];
    const float sa = o.sw[pT], sc = o.sw[pB];
    const u64 t2 = pk1(wT), b2 = pk1(wB);
    float r, g, b, z;
    // (both switched off: cT == cB == the lane's scratch cell, which takes two meaningless writes -- no branch)
    upk(fma2(rg, t2, pk(a.x, a.y)), r, g); upk(fma2(bz, t2, pk(a.z, a.w)), b, z);
    o.s4[pT] = make_float4(r, g, b, z);
    o.sw[pT] = sa + wT;
    upk(fma2(rg, b2, pk(c.x, c.y)), r, g); upk(fma2(bz, b2, pk(c.z, c.w)), b, z);
    o.s4[pB] = make_float4(r, g, b, z);
    o.sw[pB] = sc + wB;
    __syncwarp();
}
// live lanes form one run of consecutive lanes?
__device__ __forceinline__ bool win_run(unsigned m) {
    if (m == 0u) return true;
    const unsigned t = m >> (__ffs(m) - 1);
    return (t & (t + 1u)) == 0u;
}

// per-unit source pointers of a strip
struct WinSrc {
    const float* fl;      // flow12 of the unit: x plane, y plane at + hw
    const float* wd;      // weight depth plane (z' in the fused path, depth1 in the generic splat)
    const float* v[4];    // value planes (nullptr: 0)
    const float* m1; const float* fm;
};

// The eight rows of a strip in one mode.  `act` = this lane's half takes part in this pass; pitch / idbase = this lane's
// window (padded floor coordinates -> cell index of the pass allocation: id = iy * pitch + ix + idbase).
template <int MODE, bool FUSED>
__device__ __forceinline__ void splat_window_rows(const WinArgs& A, const WinSrc& S, const WinOut& o, int lane, int x0s, int y0, int y1,
                                                  bool act, int pitch, int idbase, float L, u64& hrg, u64& hbz, float& hwt, bool& any_pile) {
    constexpr unsigned FULL = 0xffffffffu;
    const int h = A.h, w = A.w;
    const int xA = x0s + 2 * lane, xB = xA + 1;
    const bool okA = xA < w && act, okB = xB < w && act;
    const unsigned hw32 = (unsigned)((size_t)h * w);
    const float fxA = (float)xA, fxB = (float)xB;
    const u64 fx2 = pk(fxA, fxB);
    float wf1 = (float)(w + 1), hf1 = (float)(h + 1);
    const bool pile_ok = w >= 2 && h >= 2;
    const int colA = xA < w ? xA : 0, colB = xB < w ? xB : 0;
    const bool vec2 = xB < w && !(w & 1) &&
        !((((size_t)S.fl) | ((size_t)S.wd) | ((size_t)S.v[0]) | ((size_t)S.v[1]) | ((size_t)S.v[2]) | ((size_t)S.v[3]) | ((size_t)S.m1) | ((size_t)S.fm)) & 7);
    const bool vec2all = __all_sync(FULL, vec2 || !(xA < w));
    // exact ld / L through a refined reciprocal (warp_pair.cuh)
    const bool Lok = L >= 9.5367431640625e-07f;
    float rL = rcp_approx(L);
    rL = fmaf(rL, fmaf(-L, rL, 1.0f), rL);
    const u64 rL2 = pk1(rL), nL2 = pk1(-L);

    float flxA_n, flxB_n, flyA_n, flyB_n, zA_n, zB_n, vA_n[4], vB_n[4], mA_n = 1.f, mB_n = 1.f;
    const unsigned hw4 = hw32 * 4u;
    auto load_row = [&](int y) {
        const unsigned q = (unsigned)y * (unsigned)w;
        if (FUSED && vec2all) {
            // the forward warp's own sources: flow x / y planes, z', three colour planes h*w apart, an optional mask; no fourth
            // value, no flow mask.  Byte offsets fit 32 bits (3 * h * w * 4 < 2^32: h * w < 2^28 is checked at submission)
            const unsigned qb = (q + colA) * 4u;
            const float2 f0 = __ldcg((const float2*)((const char*)S.fl + qb)), f1 = __ldcg((const float2*)((const char*)S.fl + (qb + hw4)));
            const float2 z2 = __ldcg((const float2*)((const char*)S.wd + qb));
            const float2 t0 = __ldg((const float2*)((const char*)S.v[0] + qb)), t1 = __ldg((const float2*)((const char*)S.v[0] + (qb + hw4)));
            const float2 t2 = __ldg((const float2*)((const char*)S.v[0] + (qb + 2u * hw4)));
            flxA_n = f0.x; flxB_n = f0.y; flyA_n = f1.x; flyB_n = f1.y; zA_n = z2.x; zB_n = z2.y;
            vA_n[0] = t0.x; vB_n[0] = t0.y; vA_n[1] = t1.x; vB_n[1] = t1.y; vA_n[2] = t2.x; vB_n[2] = t2.y; vA_n[3] = 0.f; vB_n[3] = 0.f;
            mA_n = 1.f; mB_n = 1.f;
            if (S.m1) { const float2 t = __ldg((const float2*)((const char*)S.m1 + qb)); mA_n = t.x; mB_n = t.y; }
        } else if (vec2all) {
            const unsigned qa = q + colA;
            const float2 f0 = __ldcg((const float2*)(S.fl + qa)), f1 = __ldcg((const float2*)(S.fl + hw32 + qa));
            const float2 z2 = __ldcg((const float2*)(S.wd + qa));
            flxA_n = f0.x; flxB_n = f0.y; flyA_n = f1.x; flyB_n = f1.y; zA_n = z2.x; zB_n = z2.y;
            #pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (S.v[c]) { const float2 t = __ldg((const float2*)(S.v[c] + qa)); vA_n[c] = t.x; vB_n[c] = t.y; }
                else { vA_n[c] = 0.f; vB_n[c] = 0.f; }
            }
            mA_n = 1.f; mB_n = 1.f;
            if (S.m1) { const float2 t = __ldg((const float2*)(S.m1 + qa)); mA_n = t.x; mB_n = t.y; }
            if (S.fm) { const float2 t = __ldg((const float2*)(S.fm + qa)); mA_n = __fmul_rn(mA_n, t.x); mB_n = __fmul_rn(mB_n, t.y); }
        } else {
            const unsigned qa = q + colA, qb = q + colB;
            flxA_n = __ldcg(S.fl + qa); flxB_n = __ldcg(S.fl + qb); flyA_n = __ldcg(S.fl + hw32 + qa); flyB_n = __ldcg(S.fl + hw32 + qb);
            zA_n = __ldcg(S.wd + qa); zB_n = __ldcg(S.wd + qb);
            #pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (S.v[c]) { vA_n[c] = __ldg(S.v[c] + qa); vB_n[c] = __ldg(S.v[c] + qb); }
                else { vA_n[c] = 0.f; vB_n[c] = 0.f; }
            }
            mA_n = 1.f; mB_n = 1.f;
            if (S.m1) { mA_n = __ldg(S.m1 + qa); mB_n = __ldg(S.m1 + qb); }
            if (S.fm) { mA_n = __fmul_rn(mA_n, __ldg(S.fm + qa)); mB_n = __fmul_rn(mB_n, __ldg(S.fm + qb)); }
        }
    };
    load_row(y0);

    for (int y = y0; y < y1; ++y) {
        const float flxA = flxA_n, flxB = flxB_n, flyA = flyA_n, flyB = flyB_n, zA = zA_n, zB = zB_n;
        float vA[4], vB[4];
        #pragma unroll
        for (int c = 0; c < 4; ++c) { vA[c] = vA_n[c]; vB[c] = vB_n[c]; }
        const float mA = mA_n, mB = mB_n;
        if (y + 1 < y1) load_row(y + 1);
        const float fy = (float)y;

        // ---- corner cells (models/utils.py:450-475), from the stored flow exactly as the reference forms them
        const u64 ox2r = add2(add2(pk(flxA, flxB), fx2), pk1(1.0f)), oy2r = add2(add2(pk(flyA, flyB), pk1(fy)), pk1(1.0f));
        float oxA, oxB, oyA, oyB;
        upk(ox2r, oxA, oxB);
        upk(oy2r, oyA, oyB);
        const bool pileA = okA && pile_ok && oxA == 2.0f && oyA == 2.0f, pileB = okB && pile_ok && oxB == 2.0f && oyB == 2.0f;
        bool liveA = okA && !pileA && oxA > 0.0f && oxA < wf1 && oyA > 0.0f && oyA < hf1;
        bool liveB = okB && !pileB && oxB > 0.0f && oxB < wf1 && oyB > 0.0f && oyB < hf1;
        // a pixel that adds nothing gets finite stand-in numbers: its products are exact zeros, never NaN
        if (!liveA) { oxA = 1.5f; oyA = 1.5f; }
        if (!liveB) { oxB = 1.5f; oyB = 1.5f; }
        const u64 ox2 = pk(oxA, oxB), oy2 = pk(oyA, oyB);
        const float fx0A = floorf(oxA), fx0B = floorf(oxB), fy0A = floorf(oyA), fy0B = floorf(oyB);
        const int ix0A = (int)fx0A, ix0B = (int)fx0B, iy0A = (int)fy0A, iy0B = (int)fy0B;
        const bool inA = (unsigned)(ix0A - 1) < (unsigned)(w - 1) && (unsigned)(iy0A - 1) < (unsigned)(h - 1);
        const bool inB = (unsigned)(ix0B - 1) < (unsigned)(w - 1) && (unsigned)(iy0B - 1) < (unsigned)(h - 1);
        const bool general = __any_sync(FULL, (liveA && !inA) || (liveB && !inB));

        // ---- soft-z weight mg = mask1 * flow_mask * exp(-50 * log(1 + sat(z)) / L)   (:490-513)
        u64 mg2;
        {
            const float sA = (liveA || pileA) ? clampf(zA, 0.0f, 1000.0f) : 0.0f, sB = (liveB || pileB) ? clampf(zB, 0.0f, 1000.0f) : 0.0f;
            const u64 ld = log2x(add2(pk(sA, sB), pk1(1.0f)));
            u64 qd;
            if (Lok) qd = div2_with_rcp(ld, nL2, rL2);
            else qd = pk(__fdiv_rn(plo(ld), L), __fdiv_rn(phi(ld), L));
            mg2 = exp2x(mul2(qd, pk1(-50.0f)));
            if (S.m1 || (!FUSED && S.fm)) mg2 = mul2(mg2, pk(mA, mB));
        }
        float mgA, mgB;
        upk(mg2, mgA, mgB);
        // values of the two pixels: (v0, v1), (v2, v3); the fused path splats z' itself as the fourth value
        const float z3A = FUSED ? zA : vA[3], z3B = FUSED ? zB : vB[3];
        u64 rgA = pk(vA[0], vA[1]), bzA = pk(vA[2], z3A), rgB = pk(vB[0], vB[1]), bzB = pk(vB[2], z3B);
        if (pileA || pileB) {      // exactly on padded cell (2,2): four unit-proximity hits on output pixel (1,1) (behind-camera points)
            any_pile = true;
            const float pa = pileA ? mgA : 0.0f, pb = pileB ? mgB : 0.0f;
            if (pileA) { hrg = fma2(rgA, pk1(pa), hrg); hbz = fma2(bzA, pk1(pa), hbz); }
            if (pileB) { hrg = fma2(rgB, pk1(pb), hrg); hbz = fma2(bzB, pk1(pb), hbz); }
            hwt += pa + pb;
        }
        if (!liveA) { mgA = 0.0f; rgA = 0; bzA = 0; }
        if (!liveB) { mgB = 0.0f; rgB = 0; bzB = 0; }
        mg2 = pk(mgA, mgB);

        int idNWA, idSWA, idNEA, idSEA, idNWB, idSWB, idNEB, idSEB;
        u64 wNW2, wSW2, wNE2, wSE2;
        // WM_SMEM: can the row's (corner x pixel) steps run as plain read-modify-writes?  Pixels A of all lanes are one step,
        // pixels B another.  Two lanes of a step share a cell only if they share its column: with k = 2 ix0 + (ix1 - ix0) of
        // the UNCLAMPED floor, left and right columns both grow strictly from lane l-1 to lane l  <=>  k_l - k_{l-1} >= 2
        // (clamping only moves columns into the pad ring, whose cells are never added); live lanes must be one run so that
        // neighbours are enough.  dist 1: yes.  dist 2: yes when the even and the odd lanes go separately (columns grow from
        // lane l-2 to lane l: images compressed to between a quarter and a half).  dist 0: the serialising routine.
        const bool dxA = oxA != fx0A, dxB = oxB != fx0B, dyA = oyA != fy0A, dyB = oyB != fy0B;   // ceil == floor + 1 ?
        int dist = 0;
        if (MODE == WM_SMEM) {
            const unsigned mA = __ballot_sync(FULL, liveA), mB = __ballot_sync(FULL, liveB);
            if (win_run(mA) && win_run(mB)) {
                const int kA = 2 * ix0A + (dxA ? 1 : 0), kB = 2 * ix0B + (dxB ? 1 : 0);
                const int qA = __shfl_up_sync(FULL, kA, 1), qB = __shfl_up_sync(FULL, kB, 1);
                const unsigned below = lane ? (1u << (lane - 1)) : 0u;
                const bool bad = (liveA && (mA & below) && kA - qA < 2) || (liveB && (mB & below) && kB - qB < 2);
                if (!__any_sync(FULL, bad)) dist = 1;
                else {
                    const int rA = __shfl_up_sync(FULL, kA, 2), rB = __shfl_up_sync(FULL, kB, 2);
                    const unsigned below2 = lane >= 2 ? (1u << (lane - 2)) : 0u;
                    const bool bad2 = (liveA && (mA & below2) && kA - rA < 2) || (liveB && (mB & below2) && kB - rB < 2);
                    if (!__any_sync(FULL, bad2)) dist = 2;
                }
            }
        }
        if (!general) {
            const u64 fx02 = pk(fx0A, fx0B), fy02 = pk(fy0A, fy0B);
            const u64 fx12 = pk(dxA ? fx0A + 1.0f : fx0A, dxB ? fx0B + 1.0f : fx0B);
            const u64 fy12 = pk(dyA ? fy0A + 1.0f : fy0A, dyB ? fy0B + 1.0f : fy0B);
            const u64 one2 = pk1(1.0f);
            const u64 ax = sub2(one2, sub2(ox2, fx02)), bx = sub2(one2, sub2(fx12, ox2));
            const u64 ay = sub2(one2, sub2(oy2, fy02)), by = sub2(one2, sub2(fy12, oy2));
            wNW2 = mul2(mul2(ay, ax), mg2); wSW2 = mul2(mul2(by, ax), mg2);
            wNE2 = mul2(mul2(ay, bx), mg2); wSE2 = mul2(mul2(by, bx), mg2);
            // a pixel that adds nothing gets NEGATIVE ids: never emitted, can only "merge" with other negative ids
            idNWA = liveA ? iy0A * pitch + ix0A + idbase : (int)0x80000000;
            idNWB = liveB ? iy0B * pitch + ix0B + idbase : (int)0x80000000;
            const int sA = liveA && dyA ? pitch : 0, sB = liveB && dyB ? pitch : 0;
            const int eA = liveA && dxA ? 1 : 0, eB = liveB && dxB ? 1 : 0;
            idSWA = idNWA + sA; idSWB = idNWB + sB;
            idNEA = idNWA + eA; idNEB = idNWB + eB;
            idSEA = idSWA + eA; idSEB = idSWB + eB;
        } else {
            // some live pixel touches the pad ring: the clamped formulas of the reference, corner by corner
            const Corners ca = corners_from_flow(flxA, flyA, xA, y, h, w), cb = corners_from_flow(flxB, flyB, xB, y, h, w);
            auto cid = [&](bool live, int ix, int iy) {
                return (live && ix >= 1 && ix <= w && iy >= 1 && iy <= h) ? iy * pitch + ix + idbase : (int)0x80000000;
            };
            idNWA = cid(liveA, ca.ix0, ca.iy0); idSWA = cid(liveA, ca.ix0, ca.iy1); idNEA = cid(liveA, ca.ix1, ca.iy0); idSEA = cid(liveA, ca.ix1, ca.iy1);
            idNWB = cid(liveB, cb.ix0, cb.iy0); idSWB = cid(liveB, cb.ix0, cb.iy1); idNEB = cid(liveB, cb.ix1, cb.iy0); idSEB = cid(liveB, cb.ix1, cb.iy1);
            const float pa0 = liveA ? ca.nw : 0.f, pa1 = liveA ? ca.sw : 0.f, pa2 = liveA ? ca.ne : 0.f, pa3 = liveA ? ca.se : 0.f;
            const float pb0 = liveB ? cb.nw : 0.f, pb1 = liveB ? cb.sw : 0.f, pb2 = liveB ? cb.ne : 0.f, pb3 = liveB ? cb.se : 0.f;
            wNW2 = mul2(pk(pa0, pb0), mg2); wSW2 = mul2(pk(pa1, pb1), mg2);
            wNE2 = mul2(pk(pa2, pb2), mg2); wSE2 = mul2(pk(pa3, pb3), mg2);
        }
        float wNWA, wNWB, wSWA, wSWB, wNEA, wNEB, wSEA, wSEB;
        upk(wNW2, wNWA, wNWB); upk(wSW2, wSWA, wSWB); upk(wNE2, wNEA, wNEB); upk(wSE2, wSEA, wSEB);
        if (MODE == WM_SMEM && dist) {
            // plain steps in a fixed order: NW, SW, NE, SE of the A pixels and of the B pixels (a negative id: pad ring or a
            // pixel that adds nothing)
            // top and bottom cell of a column go in one step; they are the same cell when the pixel sits exactly on a cell
            // row (or the bottom one is a pad cell): then the top takes both weights and the bottom step is switched off
            const bool tA = idSWA >= 0 && idSWA != idNWA, tB = idSWB >= 0 && idSWB != idNWB;
            const bool uA = idSEA >= 0 && idSEA != idNEA, uB = idSEB >= 0 && idSEB != idNEB;
            if (idSWA >= 0 && !tA) { wNWA += wSWA; }
            if (idSWB >= 0 && !tB) { wNWB += wSWB; }
            if (idSEA >= 0 && !uA) { wNEA += wSEA; }
            if (idSEB >= 0 && !uB) { wNEB += wSEB; }
            for (int par = 0; par < dist; ++par) {
                const bool mine = dist == 1 || (lane & 1) == par;
                win_step2(o, lane, mine && idNWA >= 0, idNWA, mine && tA, idSWA, rgA, bzA, wNWA, wSWA);
                win_step2(o, lane, mine && idNWB >= 0, idNWB, mine && tB, idSWB, rgB, bzB, wNWB, wSWB);
                win_step2(o, lane, mine && idNEA >= 0, idNEA, mine && uA, idSEA, rgA, bzA, wNEA, wSEA);
                win_step2(o, lane, mine && idNEB >= 0, idNEB, mine && uB, idSEB, rgB, bzB, wNEB, wSEB);
            }
            continue;
        }

        // ---- register-level merge (the rules of warp_pair.cuh without the row carries: the window adds across rows).
        // Destinations: Lt / Lb = pixel A's left cells, Mt / Mb = pixel B's left cells.  Sources: the left lane's B
        // right-hand cells (rotate by one lane) and A's right-hand cells; what matches no destination is added on its own.
        const int srcl = (lane + 31) & 31;
        const int idNEl = __shfl_sync(FULL, idNEB, srcl), idSEl = __shfl_sync(FULL, idSEB, srcl);
        const float wNEl = __shfl_sync(FULL, wNEB, srcl), wSEl = __shfl_sync(FULL, wSEB, srcl);
        const u64 rgl = __shfl_sync(FULL, rgB, srcl), bzl = __shfl_sync(FULL, bzB, srcl);
        u64 Ltrg = mul2(rgA, pk1(wNWA)), Ltbz = mul2(bzA, pk1(wNWA)), Lbrg = mul2(rgA, pk1(wSWA)), Lbbz = mul2(bzA, pk1(wSWA));
        u64 Mtrg = mul2(rgB, pk1(wNWB)), Mtbz = mul2(bzB, pk1(wNWB)), Mbrg = mul2(rgB, pk1(wSWB)), Mbbz = mul2(bzB, pk1(wSWB));
        u64 ANrg = mul2(rgA, pk1(wNEA)), ANbz = mul2(bzA, pk1(wNEA)), ASrg = mul2(rgA, pk1(wSEA)), ASbz = mul2(bzA, pk1(wSEA));
        float Ltw = wNWA, Lbw = wSWA, Mtw = wNWB, Mbw = wSWB, ANw = wNEA, ASw = wSEA;
        {   // the left lane's NE cell: my A's top-left, else my A's bottom-left, else my A's top-right
            const bool a = idNEl == idNWA, b = !a && idNEl == idSWA, c = !a && !b && idNEl == idNEA;
            const float wa = a ? wNEl : 0.0f, wb = b ? wNEl : 0.0f, wc = c ? wNEl : 0.0f;
            Ltrg = fma2(rgl, pk1(wa), Ltrg); Ltbz = fma2(bzl, pk1(wa), Ltbz); Ltw += wa;
            Lbrg = fma2(rgl, pk1(wb), Lbrg); Lbbz = fma2(bzl, pk1(wb), Lbbz); Lbw += wb;
            ANrg = fma2(rgl, pk1(wc), ANrg); ANbz = fma2(bzl, pk1(wc), ANbz); ANw += wc;
            emit_window_single<MODE>(o, lane, !(a || b || c), idNEl, rgl, bzl, wNEl);
        }
        {   // the left lane's SE cell: my A's bottom-left, else my A's top-left, else my A's bottom-right
            const bool a = idSEl == idSWA, b = !a && idSEl == idNWA, c = !a && !b && idSEl == idSEA;
            const float wa = a ? wSEl : 0.0f, wb = b ? wSEl : 0.0f, wc = c ? wSEl : 0.0f;
            Lbrg = fma2(rgl, pk1(wa), Lbrg); Lbbz = fma2(bzl, pk1(wa), Lbbz); Lbw += wa;
            Ltrg = fma2(rgl, pk1(wb), Ltrg); Ltbz = fma2(bzl, pk1(wb), Ltbz); Ltw += wb;
            ASrg = fma2(rgl, pk1(wc), ASrg); ASbz = fma2(bzl, pk1(wc), ASbz); ASw += wc;
            emit_window_single<MODE>(o, lane, !(a || b || c), idSEl, rgl, bzl, wSEl);
        }
        {   // A's NE sum: B's top-left, else B's bottom-left
            const bool a = idNEA == idNWB, b = !a && idNEA == idSWB;
            const float fa = a ? 1.0f : 0.0f, fb = b ? 1.0f : 0.0f;
            Mtrg = fma2(ANrg, pk1(fa), Mtrg); Mtbz = fma2(ANbz, pk1(fa), Mtbz); Mtw = fmaf(ANw, fa, Mtw);
            Mbrg = fma2(ANrg, pk1(fb), Mbrg); Mbbz = fma2(ANbz, pk1(fb), Mbbz); Mbw = fmaf(ANw, fb, Mbw);
            emit_window<MODE>(o, lane, (a || b) ? -1 : idNEA, ANrg, ANbz, ANw);
        }
        {   // A's SE sum: B's bottom-left, else B's top-left
            const bool a = idSEA == idSWB, b = !a && idSEA == idNWB;
            const float fa = a ? 1.0f : 0.0f, fb = b ? 1.0f : 0.0f;
            Mbrg = fma2(ASrg, pk1(fa), Mbrg); Mbbz = fma2(ASbz, pk1(fa), Mbbz); Mbw = fmaf(ASw, fa, Mbw);
            Mtrg = fma2(ASrg, pk1(fb), Mtrg); Mtbz = fma2(ASbz, pk1(fb), Mtbz); Mtw = fmaf(ASw, fb, Mtw);
            emit_window<MODE>(o, lane, (a || b) ? -1 : idSEA, ASrg, ASbz, ASw);
        }
        // pixels A and B in the same column: B's left cells are A's
        const bool sT = idNWB == idNWA, sB = idSWB == idSWA;
        {
            const float fa = sT ? 1.0f : 0.0f, fb = sB ? 1.0f : 0.0f;
            Ltrg = fma2(Mtrg, pk1(fa), Ltrg); Ltbz = fma2(Mtbz, pk1(fa), Ltbz); Ltw = fmaf(Mtw, fa, Ltw);
            Lbrg = fma2(Mbrg, pk1(fb), Lbrg); Lbbz = fma2(Mbbz, pk1(fb), Lbbz); Lbw = fmaf(Mbw, fb, Lbw);
        }
        // (top and bottom cells of one pixel may coincide for an exact-integer y: they are added in separate steps)
        emit_window<MODE>(o, lane, idNWA, Ltrg, Ltbz, Ltw);
        emit_window<MODE>(o, lane, idSWA, Lbrg, Lbbz, Lbw);
        emit_window<MODE>(o, lane, sT ? -1 : idNWB, Mtrg, Mtbz, Mtw);
        emit_window<MODE>(o, lane, sB ? -1 : idSWB, Mbrg, Mbbz, Mbw);
    }
}

#ifndef TCW_WIN_MINB
#define TCW_WIN_MINB (12 / TCW_WIN_WARPS)
#endif
template <bool FUSED>      // FUSED: the forward warp (sources = K1's flow / z' + the frame); else the generic splat of tcw_splat
__global__ void __launch_bounds__(WIN_WARPS * 32, TCW_WIN_MINB)
k_splat_window(WinArgs A, int unit0, int prep0, int slot0) {
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char win_smem[];
    const int lane = threadIdx.x & 31, wv = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);      // the shuffle tells the compiler the warp index is uniform
    const int strip = blockIdx.x * WIN_WARPS + wv;
    if (strip >= A.ns) return;                                           // whole warps leave together
    const int uc = blockIdx.y, u = unit0 + uc, ps = prep0 + uc, sl = slot0 + uc;
    const int h = A.h, w = A.w;
    const size_t hw = (size_t)h * w;
    const int sy = strip / A.strips_x, sx = strip - sy * A.strips_x;
    const int hs0 = sy * 2 * A.strips_x + 2 * sx;
    const int y0 = sy * 8, y1 = y0 + 8 < h ? y0 + 8 : h;

    unsigned char* base = win_smem + (size_t)wv * WIN_BYTES;
    WinOut o;
    o.s4 = (float4*)base; o.sw = (float*)(base + WIN_P4 * 16); o.tag = base + WIN_P4 * 16 + WIN_PW * 4;
    o.g4 = nullptr; o.gw = nullptr; o.scratch = 0; o.status = A.status; o.glimit = 0;

    // sources of this unit
    WinSrc S;
    if (FUSED) {
        const int s = win_src(A, u);
        S.fl = A.flow_out ? A.flow_out + (size_t)win_dst(A, u) * 2 * hw : A.flowbuf + (size_t)ps * 2 * hw;
        S.wd = A.zbuf + (size_t)ps * hw;
        S.v[0] = A.frame + (size_t)s * 3 * hw; S.v[1] = S.v[0] + hw; S.v[2] = S.v[1] + hw; S.v[3] = nullptr;
        S.m1 = A.mask1 ? A.mask1 + (size_t)s * hw : nullptr; S.fm = nullptr;
    } else {
        S.fl = A.gflow + (size_t)u * 2 * hw;
        S.wd = A.depth + (size_t)u * hw;
        #pragma unroll
        for (int c = 0; c < 4; ++c) S.v[c] = c < A.nv ? A.frame + ((size_t)u * A.nv + c) * hw : nullptr;
        S.m1 = A.mask1 ? A.mask1 + (size_t)u * hw : nullptr;
        S.fm = A.gflow_mask ? A.gflow_mask + (size_t)u * hw : nullptr;
    }
    const UnitStats* st = A.stats + (A.stat_index_global >= 0 ? A.stat_index_global : u);
    // (values every lane loads from one address go through a shuffle from lane 0: the compiler then knows they are uniform and
    // drops the convergence checks around the collectives of the row loop)
    const float L = __uint_as_float(__shfl_sync(FULL, st->lbits, 0));

    WinBox b0, b1;
    {
        const int4* __restrict__ bp = (const int4*)(A.box + (size_t)ps * A.nh + hs0);
        int4 q0 = bp[0], q1 = bp[1];
        q0.x = __shfl_sync(FULL, q0.x, 0); q0.y = __shfl_sync(FULL, q0.y, 0); q0.z = __shfl_sync(FULL, q0.z, 0);
        q1.x = __shfl_sync(FULL, q1.x, 0); q1.y = __shfl_sync(FULL, q1.y, 0); q1.z = __shfl_sync(FULL, q1.z, 0);
        b0.x0 = (short)(q0.x & 0xffff); b0.y0 = (short)(q0.x >> 16); b0.bw = (short)(q0.y & 0xffff); b0.bh = (short)(q0.y >> 16); b0.flags = q0.z; b0.pad = 0;
        b1.x0 = (short)(q1.x & 0xffff); b1.y0 = (short)(q1.x >> 16); b1.bw = (short)(q1.y & 0xffff); b1.bh = (short)(q1.y >> 16); b1.flags = q1.z; b1.pad = 0;
    }
    const int a0 = (int)b0.bw * b0.bh, a1 = (int)b1.bw * b1.bh;
    float* pilep = A.pile + ((size_t)sl * A.ns + strip) * 8;
    u64 hrg = 0, hbz = 0;
    float hwt = 0.f;
    bool any_pile = false;
    int off0 = -1, off1 = -1;

    // one slab for the whole strip when the union of the two half-strip boxes costs at most a quarter more cells than the two
    // boxes: the gather then meets one slab instead of two
    WinBox bu = b0;
    bool uni = false;
    if (a0 > 0 && a1 > 0) {
        const int ux0 = min((int)b0.x0, (int)b1.x0), uy0 = min((int)b0.y0, (int)b1.y0);
        const int ux1 = max((int)b0.x0 + b0.bw, (int)b1.x0 + b1.bw), uy1 = max((int)b0.y0 + b0.bh, (int)b1.y0 + b1.bh);
        const int ua = (ux1 - ux0) * (uy1 - uy0);
        if (ua <= WIN_CAP && 4 * ua <= 5 * (a0 + a1)) {
            uni = true;
            bu.x0 = (short)ux0; bu.y0 = (short)uy0; bu.bw = (short)(ux1 - ux0); bu.bh = (short)(uy1 - uy0);
        }
    }
    if (a0 + a1 > 0 || ((b0.flags | b1.flags) & 1)) {
        const int half = lane >> 4;
        // passes: both halves in one window if they fit together, else one half per pass
        const bool together = uni || a0 + a1 <= WIN_CAP;
        const int npass = together ? 1 : 2;
        for (int p = 0; p < npass; ++p) {
            const bool act0 = together || p == 0, act1 = together || p == 1;
            const int n0 = uni ? (int)bu.bw * bu.bh : (act0 ? a0 : 0), n1 = uni ? 0 : (act1 ? a1 : 0), cells = n0 + n1;
            const bool pile_only = cells == 0;
            if (pile_only && !(((act0 ? b0.flags : 0) | (act1 ? b1.flags : 0)) & 1)) continue;
            // window of this lane's half for this pass
            const bool act = half ? act1 : act0;
            const WinBox& bm = uni ? bu : (half ? b1 : b0);
            const int pitch = bm.bw;
            int mode = cells <= WIN_CAP ? WM_SMEM : WM_GLOBAL;
            long long off = 0;
            bool shared_plane = false;
            if (cells > 0) {
                if (mode == WM_SMEM) {
                    // a window's slab has a place of its own: slot 2 * strip + pass of WIN_CAP cells each -- no allocation, no atomic
                    off = (long long)(2 * strip + p) * WIN_CAP;
                } else {
                    // larger footprints share the rest of the pool (one atomic per such footprint)
                    if (lane == 0) off = (long long)A.ns * 2 * WIN_CAP + (long long)atomicAdd(A.counter + sl, cells);
                    off = __shfl_sync(FULL, off, 0);
                }
                if (off + cells > A.pool_cells || off < 0) { shared_plane = true; mode = WM_GLOBAL; }      // pool exhausted (the counter may wrap for absurd inputs)
            }
            int idbase;
            if (shared_plane) {
                // the unit's shared accumulation plane: global cell ids, float reductions (order not fixed) -- flagged
                o.g4 = A.spill4 + (size_t)sl * hw; o.gw = A.spillw + (size_t)sl * hw; o.glimit = (int)hw;
                idbase = -(w + 1);
                if (lane == 0) { A.spill_flag[sl] = 1; if (A.status) atomicOr(A.status, 2); }
            } else {
                idbase = (!uni && half && act0 ? n0 : 0) - ((int)bm.y0 + 1) * pitch - ((int)bm.x0 + 1);
                o.g4 = A.pool4 + (size_t)sl * A.pool_cells + off; o.gw = A.poolw + (size_t)sl * A.pool_cells + off; o.glimit = cells;
                (void)TCW_INRANGE(A.status, off >= 0 && off + cells <= A.pool_cells && cells + WIN_SCRATCH <= WIN_CAP + WIN_SCRATCH + (mode == WM_GLOBAL ? (1 << 30) : 0));
            }
            const int pitch_l = shared_plane ? w : pitch;
            mode = __shfl_sync(FULL, mode, 0);      // uniform, and the compiler can see it (no convergence checks in the row loop)
            if (mode == WM_SMEM) {
                o.scratch = cells;
                {
                    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    const int n4 = cells + (cells >> 3) + 1, nw = n4;
                    for (int i = lane; i < n4; i += 32) o.s4[i] = z4;
                    for (int i = lane; i < nw; i += 32) o.sw[i] = 0.f;
                    __syncwarp();
                }
                splat_window_rows<WM_SMEM, FUSED>(A, S, o, lane, sx * 64, y0, y1, act, pitch_l, idbase, L, hrg, hbz, hwt, any_pile);
                __syncwarp();
                // the window becomes a dense slab (cells in window order == slab order)
                // (four cells per lane and trip: the shared-memory reads of a trip are issued before its global stores)
                for (int i0 = lane; i0 < cells; i0 += 128) {
                    float4 v4[4];
                    float vw[4];
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int i = i0 + 32 * k, j = i < cells ? i : lane;
                        v4[k] = o.s4[j + (j >> 3)];
                        vw[k] = o.sw[j + (j >> 3)];
                    }
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int i = i0 + 32 * k;
                        if (i < cells) { __stcg(o.g4 + i, v4[k]); __stcg(o.gw + i, vw[k]); }
                    }
                }
                __syncwarp();
            } else {
                if (!shared_plane) {
                    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int i = lane; i < cells; i += 32) { o.g4[i] = z4; o.gw[i] = 0.f; }
                    __threadfence();      // the zeros are in L2 before this warp's first reduction arrives
                    __syncwarp();
                }
                splat_window_rows<WM_GLOBAL, FUSED>(A, S, o, lane, sx * 64, y0, y1, act, pitch_l, idbase, L, hrg, hbz, hwt, any_pile);
            }
            if (!shared_plane) {
                if (uni) off0 = (int)off;
                else {
                    if (act0 && a0 > 0) off0 = (int)off;
                    if (act1 && a1 > 0) off1 = (int)off + n0;
                }
            }
        }
    }
    if (lane < 2) {      // what the gather will find: the slab(s) of this strip
        WinBox r = lane ? b1 : b0;
        if (uni) { r = bu; if (lane) r.bw = 0; }
        const int off = lane ? off1 : off0;
        if (off < 0) r.bw = 0;
        r.pad = off;
        A.sbox[(size_t)sl * A.nh + hs0 + lane] = r;
    }
    // pixels that sat exactly on padded cell (2,2): one sum per strip, added by K3 to output pixel (1,1)
    float hr = 0.f, hg = 0.f, hb = 0.f, hz = 0.f;
    const bool anyp = __any_sync(FULL, any_pile);
    if (anyp) {
        upk(hrg, hr, hg);
        upk(hbz, hb, hz);
        #pragma unroll
        for (int ofs = 16; ofs > 0; ofs >>= 1) {
            hr += __shfl_xor_sync(FULL, hr, ofs); hg += __shfl_xor_sync(FULL, hg, ofs); hb += __shfl_xor_sync(FULL, hb, ofs);
            hz += __shfl_xor_sync(FULL, hz, ofs); hwt += __shfl_xor_sync(FULL, hwt, ofs);
        }
    }
    if (lane == 0) {
        *(float4*)pilep = make_float4(4.f * hr, 4.f * hg, 4.f * hb, 4.f * hz);
        *(float4*)(pilep + 4) = make_float4(4.f * hwt, anyp ? 1.f : 0.f, 0.f, 0.f);
    }
}

// ------------------------------------------------------------------------------------------------
// K3: CTA = 32 x 32 target tile, warp = 32 columns x 8 rows (lane = column, 8 rows of sums in registers).
// Warp 0 lists the half-strips whose box overlaps the tile in ascending order (box groups first); every warp then
// adds the listed slabs in that order, the shared plane if the unit used it, the pile sums (output pixel (1,1) only),
// normalises and stores.  `is_image`: holes = -1, range check, clamp (models/utils.py:571-582).
// ------------------------------------------------------------------------------------------------
struct WinNormArgs {
    float* out_frame; int out_c;          // (n, out_c, h, w): channels 0 .. out_c-1 of the sums / sum w
    float* out_mask; float* out_depth;    // out_depth: fourth sum / sum w (fused path), may be nullptr
    int is_image; int* range_flag;
    int mask_u8;                          // out_mask is a uint8 buffer (TCW_MASK_U8)
};

// IMG3: the forward warp's own shape -- three colour planes (holes -1, range check, clamp) and a depth plane -- with the
// branches of the generic store loop resolved at compile time and 32-bit byte offsets from one uniform base per output.
template <bool IMG3, bool SCATTER>      // SCATTER: tcw_forward_warp_scatter (its own instantiation: the default one stays as it was)
// (8 CTAs per SM = 64 registers, 32 warps: in the 12-unit launches of the timed schedule the slabs come from HBM and the kernel is
// latency-bound: 3.72 -> 3.60 ms per step; 10 CTAs spill: 4.04)
__global__ void __launch_bounds__(128, 8)
k_gather_normalise(WinArgs A, WinNormArgs N, int unit0, int prep0, int slot0, int tiles_x) {
    constexpr unsigned FULL = 0xffffffffu;
    __shared__ int4 cand_all[4][WIN_MAXC];                          // per warp: {x0, y0, bw | bh << 16, slab offset}
    const int lane = threadIdx.x & 31, wv = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);      // the shuffle tells the compiler the warp index is uniform
    int4* cand = cand_all[wv];
    __shared__ int hitl_all[4][32];
    int* hitl = hitl_all[wv];
    const int uc = blockIdx.y, u = unit0 + uc, ps = prep0 + uc, sl = slot0 + uc;
    const int h = A.h, w = A.w;
    const size_t hw = (size_t)h * w;
    const int ty_t = blockIdx.x / tiles_x, tx_t = blockIdx.x - ty_t * tiles_x;
    const int X0 = tx_t * 32, Y0 = ty_t * 32, X1 = X0 + 32 < w ? X0 + 32 : w;
    const int tx = X0 + lane, row0 = Y0 + wv * 8;
    const int R1 = row0 + 8 < h ? row0 + 8 : h;                     // this warp's tile: [X0, X1) x [row0, R1)
    const WinBox* __restrict__ box = A.sbox + (size_t)sl * A.nh;          // slabs as K2 wrote them (offset in `pad`)
    const WinGrp* __restrict__ grp = A.grp + (size_t)ps * A.ng;
    const float4* __restrict__ p4 = A.pool4 + (size_t)sl * A.pool_cells;
    const float* __restrict__ pw = A.poolw + (size_t)sl * A.pool_cells;

    float4 acc[8];
    float accw[8];
    #pragma unroll
    for (int r = 0; r < 8; ++r) { acc[r] = make_float4(0.f, 0.f, 0.f, 0.f); accw[r] = 0.f; }

    // adds the n listed slabs (ascending half-strip order) to this warp's 32 x 8 cells.  Branch-free: a cell outside
    // the slab reads the zero cell behind the pools (an L1 hit), so all sixteen loads of a slab are in flight together and every
    // accumulator takes the same sequence of additions (x + 0 changes nothing).  Predicated loads instead were slower (114 registers).
    const unsigned zero_cell = (unsigned)((long long)(A.n_slots - sl) * A.pool_cells);
    const char* __restrict__ b4 = (const char*)p4;
    const char* __restrict__ b1 = (const char*)pw;
    auto add_listed = [&](int n) {
        __syncwarp();
        for (int i = 0; i < n; ++i) {
            const int4 e = cand[i];
            const int bw = e.z & 0xffff, bh = (int)((unsigned)e.z >> 16);
            const int c = tx - e.x, r0 = row0 - e.y;
            const bool colin = (unsigned)c < (unsigned)bw;
            int ci = e.w + r0 * bw + c;
            float4 s4[8];
            float sw[8];
            #pragma unroll
            for (int r = 0; r < 8; ++r) {
                const unsigned cu = ((unsigned)(r0 + r) < (unsigned)bh && colin && TCW_INRANGE(N.range_flag, (unsigned)ci < (unsigned)A.pool_cells)) ? (unsigned)ci : zero_cell;
                s4[r] = __ldg((const float4*)(b4 + (size_t)cu * 16u));       // written by an earlier kernel, read-only here: L1 keeps the
                sw[r] = __ldg((const float*)(b1 + (size_t)cu * 4u));        // zero cell and the lines two rows share (ld.cg: 2.40 -> 1.48 ms)
                ci += bw;
            }
            #pragma unroll
            for (int r = 0; r < 8; ++r) {
                acc[r].x += s4[r].x; acc[r].y += s4[r].y; acc[r].z += s4[r].z; acc[r].w += s4[r].w; accw[r] += sw[r];
            }
        }
        __syncwarp();
    };

    if (row0 < h) {
        // ---- every warp lists the half-strips whose box overlaps its own tile, ascending: 32 box groups per round,
        // then the boxes of up to four hit groups per step (lane = group-in-step * 8 + half-strip-in-group)
        int n = 0;
        const int sel = lane >> 3, j = lane & 7;
        for (int g0 = 0; g0 < A.ng; g0 += 32) {
            const int gi = g0 + lane;
            bool hit = false;
            if (gi < A.ng) { const WinGrp q = grp[gi]; hit = q.x0 < X1 && q.x1 > X0 && q.y0 < R1 && q.y1 > row0; }
            const unsigned hits = __ballot_sync(FULL, hit);
            const int nhit = __popc(hits);
            if (nhit == 0) continue;
            __syncwarp();
            if (hit) hitl[__popc(hits & ((1u << lane) - 1u))] = gi;          // hit groups, ascending
            __syncwarp();
            for (int b = 0; b < nhit; b += 4) {
                if (n > WIN_MAXC - 32) { add_listed(n); n = 0; }
                bool in = false;
                int4 e = make_int4(0, 0, 0, 0);
                if (b + sel < nhit) {
                    const int hs = hitl[b + sel] * WIN_GROUP + j;
                    if (hs < A.nh) {
                        const WinBox bx = box[hs];
                        const int off = bx.pad;
                        in = bx.bw > 0 && off >= 0 && bx.x0 < X1 && bx.x0 + bx.bw > X0 && bx.y0 < R1 && bx.y0 + bx.bh > row0;
                        e = make_int4(bx.x0, bx.y0, (int)bx.bw | ((int)bx.bh << 16), off);
                    }
                }
                const unsigned m = __ballot_sync(FULL, in);
                if (in) cand[n + __popc(m & ((1u << lane) - 1u))] = e;
                n += __popc(m);
            }
        }
        add_listed(n);
    }
    // ---- the unit's shared plane (only when a footprint did not fit the slab pool); re-zeroed for the next unit
    if (A.spill_flag[sl]) {
        float4* s4 = A.spill4 + (size_t)sl * hw;
        float* sw = A.spillw + (size_t)sl * hw;
        #pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (row0 + r < h && tx < w) {
                const size_t p = (size_t)(row0 + r) * w + tx;
                const float4 s = __ldcg(s4 + p);
                const float t = __ldcg(sw + p);
                acc[r].x += s.x; acc[r].y += s.y; acc[r].z += s.z; acc[r].w += s.w; accw[r] += t;
                s4[p] = make_float4(0.f, 0.f, 0.f, 0.f); sw[p] = 0.f;
            }
        }
    }
    // ---- pile sums -> output pixel (1,1): strips in ascending order
    if (blockIdx.x == 0 && wv == 0 && w >= 2 && h >= 2) {
        const float* __restrict__ pile = A.pile + (size_t)sl * A.ns * 8;
        float4 t4 = make_float4(0.f, 0.f, 0.f, 0.f);
        float tw = 0.f;
        for (int s0 = 0; s0 < A.ns; s0 += 32) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), bq = a;
            if (s0 + lane < A.ns) { a = __ldcg((const float4*)(pile + (size_t)(s0 + lane) * 8)); bq = __ldcg((const float4*)(pile + (size_t)(s0 + lane) * 8 + 4)); }
            unsigned m = __ballot_sync(FULL, bq.y != 0.f);
            while (m) {
                const int k = __ffs(m) - 1;
                m &= m - 1;
                t4.x += __shfl_sync(FULL, a.x, k); t4.y += __shfl_sync(FULL, a.y, k); t4.z += __shfl_sync(FULL, a.z, k);
                t4.w += __shfl_sync(FULL, a.w, k); tw += __shfl_sync(FULL, bq.x, k);
            }
        }
        if (lane == 1) { acc[1].x += t4.x; acc[1].y += t4.y; acc[1].z += t4.z; acc[1].w += t4.w; accw[1] += tw; }
    }
    // ---- normalise, mask, clip (models/utils.py:571-583)
    bool bad = false;
    if (IMG3) {
        if (tx < w && row0 < h) {
            // byte offsets fit 32 bits: 3 * h * w * 4 < 2^32 (h * w < 2^28 is checked at submission)
            const int uo = SCATTER ? win_dst(A, u) : u;
            char* __restrict__ fb = (char*)(N.out_frame + (size_t)uo * 3 * hw);
            char* __restrict__ db = (char*)(N.out_depth + (size_t)uo * hw);
            char* __restrict__ mb = (char*)N.out_mask + (size_t)uo * hw * (N.mask_u8 ? 1 : 4);
            const unsigned hw4 = (unsigned)hw * 4u, w4 = (unsigned)w * 4u;
            unsigned off = ((unsigned)row0 * (unsigned)w + (unsigned)tx) * 4u;
            const int nr = R1 - row0;
            #pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (r >= nr) break;
                const float W = accw[r];
                const bool valid = W > 0.0f;
                float rr = rcp_approx(W);
                rr = fmaf(rr, fmaf(-W, rr, 1.0f), rr);
                const float x0 = valid ? acc[r].x * rr : -1.0f, x1 = valid ? acc[r].y * rr : -1.0f, x2 = valid ? acc[r].z * rr : -1.0f;
                bad |= !(fabsf(x0) <= 1.1f) | !(fabsf(x1) <= 1.1f) | !(fabsf(x2) <= 1.1f);      // (a NaN fails too, as in the reference's assert)
                __stcs((float*)(fb + off), clampf(x0, -1.0f, 1.0f));
                __stcs((float*)(fb + (off + hw4)), clampf(x1, -1.0f, 1.0f));
                __stcs((float*)(fb + (off + 2u * hw4)), clampf(x2, -1.0f, 1.0f));
                __stcs((float*)(db + off), valid ? acc[r].w * rr : 0.0f);
                if (N.mask_u8) *(unsigned char*)(mb + (off >> 2)) = valid ? 1 : 0;
                else __stcs((float*)(mb + off), valid ? 1.0f : 0.0f);
                off += w4;
            }
        }
    } else if (tx < w && row0 < h) {
        const size_t p0 = (size_t)row0 * w + tx;
        const int uo = SCATTER ? win_dst(A, u) : u;
        float* o0 = N.out_frame + (size_t)uo * N.out_c * hw + p0;
        float* om = N.out_mask + (size_t)uo * hw + p0;
        unsigned char* om8 = (unsigned char*)N.out_mask + (size_t)uo * hw + p0;
        float* od = N.out_depth ? N.out_depth + (size_t)uo * hw + p0 : nullptr;
        const float hole = N.is_image ? -1.0f : 0.0f;
        #pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (row0 + r >= h) break;
            const float W = accw[r];
            const bool valid = W > 0.0f;
            float rr = rcp_approx(W);
            rr = fmaf(rr, fmaf(-W, rr, 1.0f), rr);
            const float v[4] = {acc[r].x * rr, acc[r].y * rr, acc[r].z * rr, acc[r].w * rr};
            #pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (c < N.out_c) {
                    float x = valid ? v[c] : hole;
                    if (N.is_image) { bad |= !(fabsf(x) <= 1.1f); x = clampf(x, -1.0f, 1.0f); }
                    __stcs(o0 + (size_t)c * hw, x);
                }
            }
            if (N.mask_u8) *om8 = valid ? 1 : 0;
            else __stcs(om, valid ? 1.0f : 0.0f);
            if (od) { __stcs(od, valid ? v[3] : 0.0f); od += w; }
            o0 += w; om += w; om8 += w;
        }
    }
    if (bad && N.range_flag) atomicOr(N.range_flag, 1);
}
