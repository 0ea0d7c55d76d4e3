// warp_pair2.cuh — EXPERIMENT for the next round (compiled only with -DTCW_ROWS2=1, never in the default build):
// the strip code of warp_pair.cuh with TWO source rows per loop iteration.
//
// Why: k_splat_pair issues 0.58 instructions per active cycle; the largest stall is the fixed-latency dependency
// wait of ONE packed dependency chain per lane (profiles/r1_v4_notes.md).  Unrolling the row loop does not help
// because every warp-uniform guard ends a basic block.  Here stage 1 of a row (projection, flow division, soft-z
// weight, corner ids and weights) is written branch-free for the common case (pinhole intrinsics, no mask1, paired
// loads, usable soft-z normaliser, every live pixel interior, division preconditions met) and evaluated for rows y
// and y+1 back to back in one basic block, so ptxas can interleave the two chains; ONE guard after both decides
// whether the pair of rows is redone by the general stage 1.  Stage 2 (merge + reductions) runs per row, in
// order, exactly as in warp_pair.cuh (TCW_MERGE_RULES == 2).
//
// Arithmetic per pixel is the same instruction sequence as warp_pair.cuh (same helpers), so the bit-exactness
// argument of DESIGN.md §2 carries over unchanged.  NOT YET RUN ON A GPU.
#pragma once

struct Row2 {                       // stage-1 result of one row: this lane's pixels A and B
    int idNWA, idSWA, idNEA, idSEA, idNWB, idSWB, idNEB, idSEB;
    u64 wNW2, wSW2, wNE2, wSE2;     // (A, B) corner weights = proximity * soft-z weight
    u64 flx2, fly2, mg2;            // flow12 and the soft-z weight (behind-camera sums need it)
    float P2A, P2B;                 // z' (the depth channel that is splatted)
    bool bhA, bhB, ok;
};

template <int SH>
__device__ __forceinline__ void splat_pair_strip2(const PairArgs& A, int u, int slot, int strip) {
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int MODE = MODE_F32;
    const int lane = threadIdx.x & 31;
    int h = A.h, w = A.w;
    const int sy = strip / A.strips_x, sx = strip - sy * A.strips_x;
    const int xA = sx * PAIR_W + 2 * lane, xB = xA + 1;
    const int y0 = sy * SH, y1 = y0 + SH < h ? y0 + SH : h;
    const bool okA = xA < w, okB = xB < w;
    const size_t hw = (size_t)h * w;
    const float* __restrict__ Mu = A.Mall + (size_t)u * M_FLOATS;
    const int s = A.src_index ? A.src_index[u] : u;
    const UnitStats* st = A.stats + (A.stat_index_global >= 0 ? A.stat_index_global : u);
    const float L = __uint_as_float(st->lbits);
    const bool Lok = L >= 9.5367431640625e-07f;
    float m0 = __ldg(Mu + 0), m1 = __ldg(Mu + 1), m2 = __ldg(Mu + 2), m3 = __ldg(Mu + 3), m4 = __ldg(Mu + 4), m5 = __ldg(Mu + 5);
    float m6 = __ldg(Mu + 6), m7 = __ldg(Mu + 7), m8 = __ldg(Mu + 8);
    float T[12], K[9];
    #pragma unroll
    for (int i = 0; i < 12; ++i) T[i] = __ldg(Mu + 9 + i);
    #pragma unroll
    for (int i = 0; i < 9; ++i) K[i] = __ldg(Mu + 21 + i);
    const bool PIN = m1 == 0.0f && m3 == 0.0f && m6 == 0.0f && m7 == 0.0f && m8 == 1.0f &&
                     K[1] == 0.0f && K[3] == 0.0f && K[6] == 0.0f && K[7] == 0.0f && K[8] == 1.0f;
    const bool vec2 = okB && !(w & 1) && !((((size_t)A.depth) | ((size_t)A.frame) | ((size_t)A.out_flow)) & 7);
    const bool vec2all = __all_sync(FULL, vec2 || !okA);
    // anything outside the common case: the one-row code handles the whole strip (warp-uniform decision)
    if (!(PIN && Lok && A.mask1 == nullptr && vec2all)) {
        splat_pair_strip<SH, MODE_F32>(A, u, slot, strip);
        return;
    }

    PairOut o;
    o.acc_i64 = nullptr; o.wmax = nullptr; o.det_scale = A.det_scale; o.inv_zabs = 1.0f; o.hw = hw;
    o.acc4 = A.acc4 + (size_t)slot * hw - (w + 1);
    o.accw = A.accw + (size_t)slot * hw - (w + 1);
    pin_ptr(o.acc4);
    pin_ptr(o.accw);
    float rL = rcp_approx(L);
    rL = fmaf(rL, fmaf(-L, rL, 1.0f), rL);
    const u64 rL2 = pk1(rL), nL2 = pk1(-L);
    const float fxA = (float)xA, fxB = (float)xB;
    const float u0A = __fadd_rn(__fmul_rn(m0, fxA), m2), u0B = __fadd_rn(__fmul_rn(m0, fxB), m2);
    float wf1 = (float)(w + 1), hf1 = (float)(h + 1);
    pin_f(wf1);
    pin_f(hf1);
    const u64 fx2 = pk(fxA, fxB);
    const int colA = okA ? xA : 0;
    const float* __restrict__ dp = A.depth + (size_t)s * hw;
    const float* __restrict__ fp = A.frame + (size_t)s * 3 * hw;
    float* __restrict__ flp = A.out_flow ? A.out_flow + (size_t)u * 2 * hw : nullptr;
    asm volatile("" : "+l"(dp));
    asm volatile("" : "+l"(fp));
    const unsigned hw32 = (unsigned)hw;

    // ---- stage 1 of one row.  general == false: branch-free, sets R.ok; general == true: every case (division
    //      fallback, clamped corner formulas), used when the guard after the pair of rows fails
    auto stage1 = [&](int y, bool rowok, float dA, float dB, bool general, Row2& R) {
        const float fy = (float)y;
        const float u1 = __fadd_rn(__fmul_rn(m4, fy), m5);
        const u64 d2 = pk(dA, dB);
        const u64 X0 = mul2(d2, pk(u0A, u0B)), X1 = mul2(d2, pk1(u1));
        float a, b, c, e, f, g;
        upk(mul2(X0, pk1(T[0])), a, b); upk(mul2(X1, pk1(T[1])), c, e); upk(mul2(d2, pk1(T[2])), f, g);
        const float Y0A = __fadd_rn(__fadd_rn(__fadd_rn(a, c), f), T[3]), Y0B = __fadd_rn(__fadd_rn(__fadd_rn(b, e), g), T[3]);
        upk(mul2(X0, pk1(T[4])), a, b); upk(mul2(X1, pk1(T[5])), c, e); upk(mul2(d2, pk1(T[6])), f, g);
        const float Y1A = __fadd_rn(__fadd_rn(__fadd_rn(a, c), f), T[7]), Y1B = __fadd_rn(__fadd_rn(__fadd_rn(b, e), g), T[7]);
        upk(mul2(X0, pk1(T[8])), a, b); upk(mul2(X1, pk1(T[9])), c, e); upk(mul2(d2, pk1(T[10])), f, g);
        const float Y2A = __fadd_rn(__fadd_rn(__fadd_rn(a, c), f), T[11]), Y2B = __fadd_rn(__fadd_rn(__fadd_rn(b, e), g), T[11]);
        const u64 Y22 = pk(Y2A, Y2B);
        upk(mul2(pk(Y0A, Y0B), pk1(K[0])), a, b); upk(mul2(Y22, pk1(K[2])), c, e);
        float P0A = __fadd_rn(a, c), P0B = __fadd_rn(b, e);
        upk(mul2(pk(Y1A, Y1B), pk1(K[4])), a, b); upk(mul2(Y22, pk1(K[5])), c, e);
        float P1A = __fadd_rn(a, c), P1B = __fadd_rn(b, e);
        float P2A = Y2A, P2B = Y2B;
        const bool bhA = Y2A <= 0.01f, bhB = Y2B <= 0.01f;                   // behind the camera (:404-417)
        P0A = bhA ? 1000.0f : P0A; P1A = bhA ? 1000.0f : P1A; P2A = bhA ? 1000.0f : P2A;
        P0B = bhB ? 1000.0f : P0B; P1B = bhB ? 1000.0f : P1B; P2B = bhB ? 1000.0f : P2B;
        R.bhA = bhA && rowok; R.bhB = bhB && rowok; R.P2A = P2A; R.P2B = P2B;

        // flow12 = P.xy / P.z - grid: the shared-reciprocal division and its preconditions (warp_pair.cuh)
        const float zlo = fminf(P2A, P2B), zhi = fmaxf(P2A, P2B);
        const float nlo = fminf(fminf(fabsf(P0A), fabsf(P0B)), fminf(fabsf(P1A), fabsf(P1B)));
        const float nhi = fmaxf(fmaxf(fabsf(P0A), fabsf(P0B)), fmaxf(fabsf(P1A), fabsf(P1B)));
        const bool fast = zlo >= 9.5367431640625e-07f && zhi <= 1048576.0f && nlo >= 7.888609052210118e-31f && nhi <= 1.2676506002282294e30f;
        u64 qx, qy;
        if (!general || __all_sync(FULL, fast)) {
            const u64 b2 = pk(P2A, P2B), nb2 = pk(-P2A, -P2B);
            const u64 r = refine_rcp2(b2, nb2);
            qx = div2_with_rcp(pk(P0A, P0B), nb2, r);
            qy = div2_with_rcp(pk(P1A, P1B), nb2, r);
        } else {
            qx = pk(__fdiv_rn(P0A, P2A), __fdiv_rn(P0B, P2B));
            qy = pk(__fdiv_rn(P1A, P2A), __fdiv_rn(P1B, P2B));
        }
        const u64 fy2 = pk1(fy);
        R.flx2 = sub2(qx, fx2); R.fly2 = sub2(qy, fy2);

        // soft-z weight (Lok holds for this strip)
        {
            const u64 a1 = add2(pk(clampf(P2A, 0.0f, 1000.0f), clampf(P2B, 0.0f, 1000.0f)), pk1(1.0f));
            const u64 ld = log2x(a1);
            R.mg2 = exp2x(mul2(div2_with_rcp(ld, nL2, rL2), pk1(-50.0f)));
        }

        // corner cells and proximity weights
        const u64 ox2 = add2(add2(R.flx2, fx2), pk1(1.0f)), oy2 = add2(add2(R.fly2, fy2), pk1(1.0f));
        float oxA, oxB, oyA, oyB;
        upk(ox2, oxA, oxB);
        upk(oy2, oyA, oyB);
        const float fx0A = floorf(oxA), fx0B = floorf(oxB), fy0A = floorf(oyA), fy0B = floorf(oyB);
        const int ix0A = (int)fx0A, ix0B = (int)fx0B, iy0A = (int)fy0A, iy0B = (int)fy0B;
        const bool inA = (unsigned)(ix0A - 1) < (unsigned)(w - 1) && (unsigned)(iy0A - 1) < (unsigned)(h - 1);
        const bool inB = (unsigned)(ix0B - 1) < (unsigned)(w - 1) && (unsigned)(iy0B - 1) < (unsigned)(h - 1);
        const bool liveA0 = okA && !bhA && rowok, liveB0 = okB && !bhB && rowok;
        // dead = every corner in the pad ring; a live pixel that is neither dead nor interior needs the clamped formulas
        const bool deadA = !liveA0 || !(oxA > 0.0f && oxA < wf1 && oyA > 0.0f && oyA < hf1);
        const bool deadB = !liveB0 || !(oxB > 0.0f && oxB < wf1 && oyB > 0.0f && oyB < hf1);
        const bool liveA = !deadA, liveB = !deadB;
        R.ok = fast && (deadA || inA) && (deadB || inB);
        if (!general || __all_sync(FULL, (deadA || inA) && (deadB || inB))) {
            const bool dxA = oxA != fx0A, dxB = oxB != fx0B, dyA = oyA != fy0A, dyB = oyB != fy0B;
            const u64 fx02 = pk(fx0A, fx0B), fy02 = pk(fy0A, fy0B);
            const u64 fx12 = pk(dxA ? fx0A + 1.0f : fx0A, dxB ? fx0B + 1.0f : fx0B);
            const u64 fy12 = pk(dyA ? fy0A + 1.0f : fy0A, dyB ? fy0B + 1.0f : fy0B);
            const u64 one2 = pk1(1.0f);
            const u64 ax = sub2(one2, sub2(ox2, fx02)), bx = sub2(one2, sub2(fx12, ox2));
            const u64 ay = sub2(one2, sub2(oy2, fy02)), by = sub2(one2, sub2(fy12, oy2));
            R.wNW2 = mul2(mul2(ay, ax), R.mg2); R.wSW2 = mul2(mul2(by, ax), R.mg2);
            R.wNE2 = mul2(mul2(ay, bx), R.mg2); R.wSE2 = mul2(mul2(by, bx), R.mg2);
            R.idNWA = ((iy0A * w + ix0A) & 0x3fffffff) | (liveA ? 0 : 0x80000000);
            R.idNWB = ((iy0B * w + ix0B) & 0x3fffffff) | (liveB ? 0 : 0x80000000);
            R.idSWA = R.idNWA + (dyA ? w : 0);
            R.idSWB = R.idNWB + (dyB ? w : 0);
            R.idNEA = R.idNWA + (dxA ? 1 : 0);
            R.idNEB = R.idNWB + (dxB ? 1 : 0);
            R.idSEA = R.idSWA + (dxA ? 1 : 0);
            R.idSEB = R.idSWB + (dxB ? 1 : 0);
        } else {
            const PairCorners ca = pair_corners_general(plo(R.flx2), plo(R.fly2), xA, y, h, w, !liveA);
            const PairCorners cb = pair_corners_general(phi(R.flx2), phi(R.fly2), xB, y, h, w, !liveB);
            const int bias = w + 1;
            R.idNWA = ca.nw < 0 ? -1 : ca.nw + bias; R.idSWA = ca.sw < 0 ? -1 : ca.sw + bias;
            R.idNEA = ca.ne < 0 ? -1 : ca.ne + bias; R.idSEA = ca.se < 0 ? -1 : ca.se + bias;
            R.idNWB = cb.nw < 0 ? -1 : cb.nw + bias; R.idSWB = cb.sw < 0 ? -1 : cb.sw + bias;
            R.idNEB = cb.ne < 0 ? -1 : cb.ne + bias; R.idSEB = cb.se < 0 ? -1 : cb.se + bias;
            R.wNW2 = mul2(pk(ca.pnw, cb.pnw), R.mg2); R.wSW2 = mul2(pk(ca.psw, cb.psw), R.mg2);
            R.wNE2 = mul2(pk(ca.pne, cb.pne), R.mg2); R.wSE2 = mul2(pk(ca.pse, cb.pse), R.mg2);
        }
    };

    // carried bottom cells, behind-camera sums (as in warp_pair.cuh)
    int cLid = -1, cMid = -1;
    u64 cLrg = 0, cLbz = 0, cMrg = 0, cMbz = 0;
    float cLw = 0.f, cMw = 0.f;
    u64 hrg = 0, hbz = 0;
    float hwt = 0.f;
    bool any_behind = false;

    // ---- stage 2 of one row: merge + reductions (TCW_MERGE_RULES == 2 of warp_pair.cuh, statement for statement)
    auto stage2 = [&](const Row2& R, u64 rgA, u64 bA_z, u64 rgB, u64 bB_z) {
        const u64 bzA = bA_z, bzB = bB_z;
        float wNWA, wNWB, wSWA, wSWB, wNEA, wNEB, wSEA, wSEB;
        upk(R.wNW2, wNWA, wNWB); upk(R.wSW2, wSWA, wSWB); upk(R.wNE2, wNEA, wNEB); upk(R.wSE2, wSEA, wSEB);
        const int idNWA = R.idNWA, idSWA = R.idSWA, idNEA = R.idNEA, idSEA = R.idSEA;
        const int idNWB = R.idNWB, idSWB = R.idSWB, idNEB = R.idNEB, idSEB = R.idSEB;
        if (R.bhA || R.bhB) {
            any_behind = true;
            const float mga = okA && R.bhA ? plo(R.mg2) : 0.0f, mgb = okB && R.bhB ? phi(R.mg2) : 0.0f;
            hrg = fma2(rgA, pk1(mga), fma2(rgB, pk1(mgb), hrg));
            hbz = fma2(bzA, pk1(mga), fma2(bzB, pk1(mgb), hbz));
            hwt += mga + mgb;
        }
        const int src = (lane + 31) & 31;
        const int idNEl = __shfl_sync(FULL, idNEB, src), idSEl = __shfl_sync(FULL, idSEB, src);
        const float wNEl = __shfl_sync(FULL, wNEB, src), wSEl = __shfl_sync(FULL, wSEB, src);
        const u64 rgl = __shfl_sync(FULL, rgB, src), bzl = __shfl_sync(FULL, bzB, src);
        u64 Ltrg = mul2(rgA, pk1(wNWA)), Ltbz = mul2(bzA, pk1(wNWA)), Lbrg = mul2(rgA, pk1(wSWA)), Lbbz = mul2(bzA, pk1(wSWA));
        u64 Mtrg = mul2(rgB, pk1(wNWB)), Mtbz = mul2(bzB, pk1(wNWB)), Mbrg = mul2(rgB, pk1(wSWB)), Mbbz = mul2(bzB, pk1(wSWB));
        u64 ANrg = mul2(rgA, pk1(wNEA)), ANbz = mul2(bzA, pk1(wNEA)), ASrg = mul2(rgA, pk1(wSEA)), ASbz = mul2(bzA, pk1(wSEA));
        float Ltw = wNWA, Lbw = wSWA, Mtw = wNWB, Mbw = wSWB, ANw = wNEA, ASw = wSEA;
        {
            const bool a = idNEl == idNWA, b = !a && idNEl == idSWA, c = !a && !b && idNEl == idNEA;
            const float wa = a ? wNEl : 0.0f, wb = b ? wNEl : 0.0f, wc = c ? wNEl : 0.0f;
            Ltrg = fma2(rgl, pk1(wa), Ltrg); Ltbz = fma2(bzl, pk1(wa), Ltbz); Ltw += wa;
            Lbrg = fma2(rgl, pk1(wb), Lbrg); Lbbz = fma2(bzl, pk1(wb), Lbbz); Lbw += wb;
            ANrg = fma2(rgl, pk1(wc), ANrg); ANbz = fma2(bzl, pk1(wc), ANbz); ANw += wc;
            emit_single<MODE>(o, !(a || b || c), idNEl, rgl, bzl, wNEl);
        }
        {
            const bool a = idSEl == idSWA, b = !a && idSEl == idNWA, c = !a && !b && idSEl == idSEA;
            const float wa = a ? wSEl : 0.0f, wb = b ? wSEl : 0.0f, wc = c ? wSEl : 0.0f;
            Lbrg = fma2(rgl, pk1(wa), Lbrg); Lbbz = fma2(bzl, pk1(wa), Lbbz); Lbw += wa;
            Ltrg = fma2(rgl, pk1(wb), Ltrg); Ltbz = fma2(bzl, pk1(wb), Ltbz); Ltw += wb;
            ASrg = fma2(rgl, pk1(wc), ASrg); ASbz = fma2(bzl, pk1(wc), ASbz); ASw += wc;
            emit_single<MODE>(o, !(a || b || c), idSEl, rgl, bzl, wSEl);
        }
        {
            const bool a = idNEA == idNWB, b = !a && idNEA == idSWB;
            const float fa = a ? 1.0f : 0.0f, fb = b ? 1.0f : 0.0f;
            Mtrg = fma2(ANrg, pk1(fa), Mtrg); Mtbz = fma2(ANbz, pk1(fa), Mtbz); Mtw = fmaf(ANw, fa, Mtw);
            Mbrg = fma2(ANrg, pk1(fb), Mbrg); Mbbz = fma2(ANbz, pk1(fb), Mbbz); Mbw = fmaf(ANw, fb, Mbw);
            emit_cell<MODE>(o, (a || b) ? -1 : idNEA, ANrg, ANbz, ANw);
        }
        {
            const bool a = idSEA == idSWB, b = !a && idSEA == idNWB;
            const float fa = a ? 1.0f : 0.0f, fb = b ? 1.0f : 0.0f;
            Mbrg = fma2(ASrg, pk1(fa), Mbrg); Mbbz = fma2(ASbz, pk1(fa), Mbbz); Mbw = fmaf(ASw, fa, Mbw);
            Mtrg = fma2(ASrg, pk1(fb), Mtrg); Mtbz = fma2(ASbz, pk1(fb), Mtbz); Mtw = fmaf(ASw, fb, Mtw);
            emit_cell<MODE>(o, (a || b) ? -1 : idSEA, ASrg, ASbz, ASw);
        }
        {
            const bool a = cLid == idNWA, b = !a && cLid == idSWA;
            const float fa = a ? 1.0f : 0.0f, fb = b ? 1.0f : 0.0f;
            emit_cell<MODE>(o, (a || b) ? -1 : cLid, cLrg, cLbz, cLw);
            Ltrg = fma2(cLrg, pk1(fa), Ltrg); Ltbz = fma2(cLbz, pk1(fa), Ltbz); Ltw = fmaf(cLw, fa, Ltw);
            Lbrg = fma2(cLrg, pk1(fb), Lbrg); Lbbz = fma2(cLbz, pk1(fb), Lbbz); Lbw = fmaf(cLw, fb, Lbw);
        }
        {
            const bool a = cMid == idNWB, b = !a && cMid == idSWB;
            const float fa = a ? 1.0f : 0.0f, fb = b ? 1.0f : 0.0f;
            emit_cell<MODE>(o, (a || b) ? -1 : cMid, cMrg, cMbz, cMw);
            Mtrg = fma2(cMrg, pk1(fa), Mtrg); Mtbz = fma2(cMbz, pk1(fa), Mtbz); Mtw = fmaf(cMw, fa, Mtw);
            Mbrg = fma2(cMrg, pk1(fb), Mbrg); Mbbz = fma2(cMbz, pk1(fb), Mbbz); Mbw = fmaf(cMw, fb, Mbw);
        }
        const bool sT = idNWB == idNWA, sB = idSWB == idSWA;
        {
            const float fa = sT ? 1.0f : 0.0f, fb = sB ? 1.0f : 0.0f;
            Ltrg = fma2(Mtrg, pk1(fa), Ltrg); Ltbz = fma2(Mtbz, pk1(fa), Ltbz); Ltw = fmaf(Mtw, fa, Ltw);
            Lbrg = fma2(Mbrg, pk1(fb), Lbrg); Lbbz = fma2(Mbbz, pk1(fb), Lbbz); Lbw = fmaf(Mbw, fb, Lbw);
        }
        emit_cell<MODE>(o, idNWA, Ltrg, Ltbz, Ltw);
        emit_cell<MODE>(o, sT ? -1 : idNWB, Mtrg, Mtbz, Mtw);
        cLid = idSWA; cLrg = Lbrg; cLbz = Lbbz; cLw = Lbw;
        cMid = sB ? -1 : idSWB; cMrg = Mbrg; cMbz = Mbbz; cMw = Mbw;
    };

    // ---- two rows per iteration; depth is prefetched one iteration ahead, colours are loaded at the top of the
    //      iteration that uses them ~300 instructions later
    auto row_of = [&](int y) { return (unsigned)(y < y1 ? y : y1 - 1) * (unsigned)w + colA; };
    float2 dn0 = __ldg((const float2*)(dp + row_of(y0))), dn1 = __ldg((const float2*)(dp + row_of(y0 + 1)));
    for (int y = y0; y < y1; y += 2) {
        const bool row1 = y + 1 < y1;
        const float2 d0 = dn0, d1 = dn1;
        const unsigned q0 = row_of(y), q1 = row_of(y + 1);
        const float2 r0 = __ldg((const float2*)(fp + q0)), g0 = __ldg((const float2*)(fp + hw32 + q0)), b0 = __ldg((const float2*)(fp + 2 * hw32 + q0));
        const float2 r1 = __ldg((const float2*)(fp + q1)), g1 = __ldg((const float2*)(fp + hw32 + q1)), b1 = __ldg((const float2*)(fp + 2 * hw32 + q1));
        if (y + 2 < y1) { dn0 = __ldg((const float2*)(dp + row_of(y + 2))); dn1 = __ldg((const float2*)(dp + row_of(y + 3))); }

        Row2 R0, R1;
        stage1(y, true, d0.x, d0.y, false, R0);
        stage1(y + 1, row1, d1.x, d1.y, false, R1);
        if (!__all_sync(FULL, R0.ok && R1.ok)) {       // rare: redo the pair of rows with every case handled
            stage1(y, true, d0.x, d0.y, true, R0);
            stage1(y + 1, row1, d1.x, d1.y, true, R1);
        }
        if (flp && okA) {
            float fxa, fxb, fya, fyb;
            upk(R0.flx2, fxa, fxb);
            upk(R0.fly2, fya, fyb);
            __stcs((float2*)(flp + q0), make_float2(fxa, fxb));
            __stcs((float2*)(flp + hw32 + q0), make_float2(fya, fyb));
            if (row1) {
                upk(R1.flx2, fxa, fxb);
                upk(R1.fly2, fya, fyb);
                __stcs((float2*)(flp + q1), make_float2(fxa, fxb));
                __stcs((float2*)(flp + hw32 + q1), make_float2(fya, fyb));
            }
        }
        stage2(R0, pk(r0.x, g0.x), pk(b0.x, R0.P2A), pk(r0.y, g0.y), pk(b0.y, R0.P2B));
        stage2(R1, pk(r1.x, g1.x), pk(b1.x, R1.P2A), pk(r1.y, g1.y), pk(b1.y, R1.P2B));
    }
    emit_cell<MODE>(o, cLid, cLrg, cLbz, cLw);
    emit_cell<MODE>(o, cMid, cMrg, cMbz, cMw);

    if (__any_sync(FULL, any_behind)) {
        float hr, hg, hb, hz;
        upk(hrg, hr, hg);
        upk(hbz, hb, hz);
        #pragma unroll
        for (int ofs = 16; ofs > 0; ofs >>= 1) {
            hr += __shfl_xor_sync(FULL, hr, ofs); hg += __shfl_xor_sync(FULL, hg, ofs); hb += __shfl_xor_sync(FULL, hb, ofs);
            hz += __shfl_xor_sync(FULL, hz, ofs); hwt += __shfl_xor_sync(FULL, hwt, ofs);
        }
        if (lane == 0 && w >= 2 && h >= 2) emit_cell<MODE>(o, 2 * w + 2, pk(4.f * hr, 4.f * hg), pk(4.f * hb, 4.f * hz), 4.f * hwt);
    }
}
