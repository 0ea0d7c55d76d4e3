/*
 * warp_oracle.c — CPU restatement of the reference's forward re-rendering path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker ("oracle") for libtcwarp.so.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product (trajectorycrafter_b200/) never imports, links or calls anything in oracle/.
 *
 * It restates, in scalar fp32 C with the reference's exact operation order, these functions of
 * /root/reference/models/utils.py (cited per function below):
 *     Warper.compute_transformed_points  :350-420
 *     Warper.forward_warp (twice=False)  :220-293
 *     Warper.bilinear_splatting          :422-583
 *     Warper.clean_points                :585-626
 *
 * Pinning: the reference ships no golden vectors or tests for this path, so
 * the oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF, generated in the build container
 * by oracle/gen_golden.py (which imports /root/reference/models/utils.py unmodified) and committed
 * under tests/golden/.  tests/test_oracle_golden.py asserts: flow12, masks and projected points
 * bit-equal; colour/depth within 1e-5 (the residual is libm log/exp rounding).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off: no FMA contraction, no fast-math).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_EXPORT __attribute__((visibility("default")))

/* acc = 0; acc += m[k]*v[k] in order: the accumulation order of torch.matmul/bmm on these tiny
 * operands (models/utils.py:388,397,399), verified bit-exact against the reference. */
static inline float dot3(const float* m, float a, float b, float c) {
    float acc = 0.0f;
    acc = acc + m[0] * a;
    acc = acc + m[1] * b;
    acc = acc + m[2] * c;
    return acc;
}
static inline float dot4(const float* m, float a, float b, float c, float d) {
    float acc = 0.0f;
    acc = acc + m[0] * a;
    acc = acc + m[1] * b;
    acc = acc + m[2] * c;
    acc = acc + m[3] * d;
    return acc;
}

/* models/utils.py:350-420.  M = [K1inv(9) | T rows 0..2 (12) | K2 (9)] per unit.  out: (h,w,3). */
static void project_unit(const float* depth, const float* M, int h, int w, float* out) {
    const float* k1i = M;
    const float* t = M + 9;
    const float* k2 = M + 21;
    for (int y = 0; y < h; ++y) {
        for (int x = 0; x < w; ++x) {
            float fx = (float)x, fy = (float)y;
            float d = depth[(size_t)y * w + x];
            /* unnormalized_pos = K1inv @ (x, y, 1)            :388 */
            float u0 = dot3(k1i + 0, fx, fy, 1.0f);
            float u1 = dot3(k1i + 3, fx, fy, 1.0f);
            float u2 = dot3(k1i + 6, fx, fy, 1.0f);
            /* world_points = depth * unnormalized_pos         :395 */
            float X0 = d * u0, X1 = d * u1, X2 = d * u2;
            /* trans_world = (T @ (X,1))[:3]                   :396-398 */
            float Y0 = dot4(t + 0, X0, X1, X2, 1.0f);
            float Y1 = dot4(t + 4, X0, X1, X2, 1.0f);
            float Y2 = dot4(t + 8, X0, X1, X2, 1.0f);
            /* trans_norm_points = K2 @ trans_world            :399 */
            float P0 = dot3(k2 + 0, Y0, Y1, Y2);
            float P1 = dot3(k2 + 3, Y0, Y1, Y2);
            float P2 = dot3(k2 + 6, Y0, Y1, Y2);
            /* behind camera -> all three components 1000      :404-417 */
            if (Y2 <= 0.01f) { P0 = 1000.0f; P1 = 1000.0f; P2 = 1000.0f; }
            float* o = out + ((size_t)y * w + x) * 3;
            o[0] = P0; o[1] = P1; o[2] = P2;
        }
    }
}

ORC_EXPORT int orc_project(const float* depth, const int* src_index, const float* M,
                           int n_units, int n_src, int h, int w, float* out) {
    for (int u = 0; u < n_units; ++u) {
        int s = src_index ? src_index[u] : u;
        if (s < 0 || s >= n_src) return -1;
        project_unit(depth + (size_t)s * h * w, M + (size_t)u * 30, h, w, out + (size_t)u * h * w * 3);
    }
    return 0;
}

static inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* float -> index with the clamp of models/utils.py:453-475.  The reference casts to int64 first
 * and clamps afterwards; clamping in float first is equivalent for every finite value whose cell
 * survives the crop (anything outside [0, n+1] lands in the discarded pad ring either way). */
static inline int clamp_index(float v, int hi) {
    if (!(v > 0.0f)) return 0;
    if (v >= (float)hi) return hi;
    return (int)v;
}

/*
 * models/utils.py:422-583 for a batch of b images sharing one global log-depth max (the
 * reference's `log_depth1.max()` is over the whole tensor).
 * frame (b,c,h,w)  mask1 (b,h,w)|NULL  depth (b,h,w)  flow (b,2,h,w)  flow_mask (b,h,w)|NULL.
 */
ORC_EXPORT int orc_splat(const float* frame, const float* mask1, const float* depth,
                         const float* flow, const float* flow_mask,
                         int b, int c, int h, int w, int is_image,
                         float* out, float* out_mask, int* range_flag) {
    const size_t hw = (size_t)h * w;
    const int ph = h + 2, pw = w + 2;
    const size_t phw = (size_t)ph * pw;
    /* log_depth1 = log(1 + clamp(depth, 0, 1000)); global max   :490-492
     * DEVIATION (documented): torch's max() propagates a NaN (a non-finite input depth whose z' is
     * inf - inf), which invalidates the reference's whole frame; `ld > L` below ignores it and only that pixel is dropped. */
    float L = -INFINITY;
    for (size_t i = 0; i < (size_t)b * hw; ++i) {
        float ld = logf(1.0f + clampf(depth[i], 0.0f, 1000.0f));
        if (ld > L) L = ld;
    }
    float* acc = (float*)calloc((size_t)c * phw, sizeof(float));
    float* wacc = (float*)calloc(phw, sizeof(float));
    int* idx = (int*)malloc(sizeof(int) * 4 * hw);       /* fx, cx, fy, cy */
    float* wt = (float*)malloc(sizeof(float) * 4 * hw);  /* nw, sw, ne, se */
    if (!acc || !wacc || !idx || !wt) { free(acc); free(wacc); free(idx); free(wt); return -2; }
    int flag = 0;
    for (int n = 0; n < b; ++n) {
        memset(acc, 0, sizeof(float) * c * phw);
        memset(wacc, 0, sizeof(float) * phw);
        const float* fl = flow + (size_t)n * 2 * hw;
        for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
            size_t p = (size_t)y * w + x;
            /* trans_pos = flow + grid; offset = trans_pos + 1      :450-452 */
            float ox = (fl[p] + (float)x) + 1.0f;
            float oy = (fl[hw + p] + (float)y) + 1.0f;
            float fxf = floorf(ox), cxf = ceilf(ox), fyf = floorf(oy), cyf = ceilf(oy);
            int ix0 = clamp_index(fxf, w + 1), ix1 = clamp_index(cxf, w + 1);
            int iy0 = clamp_index(fyf, h + 1), iy1 = clamp_index(cyf, h + 1);
            ox = clampf(ox, 0.0f, (float)(w + 1));
            oy = clampf(oy, 0.0f, (float)(h + 1));
            /* proximity weights                                     :477-488 */
            float ax = 1.0f - (ox - (float)ix0), bx = 1.0f - ((float)ix1 - ox);
            float ay = 1.0f - (oy - (float)iy0), by = 1.0f - ((float)iy1 - oy);
            float ld = logf(1.0f + clampf(depth[(size_t)n * hw + p], 0.0f, 1000.0f));
            float dw = expf(ld / L * 50.0f);
            float m = mask1 ? mask1[(size_t)n * hw + p] : 1.0f;
            float fm = flow_mask ? flow_mask[(size_t)n * hw + p] : 1.0f;
            /* weight = prox * mask1 * flow12_mask / depth_weights   :494-513 */
            wt[4 * p + 0] = ((ay * ax) * m) * fm / dw;
            wt[4 * p + 1] = ((by * ax) * m) * fm / dw;
            wt[4 * p + 2] = ((ay * bx) * m) * fm / dw;
            wt[4 * p + 3] = ((by * bx) * m) * fm / dw;
            idx[4 * p + 0] = ix0; idx[4 * p + 1] = ix1; idx[4 * p + 2] = iy0; idx[4 * p + 3] = iy1;
        }
        /* index_put_(accumulate=True), serial (deterministic mode), in the order
         * NW(fy,fx) SW(cy,fx) NE(fy,cx) SE(cy,cx), raster order inside each     :524-564 */
        for (int k = 0; k < 4; ++k) {
            for (size_t p = 0; p < hw; ++p) {
                int cx = (k & 2) ? idx[4 * p + 1] : idx[4 * p + 0];
                int cy = (k & 1) ? idx[4 * p + 3] : idx[4 * p + 2];
                size_t cell = (size_t)cy * pw + cx;
                float wk = wt[4 * p + k];
                for (int ch = 0; ch < c; ++ch)
                    acc[(size_t)ch * phw + cell] += frame[((size_t)n * c + ch) * hw + p] * wk;
            }
        }
        for (int k = 0; k < 4; ++k) {
            for (size_t p = 0; p < hw; ++p) {
                int cx = (k & 2) ? idx[4 * p + 1] : idx[4 * p + 0];
                int cy = (k & 1) ? idx[4 * p + 3] : idx[4 * p + 2];
                wacc[(size_t)cy * pw + cx] += wt[4 * p + k];
            }
        }
        /* crop, mask, normalise, clip                                            :566-583 */
        for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
            size_t cell = (size_t)(y + 1) * pw + (x + 1);
            size_t p = (size_t)y * w + x;
            float W = wacc[cell];
            int valid = W > 0.0f;
            out_mask[(size_t)n * hw + p] = valid ? 1.0f : 0.0f;
            for (int ch = 0; ch < c; ++ch) {
                float v = valid ? acc[(size_t)ch * phw + cell] / W : (is_image ? -1.0f : 0.0f);
                if (is_image) {
                    if (!(v >= -1.1f) || !(v <= 1.1f)) flag = 1;   /* the reference asserts here */
                    v = clampf(v, -1.0f, 1.0f);
                }
                out[((size_t)n * c + ch) * hw + p] = v;
            }
        }
    }
    if (range_flag) *range_flag = flag;
    free(acc); free(wacc); free(idx); free(wt);
    return 0;
}

/*
 * models/utils.py:220-293 (twice=False).  per_unit_max != 0 : n_units independent b=1 calls
 * (what every production caller does); == 0 : one call with b = n_units (global log-depth max).
 * frame (n_src,3,h,w) mask1 (n_src,h,w)|NULL depth (n_src,h,w).
 */
ORC_EXPORT int orc_forward_warp(const float* frame, const float* mask1, const float* depth,
                                const int* src_index, const float* M,
                                int n_units, int n_src, int h, int w, int per_unit_max,
                                float* out_frame, float* out_mask, float* out_depth,
                                float* out_flow, int* range_flag) {
    const size_t hw = (size_t)h * w;
    float* pts = (float*)malloc(sizeof(float) * 3 * hw * n_units);
    float* zs = (float*)malloc(sizeof(float) * hw * n_units);
    float* flow = out_flow ? out_flow : (float*)malloc(sizeof(float) * 2 * hw * n_units);
    float* gf = (float*)malloc(sizeof(float) * 3 * hw * n_units);
    float* gm = mask1 ? (float*)malloc(sizeof(float) * hw * n_units) : NULL;
    float* dmask = (float*)malloc(sizeof(float) * hw * n_units);
    float* dtmp = out_depth ? out_depth : (float*)malloc(sizeof(float) * hw * n_units);
    if (!pts || !zs || !flow || !gf || !dmask || !dtmp || (mask1 && !gm)) return -2;
    int rc = orc_project(depth, src_index, M, n_units, n_src, h, w, pts);
    if (rc) return rc;
    for (int u = 0; u < n_units; ++u) {
        int s = src_index ? src_index[u] : u;
        memcpy(gf + (size_t)u * 3 * hw, frame + (size_t)s * 3 * hw, sizeof(float) * 3 * hw);
        if (mask1) memcpy(gm + (size_t)u * hw, mask1 + (size_t)s * hw, sizeof(float) * hw);
        for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
            size_t p = (size_t)y * w + x;
            const float* P = pts + ((size_t)u * hw + p) * 3;
            /* coords = P.xy / P.z ; trans_depth = P.z ; flow12 = coords - grid   :273-278 */
            float cx = P[0] / P[2], cy = P[1] / P[2];
            zs[(size_t)u * hw + p] = P[2];
            flow[(size_t)u * 2 * hw + p] = cx - (float)x;
            flow[(size_t)u * 2 * hw + hw + p] = cy - (float)y;
        }
    }
    int flag = 0, f1 = 0;
    if (per_unit_max) {
        for (int u = 0; u < n_units && !rc; ++u) {
            rc = orc_splat(gf + (size_t)u * 3 * hw, gm ? gm + (size_t)u * hw : NULL, zs + (size_t)u * hw,
                           flow + (size_t)u * 2 * hw, NULL, 1, 3, h, w, 1,
                           out_frame + (size_t)u * 3 * hw, out_mask + (size_t)u * hw, &f1);
            flag |= f1;
            if (!rc) rc = orc_splat(zs + (size_t)u * hw, gm ? gm + (size_t)u * hw : NULL, zs + (size_t)u * hw,
                           flow + (size_t)u * 2 * hw, NULL, 1, 1, h, w, 0,
                           dtmp + (size_t)u * hw, dmask + (size_t)u * hw, NULL);
        }
    } else {
        rc = orc_splat(gf, gm, zs, flow, NULL, n_units, 3, h, w, 1, out_frame, out_mask, &flag);
        if (!rc) rc = orc_splat(zs, gm, zs, flow, NULL, n_units, 1, h, w, 0, dtmp, dmask, NULL);
    }
    if (range_flag) *range_flag = flag;
    free(pts); free(zs); free(gf); free(gm); free(dmask);
    if (!out_flow) free(flow);
    if (!out_depth) free(dtmp);
    return rc;
}

/*
 * models/utils.py:585-626 for one (1,c,h,w) image.  cv2.dilate(5x5 ones, iterations=1) with its
 * default border (outside pixels never win the max).  Outputs are float64 like the reference's.
 */
ORC_EXPORT int orc_clean_points(const float* frame, const float* mask2, int c, int h, int w,
                                double* out_frame, double* out_mask) {
    const size_t hw = (size_t)h * w;
    unsigned char* hole = (unsigned char*)malloc(hw);
    if (!hole) return -2;
    for (size_t p = 0; p < hw; ++p) hole[p] = (1.0f - mask2[p]) >= 0.5f;   /* :589-591 */
    for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
        int m = 0;
        for (int dy = -2; dy <= 2 && !m; ++dy) {
            int yy = y + dy; if (yy < 0 || yy >= h) continue;
            for (int dx = -2; dx <= 2; ++dx) {
                int xx = x + dx; if (xx < 0 || xx >= w) continue;
                if (hole[(size_t)yy * w + xx]) { m = 1; break; }
            }
        }
        size_t p = (size_t)y * w + x;
        out_mask[p] = 1.0 - (double)m;                                      /* :626 */
        for (int ch = 0; ch < c; ++ch) {
            float half = (frame[(size_t)ch * hw + p] + 1.0f) / 2.0f;        /* :588 (f32) */
            out_frame[(size_t)ch * hw + p] = (double)half * (1.0 - (double)m) * 2.0 - 1.0;  /* :625-626 (f64) */
        }
    }
    free(hole);
    return 0;
}

/*
 * Number of scatter contributions each output pixel receives (4 corners per source pixel, the same
 * indices as orc_splat; models/utils.py:450-475, 524-564).  Test infrastructure for the parity
 * tolerance: the reference accumulates sequentially in fp32, so a cell with n contributors carries
 * a rounding error of up to ~n*2^-24 relative in the reference itself; tests widen the 1e-5
 * tolerance by that bound for cells with n > 64 (fold lines, the behind-camera pixel (1,1)).
 */
ORC_EXPORT int orc_contrib_count(const float* flow, int b, int h, int w, int* out) {
    const size_t hw = (size_t)h * w;
    memset(out, 0, sizeof(int) * hw * b);
    for (int n = 0; n < b; ++n) {
        const float* fl = flow + (size_t)n * 2 * hw;
        for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
            size_t p = (size_t)y * w + x;
            float ox = (fl[p] + (float)x) + 1.0f;
            float oy = (fl[hw + p] + (float)y) + 1.0f;
            int ix[2] = {clamp_index(floorf(ox), w + 1), clamp_index(ceilf(ox), w + 1)};
            int iy[2] = {clamp_index(floorf(oy), h + 1), clamp_index(ceilf(oy), h + 1)};
            for (int k = 0; k < 4; ++k) {
                int cx = ix[k >> 1], cy = iy[k & 1];
                if (cx >= 1 && cx <= w && cy >= 1 && cy <= h) out[(size_t)n * hw + (size_t)(cy - 1) * w + (cx - 1)]++;
            }
        }
    }
    return 0;
}

/*
 * The step right after the warp loop (reference demo.py:91-100, 117-120):
 *     cond_video = F.interpolate((cat(warped) + 1) / 2, size, mode='bilinear', align_corners=False)
 *     cond_masks = F.interpolate(cat(masks), size, mode='nearest')
 *     cond_video = cond_video.permute(1,0,2,3)[None]                  -> (1,3,F,h,w)
 *     cond_masks = (1 - cond_masks.permute(1,0,2,3)[None]) * 255      -> (1,1,F,h,w)
 * The arithmetic lives in torch (ATen UpSampleKernel.cpp, not vendored by the reference): restated
 * from its published algorithm --
 *   bilinear, align_corners=False, no scale_factor: scale = (float)in / out;
 *     src = max(fma(scale, dst + 0.5f, -0.5f), 0);  i0 = min((int)src, in - 1);  l1 = clamp(src - i0, 0, 1);
 *       (torch 2.11 CPU evaluates the source index with ONE rounding -- probed in this container: the two-rounding
 *        form moves the weights by up to ulp(src) and misses the goldens by 1.6e-6, the fused form by 1.2e-7)
 *     i1 = i0 + (i0 < in - 1);  l0 = 1 - l1;   (in == out: i0 = i1 = dst, l0 = 1, l1 = 0)
 *     out = (a*lx0 + b*lx1) * ly0 + (c*lx0 + d*lx1) * ly1
 *   nearest (legacy): src = min((int)floorf(dst * scale), in - 1)   (in == out: dst; out == 2 in: dst >> 1)
 * ATen's vectorised kernel may contract products into FMAs, so parity of the colour is 1e-6 absolute
 * (values in [0,1]); the packed mask is exact.  Pinned by tests/golden/pack_cases.npz (oracle/gen_golden_pack.py).
 */
static void lin_index(int dst, int in, int out, float scale, int* i0, int* i1, float* l0, float* l1) {
    if (in == out) { *i0 = *i1 = dst; *l0 = 1.0f; *l1 = 0.0f; return; }
    float src = fmaf(scale, (float)dst + 0.5f, -0.5f);   /* one rounding: what the ATen build pinned by the goldens does */
    if (src < 0.0f) src = 0.0f;
    int k = (int)floorf(src);
    if (k > in - 1) k = in - 1;
    float l = src - (float)k;
    l = l < 0.0f ? 0.0f : (l > 1.0f ? 1.0f : l);
    *i0 = k; *i1 = k + (k < in - 1 ? 1 : 0); *l1 = l; *l0 = 1.0f - l;
}
static int near_index(int dst, int in, int out, float scale) {
    if (in == out) return dst;
    if (out == 2 * in) return dst >> 1;
    int k = (int)floorf((float)dst * scale);
    return k < in - 1 ? k : in - 1;
}
ORC_EXPORT int orc_resize_pack(const float* frame, const float* mask, int n, int H, int W, int h, int w,
                               float* out_video, float* out_masks) {
    if (n < 0 || H <= 0 || W <= 0 || h <= 0 || w <= 0) return -1;
    const size_t HW = (size_t)H * W, hw = (size_t)h * w;
    const float sy = (float)H / (float)h, sx = (float)W / (float)w;
    for (int f = 0; f < n; ++f) for (int y = 0; y < h; ++y) {
        int y0, y1; float ly0, ly1;
        lin_index(y, H, h, sy, &y0, &y1, &ly0, &ly1);
        const int yn = near_index(y, H, h, sy);
        for (int x = 0; x < w; ++x) {
            int x0, x1; float lx0, lx1;
            lin_index(x, W, w, sx, &x0, &x1, &lx0, &lx1);
            for (int c = 0; c < 3; ++c) {
                const float* s = frame + ((size_t)f * 3 + c) * HW;
                const float a = (s[(size_t)y0 * W + x0] + 1.0f) / 2.0f, b = (s[(size_t)y0 * W + x1] + 1.0f) / 2.0f;
                const float cc = (s[(size_t)y1 * W + x0] + 1.0f) / 2.0f, d = (s[(size_t)y1 * W + x1] + 1.0f) / 2.0f;
                const float t0 = a * lx0 + b * lx1, t1 = cc * lx0 + d * lx1;
                out_video[((size_t)c * n + f) * hw + (size_t)y * w + x] = t0 * ly0 + t1 * ly1;
            }
            const float m = mask[(size_t)f * HW + (size_t)yn * W + near_index(x, W, w, sx)];
            out_masks[(size_t)f * hw + (size_t)y * w + x] = (1.0f - m) * 255.0f;
        }
    }
    return 0;
}
