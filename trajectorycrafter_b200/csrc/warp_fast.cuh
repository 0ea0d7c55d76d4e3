// warp_fast.cuh — the fp32 fast path of forward_warp (included by tcwarp.cu after warp_kernels.cuh).
//
//   K0 k_zmax_quad      soft-z normaliser pre-pass: 4 px per thread, only the z' row of the projection
//   (K2, the fused project + splat kernel, lives in warp_strip.cuh)
//   K3 k_normalise_quad normalise / mask / clip, 4 px per thread, re-zeroes the accumulators it read
//
// Reference: models/utils.py:350-420 (projection), :422-583 (splat), :566-583 (normalise).
#pragma once
#include <limits.h>


// ------------------------------------------------------------------------------------------------
// K0: per-unit max of log(1+sat(z')) and of |z'|.  Only P.z is needed; when K2's last row is
// (0, 0, k) the reference's dot product ((0 + 0*Y0) + 0*Y1) + k*Y2 equals k*Y2 (finite Y), so Y0/Y1
// are skipped.  models/utils.py:388-417, 490-492.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float project_z(const float* M, bool simple, float fx, float fy, float d) {
    const float u0 = dot3(M + 0, fx, fy, 1.0f);
    const float u1 = dot3(M + 3, fx, fy, 1.0f);
    const float u2 = dot3(M + 6, fx, fy, 1.0f);
    const float X0 = __fmul_rn(d, u0), X1 = __fmul_rn(d, u1), X2 = __fmul_rn(d, u2);
    const float Y2 = dot4(M + 17, X0, X1, X2, 1.0f);
    float P2;
    if (simple) {
        P2 = __fadd_rn(0.0f, __fmul_rn(M[29], Y2));
    } else {
        const float Y0 = dot4(M + 9, X0, X1, X2, 1.0f);
        const float Y1 = dot4(M + 13, X0, X1, X2, 1.0f);
        P2 = dot3(M + 27, Y0, Y1, Y2);
    }
    return Y2 <= 0.01f ? 1000.0f : P2;
}

__device__ float log1x_fwd(float a);      // warp_pair.cuh

__global__ void __launch_bounds__(TILE_THREADS)
k_zmax_quad(const float* __restrict__ depth, const int* __restrict__ src_index, const float* __restrict__ Mall,
            int unit0, int n_src, int h, int w, UnitStats* __restrict__ stats, int n_units_total, int poly_log) {
    __shared__ float M[M_FLOATS];
    const int u = unit0 + blockIdx.y;
    if (threadIdx.x < M_FLOATS) M[threadIdx.x] = Mall[(size_t)u * M_FLOATS + threadIdx.x];
    __syncthreads();
    const bool simple = M[27] == 0.0f && M[28] == 0.0f;
    // pinhole K1^-1 (no skew, last row (0,0,1)): u = (m0 x + m2, m4 y + m5, 1) -- the zero products of the reference's
    // dot products add +-0 to a non-zero partial sum, so dropping them changes no bit (same argument as warp_pair.cuh)
    const bool pin = simple && M[1] == 0.0f && M[3] == 0.0f && M[6] == 0.0f && M[7] == 0.0f && M[8] == 1.0f;
    const float m0 = M[0], m2 = M[2], m4 = M[4], m5 = M[5], t8 = M[17], t9 = M[18], t10 = M[19], t11 = M[20], k8 = M[29];
    int s = src_index ? src_index[u] : u;
    if ((unsigned)s >= (unsigned)n_src) s = 0;      // reported by the splat kernel (TCW_STATUS_SRC_INDEX)
    const float* dsrc = depth + (size_t)s * h * w;
    const int wq = (w + 3) >> 2, quads = h * wq;
    const bool vec = (w & 3) == 0 && ((((size_t)dsrc) & 15) == 0);
    float zs = 0.0f, za = 0.0f;      // max of sat(z') (log is monotone: take it once at the end) and of |z'|
    for (int q = blockIdx.x * TILE_THREADS + threadIdx.x; q < quads; q += gridDim.x * TILE_THREADS) {
        const int y = q / wq, x0 = (q - y * wq) << 2;
        float d[4];
        if (vec) {
            const float4 v = __ldg((const float4*)(dsrc + (size_t)y * w + x0));
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        } else {
            #pragma unroll
            for (int i = 0; i < 4; ++i) d[i] = x0 + i < w ? dsrc[(size_t)y * w + x0 + i] : 0.0f;
        }
        if (pin) {
            const float u1 = __fadd_rn(__fmul_rn(m4, (float)y), m5);
            #pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (x0 + i < w) {
                    const float u0 = __fadd_rn(__fmul_rn(m0, (float)(x0 + i)), m2);
                    const float X0 = __fmul_rn(d[i], u0), X1 = __fmul_rn(d[i], u1);
                    const float Y2 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(t8, X0), __fmul_rn(t9, X1)), __fmul_rn(t10, d[i])), t11);
                    const float P2 = Y2 <= 0.01f ? 1000.0f : __fadd_rn(0.0f, __fmul_rn(k8, Y2));
                    zs = fmaxf(zs, clampf(P2, 0.0f, 1000.0f));
                    za = fmaxf(za, fabsf(P2));
                }
            }
        } else {
            #pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (x0 + i < w) {
                    const float P2 = project_z(M, simple, (float)(x0 + i), (float)y, d[i]);
                    zs = fmaxf(zs, clampf(P2, 0.0f, 1000.0f));
                    za = fmaxf(za, fabsf(P2));
                }
            }
        }
    }
    // == max over pixels of log(1+sat(z')): log is monotone.  poly_log: the same polynomial the v4 splat kernel
    // evaluates per pixel (warp_pair.cuh log2x), so that ld / L is exactly 1 at the arg-max pixel
    const float ld = poly_log ? log1x_fwd(__fadd_rn(1.0f, zs)) : logf(__fadd_rn(1.0f, zs));
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

// ------------------------------------------------------------------------------------------------
// K3: normalise / mask / clip, 4 px per thread (models/utils.py:566-583).  Reads the five fp32 sums,
// writes colour/mask/depth with streaming stores, and re-zeroes what it read so the next chunk of
// units finds clean accumulators without a separate memset.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void normalise_px(float sr, float sg, float sb, float sz, float W,
                                             float& cr, float& cg, float& cb, float& dep, float& mk, bool& bad) {
    const bool valid = W > 0.0f;
    cr = valid ? __fdiv_rn(sr, W) : -1.0f;
    cg = valid ? __fdiv_rn(sg, W) : -1.0f;
    cb = valid ? __fdiv_rn(sb, W) : -1.0f;
    dep = valid ? __fdiv_rn(sz, W) : 0.0f;
    mk = valid ? 1.0f : 0.0f;
    bad |= !(cr >= -1.1f) || !(cr <= 1.1f) || !(cg >= -1.1f) || !(cg <= 1.1f) || !(cb >= -1.1f) || !(cb <= 1.1f);
    cr = clampf(cr, -1.0f, 1.0f); cg = clampf(cg, -1.0f, 1.0f); cb = clampf(cb, -1.0f, 1.0f);
}

// quads [q, q+1) (VEC: 4 px each) or pixels (scalar) of one unit: read sums (L2, bypassing L1), re-zero, write outputs
template <bool VEC>
__device__ __forceinline__ void normalise_item(float4* __restrict__ a4, float* __restrict__ aw, size_t hw, size_t i,
                                               float* __restrict__ of, float* __restrict__ om, float* __restrict__ od, bool& bad) {
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    if (VEC) {
        const size_t q = i;                                        // quad index, hw % 4 == 0
        const float4 W = __ldcg((const float4*)aw + q);
        const float4 S0 = __ldcg(a4 + 4 * q + 0), S1 = __ldcg(a4 + 4 * q + 1), S2 = __ldcg(a4 + 4 * q + 2), S3 = __ldcg(a4 + 4 * q + 3);
        *((float4*)aw + q) = zero;
        a4[4 * q + 0] = zero; a4[4 * q + 1] = zero; a4[4 * q + 2] = zero; a4[4 * q + 3] = zero;
        float4 R, G, B, D, K;
        normalise_px(S0.x, S0.y, S0.z, S0.w, W.x, R.x, G.x, B.x, D.x, K.x, bad);
        normalise_px(S1.x, S1.y, S1.z, S1.w, W.y, R.y, G.y, B.y, D.y, K.y, bad);
        normalise_px(S2.x, S2.y, S2.z, S2.w, W.z, R.z, G.z, B.z, D.z, K.z, bad);
        normalise_px(S3.x, S3.y, S3.z, S3.w, W.w, R.w, G.w, B.w, D.w, K.w, bad);
        __stcs((float4*)(of + 0 * hw) + q, R);
        __stcs((float4*)(of + 1 * hw) + q, G);
        __stcs((float4*)(of + 2 * hw) + q, B);
        __stcs((float4*)om + q, K);
        if (od) __stcs((float4*)od + q, D);
    } else {
        const size_t p = i;
        const float4 S = __ldcg(a4 + p);
        float cr, cg, cb, dep, mk;
        normalise_px(S.x, S.y, S.z, S.w, __ldcg(aw + p), cr, cg, cb, dep, mk, bad);
        a4[p] = zero;
        aw[p] = 0.0f;
        of[0 * hw + p] = cr; of[1 * hw + p] = cg; of[2 * hw + p] = cb;
        om[p] = mk;
        if (od) od[p] = dep;
    }
}

template <bool VEC>
__global__ void __launch_bounds__(TILE_THREADS)
k_normalise_quad(float4* __restrict__ acc4, float* __restrict__ accw, int unit0, size_t hw,
                 float* __restrict__ out_frame, float* __restrict__ out_mask, float* __restrict__ out_depth,
                 int* __restrict__ range_flag) {
    const int uc = blockIdx.y, u = unit0 + uc;
    const size_t i = (size_t)blockIdx.x * TILE_THREADS + threadIdx.x;
    if (i >= (VEC ? hw / 4 : hw)) return;
    bool bad = false;
    normalise_item<VEC>(acc4 + (size_t)uc * hw, accw + (size_t)uc * hw, hw, i, out_frame + (size_t)u * 3 * hw,
                        out_mask + (size_t)u * hw, out_depth ? out_depth + (size_t)u * hw : nullptr, bad);
    if (bad && range_flag) atomicOr(range_flag, 1);
}
