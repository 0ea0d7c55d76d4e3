// depth_post.cuh — the step right BEFORE the warp: DepthCrafter's raw output -> metric-like depth, and
// the centre-pixel radius that scales the trajectory, without the depth map ever leaving the device.
//
// Reference: models/infer.py:79-92 (numpy / torch lines at the end of DepthCrafterDemo.infer)
//     res    = res.sum(-1) / res.shape[-1]              # (n,h,w,3) -> (n,h,w), float32: ((a + b) + c) / 3
//     depths = (res - res.min()) / (res.max() - res.min())
//     depths *= 3900 ; depths[depths < 1e-5] = 1e-5 ; depths = 10000.0 / depths ; depths = depths.clip(near, far)
// and demo.py:539-543   radius = min(depths[0, 0, h // 2, w // 2] * radius_scale, 5).
// Every operation is one correctly rounded float32 operation in the reference, so the kernels below reproduce the result bit
// for bit (__fadd_rn / __fmul_rn / __fdiv_rn, never contracted).
#pragma once

__device__ __forceinline__ unsigned dp_encode(float f) {          // order-preserving float -> uint
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float dp_decode(unsigned u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ float dp_mean(const float* __restrict__ p, int c) {
    // numpy adds the channels of one pixel left to right in float32, then divides by the channel count
    float s = p[0];
    for (int k = 1; k < c; ++k) s = __fadd_rn(s, p[k]);
    return c > 1 ? __fdiv_rn(s, (float)c) : s;
}

// pass 1: channel mean of every pixel, global min / max (order-preserving uint atomics: the result does not depend on order)
__global__ void __launch_bounds__(256)
k_depth_minmax(const float* __restrict__ res, size_t n_px, int c, unsigned* __restrict__ mm) {
    float lo = __int_as_float(0x7f800000), hi = -lo;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_px; i += (size_t)gridDim.x * blockDim.x) {
        const float v = dp_mean(res + i * c, c);
        lo = fminf(lo, v);
        hi = fmaxf(hi, v);
    }
    unsigned ulo = dp_encode(lo), uhi = dp_encode(hi);
    ulo = __reduce_min_sync(0xffffffffu, ulo);
    uhi = __reduce_max_sync(0xffffffffu, uhi);
    if ((threadIdx.x & 31) == 0) {
        atomicMin(mm + 0, ulo);
        atomicMax(mm + 1, uhi);
    }
}

__device__ __forceinline__ float dp_post(float v, float lo, float span, float nearv, float farv) {
    float d = __fdiv_rn(__fsub_rn(v, lo), span);          // (res - min) / (max - min)
    d = __fmul_rn(d, 3900.0f);                            // "compatible with da output"
    if (d < 1e-5f) d = 1e-5f;
    d = __fmul_rn(__frcp_rn(d), 10000.0f);                // torch: `10000.0 / t` == t.reciprocal() * 10000.0 (two roundings)
    return fminf(fmaxf(d, nearv), farv);                  // .clip(near, far)
}

// pass 2: the post-processing itself
__global__ void __launch_bounds__(256)
k_depth_post(const float* __restrict__ res, size_t n_px, int c, const unsigned* __restrict__ mm, float nearv, float farv,
             float* __restrict__ out) {
    const float lo = dp_decode(mm[0]), hi = dp_decode(mm[1]);
    const float span = __fsub_rn(hi, lo);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_px; i += (size_t)gridDim.x * blockDim.x)
        out[i] = dp_post(dp_mean(res + i * c, c), lo, span, nearv, farv);
}

// stats[0] = min, stats[1] = max (decoded in place), stats[2] = depth[0, 0, h/2, w/2], stats[3] = min(stats[2] * radius_scale, 5)
__global__ void k_depth_stats(unsigned* __restrict__ mm, const float* __restrict__ depth, int h, int w, float radius_scale) {
    if (threadIdx.x || blockIdx.x) return;
    float* st = (float*)mm;
    const float lo = dp_decode(mm[0]), hi = dp_decode(mm[1]);
    const float centre = depth[(size_t)(h / 2) * w + (w / 2)];
    st[0] = lo; st[1] = hi; st[2] = centre;
    st[3] = fminf(__fmul_rn(centre, radius_scale), 5.0f);
}
