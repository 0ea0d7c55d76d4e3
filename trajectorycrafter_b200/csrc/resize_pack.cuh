// resize_pack.cuh — the step right after the warp loop (reference demo.py:91-100, 117-120):
//     cond_video = F.interpolate((cat(warped) + 1) / 2, size, 'bilinear', align_corners=False).permute(1,0,2,3)[None]
//     cond_masks = (1 - F.interpolate(cat(masks), size, 'nearest').permute(1,0,2,3)[None]) * 255
// as ONE pass over the warp's outputs: reads colour (n,3,H,W) + mask (n,1,H,W), writes the two tensors the
// diffusion pipeline consumes, (3,n,h,w) and (1,n,h,w).  HBM-bound: 16 B per source pixel in, 16 B per target
// pixel out; one thread per target pixel, a warp covers 32 consecutive x (128 B stores, taps served by L1).
// Index / weight arithmetic follows ATen's UpSampleKernel (see oracle/warp_oracle.c::orc_resize_pack): the
// source index is evaluated with ONE rounding (fmaf) like the torch build the goldens come from.
#pragma once

struct LinTap { int i0, i1; float l0, l1; };

__device__ __forceinline__ LinTap lin_tap(int dst, int in, int out, float scale) {
    LinTap t;
    if (in == out) { t.i0 = t.i1 = dst; t.l0 = 1.0f; t.l1 = 0.0f; return t; }
    float src = fmaf(scale, (float)dst + 0.5f, -0.5f);
    src = src < 0.0f ? 0.0f : src;
    int k = (int)floorf(src);
    k = min(k, in - 1);
    const float l = fminf(fmaxf(src - (float)k, 0.0f), 1.0f);
    t.i0 = k; t.i1 = k + (k < in - 1 ? 1 : 0); t.l1 = l; t.l0 = 1.0f - l;
    return t;
}
__device__ __forceinline__ int near_tap(int dst, int in, int out, float scale) {
    if (in == out) return dst;
    if (out == 2 * in) return dst >> 1;
    return min((int)floorf(__fmul_rn((float)dst, scale)), in - 1);
}

__global__ void __launch_bounds__(256)
k_resize_pack(const float* __restrict__ frame, const float* __restrict__ mask, int n, int H, int W, int h, int w,
              float sy, float sx, float* __restrict__ out_video, float* __restrict__ out_masks, int n_total, int f0) {
    // the n frames given are frames f0 .. f0+n-1 of an output clip of n_total frames (the fused epilogue of tcw_forward_warp_pack
    // hands over one chunk of units at a time; the standalone entry passes n_total = n, f0 = 0)
    const size_t hw = (size_t)h * w, HW = (size_t)H * W;
    const size_t total = (size_t)n * hw;
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (size_t)gridDim.x * blockDim.x) {
        const int f = (int)(p / hw);
        const unsigned q = (unsigned)(p - (size_t)f * hw);
        const int y = q / (unsigned)w, x = q - y * (unsigned)w;
        const LinTap ty = lin_tap(y, H, h, sy), tx = lin_tap(x, W, w, sx);
        const size_t r0 = (size_t)ty.i0 * W, r1 = (size_t)ty.i1 * W;
        #pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float* s = frame + ((size_t)f * 3 + c) * HW;
            // (x + 1) / 2 per tap first, like the reference (demo.py:91), then the separable blend
            const float a = __fmul_rn(__fadd_rn(__ldg(s + r0 + tx.i0), 1.0f), 0.5f), b = __fmul_rn(__fadd_rn(__ldg(s + r0 + tx.i1), 1.0f), 0.5f);
            const float cc = __fmul_rn(__fadd_rn(__ldg(s + r1 + tx.i0), 1.0f), 0.5f), d = __fmul_rn(__fadd_rn(__ldg(s + r1 + tx.i1), 1.0f), 0.5f);
            const float t0 = __fadd_rn(__fmul_rn(a, tx.l0), __fmul_rn(b, tx.l1));
            const float t1 = __fadd_rn(__fmul_rn(cc, tx.l0), __fmul_rn(d, tx.l1));
            __stcs(out_video + ((size_t)c * n_total + f0 + f) * hw + q, __fadd_rn(__fmul_rn(t0, ty.l0), __fmul_rn(t1, ty.l1)));
        }
        const float m = __ldg(mask + (size_t)f * HW + (size_t)near_tap(y, H, h, sy) * W + near_tap(x, W, w, sx));
        __stcs(out_masks + (size_t)f0 * hw + p, __fmul_rn(__fsub_rn(1.0f, m), 255.0f));
    }
}
