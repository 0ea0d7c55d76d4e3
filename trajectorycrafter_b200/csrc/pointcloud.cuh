// pointcloud.cuh — SURVEY.md 8f N2: the fork's global point-cloud renderer (reference
// notebooks/06_10_25_vggt/warper_point_cloud.py, class GlobalPointCloudWarper) as CUDA kernels.
//
//   k_pc_lift       lift_to_3d_pointcloud :12-77     pixels + depth -> world points (b,h,w,3) and colours (b,h,w,3)
//   k_pc_zbuffer    render_pointcloud_zbuffer_vectorized_point_size :522-646, first half: project every point, expand
//                   its point_size x point_size footprint, keep the nearest entry per pixel
//   k_pc_resolve    second half: winner -> colour / mask / depth
//   k_pc_range, k_pc_bilinear, k_pc_normalise      render_from_camera :158-256 + _splat_points_bilinear_efficient :351-443
//   k_pc_disc       render_from_camera(use_bilinear=False): _splat_points_to_image_with_size :258-349
//   (k_pc_zbuffer with first_wins: render_pointcloud_zbuffer_vectorized_fixed :650-765)
//
// Arithmetic.  The reference's per-pixel products are broadcast matmuls of (3x3)(3x1) / (4x4)(4x1) blocks: sequential un-fused
// fp32 like the main path (models/utils.py:388-399).  Its per-POINT products are plain GEMMs `T @ points.T` (4 x N) and
// `K @ cam.T` (3 x N); torch's CPU GEMM evaluates each element as a chain of fused multiply-adds in index order starting
// from 0 (probed on torch 2.11: 100 % of 1 M points bit-equal, tests/golden/pointcloud_cases.npz pins it), so those use
// fmaf.  Pixel coordinates are IEEE divisions, rounding is torch.round (half to even = rintf).
//
// Z-buffer.  One 64-bit atomicMin per footprint entry on key = depth bits << 32 | (0xffffffff - entry index): the smallest
// depth wins, among equal depths the LAST entry in the reference's (point, offset) order -- what the reference's
// sequential `color_buffer[:, winner_indices] = ...` leaves on the CPU.  The result is independent of the order in which
// the atomics arrive.
#pragma once

struct PcCamera { float T[12]; float K[9]; };      // rows 0..2 of the extrinsic, the intrinsic

// ---- lift_to_3d_pointcloud (:34-77): u = K^-1 (x, y, 1); cam = d * u; world = T^-1 (cam, 1); both times mask1
__global__ void __launch_bounds__(256)
k_pc_lift(const float* __restrict__ frame, const float* __restrict__ mask1, const float* __restrict__ depth,
          const float* __restrict__ Mall /* per image: K^-1 (9) | rows 0..2 of T^-1 (12) */, int h, int w, int step,
          float* __restrict__ points, float* __restrict__ colors) {
    const int bi = blockIdx.y;
    const size_t hw = (size_t)h * w;
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= hw) return;
    const float* M = Mall + (size_t)bi * 21;
    const int y = (int)(p / w), x = (int)(p - (size_t)y * w);
    const float fx = (float)(x * step), fy = (float)(y * step);      // step > 1: the inputs are a ::step subsample (core.py:30-45)
    const float u0 = dot3(M + 0, fx, fy, 1.0f), u1 = dot3(M + 3, fx, fy, 1.0f), u2 = dot3(M + 6, fx, fy, 1.0f);
    const float d = depth[(size_t)bi * hw + p];
    const float X0 = __fmul_rn(d, u0), X1 = __fmul_rn(d, u1), X2 = __fmul_rn(d, u2);
    const float m = mask1 ? mask1[(size_t)bi * hw + p] : 1.0f;
    float* o = points + ((size_t)bi * hw + p) * 3;
    o[0] = __fmul_rn(dot4(M + 9, X0, X1, X2, 1.0f), m);
    o[1] = __fmul_rn(dot4(M + 13, X0, X1, X2, 1.0f), m);
    o[2] = __fmul_rn(dot4(M + 17, X0, X1, X2, 1.0f), m);
    float* c = colors + ((size_t)bi * hw + p) * 3;
    const float* f = frame + (size_t)bi * 3 * hw + p;
    c[0] = __fmul_rn(f[0], m); c[1] = __fmul_rn(f[hw], m); c[2] = __fmul_rn(f[2 * hw], m);
}

// camera-space point and pixel coordinates of one world point, the GEMM way (fma chains from 0, index order)
__device__ __forceinline__ void pc_project(const PcCamera& C, const float* __restrict__ pt, float& cx, float& cy, float& cz,
                                           float& px, float& py, float& pz) {
    const float X = pt[0], Y = pt[1], Z = pt[2];
    cx = fmaf(C.T[3], 1.0f, fmaf(C.T[2], Z, fmaf(C.T[1], Y, fmaf(C.T[0], X, 0.0f))));
    cy = fmaf(C.T[7], 1.0f, fmaf(C.T[6], Z, fmaf(C.T[5], Y, fmaf(C.T[4], X, 0.0f))));
    cz = fmaf(C.T[11], 1.0f, fmaf(C.T[10], Z, fmaf(C.T[9], Y, fmaf(C.T[8], X, 0.0f))));
    const float q0 = fmaf(C.K[2], cz, fmaf(C.K[1], cy, fmaf(C.K[0], cx, 0.0f)));
    const float q1 = fmaf(C.K[5], cz, fmaf(C.K[4], cy, fmaf(C.K[3], cx, 0.0f)));
    pz = fmaf(C.K[8], cz, fmaf(C.K[7], cy, fmaf(C.K[6], cx, 0.0f)));
    px = __fdiv_rn(q0, pz);
    py = __fdiv_rn(q1, pz);
}

constexpr unsigned long long PC_EMPTY = 0xffffffffffffffffull;

__global__ void __launch_bounds__(256)
k_pc_zbuffer(const float* __restrict__ points, PcCamera C, long long n_points, int h, int w, int point_size,
             int first_wins, unsigned long long* __restrict__ zbuf) {
    // first_wins: render_pointcloud_zbuffer_vectorized_fixed (:650-765), a sequential loop with a strict `<`: among equal depths
    // the FIRST entry stays.  point_size 0: the single-pixel variant render_pointcloud_zbuffer_vectorized (:445-519), which CLAMPS the rounded pixel
    // into the image (:483-484) instead of dropping it
    const bool clamp_mode = point_size == 0;
    const int radius = point_size / 2, side = 2 * radius + 1, n_splats = side * side;
    const float wf = (float)w, hf = (float)h;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_points; i += (long long)gridDim.x * blockDim.x) {
        float cx, cy, cz, px, py, pz;
        pc_project(C, points + i * 3, cx, cy, cz, px, py, pz);
        // :545-548: depth = camera z; valid = depth > 0.01 and the un-rounded pixel inside the image
        if (!(cz > 0.01f && px >= 0.0f && px < wf && py >= 0.0f && py < hf)) continue;
        const unsigned long long hi = (unsigned long long)__float_as_uint(cz) << 32;
        int j = 0;
        for (int dy = -radius; dy <= radius; ++dy) {
            for (int dx = -radius; dx <= radius; ++dx, ++j) {
                // :592-600: coords + offsets in float32, torch.round, bounds
                float xr = rintf(__fadd_rn(px, (float)dx)), yr = rintf(__fadd_rn(py, (float)dy));
                if (clamp_mode) { xr = fminf(fmaxf(xr, 0.0f), wf - 1.0f); yr = fminf(fmaxf(yr, 0.0f), hf - 1.0f); }
                if (!(xr >= 0.0f && xr < wf && yr >= 0.0f && yr < hf)) continue;
                const unsigned entry = (unsigned)(i * n_splats + j);
                atomicMin(zbuf + (size_t)((int)yr) * w + (int)xr, hi | (unsigned long long)(first_wins ? entry : 0xffffffffu - entry));
            }
        }
    }
}

// :622-641: colour of the winner (empty pixels -1), mask = depth > 0, depth buffer (empty 0)
__global__ void __launch_bounds__(256)
k_pc_resolve(const unsigned long long* __restrict__ zbuf, const float* __restrict__ colors, int hw, int n_splats, int first_wins,
             float* __restrict__ out_frame, float* __restrict__ out_mask, float* __restrict__ out_depth) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= hw) return;
    const unsigned long long k = zbuf[p];
    float r = -1.0f, g = -1.0f, b = -1.0f, d = 0.0f;
    if (k != PC_EMPTY) {
        const unsigned low = (unsigned)(k & 0xffffffffull);
        const unsigned entry = first_wins ? low : 0xffffffffu - low;
        const float* c = colors + (size_t)(entry / (unsigned)n_splats) * 3;
        r = c[0]; g = c[1]; b = c[2];
        d = __uint_as_float((unsigned)(k >> 32));
    }
    out_frame[p] = r; out_frame[hw + p] = g; out_frame[2 * hw + p] = b;
    out_mask[p] = d > 0.0f ? 1.0f : 0.0f;
    if (out_depth) out_depth[p] = d;
}

// ---- render_from_camera (:158-256) with use_bilinear=True: _splat_points_bilinear_efficient (:351-443)
// pass 1: range of the projected depths (projected z, :226) over the points that survive both filters (:210-236)
__device__ __forceinline__ bool pc_visible(const PcCamera& C, const float* __restrict__ pt, float depth_threshold, float wf, float hf,
                                           float& px, float& py, float& pz) {
    float cx, cy, cz;
    pc_project(C, pt, cx, cy, cz, px, py, pz);
    return cz > 0.1f && cz < depth_threshold && px >= 0.0f && px < wf && py >= 0.0f && py < hf;
}

__global__ void __launch_bounds__(256)
k_pc_range(const float* __restrict__ points, PcCamera C, long long n_points, int h, int w, float depth_threshold,
           unsigned* __restrict__ mm /* [0] min code, [1] max code, [2] count of visible points */) {
    const float wf = (float)w, hf = (float)h;
    unsigned lo = 0xffffffffu, hi = 0u, cnt = 0u;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_points; i += (long long)gridDim.x * blockDim.x) {
        float px, py, pz;
        if (!pc_visible(C, points + i * 3, depth_threshold, wf, hf, px, py, pz)) continue;
        const unsigned code = dp_encode(pz);
        lo = min(lo, code); hi = max(hi, code); cnt++;
    }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0 && cnt) { atomicMin(mm, lo); atomicMax(mm + 1, hi); atomicAdd(mm + 2, cnt); }
}

// pass 2: bilinear scatter of colour * weight and weight into acc (h*w float4 = sum r, sum g, sum b, sum w).
// Float reductions in L2: the order of the additions is not fixed (the reference's own scatter_add_ is sequential on the
// CPU and unordered on CUDA); parity is a tolerance, not bit equality.
__global__ void __launch_bounds__(256)
k_pc_bilinear(const float* __restrict__ points, const float* __restrict__ colors, const float* __restrict__ weights,
              PcCamera C, long long n_points, int h, int w, float depth_threshold, const unsigned* __restrict__ mm,
              float4* __restrict__ acc) {
    const float wf = (float)w, hf = (float)h;
    if (mm[2] == 0u) return;
    const float dmin = dp_decode(mm[0]), dmax = dp_decode(mm[1]);
    const float span = __fadd_rn(__fsub_rn(dmax, dmin), 1e-8f);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_points; i += (long long)gridDim.x * blockDim.x) {
        float px, py, pz;
        if (!pc_visible(C, points + i * 3, depth_threshold, wf, hf, px, py, pz)) continue;
        // :377-379  gentle depth weighting
        const float nd = __fdiv_rn(__fsub_rn(pz, dmin), span);
        const float pw = __fmul_rn(weights[i], expf(__fmul_rn(-nd, 2.0f)));
        // :383-396  corners and bilinear weights, products left to right
        const float fx0 = floorf(px), fy0 = floorf(py);
        const float dx = __fsub_rn(px, fx0), dy = __fsub_rn(py, fy0);
        const float ax = __fsub_rn(1.0f, dx), ay = __fsub_rn(1.0f, dy);
        const int x0 = (int)fx0, y0 = (int)fy0;
        const float wq[4] = {__fmul_rn(__fmul_rn(ax, ay), pw), __fmul_rn(__fmul_rn(dx, ay), pw),
                             __fmul_rn(__fmul_rn(ax, dy), pw), __fmul_rn(__fmul_rn(dx, dy), pw)};
        const float r = colors[i * 3], g = colors[i * 3 + 1], b = colors[i * 3 + 2];
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = x0 + (k & 1), y = y0 + (k >> 1);
            if (x < 0 || x >= w || y < 0 || y >= h || !(wq[k] > 1e-8f)) continue;          // :405-409
            float4* a = acc + (size_t)y * w + x;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                         :: "l"(a), "f"(__fmul_rn(r, wq[k])), "f"(__fmul_rn(g, wq[k])), "f"(__fmul_rn(b, wq[k])), "f"(wq[k]) : "memory");
        }
    }
}

// ---- render_from_camera with use_bilinear=False: _splat_points_to_image_with_size (:258-349).  Every visible point is
// stamped at round(pixel + offset) for the offsets of a disc of diameter point_size, with weight
// point weight * exp(-2 * depth / (max depth + 1e-8)) * Gaussian falloff; the disc's offsets and falloff weights are built on
// the host exactly as the reference builds them (Python double arithmetic, then float32 exp) and arrive as a table.
struct PcDisc { int n; signed char dx[128]; signed char dy[128]; float w[128]; };

__global__ void __launch_bounds__(256)
k_pc_disc(const float* __restrict__ points, const float* __restrict__ colors, const float* __restrict__ weights,
          PcCamera C, PcDisc D, long long n_points, int h, int w, float depth_threshold, const unsigned* __restrict__ mm,
          float4* __restrict__ acc) {
    const float wf = (float)w, hf = (float)h;
    if (mm[2] == 0u) return;
    const float denom = __fadd_rn(dp_decode(mm[1]), 1e-8f);                                  // :277 depths.max() + 1e-8
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_points; i += (long long)gridDim.x * blockDim.x) {
        float px, py, pz;
        if (!pc_visible(C, points + i * 3, depth_threshold, wf, hf, px, py, pz)) continue;
        const float nd = __fdiv_rn(pz, denom);
        const float tw = __fmul_rn(weights[i], expf(__fmul_rn(-nd, 2.0f)));                   // :278-279
        const float r = colors[i * 3], g = colors[i * 3 + 1], b = colors[i * 3 + 2];
        for (int j = 0; j < D.n; ++j) {
            const float xr = rintf(__fadd_rn(px, (float)D.dx[j])), yr = rintf(__fadd_rn(py, (float)D.dy[j]));   // :302, :315
            if (!(xr >= 0.0f && xr < wf && yr >= 0.0f && yr < hf)) continue;
            const float fw = __fmul_rn(tw, D.w[j]);                                           // :312
            float4* a = acc + (size_t)((int)yr) * w + (int)xr;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                         :: "l"(a), "f"(__fmul_rn(r, fw)), "f"(__fmul_rn(g, fw)), "f"(__fmul_rn(b, fw)), "f"(fw) : "memory");
        }
    }
}

// :433-441 / :343-347: colour = sum / weight where weight > threshold (1e-8 bilinear, 1e-6 disc; else the zero the buffer
// started with), mask = that test
__global__ void __launch_bounds__(256)
k_pc_normalise(const float4* __restrict__ acc, int hw, float threshold, float* __restrict__ out_frame, float* __restrict__ out_mask) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= hw) return;
    const float4 a = acc[p];
    const bool v = a.w > threshold;
    out_frame[p] = v ? __fdiv_rn(a.x, a.w) : 0.0f;
    out_frame[hw + p] = v ? __fdiv_rn(a.y, a.w) : 0.0f;
    out_frame[2 * hw + p] = v ? __fdiv_rn(a.z, a.w) : 0.0f;
    out_mask[p] = v ? 1.0f : 0.0f;
}
