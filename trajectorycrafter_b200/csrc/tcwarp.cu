// tcwarp.cu — sm_100a kernels + C ABI (include/tcwarp.h) for the forward re-rendering path.
//
// Reference behaviour restated here: /root/reference/models/utils.py
//   compute_transformed_points :350-420, forward_warp :220-293, bilinear_splatting :422-583,
//   clean_points :585-626.  See DESIGN.md for the data layout and the per-kernel rooflines.
//
// Numerical contract: the per-pixel projection is evaluated with explicitly
// un-fused, round-to-nearest fp32 operations in the reference's order (__fmul_rn/__fadd_rn/
// __fdiv_rn are never contracted into FMAs), so flow12 and the integer corner indices are
// bit-identical to the reference's CPU torch result.  This TU is also compiled with -fmad=false.
#include "tcwarp.h"

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define TCW_API extern "C" __attribute__((visibility("default")))

#ifndef TCW_CONST_M
#define TCW_CONST_M 0         // see warp_pair.cuh
#endif
#if TCW_CONST_M
#include <atomic>
struct tcw_ctx;
static std::atomic<tcw_ctx*> g_const_owner{nullptr};      // the one context that may use the __constant__ matrix bank
#endif

// ----------------------------------------------------------------------------------------------
// errors / context
// ----------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, const char* a = "", const char* b = "") {
    snprintf(g_err, sizeof(g_err), fmt, a, b);
    return code;
}
#define CUDA_TRY(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess) return fail(TCW_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); \
    } while (0)

struct tcw_ctx {
    int device;
    int sm_count;
    size_t l2_bytes;
    long long launches;
    int chunk_units_override;   // TCW_CHUNK_UNITS env (0 = auto)
    int pipeline_off;           // TCW_NO_PIPELINE env: use the per-chunk launches for the fp32 path too
    int pipe_lag_override;      // TCW_PIPE_LAG env (0 = auto)
    int pipe_blocks;            // resident CTAs of k_warp_pipeline on this device (0 = not queried yet)
    int pipe2_blocks;           // resident CTAs of k_warp_pipeline2 (0 = not queried yet)
    int use_v3;                 // TCW_V3 env: keep the one-pixel-per-lane kernels (warp_strip.cuh) for A/B runs
    // resources of the host-buffer entry point (grow-only)
    cudaStream_t s_in, s_run, s_out;
    cudaStream_t s_aux;         // second stream of the two-stream fp32 path (normalise of chunk k under splat of chunk k+1)
    cudaEvent_t ev_splat[2], ev_norm[2];
    cudaStream_t s_lane[7];     // extra streams of the multi-lane schedule (lane 0 is the caller's stream)
    cudaEvent_t ev_lane[7];
    int lanes;                  // TCW_LANES env (0 = default)
    // CUDA graph of one whole tcw_forward_warp call (replayed while the arguments stay the same)
    int graphs_off;             // TCW_NO_GRAPH env
    struct GraphKey {
        const void *frame, *mask1, *depth, *src_index, *M, *out_frame, *out_mask, *out_depth, *out_flow, *range_flag, *workspace;
        size_t workspace_bytes;
        int n_units, n_src, h, w, flags;
    } last_key, graph_key;
    int last_key_valid;
    cudaGraphExec_t graph_exec;
    long long graph_launches;   // kernels one replay launches
    cudaStream_t s_cap;         // origin stream of the capture (the caller's stream may be the legacy default stream)
    int serial;                 // TCW_SERIAL env: schedule of the chunked fp32 path (see tcw_create)
    int persist;                // TCW_PERSIST env: use the single persistent kernel (warp_pipeline2.cuh) instead
    void* pool;
    size_t pool_bytes;
    // optional per-kernel timing (tcw_profile_*): event pairs around every launch
    int prof_on;
    struct ProfRec { int kernel; double units; cudaEvent_t e0, e1; }* prof;
    int prof_n, prof_cap;
};

enum { PK_ZMAX = 0, PK_SPLAT = 1, PK_NORMALISE = 2, PK_WMAX = 3, PK_COUNT = 4 };

static void prof_begin(tcw_ctx* ctx, int kernel, double units, cudaStream_t st) {
    if (!ctx->prof_on) return;
    if (ctx->prof_n == ctx->prof_cap) {
        int cap = ctx->prof_cap ? ctx->prof_cap * 2 : 1024;
        if (cap > (1 << 17)) { ctx->prof_on = 2; return; }      // cap reached: stop recording
        auto* p = (tcw_ctx::ProfRec*)realloc(ctx->prof, sizeof(tcw_ctx::ProfRec) * cap);
        if (!p) return;
        for (int i = ctx->prof_cap; i < cap; ++i) {
            cudaEventCreate(&p[i].e0);
            cudaEventCreate(&p[i].e1);
        }
        ctx->prof = p;
        ctx->prof_cap = cap;
    }
    if (ctx->prof_on != 1) return;
    auto& r = ctx->prof[ctx->prof_n];
    r.kernel = kernel;
    r.units = units;
    cudaEventRecord(r.e0, st);
}
static void prof_end(tcw_ctx* ctx, cudaStream_t st) {
    if (ctx->prof_on != 1) return;
    cudaEventRecord(ctx->prof[ctx->prof_n].e1, st);
    ctx->prof_n++;
}

TCW_API int tcw_profile_enable(tcw_ctx* ctx, int enable) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_profile_enable: ctx is NULL");
    ctx->prof_on = enable ? 1 : 0;
    ctx->prof_n = 0;
    return TCW_OK;
}

TCW_API int tcw_profile_read(tcw_ctx* ctx, double* ms, long long* launches, double* units) {
    if (!ctx || !ms || !launches || !units) return fail(TCW_E_ARG, "tcw_profile_read: NULL argument");
    for (int k = 0; k < PK_COUNT; ++k) { ms[k] = 0; launches[k] = 0; units[k] = 0; }
    for (int i = 0; i < ctx->prof_n; ++i) {
        auto& r = ctx->prof[i];
        CUDA_TRY(cudaEventSynchronize(r.e1));
        float t = 0;
        CUDA_TRY(cudaEventElapsedTime(&t, r.e0, r.e1));
        ms[r.kernel] += t;
        launches[r.kernel] += 1;
        units[r.kernel] += r.units;
    }
    ctx->prof_n = 0;
    if (ctx->prof_on == 2) ctx->prof_on = 1;
    return TCW_OK;
}

TCW_API int tcw_version(void) { return TCW_VERSION; }
TCW_API const char* tcw_last_error(void) { return g_err; }
TCW_API long long tcw_launch_count(const tcw_ctx* ctx) { return ctx ? ctx->launches : 0; }

TCW_API int tcw_create(int device, tcw_ctx** out) {
    if (!out) return fail(TCW_E_ARG, "tcw_create: out is NULL");
    int n = 0;
    CUDA_TRY(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(TCW_E_ARG, "tcw_create: no such CUDA device");
    cudaDeviceProp p;
    CUDA_TRY(cudaGetDeviceProperties(&p, device));
    tcw_ctx* c = (tcw_ctx*)calloc(1, sizeof(tcw_ctx));
    if (!c) return fail(TCW_E_NOMEM, "tcw_create: out of host memory");
    c->device = device;
    c->sm_count = p.multiProcessorCount;
    c->l2_bytes = (size_t)p.l2CacheSize;
    const char* env = getenv("TCW_CHUNK_UNITS");
    c->chunk_units_override = env ? atoi(env) : 0;
    env = getenv("TCW_NO_PIPELINE");
    c->pipeline_off = env ? atoi(env) : 0;
    env = getenv("TCW_PIPE_LAG");
    c->pipe_lag_override = env ? atoi(env) : 0;
    env = getenv("TCW_V3");
    c->use_v3 = env ? atoi(env) : 0;
    env = getenv("TCW_NO_GRAPH");
    c->graphs_off = env ? atoi(env) : 0;
    env = getenv("TCW_LANES");
    c->lanes = env ? atoi(env) : 0;
    env = getenv("TCW_SERIAL");     // 2 (default): independent lanes; 1: one stream; 0: splat stream + normalise stream
    c->serial = env ? atoi(env) : 2;
    env = getenv("TCW_PERSIST");
    c->persist = env ? atoi(env) : 0;
    *out = c;
    return TCW_OK;
}

TCW_API void tcw_destroy(tcw_ctx* c) {
    if (!c) return;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(c->device);
#if TCW_CONST_M
    { tcw_ctx* me = c; g_const_owner.compare_exchange_strong(me, nullptr); }
#endif
    if (c->pool) cudaFree(c->pool);
    for (int i = 0; i < c->prof_cap; ++i) { cudaEventDestroy(c->prof[i].e0); cudaEventDestroy(c->prof[i].e1); }
    free(c->prof);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_run) cudaStreamDestroy(c->s_run);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    if (c->graph_exec) cudaGraphExecDestroy(c->graph_exec);
    if (c->s_cap) cudaStreamDestroy(c->s_cap);
    for (int i = 0; i < 7; ++i) if (c->s_lane[i]) { cudaStreamDestroy(c->s_lane[i]); cudaEventDestroy(c->ev_lane[i]); }
    if (c->s_aux) {
        cudaStreamDestroy(c->s_aux);
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(c->ev_splat[i]); cudaEventDestroy(c->ev_norm[i]); }
    }
    cudaSetDevice(prev);
    free(c);
}

// ----------------------------------------------------------------------------------------------
// device helpers
// ----------------------------------------------------------------------------------------------
constexpr int M_FLOATS = 30;
constexpr int TILE_W = 32, TILE_H = 8, TILE_THREADS = TILE_W * TILE_H;
constexpr int NCH = 5;   // accumulator planes: sum w*r, w*g, w*b, w*z, w

__device__ __forceinline__ float dot3(const float* m, float a, float b, float c) {
    // acc = 0; acc += m[k]*v[k]  — the reference's matmul order, never fused (models/utils.py:388,399)
    float acc = __fadd_rn(0.0f, __fmul_rn(m[0], a));
    acc = __fadd_rn(acc, __fmul_rn(m[1], b));
    acc = __fadd_rn(acc, __fmul_rn(m[2], c));
    return acc;
}
__device__ __forceinline__ float dot4(const float* m, float a, float b, float c, float d) {
    float acc = __fadd_rn(0.0f, __fmul_rn(m[0], a));
    acc = __fadd_rn(acc, __fmul_rn(m[1], b));
    acc = __fadd_rn(acc, __fmul_rn(m[2], c));
    acc = __fadd_rn(acc, __fmul_rn(m[3], d));
    return acc;
}

// compute_transformed_points for one pixel (models/utils.py:388-417). M = K1inv | T[:3] | K2.
__device__ __forceinline__ void project_px(const float* M, float fx, float fy, float d,
                                           float& P0, float& P1, float& P2) {
    float u0 = dot3(M + 0, fx, fy, 1.0f);
    float u1 = dot3(M + 3, fx, fy, 1.0f);
    float u2 = dot3(M + 6, fx, fy, 1.0f);
    float X0 = __fmul_rn(d, u0), X1 = __fmul_rn(d, u1), X2 = __fmul_rn(d, u2);
    float Y0 = dot4(M + 9, X0, X1, X2, 1.0f);
    float Y1 = dot4(M + 13, X0, X1, X2, 1.0f);
    float Y2 = dot4(M + 17, X0, X1, X2, 1.0f);
    P0 = dot3(M + 21, Y0, Y1, Y2);
    P1 = dot3(M + 24, Y0, Y1, Y2);
    P2 = dot3(M + 27, Y0, Y1, Y2);
    if (Y2 <= 0.01f) { P0 = 1000.0f; P1 = 1000.0f; P2 = 1000.0f; }   // behind the camera :404-417
}

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// log(1 + clamp(z, 0, 1000))   (models/utils.py:490-491)
__device__ __forceinline__ float log_depth(float z) { return logf(__fadd_rn(1.0f, clampf(z, 0.0f, 1000.0f))); }

struct Corners {
    int ix0, ix1, iy0, iy1;       // padded-grid indices in [0, w+1] x [0, h+1]
    float nw, sw, ne, se;          // proximity weights
};

// indices and proximity weights from a flow vector (models/utils.py:450-488)
__device__ __forceinline__ Corners corners_from_flow(float flx, float fly, int x, int y, int h, int w) {
    float ox = __fadd_rn(__fadd_rn(flx, (float)x), 1.0f);
    float oy = __fadd_rn(__fadd_rn(fly, (float)y), 1.0f);
    const float wx = (float)(w + 1), hy = (float)(h + 1);
    float fx0 = clampf(floorf(ox), 0.0f, wx), fx1 = clampf(ceilf(ox), 0.0f, wx);
    float fy0 = clampf(floorf(oy), 0.0f, hy), fy1 = clampf(ceilf(oy), 0.0f, hy);
    ox = clampf(ox, 0.0f, wx);
    oy = clampf(oy, 0.0f, hy);
    float ax = __fsub_rn(1.0f, __fsub_rn(ox, fx0)), bx = __fsub_rn(1.0f, __fsub_rn(fx1, ox));
    float ay = __fsub_rn(1.0f, __fsub_rn(oy, fy0)), by = __fsub_rn(1.0f, __fsub_rn(fy1, oy));
    Corners c;
    c.ix0 = (int)fx0; c.ix1 = (int)fx1; c.iy0 = (int)fy0; c.iy1 = (int)fy1;
    c.nw = __fmul_rn(ay, ax); c.sw = __fmul_rn(by, ax); c.ne = __fmul_rn(ay, bx); c.se = __fmul_rn(by, bx);
    return c;
}

__device__ __forceinline__ float block_max_256(float v) {
    __shared__ float red[TILE_THREADS / 32];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = threadIdx.x < TILE_THREADS / 32 ? red[threadIdx.x] : 0.0f;
        #pragma unroll
        for (int o = 4; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    }
    return v;   // valid in thread 0
}

#include "warp_kernels.cuh"
#include "warp_fast.cuh"
#include "warp_strip.cuh"
#include "warp_pipeline.cuh"
#include "warp_pair.cuh"
#ifndef TCW_ROWS2
#define TCW_ROWS2 0           // 1: next-round experiment, two source rows per iteration (warp_pair2.cuh); not in the default build
#endif
#if TCW_ROWS2
#include "warp_pair2.cuh"
#endif
#include "warp_pipeline2.cuh"
#include "resize_pack.cuh"

// ----------------------------------------------------------------------------------------------
// compute_transformed_points as a standalone op
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TILE_THREADS)
k_project(const float* __restrict__ depth, const int* __restrict__ src_index, const float* __restrict__ Mall,
          int h, int w, int tiles_x, float* __restrict__ out) {
    __shared__ float M[M_FLOATS];
    const int u = blockIdx.y;
    if (threadIdx.x < M_FLOATS) M[threadIdx.x] = Mall[(size_t)u * M_FLOATS + threadIdx.x];
    __syncthreads();
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    if (x >= w || y >= h) return;
    const int s = src_index ? src_index[u] : u;
    const size_t hw = (size_t)h * w, p = (size_t)y * w + x;
    float P0, P1, P2;
    project_px(M, (float)x, (float)y, depth[(size_t)s * hw + p], P0, P1, P2);
    float* o = out + ((size_t)u * hw + p) * 3;
    o[0] = P0; o[1] = P1; o[2] = P2;
}

// ----------------------------------------------------------------------------------------------
// generic bilinear_splatting (models/utils.py:422-583), c in 1..4
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TILE_THREADS)
k_logmax(const float* __restrict__ depth, size_t hw, unsigned* __restrict__ Lbits, int b_total) {
    const int n = blockIdx.y;
    float ld = 0.0f;
    for (size_t i = (size_t)blockIdx.x * TILE_THREADS + threadIdx.x; i < hw; i += (size_t)gridDim.x * TILE_THREADS)
        ld = fmaxf(ld, log_depth(depth[(size_t)n * hw + i]));
    float m = block_max_256(ld);
    if (threadIdx.x == 0) {
        atomicMax(&Lbits[n], __float_as_uint(m));
        atomicMax(&Lbits[b_total], __float_as_uint(m));
    }
}

template <int C>
__global__ void __launch_bounds__(TILE_THREADS)
k_splat_f32(const float* __restrict__ frame, const float* __restrict__ mask1, const float* __restrict__ depth,
            const float* __restrict__ flow, const float* __restrict__ flow_mask, int unit0, int h, int w,
            int tiles_x, const unsigned* __restrict__ Lbits, int l_index_global, float* __restrict__ acc) {
    const int uc = blockIdx.y, n = unit0 + uc;
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    if (x >= w || y >= h) return;
    const size_t hw = (size_t)h * w, p = (size_t)y * w + x;
    const Corners c = corners_from_flow(flow[(size_t)n * 2 * hw + p], flow[(size_t)n * 2 * hw + hw + p], x, y, h, w);
    const float L = __uint_as_float(Lbits[l_index_global >= 0 ? l_index_global : n]);
    const float g = expf(__fmul_rn(__fdiv_rn(log_depth(depth[(size_t)n * hw + p]), L), 50.0f));
    const float m = mask1 ? mask1[(size_t)n * hw + p] : 1.0f;
    const float fm = flow_mask ? flow_mask[(size_t)n * hw + p] : 1.0f;
    float v[C];
    #pragma unroll
    for (int ch = 0; ch < C; ++ch) v[ch] = frame[((size_t)n * C + ch) * hw + p];
    float* a = acc + (size_t)uc * (C + 1) * hw;
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int cx = (k & 2) ? c.ix1 : c.ix0;
        const int cy = (k & 1) ? c.iy1 : c.iy0;
        if (cx < 1 || cx > w || cy < 1 || cy > h) continue;
        const float prox = k == 0 ? c.nw : (k == 1 ? c.sw : (k == 2 ? c.ne : c.se));
        const float wk = __fdiv_rn(__fmul_rn(__fmul_rn(prox, m), fm), g);
        const size_t cell = (size_t)(cy - 1) * w + (cx - 1);
        #pragma unroll
        for (int ch = 0; ch < C; ++ch) atomicAdd(a + ch * hw + cell, __fmul_rn(v[ch], wk));
        atomicAdd(a + C * hw + cell, wk);
    }
}

template <int C>
__global__ void __launch_bounds__(TILE_THREADS)
k_splat_normalise(const float* __restrict__ acc, int unit0, size_t hw, int is_image,
                  float* __restrict__ out, float* __restrict__ out_mask, int* __restrict__ range_flag) {
    const int uc = blockIdx.y, n = unit0 + uc;
    const size_t p = (size_t)blockIdx.x * TILE_THREADS + threadIdx.x;
    if (p >= hw) return;
    const float* a = acc + (size_t)uc * (C + 1) * hw;
    const float W = a[C * hw + p];
    const bool valid = W > 0.0f;
    bool bad = false;
    #pragma unroll
    for (int ch = 0; ch < C; ++ch) {
        float v = valid ? __fdiv_rn(a[ch * hw + p], W) : (is_image ? -1.0f : 0.0f);
        if (is_image) {
            bad |= !(v >= -1.1f) || !(v <= 1.1f);
            v = clampf(v, -1.0f, 1.0f);
        }
        out[((size_t)n * C + ch) * hw + p] = v;
    }
    if (bad && range_flag) atomicOr(range_flag, 1);
    out_mask[(size_t)n * hw + p] = valid ? 1.0f : 0.0f;
}

// standalone clean_points (models/utils.py:585-626)
__global__ void __launch_bounds__(TILE_THREADS)
k_clean_points(const float* __restrict__ frame, const float* __restrict__ mask2, int c, int h, int w, int tiles_x,
               double* __restrict__ out_frame, double* __restrict__ out_mask) {
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    if (x >= w || y >= h) return;
    const size_t hw = (size_t)h * w, p = (size_t)y * w + x;
    bool hole = false;
    for (int dy = -2; dy <= 2; ++dy) {
        const int yy = y + dy;
        if (yy < 0 || yy >= h) continue;
        for (int dx = -2; dx <= 2; ++dx) {
            const int xx = x + dx;
            if (xx < 0 || xx >= w) continue;
            hole |= __fsub_rn(1.0f, mask2[(size_t)yy * w + xx]) >= 0.5f;
        }
    }
    const double keep = hole ? 0.0 : 1.0;
    for (int ch = 0; ch < c; ++ch) {
        const float half = __fdiv_rn(__fadd_rn(frame[(size_t)ch * hw + p], 1.0f), 2.0f);
        out_frame[(size_t)ch * hw + p] = __dsub_rn(__dmul_rn(__dmul_rn((double)half, keep), 2.0), 1.0);
    }
    out_mask[p] = keep;
}

// ----------------------------------------------------------------------------------------------
// host side of the C ABI
// ----------------------------------------------------------------------------------------------
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
// workspace header: 256 B of control words (ticket counter of the pipeline kernel) + per-unit stats (+1 for the whole call)
constexpr size_t CTRL_BYTES = 256;
static size_t header_bytes(int n_units) { return CTRL_BYTES + align_up((size_t)(n_units + 1) * sizeof(UnitStats), 256); }

// accumulator bytes per pixel of one unit: fp32 path 5 planes x 4 B; deterministic path 5 x int64 + max-weight plane
static size_t acc_bytes_per_px(int flags) { return (flags & TCW_DETERMINISTIC) ? (size_t)NCH * 8 + 4 : (size_t)NCH * 4; }

// units per chunk: keep one chunk's accumulators + inputs + outputs inside a fraction of L2
static int auto_chunk_units(const tcw_ctx* ctx, int n_units, int h, int w, size_t acc_per_px) {
    if (ctx->chunk_units_override > 0) return ctx->chunk_units_override < n_units ? ctx->chunk_units_override : n_units;
    const size_t per_unit = (size_t)h * w * (acc_per_px + 44);
    size_t budget = ctx->l2_bytes ? ctx->l2_bytes * 6 / 10 : (size_t)64 << 20;
    int g = (int)(budget / per_unit);
    if (g < 1) g = 1;
    return g < n_units ? g : n_units;
}

// lanes x units-per-chunk of the multi-stream fp32 path.  Units in flight = lanes * chunk: their accumulators
// (20 B/px) should stay L2-resident next to the streaming traffic; measured optimum on B200 at 384x672:
// 4 lanes x 3 units = 62 MB of the 126 MB L2 (profiles/r1_v4_notes.md).
static void lanes_plan(const tcw_ctx* ctx, int n_units, size_t hw, int* lanes, int* chunk) {
    int NL = ctx->lanes > 0 ? ctx->lanes : 4;
    if (NL > 8) NL = 8;
    int g2 = ctx->chunk_units_override;
    if (g2 <= 0) {
        const size_t l2 = ctx->l2_bytes ? ctx->l2_bytes : ((size_t)64 << 20);
        int inflight = (int)((l2 / 2) / (hw * NCH * sizeof(float)));
        if (inflight < 2) inflight = 2;
        if (ctx->lanes <= 0 && NL > inflight) NL = inflight;
        g2 = inflight / NL;
        if (g2 < 1) g2 = 1;
        if (g2 > 4096) g2 = 4096;
    }
    if (ctx->serial != 2) NL = 2;
    const int per_lane = (n_units + NL - 1) / NL;
    if (g2 > per_lane) g2 = per_lane;
    *lanes = NL; *chunk = g2;
}

TCW_API int tcw_workspace_bytes(const tcw_ctx* ctx, int n_units, int h, int w, int flags, size_t* bytes) {
    if (!ctx || !bytes || n_units <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_workspace_bytes: bad argument");
    const size_t per_px = acc_bytes_per_px(flags);
    int g = auto_chunk_units(ctx, n_units, h, w, per_px);
    int ring_max = PIPE_RING > PIPE2_RING ? PIPE_RING : PIPE2_RING;                   // accumulator ring of the pipeline kernels
    {   // lanes x chunk units of the multi-stream path
        int nl = 1, g2 = 1;
        lanes_plan(ctx, n_units, (size_t)h * w, &nl, &g2);
        if (ring_max < nl * g2) ring_max = nl * g2;
    }
    if (!(flags & (TCW_DETERMINISTIC | TCW_CLEAN)) && g < ring_max) g = ring_max;
    *bytes = header_bytes(n_units) + align_up((size_t)g * per_px * h * w, 256);
    return TCW_OK;
}

static int check_launch(tcw_ctx* ctx, const char* what) {
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(TCW_E_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
    return TCW_OK;
}
#define LAUNCH_CHECK(what) do { int rc__ = check_launch(ctx, what); if (rc__) return rc__; } while (0)

static int forward_warp_enqueue(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                                const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                                int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                                int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_forward_warp: ctx is NULL");
    if (!frame || !depth || !M || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_forward_warp: NULL buffer");
    if (n_units <= 0 || n_src <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_forward_warp: bad shape");
    if (!src_index && n_src != n_units) return fail(TCW_E_ARG, "tcw_forward_warp: src_index is NULL but n_src != n_units");
    const bool det = flags & TCW_DETERMINISTIC;
    const size_t hw = (size_t)h * w;
    const size_t hdr = header_bytes(n_units);
    const size_t unit_bytes = acc_bytes_per_px(flags) * hw;
    if (!workspace || workspace_bytes < hdr + unit_bytes)
        return fail(TCW_E_WORKSPACE, "tcw_forward_warp: workspace too small (see tcw_workspace_bytes)");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    int g = (int)((workspace_bytes - hdr) / unit_bytes);
    const bool ring_fits = g >= PIPE_RING;
    const bool ring2_fits = g >= PIPE2_RING;
    const int want = auto_chunk_units(ctx, n_units, h, w, acc_bytes_per_px(flags));
    if (g > want) g = want;
    UnitStats* stats = (UnitStats*)((char*)workspace + CTRL_BYTES);
    int* ticket = (int*)workspace;
    char* accbase = (char*)workspace + hdr;
    // chunk layout: fp32: [g][hw] float4 (sum w*r,g,b,z) then [g][hw] float (sum w);
    //               deterministic: [g][5][hw] int64 then [g][hw] uint32 max-weight bits
    float4* acc4 = det ? nullptr : (float4*)accbase;
    float* accw = det ? nullptr : (float*)(accbase + (size_t)g * hw * sizeof(float4));
    long long* acc_i64 = det ? (long long*)accbase : nullptr;
    unsigned* wmax = det ? (unsigned*)(accbase + (size_t)g * NCH * hw * sizeof(long long)) : nullptr;
    const float det_scale = det ? ldexpf(1.0f, det_shift(h, w)) : 0.0f;
    const int tiles_x = (w + TILE_W - 1) / TILE_W, tiles_y = (h + TILE_H - 1) / TILE_H;
    const int strips_x = (w + STRIP_W - 1) / STRIP_W, strips_per_unit = strips_x * ((h + STRIP_H - 1) / STRIP_H);
    const int strip_blocks = (strips_per_unit + TILE_THREADS / 32 - 1) / (TILE_THREADS / 32);
    const bool per_unit = flags & TCW_PER_UNIT_MAX;
    const bool clean = flags & TCW_CLEAN;
    const int sidx = per_unit ? -1 : n_units;
    // v4 (two pixels per lane, warp-granular pipeline) serves the plain fp32 path; the other modes keep warp_strip.cuh
    const bool rezero = !det && !clean;
    const bool use_v4 = rezero && ctx->persist && !ctx->use_v3 && !ctx->pipeline_off && ring2_fits && hw < 0x7fffffffull / 4;
    const bool pair_ok = !ctx->use_v3 && hw < 0x7fffffffull / 4;                           // v4 strip code (warp_pair.cuh) usable
    const bool pair_chunk = rezero && pair_ok && !use_v4;                                  // ... in per-chunk launches, fp32 path
    CUDA_TRY(cudaMemsetAsync(workspace, 0, hdr, st));   // ticket + per-unit stats and counters
    int k0_head = n_units, k0_bx = 1;      // units whose normaliser is computed up front
    {   // K0: soft-z normaliser of every unit (and of the whole call) before any splat.  (Launching it per chunk
        // inside the lanes was measured: 98 more small launches, 2.51 -> 2.97 ms/step.)
        const int quads = h * ((w + 3) / 4);
        int bx = (quads + TILE_THREADS * 2 - 1) / (TILE_THREADS * 2);
        if (bx < 1) bx = 1;
        k0_bx = bx;
        // multi-lane schedule with per-unit normalisers: only the units of the first round of chunks here, the rest
        // runs on its own stream under the first round (see the lanes loop)
        if (pair_chunk && !ctx->pipeline_off && ctx->serial == 2 && per_unit && !ctx->prof_on) {
            int nl = 1, g2 = 1;
            lanes_plan(ctx, n_units, hw, &nl, &g2);
            if (nl * g2 < n_units && n_units < 65535) k0_head = nl * g2;
        }
        for (int u0 = 0; u0 < k0_head; u0 += 65535) {
            const int nu = k0_head - u0 < 65535 ? k0_head - u0 : 65535;
            prof_begin(ctx, PK_ZMAX, nu, st);
            k_zmax_quad<<<dim3(bx, nu), TILE_THREADS, 0, st>>>(depth, src_index, M, u0, h, w, stats, n_units, (use_v4 || pair_ok) ? 1 : 0);
            prof_end(ctx, st);
            LAUNCH_CHECK("k_zmax_quad");
        }
    }
    // the plain fp32 normalise kernel re-zeroes what it read: one memset per call instead of one per chunk
    const bool vec = (hw % 4 == 0) && ((((size_t)accbase) | ((size_t)out_frame) | ((size_t)out_mask) | ((size_t)out_depth)) % 16 == 0);
    SplatArgs A;
    A.frame = frame; A.mask1 = mask1; A.depth = depth; A.src_index = src_index; A.Mall = M;
    A.h = h; A.w = w; A.strips_x = strips_x; A.strips_per_unit = strips_per_unit;
    A.stats = stats; A.stat_index_global = sidx;
    A.acc4 = acc4; A.accw = accw; A.acc_i64 = acc_i64; A.wmax = wmax; A.det_scale = det_scale; A.out_flow = out_flow;
    if (pair_chunk && !ctx->pipeline_off) {
        // fp32 path: chunks of g2 units, each with its own accumulator buffer, spread over NL streams ("lanes")
        if (!ctx->s_aux) {
            CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_aux, cudaStreamNonBlocking));
            for (int i = 0; i < 2; ++i) {
                CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_splat[i], cudaEventDisableTiming));
                CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_norm[i], cudaEventDisableTiming));
            }
        }
        const int cap = (int)((workspace_bytes - hdr) / unit_bytes);          // units the workspace can hold
        int NL = 1, g2 = 1;
        lanes_plan(ctx, n_units, hw, &NL, &g2);
        if (NL > cap) NL = cap > 0 ? cap : 1;
        if (g2 > cap / NL) g2 = cap / NL;
        if (g2 >= 1) {
            if (g2 > n_units) g2 = n_units;
            PairArgs B;
            B.frame = frame; B.mask1 = mask1; B.depth = depth; B.src_index = src_index; B.Mall = M;
            B.h = h; B.w = w; B.strips_x = (w + PAIR_W - 1) / PAIR_W;
            B.strips_per_unit = B.strips_x * ((h + PAIR_H_CHUNK - 1) / PAIR_H_CHUNK);
            B.stats = stats; B.stat_index_global = sidx; B.out_flow = out_flow;
            const int pb = (B.strips_per_unit + PIPE2_THREADS / 32 - 1) / (PIPE2_THREADS / 32);
            // buffer b: [g2][hw] float4 then [g2][hw] float
            const size_t buf_bytes = (size_t)g2 * NCH * hw * sizeof(float);
            if (ctx->serial != 2) CUDA_TRY(cudaMemsetAsync(accbase, 0, 2 * buf_bytes, st));
            int k = 0;
            if (ctx->serial == 2) {
                // independent lanes: chunk k runs splat + normalise on stream k % NL with buffer k % NL; the tail of
                // one lane's kernel overlaps the other lanes' kernels, and nothing waits across lanes until the join
                for (int i = 0; i + 1 < NL; ++i) if (!ctx->s_lane[i]) {
                    CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_lane[i], cudaStreamNonBlocking));
                    CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_lane[i], cudaEventDisableTiming));
                }
                CUDA_TRY(cudaMemsetAsync(accbase, 0, (size_t)NL * buf_bytes, st));
                const bool k0_rest = k0_head < n_units && k0_head == NL * g2;
                if (k0_head < n_units && !k0_rest) {      // the workspace forced another lane plan: finish the pre-pass here
                    k_zmax_quad<<<dim3(k0_bx, n_units - k0_head), TILE_THREADS, 0, st>>>(depth, src_index, M, k0_head, h, w, stats, n_units, 1);
                    LAUNCH_CHECK("k_zmax_quad");
                }
#if TCW_CONST_M
                // one context per process may use the constant bank (a second one would overwrite it under running kernels)
                bool use_cm = false;
                if (n_units <= CONST_M_UNITS) {
                    tcw_ctx* none = nullptr;
                    use_cm = g_const_owner.compare_exchange_strong(none, ctx) || none == ctx;
                }
                if (use_cm) CUDA_TRY(cudaMemcpyToSymbolAsync(c_tcw_M, M, (size_t)n_units * M_FLOATS * sizeof(float), 0,
                                                             cudaMemcpyDeviceToDevice, st));
#endif
                CUDA_TRY(cudaEventRecord(ctx->ev_splat[0], st));
                for (int i = 0; i + 1 < NL; ++i) CUDA_TRY(cudaStreamWaitEvent(ctx->s_lane[i], ctx->ev_splat[0], 0));
                if (k0_rest) {      // normalisers of the later rounds: own stream, concurrent with the first round of chunks
                    CUDA_TRY(cudaStreamWaitEvent(ctx->s_aux, ctx->ev_splat[0], 0));
                    k_zmax_quad<<<dim3(k0_bx, n_units - k0_head), TILE_THREADS, 0, ctx->s_aux>>>(depth, src_index, M, k0_head, h, w, stats, n_units, 1);
                    LAUNCH_CHECK("k_zmax_quad");
                    CUDA_TRY(cudaEventRecord(ctx->ev_norm[0], ctx->s_aux));
                }
                for (int u0 = 0; u0 < n_units; u0 += g2, ++k) {
                    const int nu = n_units - u0 < g2 ? n_units - u0 : g2, b = k % NL;
                    cudaStream_t sk = b ? ctx->s_lane[b - 1] : st;
                    if (k0_rest && k >= NL && k < 2 * NL) CUDA_TRY(cudaStreamWaitEvent(sk, ctx->ev_norm[0], 0));
                    char* buf = accbase + (size_t)b * buf_bytes;
                    B.acc4 = (float4*)buf;
                    B.accw = (float*)(buf + (size_t)g2 * hw * sizeof(float4));
                    prof_begin(ctx, PK_SPLAT, nu, sk);
#if TCW_CONST_M
                    if (use_cm) k_splat_pair_cm<<<dim3(pb, nu), PIPE2_THREADS, 0, sk>>>(B, u0);
                    else
#endif
                    k_splat_pair<<<dim3(pb, nu), PIPE2_THREADS, 0, sk>>>(B, u0);
                    prof_end(ctx, sk);
                    LAUNCH_CHECK("k_splat_pair");
                    prof_begin(ctx, PK_NORMALISE, nu, sk);
                    k_normalise_pair<<<dim3((unsigned)((hw + 1023) / 1024), nu), 256, 0, sk>>>(
                        B.acc4, B.accw, u0, hw, (float*)out_frame, (float*)out_mask, out_depth, range_flag);
                    prof_end(ctx, sk);
                    LAUNCH_CHECK("k_normalise_pair");
                }
                for (int i = 0; i + 1 < NL; ++i) {
                    CUDA_TRY(cudaEventRecord(ctx->ev_lane[i], ctx->s_lane[i]));
                    CUDA_TRY(cudaStreamWaitEvent(st, ctx->ev_lane[i], 0));
                }
                return TCW_OK;
            }
            for (int u0 = 0; u0 < n_units; u0 += g2, ++k) {
                const int nu = n_units - u0 < g2 ? n_units - u0 : g2, b = k & 1;
                char* buf = accbase + (size_t)b * buf_bytes;
                B.acc4 = (float4*)buf;
                B.accw = (float*)(buf + (size_t)g2 * hw * sizeof(float4));
                cudaStream_t sn = ctx->serial ? st : ctx->s_aux;
                if (k >= 2 && !ctx->serial) CUDA_TRY(cudaStreamWaitEvent(st, ctx->ev_norm[b], 0));      // buffer b normalised + re-zeroed
                prof_begin(ctx, PK_SPLAT, nu, st);
                k_splat_pair<<<dim3(pb, nu), PIPE2_THREADS, 0, st>>>(B, u0);
                prof_end(ctx, st);
                LAUNCH_CHECK("k_splat_pair");
                if (!ctx->serial) {
                    CUDA_TRY(cudaEventRecord(ctx->ev_splat[b], st));
                    CUDA_TRY(cudaStreamWaitEvent(sn, ctx->ev_splat[b], 0));
                }
                prof_begin(ctx, PK_NORMALISE, nu, sn);
                k_normalise_pair<<<dim3((unsigned)((hw + 1023) / 1024), nu), 256, 0, sn>>>(
                    B.acc4, B.accw, u0, hw, (float*)out_frame, (float*)out_mask, out_depth, range_flag);
                prof_end(ctx, sn);
                LAUNCH_CHECK("k_normalise_pair");
                if (!ctx->serial) CUDA_TRY(cudaEventRecord(ctx->ev_norm[b], sn));
            }
            // join: the caller's stream continues only after the last normalise of either buffer
            if (!ctx->serial) {
                CUDA_TRY(cudaStreamWaitEvent(st, ctx->ev_norm[(k - 1) & 1], 0));
                if (k >= 2) CUDA_TRY(cudaStreamWaitEvent(st, ctx->ev_norm[k & 1], 0));
            }
            return TCW_OK;
        }
    }
    if (use_v4) {
        // fp32 path: one persistent kernel, warp-granular items (warp_pipeline2.cuh)
        if (!ctx->pipe2_blocks) {
            int per_sm = 0;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_warp_pipeline2, PIPE2_THREADS, 0));
            ctx->pipe2_blocks = (per_sm > 0 ? per_sm : 1) * ctx->sm_count;
        }
        Pipe2Args P;
        P.A.frame = frame; P.A.mask1 = mask1; P.A.depth = depth; P.A.src_index = src_index; P.A.Mall = M;
        P.A.h = h; P.A.w = w; P.A.strips_x = (w + PAIR_W - 1) / PAIR_W;
        P.A.strips_per_unit = P.A.strips_x * ((h + PAIR_H - 1) / PAIR_H);
        P.A.stats = stats; P.A.stat_index_global = sidx;
        // ring layout: [PIPE2_RING][hw] float4 then [PIPE2_RING][hw] float
        P.A.acc4 = (float4*)accbase;
        P.A.accw = (float*)(accbase + (size_t)PIPE2_RING * hw * sizeof(float4));
        P.A.out_flow = out_flow;
        CUDA_TRY(cudaMemsetAsync(accbase, 0, (size_t)PIPE2_RING * NCH * hw * sizeof(float), st));
        P.stats_rw = stats; P.ticket = ticket;
        { const char* dg = getenv("TCW_DIAG"); P.diag = dg ? atoi(dg) : 0; }
        P.n_units = n_units; P.ns = P.A.strips_per_unit; P.vec = vec ? 1 : 0;
        P.nn = (int)(((vec ? hw / 4 : hw) + NORM2_QUADS - 1) / NORM2_QUADS);
        P.out_frame = (float*)out_frame; P.out_mask = (float*)out_mask; P.out_depth = out_depth; P.range_flag = range_flag;
        const long long total_s = (long long)n_units * P.ns, total_n = (long long)n_units * P.nn;
        if (total_s + total_n > 0x7fffff00ll) return fail(TCW_E_ARG, "tcw_forward_warp: too many work items for one call");
        const long long most = total_s > total_n ? total_s : total_n;
        const long long want_blocks = (most + PIPE2_THREADS / 32 - 1) / (PIPE2_THREADS / 32);
        const int blocks = (long long)ctx->pipe2_blocks < want_blocks ? ctx->pipe2_blocks : (int)want_blocks;
        {   // ring: a unit stays in the ring for about two strip times (its strips run, then its owners pick the
            // normalise items up after their next strip); the grid starts warps / ns units per strip time
            const int warps = blocks * (PIPE2_THREADS / 32);
            int ring = (int)((2.2 * warps) / P.ns) + 3;
            if (ctx->pipe_lag_override > 0) ring = ctx->pipe_lag_override + 2;
            if (ring < 2) ring = 2;
            if (ring > PIPE2_RING) ring = PIPE2_RING;
            P.lag = ring - 2; P.ring = ring;
        }
        prof_begin(ctx, PK_SPLAT, n_units, st);
        void* kargs[] = {&P};      // cooperative: the static schedule needs every warp of the grid resident
        CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_warp_pipeline2, dim3(blocks), dim3(PIPE2_THREADS), kargs, 0, st));
        prof_end(ctx, st);
        LAUNCH_CHECK("k_warp_pipeline2");
        return TCW_OK;
    }
    if (rezero && !ctx->pipeline_off && ring_fits) {
        // fp32 path: one persistent kernel runs splat + normalise of every unit (warp_pipeline.cuh)
        if (!ctx->pipe_blocks) {
            int per_sm = 0;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_warp_pipeline, TILE_THREADS, 0));
            ctx->pipe_blocks = (per_sm > 0 ? per_sm : 1) * ctx->sm_count;
        }
        // ring layout: [PIPE_RING][hw] float4 then [PIPE_RING][hw] float
        A.accw = (float*)(accbase + (size_t)PIPE_RING * hw * sizeof(float4));
        CUDA_TRY(cudaMemsetAsync(accbase, 0, (size_t)PIPE_RING * NCH * hw * sizeof(float), st));
        PipeArgs P;
        P.A = A; P.stats_rw = stats; P.ticket = ticket;
        P.n_units = n_units; P.ns = strip_blocks; P.vec = vec ? 1 : 0;
        P.nn = (int)(((vec ? hw / 4 : hw) + NORM_ITEM - 1) / NORM_ITEM);
        {   // lag: a full wave of resident CTAs between a unit's splat items and its normalise items
            const int per_unit_items = P.ns + P.nn;
            int lag = (ctx->pipe_blocks + per_unit_items - 1) / per_unit_items;
            if (ctx->pipe_lag_override > 0) lag = ctx->pipe_lag_override;
            if (lag < 1) lag = 1;
            if (lag > PIPE_RING - 2) lag = PIPE_RING - 2;
            P.lag = lag; P.ring = lag + 2;
        }
        P.out_frame = (float*)out_frame; P.out_mask = (float*)out_mask; P.out_depth = out_depth; P.range_flag = range_flag;
        const long long items = (long long)n_units * (P.ns + P.nn);
        if (items > 0x7fffff00ll) return fail(TCW_E_ARG, "tcw_forward_warp: too many work items for one call");
        const int blocks = (long long)ctx->pipe_blocks < items ? ctx->pipe_blocks : (int)items;
        prof_begin(ctx, PK_SPLAT, n_units, st);
        k_warp_pipeline<<<blocks, TILE_THREADS, 0, st>>>(P);
        prof_end(ctx, st);
        LAUNCH_CHECK("k_warp_pipeline");
        return TCW_OK;
    }
    if (rezero) CUDA_TRY(cudaMemsetAsync(accbase, 0, (size_t)(g < n_units ? g : n_units) * NCH * hw * sizeof(float), st));
    for (int u0 = 0; u0 < n_units; u0 += g) {
        const int nu = n_units - u0 < g ? n_units - u0 : g;
        const dim3 sgrid(strip_blocks, nu);
        PairArgs B;
        B.frame = frame; B.mask1 = mask1; B.depth = depth; B.src_index = src_index; B.Mall = M;
        B.h = h; B.w = w; B.strips_x = (w + PAIR_W - 1) / PAIR_W;
        B.strips_per_unit = B.strips_x * ((h + PAIR_H_CHUNK - 1) / PAIR_H_CHUNK);
        B.stats = stats; B.stat_index_global = sidx; B.acc4 = acc4; B.accw = accw; B.out_flow = out_flow;
        B.acc_i64 = acc_i64; B.wmax = wmax; B.det_scale = det_scale;
        const dim3 pgrid((B.strips_per_unit + PIPE2_THREADS / 32 - 1) / (PIPE2_THREADS / 32), nu);
        if (det) {
            CUDA_TRY(cudaMemsetAsync(acc_i64, 0, (size_t)nu * NCH * hw * sizeof(long long), st));
            CUDA_TRY(cudaMemsetAsync(wmax, 0, (size_t)nu * hw * sizeof(unsigned), st));
            A.out_flow = nullptr;
            prof_begin(ctx, PK_WMAX, nu, st);
            if (pair_ok) k_splat_pair_det<MODE_WMAX><<<pgrid, PIPE2_THREADS, 0, st>>>(B, u0);
            else k_splat_strip<MODE_WMAX><<<sgrid, TILE_THREADS, 0, st>>>(A, u0);
            prof_end(ctx, st);
            LAUNCH_CHECK("k_splat<wmax>");
            A.out_flow = out_flow;
            prof_begin(ctx, PK_SPLAT, nu, st);
            if (pair_ok) k_splat_pair_det<MODE_I64><<<pgrid, PIPE2_THREADS, 0, st>>>(B, u0);
            else k_splat_strip<MODE_I64><<<sgrid, TILE_THREADS, 0, st>>>(A, u0);
            prof_end(ctx, st);
            LAUNCH_CHECK("k_splat<i64>");
        } else {
            if (!rezero) CUDA_TRY(cudaMemsetAsync(accbase, 0, (size_t)g * NCH * hw * sizeof(float), st));
            prof_begin(ctx, PK_SPLAT, nu, st);
            if (pair_ok) k_splat_pair<<<pgrid, PIPE2_THREADS, 0, st>>>(B, u0);
            else k_splat_strip<MODE_F32><<<sgrid, TILE_THREADS, 0, st>>>(A, u0);
            prof_end(ctx, st);
            LAUNCH_CHECK("k_splat<f32>");
        }
        prof_begin(ctx, PK_NORMALISE, nu, st);
        if (rezero) {
            if (vec) {
                k_normalise_quad<true><<<dim3((unsigned)((hw / 4 + TILE_THREADS - 1) / TILE_THREADS), nu), TILE_THREADS, 0, st>>>(
                    acc4, accw, u0, hw, (float*)out_frame, (float*)out_mask, out_depth, range_flag);
            } else {
                k_normalise_quad<false><<<dim3((unsigned)((hw + TILE_THREADS - 1) / TILE_THREADS), nu), TILE_THREADS, 0, st>>>(
                    acc4, accw, u0, hw, (float*)out_frame, (float*)out_mask, out_depth, range_flag);
            }
        } else {
            const dim3 grid(tiles_x * tiles_y, nu);
            #define NORM_ARGS acc4, accw, acc_i64, stats, sidx, u0, h, w, tiles_x, out_frame, out_mask, out_depth, range_flag
            if (clean && det) k_normalise<true, true><<<grid, TILE_THREADS, 0, st>>>(NORM_ARGS);
            else if (clean)   k_normalise<true, false><<<grid, TILE_THREADS, 0, st>>>(NORM_ARGS);
            else              k_normalise<false, true><<<grid, TILE_THREADS, 0, st>>>(NORM_ARGS);
            #undef NORM_ARGS
        }
        prof_end(ctx, st);
        LAUNCH_CHECK("k_normalise");
    }
    return TCW_OK;
}

// Public entry: enqueue, or replay the CUDA graph of the previous identical call.  One call is ~200 kernel
// launches on four streams; enqueueing them costs the host as long as the GPU needs to run them (2.2 ms measured),
// so a call whose arguments equal those of the call before is captured once and replayed afterwards.  Kernels
// read the matrices, indices and images through the pointers at run time, so new CONTENTS need no new graph.
TCW_API int tcw_forward_warp(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                             const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                             int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                             int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_forward_warp: ctx is NULL");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    bool eligible = !ctx->graphs_off && !ctx->prof_on && !(flags & (TCW_DETERMINISTIC | TCW_CLEAN)) && !ctx->use_v3 &&
                    !ctx->persist && !ctx->pipeline_off && ctx->serial == 2;
    if (eligible) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) { cudaGetLastError(); eligible = false; }
    }
    if (!eligible)
        return forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream);
    tcw_ctx::GraphKey key;
    memset(&key, 0, sizeof(key));
    key.frame = frame; key.mask1 = mask1; key.depth = depth; key.src_index = src_index; key.M = M;
    key.out_frame = out_frame; key.out_mask = out_mask; key.out_depth = out_depth; key.out_flow = out_flow;
    key.range_flag = range_flag; key.workspace = workspace; key.workspace_bytes = workspace_bytes;
    key.n_units = n_units; key.n_src = n_src; key.h = h; key.w = w; key.flags = flags;
    if (ctx->graph_exec && !memcmp(&key, &ctx->graph_key, sizeof(key))) {
        CUDA_TRY(cudaGraphLaunch(ctx->graph_exec, st));
        ctx->launches += ctx->graph_launches;
        return TCW_OK;
    }
    const bool repeat = ctx->last_key_valid && !memcmp(&key, &ctx->last_key, sizeof(key));
    ctx->last_key = key;
    ctx->last_key_valid = 1;
    if (!repeat)      // first sight of these arguments: plain enqueue (a caller that allocates fresh outputs never repeats)
        return forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream);
    // second identical call in a row: capture it on an internal stream, instantiate, launch on the caller's stream
    if (!ctx->s_cap) CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_cap, cudaStreamNonBlocking));
    if (ctx->graph_exec) { cudaGraphExecDestroy(ctx->graph_exec); ctx->graph_exec = nullptr; }
    const long long before = ctx->launches;
    CUDA_TRY(cudaStreamBeginCapture(ctx->s_cap, cudaStreamCaptureModeThreadLocal));
    const int rc = forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                        out_depth, out_flow, range_flag, workspace, workspace_bytes, ctx->s_cap);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(ctx->s_cap, &graph);
    ctx->graph_launches = ctx->launches - before;
    ctx->launches = before;
    if (rc || ce != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        ctx->graphs_off = 1;      // do not try again in this context; fall back to plain enqueue
        if (rc) return rc;
        return forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream);
    }
    cudaError_t ie = cudaGraphInstantiate(&ctx->graph_exec, graph, 0);
    if (ie == cudaSuccess) ie = cudaGraphUpload(ctx->graph_exec, st);      // pre-pay the first-launch upload
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) {
        cudaGetLastError();
        if (ctx->graph_exec) { cudaGraphExecDestroy(ctx->graph_exec); ctx->graph_exec = nullptr; }
        ctx->graphs_off = 1;
        return forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream);
    }
    ctx->graph_key = key;
    CUDA_TRY(cudaGraphLaunch(ctx->graph_exec, st));
    ctx->launches += ctx->graph_launches;
    return TCW_OK;
}

TCW_API int tcw_project(tcw_ctx* ctx, const float* depth, const int* src_index, const float* M,
                        int n_units, int n_src, int h, int w, float* out_points, void* cuda_stream) {
    if (!ctx || !depth || !M || !out_points) return fail(TCW_E_ARG, "tcw_project: NULL argument");
    if (n_units <= 0 || n_units > 65535 || n_src <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_project: bad shape");
    if (!src_index && n_src != n_units) return fail(TCW_E_ARG, "tcw_project: src_index is NULL but n_src != n_units");
    const int tiles_x = (w + TILE_W - 1) / TILE_W, tiles_y = (h + TILE_H - 1) / TILE_H;
    k_project<<<dim3(tiles_x * tiles_y, n_units), TILE_THREADS, 0, (cudaStream_t)cuda_stream>>>(depth, src_index, M, h, w, tiles_x, out_points);
    LAUNCH_CHECK("k_project");
    return TCW_OK;
}

template <int C>
static int splat_c(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth, const float* flow,
                   const float* flow_mask, int b, int h, int w, int flags, float* out, float* out_mask,
                   int* range_flag, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    const size_t hw = (size_t)h * w;
    const size_t hdr = header_bytes(b);      // used here as an array of (b+1) uint32 log-depth maxima
    if (!workspace || workspace_bytes < hdr + (C + 1) * hw * sizeof(float))
        return fail(TCW_E_WORKSPACE, "tcw_splat: workspace too small (see tcw_workspace_bytes)");
    int g = (int)((workspace_bytes - hdr) / ((C + 1) * hw * sizeof(float)));
    const int want = auto_chunk_units(ctx, b, h, w, (size_t)(C + 1) * 4);
    if (g > want) g = want;
    unsigned* Lbits = (unsigned*)workspace;
    float* acc = (float*)((char*)workspace + hdr);
    const int tiles_x = (w + TILE_W - 1) / TILE_W, tiles_y = (h + TILE_H - 1) / TILE_H;
    const bool per_unit = flags & TCW_PER_UNIT_MAX;
    CUDA_TRY(cudaMemsetAsync(Lbits, 0, hdr, st));
    int red_blocks = (int)((hw + TILE_THREADS * 8 - 1) / (TILE_THREADS * 8));
    if (red_blocks < 1) red_blocks = 1;
    k_logmax<<<dim3(red_blocks, b), TILE_THREADS, 0, st>>>(depth, hw, Lbits, b);
    LAUNCH_CHECK("k_logmax");
    for (int u0 = 0; u0 < b; u0 += g) {
        const int nu = b - u0 < g ? b - u0 : g;
        CUDA_TRY(cudaMemsetAsync(acc, 0, (size_t)nu * (C + 1) * hw * sizeof(float), st));
        k_splat_f32<C><<<dim3(tiles_x * tiles_y, nu), TILE_THREADS, 0, st>>>(frame, mask1, depth, flow, flow_mask, u0, h, w,
                                                                           tiles_x, Lbits, per_unit ? -1 : b, acc);
        LAUNCH_CHECK("k_splat_f32");
        k_splat_normalise<C><<<dim3((unsigned)((hw + TILE_THREADS - 1) / TILE_THREADS), nu), TILE_THREADS, 0, st>>>(
            acc, u0, hw, (flags & TCW_IS_IMAGE) ? 1 : 0, out, out_mask, range_flag);
        LAUNCH_CHECK("k_splat_normalise");
    }
    return TCW_OK;
}

TCW_API int tcw_splat(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                      const float* flow, const float* flow_mask, int b, int c, int h, int w, int flags,
                      float* out, float* out_mask, int* range_flag, void* workspace, size_t workspace_bytes,
                      void* cuda_stream) {
    if (!ctx || !frame || !depth || !flow || !out || !out_mask) return fail(TCW_E_ARG, "tcw_splat: NULL argument");
    if (b <= 0 || b > 65535 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_splat: bad shape");
    if (flags & TCW_DETERMINISTIC) return fail(TCW_E_ARG, "tcw_splat: TCW_DETERMINISTIC not available in this build");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    switch (c) {
        case 1: return splat_c<1>(ctx, frame, mask1, depth, flow, flow_mask, b, h, w, flags, out, out_mask, range_flag, workspace, workspace_bytes, st);
        case 2: return splat_c<2>(ctx, frame, mask1, depth, flow, flow_mask, b, h, w, flags, out, out_mask, range_flag, workspace, workspace_bytes, st);
        case 3: return splat_c<3>(ctx, frame, mask1, depth, flow, flow_mask, b, h, w, flags, out, out_mask, range_flag, workspace, workspace_bytes, st);
        case 4: return splat_c<4>(ctx, frame, mask1, depth, flow, flow_mask, b, h, w, flags, out, out_mask, range_flag, workspace, workspace_bytes, st);
        default: return fail(TCW_E_ARG, "tcw_splat: c must be 1..4");
    }
}

TCW_API int tcw_clean_points(tcw_ctx* ctx, const float* frame, const float* mask2, int c, int h, int w,
                             double* out_frame, double* out_mask, void* cuda_stream) {
    if (!ctx || !frame || !mask2 || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_clean_points: NULL argument");
    if (c < 1 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_clean_points: bad shape");
    const int tiles_x = (w + TILE_W - 1) / TILE_W, tiles_y = (h + TILE_H - 1) / TILE_H;
    k_clean_points<<<tiles_x * tiles_y, TILE_THREADS, 0, (cudaStream_t)cuda_stream>>>(frame, mask2, c, h, w, tiles_x, out_frame, out_mask);
    LAUNCH_CHECK("k_clean_points");
    return TCW_OK;
}

TCW_API int tcw_resize_pack(tcw_ctx* ctx, const float* frame, const float* mask, int n, int H, int W, int h, int w,
                            float* out_video, float* out_masks, void* cuda_stream) {
    if (!ctx || !frame || !mask || !out_video || !out_masks) return fail(TCW_E_ARG, "tcw_resize_pack: NULL argument");
    if (n <= 0 || H <= 0 || W <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_resize_pack: bad shape");
    if ((size_t)h * w >= (1u << 31) || (size_t)H * W >= (1u << 31)) return fail(TCW_E_ARG, "tcw_resize_pack: image too large");
    const size_t total = (size_t)n * h * w;
    const size_t want = (total + 255) / 256;
    const int blocks = (int)(want < (size_t)ctx->sm_count * 8 ? want : (size_t)ctx->sm_count * 8);
    k_resize_pack<<<blocks, 256, 0, (cudaStream_t)cuda_stream>>>(frame, mask, n, H, W, h, w, (float)H / (float)h, (float)W / (float)w,
                                                               out_video, out_masks);
    LAUNCH_CHECK("k_resize_pack");
    return TCW_OK;
}

// ----------------------------------------------------------------------------------------------
// host-buffer entry: upload -> warp -> download, pipelined over three streams
// ----------------------------------------------------------------------------------------------
static int ensure_host_resources(tcw_ctx* ctx, size_t need) {
    if (!ctx->s_in) {
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_run, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
    }
    if (ctx->pool_bytes < need) {
        if (ctx->pool) { CUDA_TRY(cudaFree(ctx->pool)); ctx->pool = nullptr; ctx->pool_bytes = 0; }
        cudaError_t e = cudaMalloc(&ctx->pool, need);
        if (e != cudaSuccess) return fail(TCW_E_NOMEM, "tcw_forward_warp_host: cudaMalloc failed: %s", cudaGetErrorString(e));
        ctx->pool_bytes = need;
    }
    return TCW_OK;
}

TCW_API int tcw_forward_warp_host(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                                  const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                                  int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                                  int* range_flag) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_forward_warp_host: ctx is NULL");
    if (!frame || !depth || !M || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_forward_warp_host: NULL buffer");
    if (n_units <= 0 || n_src <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_forward_warp_host: bad shape");
    if (!src_index && n_src != n_units) return fail(TCW_E_ARG, "tcw_forward_warp_host: src_index is NULL but n_src != n_units");
    if (src_index) for (int u = 0; u < n_units; ++u)
        if (src_index[u] < 0 || src_index[u] >= n_src) return fail(TCW_E_ARG, "tcw_forward_warp_host: src_index out of range");
    int prev = 0;
    CUDA_TRY(cudaGetDevice(&prev));
    CUDA_TRY(cudaSetDevice(ctx->device));
    const size_t hw = (size_t)h * w, fb = sizeof(float);
    const size_t esz = (flags & TCW_CLEAN) ? sizeof(double) : sizeof(float);
    size_t ws_bytes = 0;
    int rc = tcw_workspace_bytes(ctx, n_units, h, w, flags, &ws_bytes);
    if (rc) return rc;
    // pool layout
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    const size_t o_frame = take((size_t)n_src * 3 * hw * fb), o_depth = take((size_t)n_src * hw * fb);
    const size_t o_mask1 = mask1 ? take((size_t)n_src * hw * fb) : 0;
    const size_t o_M = take((size_t)n_units * M_FLOATS * fb), o_si = take((size_t)n_units * sizeof(int));
    const size_t o_flag = take(256);
    const size_t o_of = take((size_t)n_units * 3 * hw * esz), o_om = take((size_t)n_units * hw * esz);
    const size_t o_od = out_depth ? take((size_t)n_units * hw * fb) : 0;
    const size_t o_ofl = out_flow ? take((size_t)n_units * 2 * hw * fb) : 0;
    const size_t o_ws = take(ws_bytes);
    rc = ensure_host_resources(ctx, off);
    if (rc) return rc;
    char* base = (char*)ctx->pool;
    float* d_frame = (float*)(base + o_frame); float* d_depth = (float*)(base + o_depth);
    float* d_mask1 = mask1 ? (float*)(base + o_mask1) : nullptr;
    float* d_M = (float*)(base + o_M); int* d_si = src_index ? (int*)(base + o_si) : nullptr;
    int* d_flag = (int*)(base + o_flag);
    char* d_of = base + o_of; char* d_om = base + o_om;
    float* d_od = out_depth ? (float*)(base + o_od) : nullptr;
    float* d_ofl = out_flow ? (float*)(base + o_ofl) : nullptr;

    // slices of units: upload the sources a slice needs, run it, download its outputs
    const int slice = auto_chunk_units(ctx, n_units, h, w, acc_bytes_per_px(flags)) * 2;
    const int n_slices = (n_units + slice - 1) / slice;
    cudaEvent_t* ev_in = (cudaEvent_t*)calloc((size_t)n_slices * 2, sizeof(cudaEvent_t));
    if (!ev_in) return fail(TCW_E_NOMEM, "tcw_forward_warp_host: out of host memory");
    cudaEvent_t* ev_run = ev_in + n_slices;
    for (int i = 0; i < 2 * n_slices; ++i) CUDA_TRY(cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming));

    CUDA_TRY(cudaMemcpyAsync(d_M, M, (size_t)n_units * M_FLOATS * fb, cudaMemcpyHostToDevice, ctx->s_in));
    if (src_index) CUDA_TRY(cudaMemcpyAsync(d_si, src_index, (size_t)n_units * sizeof(int), cudaMemcpyHostToDevice, ctx->s_in));
    CUDA_TRY(cudaMemsetAsync(d_flag, 0, sizeof(int), ctx->s_in));
    int uploaded = 0;   // sources [0, uploaded) are queued on s_in
    for (int i = 0; i < n_slices; ++i) {
        const int u0 = i * slice, nu = n_units - u0 < slice ? n_units - u0 : slice;
        int need = 0;
        for (int u = u0; u < u0 + nu; ++u) { int s = src_index ? src_index[u] : u; if (s + 1 > need) need = s + 1; }
        if (need > uploaded) {
            const size_t a = uploaded, n = need - uploaded;
            CUDA_TRY(cudaMemcpyAsync(d_frame + a * 3 * hw, frame + a * 3 * hw, n * 3 * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            CUDA_TRY(cudaMemcpyAsync(d_depth + a * hw, depth + a * hw, n * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            if (mask1) CUDA_TRY(cudaMemcpyAsync(d_mask1 + a * hw, mask1 + a * hw, n * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            uploaded = need;
        }
        CUDA_TRY(cudaEventRecord(ev_in[i], ctx->s_in));
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_run, ev_in[i], 0));
        if (!(flags & TCW_PER_UNIT_MAX) && i == 0 && uploaded < n_src) {
            // call-wide max needs every source before the first splat: upload the rest now
            const size_t a = uploaded, n = n_src - uploaded;
            CUDA_TRY(cudaMemcpyAsync(d_frame + a * 3 * hw, frame + a * 3 * hw, n * 3 * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            CUDA_TRY(cudaMemcpyAsync(d_depth + a * hw, depth + a * hw, n * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            if (mask1) CUDA_TRY(cudaMemcpyAsync(d_mask1 + a * hw, mask1 + a * hw, n * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            uploaded = n_src;
            CUDA_TRY(cudaEventRecord(ev_in[i], ctx->s_in));
            CUDA_TRY(cudaStreamWaitEvent(ctx->s_run, ev_in[i], 0));
        }
        if (flags & TCW_PER_UNIT_MAX) {
            const float *sf = d_frame, *sd = d_depth, *sm = d_mask1;
            int ns = n_src;
            if (!d_si) {   // identity mapping: unit u reads source u, so offset the source pointers
                sf += (size_t)u0 * 3 * hw; sd += (size_t)u0 * hw; if (sm) sm += (size_t)u0 * hw; ns = nu;
            }
            rc = tcw_forward_warp(ctx, sf, sm, sd, d_si ? d_si + u0 : nullptr, d_M + (size_t)u0 * M_FLOATS,
                                  nu, ns, h, w, flags,
                                  d_of + (size_t)u0 * 3 * hw * esz, d_om + (size_t)u0 * hw * esz,
                                  d_od ? d_od + (size_t)u0 * hw : nullptr, d_ofl ? d_ofl + (size_t)u0 * 2 * hw : nullptr,
                                  d_flag, base + o_ws, ws_bytes, ctx->s_run);
        } else if (i == 0) {
            rc = tcw_forward_warp(ctx, d_frame, d_mask1, d_depth, d_si, d_M, n_units, n_src, h, w, flags,
                                  d_of, d_om, d_od, d_ofl, d_flag, base + o_ws, ws_bytes, ctx->s_run);
        }
        if (rc) break;
        CUDA_TRY(cudaEventRecord(ev_run[i], ctx->s_run));
        if ((flags & TCW_PER_UNIT_MAX) || i == n_slices - 1) {
            const int d0 = (flags & TCW_PER_UNIT_MAX) ? u0 : 0;
            const int dn = (flags & TCW_PER_UNIT_MAX) ? nu : n_units;
            CUDA_TRY(cudaStreamWaitEvent(ctx->s_out, ev_run[i], 0));
            CUDA_TRY(cudaMemcpyAsync((char*)out_frame + (size_t)d0 * 3 * hw * esz, d_of + (size_t)d0 * 3 * hw * esz, (size_t)dn * 3 * hw * esz, cudaMemcpyDeviceToHost, ctx->s_out));
            CUDA_TRY(cudaMemcpyAsync((char*)out_mask + (size_t)d0 * hw * esz, d_om + (size_t)d0 * hw * esz, (size_t)dn * hw * esz, cudaMemcpyDeviceToHost, ctx->s_out));
            if (out_depth) CUDA_TRY(cudaMemcpyAsync(out_depth + (size_t)d0 * hw, d_od + (size_t)d0 * hw, (size_t)dn * hw * fb, cudaMemcpyDeviceToHost, ctx->s_out));
            if (out_flow) CUDA_TRY(cudaMemcpyAsync(out_flow + (size_t)d0 * 2 * hw, d_ofl + (size_t)d0 * 2 * hw, (size_t)dn * 2 * hw * fb, cudaMemcpyDeviceToHost, ctx->s_out));
        }
    }
    int flag_host = 0;
    if (!rc) {
        CUDA_TRY(cudaStreamSynchronize(ctx->s_run));
        CUDA_TRY(cudaMemcpyAsync(&flag_host, d_flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->s_out));
        CUDA_TRY(cudaStreamSynchronize(ctx->s_out));
        CUDA_TRY(cudaStreamSynchronize(ctx->s_in));
    } else {
        cudaDeviceSynchronize();
    }
    for (int i = 0; i < 2 * n_slices; ++i) cudaEventDestroy(ev_in[i]);
    free(ev_in);
    if (range_flag) *range_flag = flag_host;
    cudaSetDevice(prev);
    return rc;
}
