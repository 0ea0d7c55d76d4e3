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
    int chunk_units_override;   // TCW_CHUNK_UNITS env (0 = auto): units per chunk
    int lanes;                  // TCW_LANES env (0 = auto): streams of the multi-lane schedule
    int graphs_off;             // TCW_NO_GRAPH env
    // resources of the host-buffer entry point (grow-only)
    cudaStream_t s_in, s_run, s_out;
    cudaEvent_t* host_ev; int host_ev_n;       // events of the host entry, created once and reused
    cudaStream_t s_aux;         // pre-pass of the later rounds (atomic path)
    cudaEvent_t ev_fork, ev_aux;
    cudaStream_t s_lane[7];     // extra streams of the multi-lane schedule (lane 0 is the caller's stream)
    cudaEvent_t ev_lane[7];
    // CUDA graph of one whole tcw_forward_warp call (replayed while the arguments stay the same)
    struct GraphKey {
        const void *frame, *mask1, *depth, *src_index, *M, *out_frame, *out_mask, *out_depth, *out_flow, *range_flag, *workspace;
        size_t workspace_bytes;
        const void* dst_index;
        int n_units, n_src, h, w, flags, n_dst;
    } last_key, graph_key;
    int last_key_valid;
    cudaGraphExec_t graph_exec;
    long long graph_launches;   // kernels one replay launches
    cudaStream_t s_cap;         // origin stream of the capture (the caller's stream may be the legacy default stream)
    int win_smem_set;           // dynamic shared memory opt-in of k_splat_window done
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
        auto* p = cap > (1 << 17) ? nullptr : (tcw_ctx::ProfRec*)realloc(ctx->prof, sizeof(tcw_ctx::ProfRec) * cap);
        if (!p) { ctx->prof_on = 2; return; }      // cap reached or out of memory: stop recording; tcw_profile_read reports it
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
    if (ctx->prof_on == 2) {
        ctx->prof_on = 1;
        return fail(TCW_E_NOMEM, "tcw_profile_read: the launch log overflowed, the sums are incomplete");
    }
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
    // tuning knobs for measurements (schedule shape only; every setting runs the same kernels)
    const char* env = getenv("TCW_CHUNK_UNITS");
    c->chunk_units_override = env ? atoi(env) : 0;
    env = getenv("TCW_LANES");
    c->lanes = env ? atoi(env) : 0;
    env = getenv("TCW_NO_GRAPH");
    c->graphs_off = env ? atoi(env) : 0;
    *out = c;
    return TCW_OK;
}

TCW_API void tcw_destroy(tcw_ctx* c) {
    if (!c) return;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(c->device);
    if (c->pool) cudaFree(c->pool);
    for (int i = 0; i < c->prof_cap; ++i) { cudaEventDestroy(c->prof[i].e0); cudaEventDestroy(c->prof[i].e1); }
    free(c->prof);
    for (int i = 0; i < c->host_ev_n; ++i) cudaEventDestroy(c->host_ev[i]);
    free(c->host_ev);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_run) cudaStreamDestroy(c->s_run);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    if (c->graph_exec) cudaGraphExecDestroy(c->graph_exec);
    if (c->s_cap) cudaStreamDestroy(c->s_cap);
    for (int i = 0; i < 7; ++i) if (c->s_lane[i]) { cudaStreamDestroy(c->s_lane[i]); cudaEventDestroy(c->ev_lane[i]); }
    if (c->s_aux) { cudaStreamDestroy(c->s_aux); cudaEventDestroy(c->ev_fork); cudaEventDestroy(c->ev_aux); }
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
#include "warp_pair.cuh"
#include "warp_window.cuh"
#include "resize_pack.cuh"
#include "depth_post.cuh"
#include "pointcloud.cuh"

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

// standalone clean_points (models/utils.py:585-626)
__global__ void __launch_bounds__(TILE_THREADS)
k_clean_points(const float* __restrict__ frame_all, const float* __restrict__ mask_all, int c, int h, int w, int tiles_x,
               double* __restrict__ out_frame_all, double* __restrict__ out_mask_all) {
    const int x = (blockIdx.x % tiles_x) * TILE_W + (threadIdx.x % TILE_W);
    const int y = (blockIdx.x / tiles_x) * TILE_H + (threadIdx.x / TILE_W);
    if (x >= w || y >= h) return;
    const size_t hw = (size_t)h * w, p = (size_t)y * w + x, n = blockIdx.y;      // image n of the batch
    const float* frame = frame_all + n * c * hw;
    const float* mask2 = mask_all + n * hw;
    double* out_frame = out_frame_all + n * c * hw;
    double* out_mask = out_mask_all + n * hw;
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
// workspace header: 256 B of control words + per-unit stats (+1 for the whole call)
constexpr size_t CTRL_BYTES = 256;
static size_t header_bytes(int n_units) { return CTRL_BYTES + align_up((size_t)(n_units + 1) * sizeof(UnitStats), 256); }

// accumulator bytes per pixel of one unit (float-reduction mode): 5 planes x 4 B
static size_t acc_bytes_per_px(int) { return (size_t)NCH * 4; }

// lanes x units-per-chunk of the multi-stream schedules.  Units in flight = lanes * chunk: what they write and read
// again (accumulators or slabs, flow, z') should stay L2-resident next to the streaming traffic.  `bytes_per_px` is
// that working set per unit pixel.  Measured optimum of the atomic path on B200 at 384x672: 4 lanes x 3 units.
static void lanes_plan(const tcw_ctx* ctx, int n_units, size_t hw, size_t bytes_per_px, int* lanes, int* chunk) {
    int NL = ctx->lanes > 0 ? ctx->lanes : 4;
    if (NL > 8) NL = 8;
    int g2 = ctx->chunk_units_override;
    if (g2 <= 0) {
        const size_t l2 = ctx->l2_bytes ? ctx->l2_bytes : ((size_t)64 << 20);
        int inflight = (int)((l2 / 2) / (hw * bytes_per_px));
        if (inflight < 2) inflight = 2;
        if (ctx->lanes <= 0 && NL > inflight) NL = inflight;
        g2 = inflight / NL;
        if (g2 < 1) g2 = 1;
        if (g2 > 4096) g2 = 4096;
    }
    if (NL > n_units) NL = n_units;
    const int per_lane = (n_units + NL - 1) / NL;
    if (g2 > per_lane) g2 = per_lane;
    *lanes = NL; *chunk = g2;
}
// The window path's schedule shape (profiles/r2_schedule_sweep.txt, second table): what pays is full launches and enough
// rounds per lane to hide the tails, not L2 residency of the slabs -- 392 units at 384x672: 4 x 2 4.58 ms, 4 x 12 3.70, plateau
// 3.65 - 3.70 from 12 units per launch on; 196 units: 4 x 7 1.82 (4 x 2 2.10); 98 and 49 units: 4 x 4 (the expensive orbit presets
// gain 27 - 36 % over 4 x 2, the cheapest one loses 3 %).
// Rule: four lanes, about eight rounds per lane, at least four units per launch, at most ~3.7 Mpixel per launch (beyond that a
// launch is many waves anyway: one 1080p unit is eight 384x672 units' worth of CTAs).
static void window_plan(const tcw_ctx* ctx, int n_units, size_t hw, int* lanes, int* chunk) {
    int NL = ctx->lanes > 0 ? ctx->lanes : 4;
    if (NL > 8) NL = 8;
    if (NL > n_units) NL = n_units;
    int g2 = ctx->chunk_units_override;
    if (g2 <= 0) {
        const int cap = (int)(3700000 / hw) > 1 ? (int)(3700000 / hw) : 1;
        g2 = n_units / (NL * 8);
        if (g2 < 4) g2 = 4;                           // (whole 49-unit presets: 4 x 4 0.47 / 0.54 / 0.62 ms, 4 x 2 0.45 / 0.69 / 0.84)
        if (g2 > cap) g2 = cap;
        if (g2 > 4096) g2 = 4096;                     // (a launch's grid.y)
    }
    const int per_lane = (n_units + NL - 1) / NL;
    if (g2 > per_lane) g2 = per_lane;
    if (g2 < 1) g2 = 1;
    *lanes = NL; *chunk = g2;
}
constexpr size_t ATOMIC_BYTES_PER_PX = NCH * sizeof(float);

static int check_launch(tcw_ctx* ctx, const char* what) {
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(TCW_E_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
    return TCW_OK;
}
#define LAUNCH_CHECK(what) do { int rc__ = check_launch(ctx, what); if (rc__) return rc__; } while (0)

static int ensure_lanes(tcw_ctx* ctx, int NL) {
    if (!ctx->s_aux) {
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_aux, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_aux, cudaEventDisableTiming));
    }
    for (int i = 0; i + 1 < NL; ++i) if (!ctx->s_lane[i]) {
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_lane[i], cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_lane[i], cudaEventDisableTiming));
    }
    return TCW_OK;
}

// ----------------------------------------------------------------------------------------------
// window path (default): workspace layout and schedule
// ----------------------------------------------------------------------------------------------
struct WinLayout {
    int NL, g2, S, P;                 // lanes, units per chunk, splat slots, prep slots
    int strips_x, ns, nh, ng;
    long long pool_cells;
    size_t hdr, o_counter, o_flag, o_box, o_grp, o_z, o_flow, o_sbox, o_pile, o_pool4, o_poolw, o_spill4, o_spillw, o_tmp, total;
};
static void win_layout(const tcw_ctx* ctx, int n_units, int h, int w, int flags, WinLayout* L) {
    const size_t hw = (size_t)h * w;
    window_plan(ctx, n_units, hw, &L->NL, &L->g2);
    L->S = L->NL * L->g2;
    L->P = (flags & TCW_PER_UNIT_MAX) ? L->S : n_units;      // a call-wide normaliser needs every unit's pre-pass first
    L->strips_x = (w + 63) / 64;
    L->ns = L->strips_x * ((h + 7) / 8);
    L->nh = 2 * L->ns;
    L->ng = (L->nh + WIN_GROUP - 1) / WIN_GROUP;
    // slab pool of a unit in flight: one WIN_CAP slot per (strip, pass) for the shared-memory windows, then WIN_POOL_MULT * h * w
    // cells shared by the footprints that are too large for a window
    L->pool_cells = (long long)L->ns * 2 * WIN_CAP + (long long)WIN_POOL_MULT * (long long)hw + 4096;
    size_t off = header_bytes(n_units);
    L->hdr = off;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    L->o_counter = take((size_t)L->S * sizeof(int));
    L->o_flag = take((size_t)L->S * sizeof(int));
    L->o_box = take((size_t)L->P * L->nh * sizeof(WinBox));
    L->o_grp = take((size_t)L->P * L->ng * sizeof(WinGrp));
    L->o_z = take((size_t)L->P * hw * sizeof(float));
    L->o_flow = take((size_t)L->P * 2 * hw * sizeof(float));
    L->o_sbox = take((size_t)L->S * L->nh * sizeof(WinBox));
    L->o_pile = take((size_t)L->S * L->ns * 8 * sizeof(float));
    // each pool array is followed by 256 B of zeros: cell S * pool_cells is the "nothing here" cell of the branch-free gather
    L->o_pool4 = take((size_t)L->S * L->pool_cells * sizeof(float4) + 256);
    L->o_poolw = take((size_t)L->S * L->pool_cells * sizeof(float) + 256);
    L->o_spill4 = take((size_t)L->S * hw * sizeof(float4));
    L->o_spillw = take((size_t)L->S * hw * sizeof(float));
    // fp32 colour + mask ahead of clean_points; colour + mask + depth ahead of the fused resize + pack (TCW_PACK)
    L->o_tmp = (flags & (TCW_CLEAN | TCW_PACK)) ? take((size_t)L->S * ((flags & TCW_PACK) ? 5 : 4) * hw * sizeof(float)) : 0;
    L->total = off;
}

// What one window submission does besides the WinArgs sources: where the results go.
struct WinJob {
    int n_units, flags;
    float* out; int out_c;             // (n, out_c, h, w) fp32, or (with TCW_CLEAN) the float64 frame
    void* out_mask; float* out_depth;
    int is_image; int* range_flag;
    // TCW_PACK: the chunk's colour / mask go to scratch and k_resize_pack writes (3, n, pack_h, pack_w) / (1, n, pack_h, pack_w)
    int pack_h, pack_w; float* pack_video; float* pack_masks;
};

static int window_enqueue(tcw_ctx* ctx, WinArgs A, const WinJob& J, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    const int n_units = J.n_units, h = A.h, w = A.w;
    const size_t hw = (size_t)h * w;
    if (hw >= ((size_t)1 << 28)) return fail(TCW_E_ARG, "tcw: image too large for this build (h*w < 2^28)");
    if (h > 32000 || w > 32000) return fail(TCW_E_ARG, "tcw: image side too large for this build (< 32000)");
    WinLayout L;
    win_layout(ctx, n_units, h, w, J.flags, &L);
    if (!workspace || workspace_bytes < L.total) return fail(TCW_E_WORKSPACE, "tcw: workspace too small (see tcw_workspace_bytes)");
    if (!ctx->win_smem_set) {
        CUDA_TRY(cudaFuncSetAttribute(k_splat_window<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, WIN_WARPS * WIN_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(k_splat_window<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, WIN_WARPS * WIN_BYTES));
        ctx->win_smem_set = 1;
    }
    char* ws = (char*)workspace;
    const bool per_unit = J.flags & TCW_PER_UNIT_MAX, clean = J.flags & TCW_CLEAN, pack = J.flags & TCW_PACK;
    A.strips_x = L.strips_x; A.ns = L.ns; A.nh = L.nh; A.ng = L.ng;
    A.stats = (const UnitStats*)(ws + CTRL_BYTES); A.stats_rw = (UnitStats*)(ws + CTRL_BYTES);
    A.stat_index_global = per_unit ? -1 : n_units; A.n_units_total = n_units;
    A.counter = (int*)(ws + L.o_counter); A.spill_flag = (int*)(ws + L.o_flag);
    A.box = (WinBox*)(ws + L.o_box); A.grp = (WinGrp*)(ws + L.o_grp);
    A.zbuf = (float*)(ws + L.o_z); A.flowbuf = (float*)(ws + L.o_flow);
    A.sbox = (WinBox*)(ws + L.o_sbox); A.pile = (float*)(ws + L.o_pile);
    A.pool4 = (float4*)(ws + L.o_pool4); A.poolw = (float*)(ws + L.o_poolw); A.pool_cells = L.pool_cells; A.n_slots = L.S;
    A.spill4 = (float4*)(ws + L.o_spill4); A.spillw = (float*)(ws + L.o_spillw);
    A.status = J.range_flag;
    float* tmp = (clean || pack) ? (float*)(ws + L.o_tmp) : nullptr;
    const int NL = L.NL, g2 = L.g2;
    int rc = ensure_lanes(ctx, NL);
    if (rc) return rc;

    CUDA_TRY(cudaMemsetAsync(workspace, 0, L.hdr, st));                                       // per-unit stats
    CUDA_TRY(cudaMemsetAsync(ws + L.o_pool4 + (size_t)L.S * L.pool_cells * sizeof(float4), 0, 256, st));      // the zero cells behind the pools
    CUDA_TRY(cudaMemsetAsync(ws + L.o_poolw + (size_t)L.S * L.pool_cells * sizeof(float), 0, 256, st));
    // shared planes: all zero on entry (k_gather_normalise re-zeroes what it read)
    const int S_used = L.S < n_units ? L.S : n_units;
    CUDA_TRY(cudaMemsetAsync(ws + L.o_spill4, 0, (size_t)S_used * hw * sizeof(float4), st));
    CUDA_TRY(cudaMemsetAsync(ws + L.o_spillw, 0, (size_t)S_used * hw * sizeof(float), st));
    const int tiles_x = (w + 31) / 32, tiles_y = (h + 31) / 32;
    const dim3 g1(L.ng), g2d((L.ns + WIN_WARPS - 1) / WIN_WARPS), g3(tiles_x * tiles_y);
    auto prepass = [&](int u0, int nu, int prep0, int slot0, cudaStream_t s) -> int {
        for (int a = 0; a < nu; a += 65535) {
            const int na = nu - a < 65535 ? nu - a : 65535;
            prof_begin(ctx, PK_ZMAX, na, s);
            if (A.fused) k_win_prepass<true><<<dim3(g1.x, na), WIN_PRE_WARPS * 32, 0, s>>>(A, u0 + a, prep0 + a, slot0 < 0 ? -1 : slot0 + a);
            else k_win_prepass<false><<<dim3(g1.x, na), WIN_PRE_WARPS * 32, 0, s>>>(A, u0 + a, prep0 + a, slot0 < 0 ? -1 : slot0 + a);
            prof_end(ctx, s);
            LAUNCH_CHECK("k_win_prepass");
        }
        return TCW_OK;
    };
    if (!per_unit) {      // call-wide normaliser (models/utils.py:492 with b > 1): every unit's pre-pass comes first
        rc = prepass(0, n_units, 0, -1, st);
        if (rc) return rc;
    }
    CUDA_TRY(cudaEventRecord(ctx->ev_fork, st));
    for (int i = 0; i + 1 < NL; ++i) CUDA_TRY(cudaStreamWaitEvent(ctx->s_lane[i], ctx->ev_fork, 0));
    int k = 0;
    for (int u0 = 0; u0 < n_units; u0 += g2, ++k) {
        const int nu = n_units - u0 < g2 ? n_units - u0 : g2, b = k % NL;
        cudaStream_t sk = b ? ctx->s_lane[b - 1] : st;
        const int slot0 = b * g2, prep0 = per_unit ? slot0 : u0;
        if (per_unit) {
            rc = prepass(u0, nu, prep0, slot0, sk);
            if (rc) return rc;
        } else {
            CUDA_TRY(cudaMemsetAsync(A.counter + slot0, 0, (size_t)nu * sizeof(int), sk));
            CUDA_TRY(cudaMemsetAsync(A.spill_flag + slot0, 0, (size_t)nu * sizeof(int), sk));
        }
        prof_begin(ctx, PK_SPLAT, nu, sk);
        if (A.fused) k_splat_window<true><<<dim3(g2d.x, nu), WIN_WARPS * 32, WIN_WARPS * WIN_BYTES, sk>>>(A, u0, prep0, slot0);
        else k_splat_window<false><<<dim3(g2d.x, nu), WIN_WARPS * 32, WIN_WARPS * WIN_BYTES, sk>>>(A, u0, prep0, slot0);
        prof_end(ctx, sk);
        LAUNCH_CHECK("k_splat_window");
        WinNormArgs N;
        N.out_c = J.out_c; N.is_image = J.is_image; N.range_flag = J.range_flag; N.out_depth = J.out_depth;
        N.mask_u8 = (J.flags & TCW_MASK_U8) ? 1 : 0;
        if (clean) {
            // fp32 results of the chunk go to scratch (laid out as units u0.. of a (n,3,h,w) / (n,1,h,w) pair), clean_points makes them float64
            N.out_frame = tmp + (size_t)slot0 * 4 * hw - (size_t)u0 * J.out_c * hw;
            N.out_mask = tmp + (size_t)slot0 * 4 * hw + (size_t)nu * J.out_c * hw - (size_t)u0 * hw;
        } else if (pack) {
            // the chunk's results stay in scratch (L2-resident: a slot is reused every NL chunks), laid out as units u0.. of
            // (n,3,h,w) | (n,1,h,w) | (n,1,h,w); the depth plane only keeps the gather on its specialised store loop
            float* base = tmp + (size_t)slot0 * 5 * hw;
            N.out_frame = base - (size_t)u0 * 3 * hw;
            N.out_mask = base + (size_t)nu * 3 * hw - (size_t)u0 * hw;
            N.out_depth = base + (size_t)nu * 4 * hw - (size_t)u0 * hw;
        } else {
            N.out_frame = J.out; N.out_mask = (float*)J.out_mask;
        }
        prof_begin(ctx, PK_NORMALISE, nu, sk);
        const bool img3 = N.out_c == 3 && N.is_image && N.out_depth;
        if (A.dst_index) {
            if (img3) k_gather_normalise<true, true><<<dim3(g3.x, nu), 128, 0, sk>>>(A, N, u0, prep0, slot0, tiles_x);
            else k_gather_normalise<false, true><<<dim3(g3.x, nu), 128, 0, sk>>>(A, N, u0, prep0, slot0, tiles_x);
        } else if (img3) k_gather_normalise<true, false><<<dim3(g3.x, nu), 128, 0, sk>>>(A, N, u0, prep0, slot0, tiles_x);
        else k_gather_normalise<false, false><<<dim3(g3.x, nu), 128, 0, sk>>>(A, N, u0, prep0, slot0, tiles_x);
        prof_end(ctx, sk);
        LAUNCH_CHECK("k_gather_normalise");
        if (clean) {
            const int ctx_x = (w + TILE_W - 1) / TILE_W, cty = (h + TILE_H - 1) / TILE_H;
            const float* tf = tmp + (size_t)slot0 * 4 * hw;
            k_clean_points<<<dim3(ctx_x * cty, nu), TILE_THREADS, 0, sk>>>(tf, tf + (size_t)nu * J.out_c * hw, J.out_c, h, w, ctx_x,
                                                                          (double*)J.out + (size_t)u0 * J.out_c * hw, (double*)J.out_mask + (size_t)u0 * hw);
            LAUNCH_CHECK("k_clean_points");
        }
        if (pack) {      // demo.py:91-100,117-120 on the chunk's units while their colour / mask are L2-hot
            const float* tf = tmp + (size_t)slot0 * 5 * hw;
            const size_t total = (size_t)nu * J.pack_h * J.pack_w, want = (total + 255) / 256;
            const int blocks = (int)(want < (size_t)ctx->sm_count * 8 ? want : (size_t)ctx->sm_count * 8);
            k_resize_pack<<<blocks, 256, 0, sk>>>(tf, tf + (size_t)nu * 3 * hw, nu, h, w, J.pack_h, J.pack_w,
                                                  (float)h / (float)J.pack_h, (float)w / (float)J.pack_w,
                                                  J.pack_video, J.pack_masks, n_units, u0);
            LAUNCH_CHECK("k_resize_pack");
        }
    }
    for (int i = 0; i + 1 < NL; ++i) {
        CUDA_TRY(cudaEventRecord(ctx->ev_lane[i], ctx->s_lane[i]));
        CUDA_TRY(cudaStreamWaitEvent(st, ctx->ev_lane[i], 0));
    }
    return TCW_OK;
}

TCW_API int tcw_workspace_bytes(const tcw_ctx* ctx, int n_units, int h, int w, int flags, size_t* bytes) {
    if (!ctx || !bytes || n_units <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_workspace_bytes: bad argument");
    const size_t hw = (size_t)h * w;
    size_t need;
    if (flags & TCW_ATOMIC_F32) {
        int nl = 1, g2 = 1;
        lanes_plan(ctx, n_units, hw, ATOMIC_BYTES_PER_PX, &nl, &g2);
        need = header_bytes(n_units) + (size_t)nl * align_up((size_t)g2 * acc_bytes_per_px(flags) * hw, 256);
    } else {
        WinLayout L;
        win_layout(ctx, n_units, h, w, flags, &L);
        need = L.total;
    }
    *bytes = need;
    return TCW_OK;
}

// atomic fp32 path (TCW_ATOMIC_F32, opt-in): k_splat_pair reduces into L2 accumulators, k_normalise_pair reads them back
static int atomic_enqueue(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                          const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                          int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                          int* range_flag, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    const size_t hw = (size_t)h * w;
    const size_t hdr = header_bytes(n_units);
    const size_t unit_bytes = acc_bytes_per_px(flags) * hw;
    UnitStats* stats = (UnitStats*)((char*)workspace + CTRL_BYTES);
    char* accbase = (char*)workspace + hdr;
    const bool per_unit = flags & TCW_PER_UNIT_MAX;
    const int sidx = per_unit ? -1 : n_units;
    int NL = 1, g2 = 1;
    lanes_plan(ctx, n_units, hw, ATOMIC_BYTES_PER_PX, &NL, &g2);
    // shrink the plan until its lane buffers (each padded to 256 B) fit the caller's workspace
    auto fits = [&](int nl, int g) { return hdr + (size_t)nl * align_up((size_t)g * unit_bytes, 256) <= workspace_bytes; };
    while (g2 > 1 && !fits(NL, g2)) --g2;
    while (NL > 1 && !fits(NL, g2)) --NL;
    if (!fits(NL, g2)) return fail(TCW_E_WORKSPACE, "tcw_forward_warp: workspace too small (see tcw_workspace_bytes)");
    int rc = ensure_lanes(ctx, NL);
    if (rc) return rc;
    CUDA_TRY(cudaMemsetAsync(workspace, 0, hdr, st));
    // K0: soft-z normaliser of the first round of chunks here, the rest on its own stream under the first round
    const int quads = h * ((w + 3) / 4);
    int k0_bx = (quads + TILE_THREADS * 2 - 1) / (TILE_THREADS * 2);
    if (k0_bx < 1) k0_bx = 1;
    int k0_head = n_units;
    if (per_unit && !ctx->prof_on && NL * g2 < n_units && n_units < 65535) k0_head = NL * g2;
    for (int u0 = 0; u0 < k0_head; u0 += 65535) {
        const int nu = k0_head - u0 < 65535 ? k0_head - u0 : 65535;
        prof_begin(ctx, PK_ZMAX, nu, st);
        k_zmax_quad<<<dim3(k0_bx, nu), TILE_THREADS, 0, st>>>(depth, src_index, M, u0, n_src, h, w, stats, n_units, 1);
        prof_end(ctx, st);
        LAUNCH_CHECK("k_zmax_quad");
    }
    PairArgs B;
    B.frame = frame; B.mask1 = mask1; B.depth = depth; B.src_index = src_index; B.Mall = M; B.n_src = n_src; B.status = range_flag;
    B.h = h; B.w = w; B.strips_x = (w + PAIR_W - 1) / PAIR_W;
    B.strips_per_unit = B.strips_x * ((h + PAIR_H_CHUNK - 1) / PAIR_H_CHUNK);
    B.stats = stats; B.stat_index_global = sidx; B.out_flow = out_flow;
    const int pb = (B.strips_per_unit + PAIR_THREADS / 32 - 1) / (PAIR_THREADS / 32);
    const size_t buf_bytes = align_up((size_t)g2 * NCH * hw * sizeof(float), 256);      // buffer b: [g2][hw] float4 then [g2][hw] float; 16-byte aligned for odd h*w too
    // independent lanes: chunk k runs splat + normalise on stream k % NL with buffer k % NL; the normalise kernel
    // re-zeroes what it read, so one memset per call is enough
    CUDA_TRY(cudaMemsetAsync(accbase, 0, (size_t)NL * buf_bytes, st));
    CUDA_TRY(cudaEventRecord(ctx->ev_fork, st));
    for (int i = 0; i + 1 < NL; ++i) CUDA_TRY(cudaStreamWaitEvent(ctx->s_lane[i], ctx->ev_fork, 0));
    const bool k0_rest = k0_head < n_units;
    if (k0_rest) {
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_aux, ctx->ev_fork, 0));
        k_zmax_quad<<<dim3(k0_bx, n_units - k0_head), TILE_THREADS, 0, ctx->s_aux>>>(depth, src_index, M, k0_head, n_src, h, w, stats, n_units, 1);
        LAUNCH_CHECK("k_zmax_quad");
        CUDA_TRY(cudaEventRecord(ctx->ev_aux, ctx->s_aux));
    }
    int k = 0;
    for (int u0 = 0; u0 < n_units; u0 += g2, ++k) {
        const int nu = n_units - u0 < g2 ? n_units - u0 : g2, b = k % NL;
        cudaStream_t sk = b ? ctx->s_lane[b - 1] : st;
        if (k0_rest && k >= NL && k < 2 * NL) CUDA_TRY(cudaStreamWaitEvent(sk, ctx->ev_aux, 0));
        char* buf = accbase + (size_t)b * buf_bytes;
        B.acc4 = (float4*)buf;
        B.accw = (float*)(buf + (size_t)g2 * hw * sizeof(float4));
        prof_begin(ctx, PK_SPLAT, nu, sk);
        k_splat_pair<<<dim3(pb, nu), PAIR_THREADS, 0, sk>>>(B, u0);
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

static int forward_warp_enqueue(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                                const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                                int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                                int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream,
                                const int* dst_index = nullptr, int n_dst = 0) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_forward_warp: ctx is NULL");
    if (!frame || !depth || !M || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_forward_warp: NULL buffer");
    if (n_units <= 0 || n_src <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_forward_warp: bad shape");
    if (!src_index && n_src != n_units) return fail(TCW_E_ARG, "tcw_forward_warp: src_index is NULL but n_src != n_units");
    if ((flags & TCW_ATOMIC_F32) && (flags & TCW_CLEAN)) return fail(TCW_E_ARG, "tcw_forward_warp: TCW_ATOMIC_F32 has no fused clean_points");
    if (flags & TCW_PACK) return fail(TCW_E_ARG, "tcw_forward_warp: TCW_PACK belongs to tcw_workspace_bytes / tcw_forward_warp_pack");
    if (dst_index && (flags & (TCW_CLEAN | TCW_ATOMIC_F32))) return fail(TCW_E_ARG, "tcw_forward_warp_scatter: default mode without TCW_CLEAN only");
    if (dst_index && n_dst <= 0) return fail(TCW_E_ARG, "tcw_forward_warp_scatter: n_dst must be positive");
    if ((flags & TCW_MASK_U8) && (flags & (TCW_CLEAN | TCW_ATOMIC_F32)))
        return fail(TCW_E_ARG, "tcw_forward_warp: TCW_MASK_U8 needs the default mode without TCW_CLEAN");
    const size_t hw = (size_t)h * w;
    if (hw >= 0x7fffffffull / 4) return fail(TCW_E_ARG, "tcw_forward_warp: image too large for this build");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (flags & TCW_ATOMIC_F32) {
        if (!workspace || workspace_bytes < header_bytes(n_units) + acc_bytes_per_px(flags) * hw)
            return fail(TCW_E_WORKSPACE, "tcw_forward_warp: workspace too small (see tcw_workspace_bytes)");
        return atomic_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask, out_depth, out_flow,
                              range_flag, workspace, workspace_bytes, st);
    }
    WinArgs A;
    memset(&A, 0, sizeof(A));
    A.frame = frame; A.mask1 = mask1; A.depth = depth; A.src_index = src_index; A.Mall = M;
    A.fused = 1; A.nv = 3; A.n_src = n_src; A.h = h; A.w = w;
    A.flow_out = out_flow;
    A.dst_index = dst_index; A.n_dst = n_dst;
    WinJob J;
    memset(&J, 0, sizeof(J));
    J.n_units = n_units; J.flags = flags; J.out = (float*)out_frame; J.out_c = 3; J.out_mask = out_mask; J.out_depth = out_depth;
    J.is_image = 1; J.range_flag = range_flag;
    return window_enqueue(ctx, A, J, workspace, workspace_bytes, st);
}

// Public entry: enqueue, or replay the CUDA graph of the previous identical call.  One call is a few hundred kernel
// launches on four streams; enqueueing them costs the host about as long as the GPU needs to run them, so a call
// whose arguments equal those of the call before is captured once and replayed afterwards.  Kernels read the
// matrices, indices and images through the pointers at run time, so new CONTENTS need no new graph.
static int forward_warp_entry(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                              const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                              int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                              int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream,
                              const int* dst_index, int n_dst) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_forward_warp: ctx is NULL");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    bool eligible = !ctx->graphs_off && !ctx->prof_on;
    if (eligible) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) { cudaGetLastError(); eligible = false; }
    }
    if (!eligible)
        return forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream, dst_index, n_dst);
    tcw_ctx::GraphKey key;
    memset(&key, 0, sizeof(key));
    key.frame = frame; key.mask1 = mask1; key.depth = depth; key.src_index = src_index; key.M = M;
    key.out_frame = out_frame; key.out_mask = out_mask; key.out_depth = out_depth; key.out_flow = out_flow;
    key.range_flag = range_flag; key.workspace = workspace; key.workspace_bytes = workspace_bytes;
    key.n_units = n_units; key.n_src = n_src; key.h = h; key.w = w; key.flags = flags; key.dst_index = dst_index; key.n_dst = n_dst;
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
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream, dst_index, n_dst);
    // second identical call in a row: capture it on an internal stream, instantiate, launch on the caller's stream
    if (!ctx->s_cap) CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_cap, cudaStreamNonBlocking));
    if (ctx->graph_exec) { cudaGraphExecDestroy(ctx->graph_exec); ctx->graph_exec = nullptr; }
    const long long before = ctx->launches;
    CUDA_TRY(cudaStreamBeginCapture(ctx->s_cap, cudaStreamCaptureModeThreadLocal));
    const int rc = forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                        out_depth, out_flow, range_flag, workspace, workspace_bytes, ctx->s_cap, dst_index, n_dst);
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
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream, dst_index, n_dst);
    }
    cudaError_t ie = cudaGraphInstantiate(&ctx->graph_exec, graph, 0);
    if (ie == cudaSuccess) ie = cudaGraphUpload(ctx->graph_exec, st);      // pre-pay the first-launch upload
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) {
        cudaGetLastError();
        if (ctx->graph_exec) { cudaGraphExecDestroy(ctx->graph_exec); ctx->graph_exec = nullptr; }
        ctx->graphs_off = 1;
        return forward_warp_enqueue(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask,
                                    out_depth, out_flow, range_flag, workspace, workspace_bytes, cuda_stream, dst_index, n_dst);
    }
    ctx->graph_key = key;
    CUDA_TRY(cudaGraphLaunch(ctx->graph_exec, st));
    ctx->launches += ctx->graph_launches;
    return TCW_OK;
}

TCW_API int tcw_forward_warp(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                             const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                             int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                             int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    return forward_warp_entry(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask, out_depth,
                              out_flow, range_flag, workspace, workspace_bytes, cuda_stream, nullptr, 0);
}

// tcw_forward_warp whose unit u writes unit slot dst_index[u] of output tensors that hold n_dst units: a rank of a multi-GPU job
// can take ANY subset of a clip's units (e.g. a few frames of every trajectory, so that all ranks cost the same) and still store
// each result where it belongs in the assembled clip -- its own or, through tcw_ipc_open, another GPU's memory.
TCW_API int tcw_forward_warp_scatter(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                                     const int* src_index, const float* M, const int* dst_index, int n_units, int n_src, int n_dst,
                                     int h, int w, int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                                     int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!dst_index) return fail(TCW_E_ARG, "tcw_forward_warp_scatter: dst_index is NULL");
    return forward_warp_entry(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask, out_depth,
                              out_flow, range_flag, workspace, workspace_bytes, cuda_stream, dst_index, n_dst);
}

// The warp and the resize + pack lines behind it, chunk by chunk (demo.py:77-100, 117-120).  Plain enqueue (no graph replay).
TCW_API int tcw_forward_warp_pack(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                                  const int* src_index, const float* M, int n_units, int n_src, int H, int W, int h, int w,
                                  int flags, float* out_video, float* out_masks, int* range_flag,
                                  void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_forward_warp_pack: ctx is NULL");
    if (!frame || !depth || !M || !out_video || !out_masks) return fail(TCW_E_ARG, "tcw_forward_warp_pack: NULL buffer");
    if (n_units <= 0 || n_src <= 0 || H <= 0 || W <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_forward_warp_pack: bad shape");
    if (!src_index && n_src != n_units) return fail(TCW_E_ARG, "tcw_forward_warp_pack: src_index is NULL but n_src != n_units");
    if (flags & ~TCW_PER_UNIT_MAX) return fail(TCW_E_ARG, "tcw_forward_warp_pack: flags may only hold TCW_PER_UNIT_MAX");
    if ((size_t)H * W >= 0x7fffffffull / 4 || (size_t)h * w >= (1u << 31)) return fail(TCW_E_ARG, "tcw_forward_warp_pack: image too large for this build");
    WinArgs A;
    memset(&A, 0, sizeof(A));
    A.frame = frame; A.mask1 = mask1; A.depth = depth; A.src_index = src_index; A.Mall = M;
    A.fused = 1; A.nv = 3; A.n_src = n_src; A.h = H; A.w = W;
    A.flow_out = nullptr;                       // flow12 stays in the workspace
    WinJob J;
    memset(&J, 0, sizeof(J));
    J.n_units = n_units; J.flags = flags | TCW_PACK; J.out_c = 3; J.is_image = 1; J.range_flag = range_flag;
    J.pack_h = h; J.pack_w = w; J.pack_video = out_video; J.pack_masks = out_masks;
    return window_enqueue(ctx, A, J, workspace, workspace_bytes, (cudaStream_t)cuda_stream);
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

// bilinear_splatting (models/utils.py:422-583) on the window kernels: the pre-pass only derives the boxes and the
// normaliser from the given flow / depth, K2 / K3 are the ones the forward warp uses.
TCW_API int tcw_splat(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                      const float* flow, const float* flow_mask, int b, int c, int h, int w, int flags,
                      float* out, float* out_mask, int* range_flag, void* workspace, size_t workspace_bytes,
                      void* cuda_stream) {
    if (!ctx || !frame || !depth || !flow || !out || !out_mask) return fail(TCW_E_ARG, "tcw_splat: NULL argument");
    if (b <= 0 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_splat: bad shape");
    if (c < 1 || c > 4) return fail(TCW_E_ARG, "tcw_splat: c must be 1..4");
    if (flags & (TCW_CLEAN | TCW_ATOMIC_F32)) return fail(TCW_E_ARG, "tcw_splat: unsupported flag");
    WinArgs A;
    memset(&A, 0, sizeof(A));
    A.frame = frame; A.mask1 = mask1; A.depth = depth; A.gflow = flow; A.gflow_mask = flow_mask;
    A.fused = 0; A.nv = c; A.n_src = b; A.h = h; A.w = w;
    WinJob J;
    memset(&J, 0, sizeof(J));
    J.n_units = b; J.flags = flags & TCW_PER_UNIT_MAX; J.out = out; J.out_c = c; J.out_mask = out_mask; J.out_depth = nullptr;
    J.is_image = (flags & TCW_IS_IMAGE) ? 1 : 0; J.range_flag = range_flag;
    return window_enqueue(ctx, A, J, workspace, workspace_bytes, (cudaStream_t)cuda_stream);
}

TCW_API int tcw_clean_points(tcw_ctx* ctx, const float* frame, const float* mask2, int c, int h, int w,
                             double* out_frame, double* out_mask, void* cuda_stream) {
    if (!ctx || !frame || !mask2 || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_clean_points: NULL argument");
    if (c < 1 || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_clean_points: bad shape");
    const int tiles_x = (w + TILE_W - 1) / TILE_W, tiles_y = (h + TILE_H - 1) / TILE_H;
    k_clean_points<<<dim3(tiles_x * tiles_y, 1), TILE_THREADS, 0, (cudaStream_t)cuda_stream>>>(frame, mask2, c, h, w, tiles_x, out_frame, out_mask);
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
                                                               out_video, out_masks, n, 0);
    LAUNCH_CHECK("k_resize_pack");
    return TCW_OK;
}

// ----------------------------------------------------------------------------------------------
// global point-cloud renderer (reference notebooks/06_10_25_vggt/warper_point_cloud.py)
// ----------------------------------------------------------------------------------------------
TCW_API int tcw_pc_lift(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth, const float* Minv,
                        int b, int h, int w, int step, float* out_points, float* out_colors, void* cuda_stream) {
    if (!ctx || !frame || !depth || !Minv || !out_points || !out_colors) return fail(TCW_E_ARG, "tcw_pc_lift: NULL argument");
    if (b <= 0 || b > 65535 || h <= 0 || w <= 0 || step < 1) return fail(TCW_E_ARG, "tcw_pc_lift: bad shape");
    const size_t hw = (size_t)h * w;
    k_pc_lift<<<dim3((unsigned)((hw + 255) / 256), b), 256, 0, (cudaStream_t)cuda_stream>>>(frame, mask1, depth, Minv, h, w, step, out_points, out_colors);
    LAUNCH_CHECK("k_pc_lift");
    return TCW_OK;
}

static int pc_grid(const tcw_ctx* ctx, long long n) {
    const long long want = (n + 255) / 256, cap = (long long)ctx->sm_count * 8;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

static int pc_zbuffer_impl(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                           long long n_points, int h, int w, int point_size, int first_wins,
                           float* out_frame, float* out_mask, float* out_depth,
                           void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!ctx || !camera || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer: NULL argument");
    if (n_points < 0 || (n_points > 0 && (!points || !colors)) || h <= 0 || w <= 0 || point_size < 0 || point_size > 64)
        return fail(TCW_E_ARG, "tcw_pc_render_zbuffer: bad argument");
    const int radius = point_size / 2, n_splats = (2 * radius + 1) * (2 * radius + 1);
    if (n_points * (long long)n_splats >= 0xffffffffll) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer: too many footprint entries (< 2^32)");
    const size_t hw = (size_t)h * w;
    if (hw >= 0x7fffffffull) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer: image too large");
    if (!workspace || workspace_bytes < hw * sizeof(unsigned long long)) return fail(TCW_E_WORKSPACE, "tcw_pc_render_zbuffer: workspace needs h*w*8 bytes");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    PcCamera C;
    memcpy(C.T, camera, sizeof(C.T));
    memcpy(C.K, camera + 12, sizeof(C.K));
    unsigned long long* zbuf = (unsigned long long*)workspace;
    CUDA_TRY(cudaMemsetAsync(zbuf, 0xff, hw * sizeof(unsigned long long), st));
    if (n_points > 0) {
        k_pc_zbuffer<<<pc_grid(ctx, n_points), 256, 0, st>>>(points, C, n_points, h, w, point_size, first_wins, zbuf);
        LAUNCH_CHECK("k_pc_zbuffer");
    }
    k_pc_resolve<<<(unsigned)((hw + 255) / 256), 256, 0, st>>>(zbuf, colors, (int)hw, n_splats, first_wins, out_frame, out_mask, out_depth);
    LAUNCH_CHECK("k_pc_resolve");
    return TCW_OK;
}

TCW_API int tcw_pc_render_zbuffer(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                                  long long n_points, int h, int w, int point_size,
                                  float* out_frame, float* out_mask, float* out_depth,
                                  void* workspace, size_t workspace_bytes, void* cuda_stream) {
    return pc_zbuffer_impl(ctx, points, colors, camera, n_points, h, w, point_size, 0, out_frame, out_mask, out_depth,
                           workspace, workspace_bytes, cuda_stream);
}

// render_pointcloud_zbuffer_vectorized_fixed (:650-765): among equal depths the FIRST entry stays; point_size <= 1 clamps
TCW_API int tcw_pc_render_zbuffer_first(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                                        long long n_points, int h, int w, int point_size,
                                        float* out_frame, float* out_mask, float* out_depth,
                                        void* workspace, size_t workspace_bytes, void* cuda_stream) {
    return pc_zbuffer_impl(ctx, points, colors, camera, n_points, h, w, point_size <= 1 ? 0 : point_size, 1, out_frame, out_mask,
                           out_depth, workspace, workspace_bytes, cuda_stream);
}

TCW_API int tcw_pc_render_bilinear(tcw_ctx* ctx, const float* points, const float* colors, const float* weights, const float* camera,
                                   long long n_points, int h, int w, float depth_threshold,
                                   float* out_frame, float* out_mask, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!ctx || !camera || !out_frame || !out_mask) return fail(TCW_E_ARG, "tcw_pc_render_bilinear: NULL argument");
    if (n_points < 0 || (n_points > 0 && (!points || !colors || !weights)) || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_pc_render_bilinear: bad argument");
    const size_t hw = (size_t)h * w;
    if (hw >= 0x7fffffffull) return fail(TCW_E_ARG, "tcw_pc_render_bilinear: image too large");
    if (!workspace || workspace_bytes < 256 + hw * sizeof(float4)) return fail(TCW_E_WORKSPACE, "tcw_pc_render_bilinear: workspace needs 256 + h*w*16 bytes");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    PcCamera C;
    memcpy(C.T, camera, sizeof(C.T));
    memcpy(C.K, camera + 12, sizeof(C.K));
    unsigned* mm = (unsigned*)workspace;
    float4* acc = (float4*)((char*)workspace + 256);
    CUDA_TRY(cudaMemsetAsync(mm, 0xff, sizeof(unsigned), st));
    CUDA_TRY(cudaMemsetAsync(mm + 1, 0, 2 * sizeof(unsigned), st));
    CUDA_TRY(cudaMemsetAsync(acc, 0, hw * sizeof(float4), st));
    if (n_points > 0) {
        k_pc_range<<<pc_grid(ctx, n_points), 256, 0, st>>>(points, C, n_points, h, w, depth_threshold, mm);
        LAUNCH_CHECK("k_pc_range");
        k_pc_bilinear<<<pc_grid(ctx, n_points), 256, 0, st>>>(points, colors, weights, C, n_points, h, w, depth_threshold, mm, acc);
        LAUNCH_CHECK("k_pc_bilinear");
    }
    k_pc_normalise<<<(unsigned)((hw + 255) / 256), 256, 0, st>>>(acc, (int)hw, 1e-8f, out_frame, out_mask);
    LAUNCH_CHECK("k_pc_normalise");
    return TCW_OK;
}

// render_from_camera(use_bilinear=False) for one camera: disc stamps with Gaussian falloff (:258-349)
TCW_API int tcw_pc_render_disc(tcw_ctx* ctx, const float* points, const float* colors, const float* weights, const float* camera,
                               long long n_points, int h, int w, float depth_threshold,
                               const int* offsets_xy, const float* splat_weights, int n_splats,
                               float* out_frame, float* out_mask, void* workspace, size_t workspace_bytes, void* cuda_stream) {
    if (!ctx || !camera || !out_frame || !out_mask || !offsets_xy || !splat_weights) return fail(TCW_E_ARG, "tcw_pc_render_disc: NULL argument");
    if (n_points < 0 || (n_points > 0 && (!points || !colors || !weights)) || h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_pc_render_disc: bad argument");
    if (n_splats < 1 || n_splats > 128) return fail(TCW_E_ARG, "tcw_pc_render_disc: 1..128 disc offsets (point_size < 12)");
    const size_t hw = (size_t)h * w;
    if (hw >= 0x7fffffffull) return fail(TCW_E_ARG, "tcw_pc_render_disc: image too large");
    if (!workspace || workspace_bytes < 256 + hw * sizeof(float4)) return fail(TCW_E_WORKSPACE, "tcw_pc_render_disc: workspace needs 256 + h*w*16 bytes");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    PcCamera C;
    memcpy(C.T, camera, sizeof(C.T));
    memcpy(C.K, camera + 12, sizeof(C.K));
    PcDisc D;
    memset(&D, 0, sizeof(D));
    D.n = n_splats;
    for (int j = 0; j < n_splats; ++j) {
        const int dx = offsets_xy[2 * j], dy = offsets_xy[2 * j + 1];
        if (dx < -127 || dx > 127 || dy < -127 || dy > 127) return fail(TCW_E_ARG, "tcw_pc_render_disc: offset out of range");
        D.dx[j] = (signed char)dx; D.dy[j] = (signed char)dy; D.w[j] = splat_weights[j];
    }
    unsigned* mm = (unsigned*)workspace;
    float4* acc = (float4*)((char*)workspace + 256);
    CUDA_TRY(cudaMemsetAsync(mm, 0xff, sizeof(unsigned), st));
    CUDA_TRY(cudaMemsetAsync(mm + 1, 0, 2 * sizeof(unsigned), st));
    CUDA_TRY(cudaMemsetAsync(acc, 0, hw * sizeof(float4), st));
    if (n_points > 0) {
        k_pc_range<<<pc_grid(ctx, n_points), 256, 0, st>>>(points, C, n_points, h, w, depth_threshold, mm);
        LAUNCH_CHECK("k_pc_range");
        k_pc_disc<<<pc_grid(ctx, n_points), 256, 0, st>>>(points, colors, weights, C, D, n_points, h, w, depth_threshold, mm, acc);
        LAUNCH_CHECK("k_pc_disc");
    }
    k_pc_normalise<<<(unsigned)((hw + 255) / 256), 256, 0, st>>>(acc, (int)hw, 1e-6f, out_frame, out_mask);
    LAUNCH_CHECK("k_pc_normalise");
    return TCW_OK;
}

// depth post-processing + trajectory radius on the device (models/infer.py:79-92, demo.py:539-543)
TCW_API int tcw_depth_post(tcw_ctx* ctx, const float* res, int n, int h, int w, int c, float near_v, float far_v,
                           float radius_scale, float* out_depth, float* stats, void* cuda_stream) {
    if (!ctx || !res || !out_depth || !stats) return fail(TCW_E_ARG, "tcw_depth_post: NULL argument");
    if (n <= 0 || h <= 0 || w <= 0 || c < 1 || c > 4) return fail(TCW_E_ARG, "tcw_depth_post: bad shape");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const size_t n_px = (size_t)n * h * w;
    unsigned* mm = (unsigned*)stats;
    CUDA_TRY(cudaMemsetAsync(mm, 0xff, sizeof(unsigned), st));          // running minimum starts at the largest code
    CUDA_TRY(cudaMemsetAsync(mm + 1, 0, sizeof(unsigned), st));
    const size_t want = (n_px + 255) / 256;
    const int blocks = (int)(want < (size_t)ctx->sm_count * 8 ? want : (size_t)ctx->sm_count * 8);
    k_depth_minmax<<<blocks, 256, 0, st>>>(res, n_px, c, mm);
    LAUNCH_CHECK("k_depth_minmax");
    k_depth_post<<<blocks, 256, 0, st>>>(res, n_px, c, mm, near_v, far_v, out_depth);
    LAUNCH_CHECK("k_depth_post");
    k_depth_stats<<<1, 32, 0, st>>>(mm, out_depth, h, w, radius_scale);
    LAUNCH_CHECK("k_depth_stats");
    return TCW_OK;
}

// ----------------------------------------------------------------------------------------------
// peer-visible buffers (multi-GPU exchange)
// ----------------------------------------------------------------------------------------------
struct DevGuard {
    int prev; bool ok;
    explicit DevGuard(int dev) : prev(0), ok(false) { ok = cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess; }
    ~DevGuard() { cudaSetDevice(prev); }
};

TCW_API int tcw_ipc_alloc(tcw_ctx* ctx, size_t bytes, void** ptr, unsigned char* handle) {
    if (!ctx || !ptr || !handle || bytes == 0) return fail(TCW_E_ARG, "tcw_ipc_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == TCW_IPC_HANDLE_BYTES, "IPC handle size");
    DevGuard g(ctx->device);
    if (!g.ok) return fail(TCW_E_CUDA, "tcw_ipc_alloc: cannot select the device");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return fail(TCW_E_NOMEM, "tcw_ipc_alloc: cudaMalloc failed: %s", cudaGetErrorString(e));
    cudaIpcMemHandle_t hnd;
    e = cudaIpcGetMemHandle(&hnd, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(TCW_E_CUDA, "tcw_ipc_alloc: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e)); }
    memcpy(handle, &hnd, sizeof(hnd));
    *ptr = p;
    return TCW_OK;
}
TCW_API int tcw_ipc_free(tcw_ctx* ctx, void* ptr) {
    if (!ctx || !ptr) return fail(TCW_E_ARG, "tcw_ipc_free: bad argument");
    DevGuard g(ctx->device);
    CUDA_TRY(cudaFree(ptr));
    return TCW_OK;
}
TCW_API int tcw_ipc_open(tcw_ctx* ctx, const unsigned char* handle, void** ptr) {
    if (!ctx || !handle || !ptr) return fail(TCW_E_ARG, "tcw_ipc_open: bad argument");
    DevGuard g(ctx->device);
    if (!g.ok) return fail(TCW_E_CUDA, "tcw_ipc_open: cannot select the device");
    cudaIpcMemHandle_t hnd;
    memcpy(&hnd, handle, sizeof(hnd));
    CUDA_TRY(cudaIpcOpenMemHandle(ptr, hnd, cudaIpcMemLazyEnablePeerAccess));
    return TCW_OK;
}
TCW_API int tcw_ipc_close(tcw_ctx* ctx, void* ptr) {
    if (!ctx || !ptr) return fail(TCW_E_ARG, "tcw_ipc_close: bad argument");
    DevGuard g(ctx->device);
    CUDA_TRY(cudaIpcCloseMemHandle(ptr));
    return TCW_OK;
}

// ----------------------------------------------------------------------------------------------
// host-buffer entry: upload -> warp -> download, pipelined over three streams
// ----------------------------------------------------------------------------------------------
__global__ void k_mask_to_u8(const float* __restrict__ m, unsigned char* __restrict__ o, size_t n) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        const float4 v = *(const float4*)(m + i);
        *(uchar4*)(o + i) = make_uchar4(v.x > 0.f, v.y > 0.f, v.z > 0.f, v.w > 0.f);
    } else {
        for (size_t j = i; j < n; ++j) o[j] = m[j] > 0.f;
    }
}

static int ensure_host_resources(tcw_ctx* ctx, size_t need, int n_events) {
    if (!ctx->s_in) {
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_run, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
    }
    if (ctx->host_ev_n < n_events) {      // created once, reused by every later call
        cudaEvent_t* p = (cudaEvent_t*)realloc(ctx->host_ev, sizeof(cudaEvent_t) * n_events);
        if (!p) return fail(TCW_E_NOMEM, "tcw_forward_warp_host: out of host memory");
        ctx->host_ev = p;
        while (ctx->host_ev_n < n_events) {
            CUDA_TRY(cudaEventCreateWithFlags(&ctx->host_ev[ctx->host_ev_n], cudaEventDisableTiming));
            ctx->host_ev_n++;
        }
    }
    if (ctx->pool_bytes < need) {
        if (ctx->pool) { CUDA_TRY(cudaFree(ctx->pool)); ctx->pool = nullptr; ctx->pool_bytes = 0; }
        cudaError_t e = cudaMalloc(&ctx->pool, need);
        if (e != cudaSuccess) return fail(TCW_E_NOMEM, "tcw_forward_warp_host: cudaMalloc failed: %s", cudaGetErrorString(e));
        ctx->pool_bytes = need;
    }
    return TCW_OK;
}

static int host_body(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                     const int* src_index, const float* M, int n_units, int n_src, int h, int w,
                     int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow, int* flag_host) {
    const size_t hw = (size_t)h * w, fb = sizeof(float);
    const bool mask_u8 = flags & TCW_HOST_MASK_U8;
    const int kflags = flags & ~TCW_HOST_MASK_U8;
    const size_t esz = (flags & TCW_CLEAN) ? sizeof(double) : sizeof(float);
    // slices of units: upload the sources a slice needs, run it, download its outputs
    int slice = n_units;
    if (flags & TCW_PER_UNIT_MAX) {
        slice = (n_units + 11) / 12;
        if (slice < 8) slice = 8;
        if (slice > n_units) slice = n_units;
    }
    const int n_slices = (n_units + slice - 1) / slice;
    size_t ws_bytes = 0;
    int rc = tcw_workspace_bytes(ctx, slice, h, w, kflags, &ws_bytes);
    if (rc) return rc;
    {   // the last slice may be shorter and plan its lanes differently: take the larger requirement
        size_t b2 = 0;
        const int last = n_units - (n_slices - 1) * slice;
        rc = tcw_workspace_bytes(ctx, last, h, w, kflags, &b2);
        if (rc) return rc;
        if (b2 > ws_bytes) ws_bytes = b2;
    }
    // pool layout
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    const size_t o_frame = take((size_t)n_src * 3 * hw * fb), o_depth = take((size_t)n_src * hw * fb);
    const size_t o_mask1 = mask1 ? take((size_t)n_src * hw * fb) : 0;
    const size_t o_M = take((size_t)n_units * M_FLOATS * fb), o_si = take((size_t)n_units * sizeof(int));
    const size_t o_flag = take(256);
    const size_t o_of = take((size_t)n_units * 3 * hw * esz), o_om = take((size_t)n_units * hw * esz);
    const size_t o_om8 = mask_u8 ? take((size_t)n_units * hw) : 0;
    const size_t o_od = out_depth ? take((size_t)n_units * hw * fb) : 0;
    const size_t o_ofl = out_flow ? take((size_t)n_units * 2 * hw * fb) : 0;
    const size_t o_ws = take(ws_bytes);
    rc = ensure_host_resources(ctx, off, 2 * n_slices);
    if (rc) return rc;
    char* base = (char*)ctx->pool;
    float* d_frame = (float*)(base + o_frame); float* d_depth = (float*)(base + o_depth);
    float* d_mask1 = mask1 ? (float*)(base + o_mask1) : nullptr;
    float* d_M = (float*)(base + o_M); int* d_si = src_index ? (int*)(base + o_si) : nullptr;
    int* d_flag = (int*)(base + o_flag);
    char* d_of = base + o_of; char* d_om = base + o_om;
    unsigned char* d_om8 = mask_u8 ? (unsigned char*)(base + o_om8) : nullptr;
    float* d_od = out_depth ? (float*)(base + o_od) : nullptr;
    float* d_ofl = out_flow ? (float*)(base + o_ofl) : nullptr;
    cudaEvent_t* ev_in = ctx->host_ev;
    cudaEvent_t* ev_run = ctx->host_ev + n_slices;

    CUDA_TRY(cudaMemcpyAsync(d_M, M, (size_t)n_units * M_FLOATS * fb, cudaMemcpyHostToDevice, ctx->s_in));
    if (src_index) CUDA_TRY(cudaMemcpyAsync(d_si, src_index, (size_t)n_units * sizeof(int), cudaMemcpyHostToDevice, ctx->s_in));
    CUDA_TRY(cudaMemsetAsync(d_flag, 0, sizeof(int), ctx->s_in));
    int uploaded = 0;   // sources [0, uploaded) are queued on s_in
    auto upload_to = [&](int need) -> int {
        if (need > uploaded) {
            const size_t a = uploaded, n = need - uploaded;
            CUDA_TRY(cudaMemcpyAsync(d_frame + a * 3 * hw, frame + a * 3 * hw, n * 3 * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            CUDA_TRY(cudaMemcpyAsync(d_depth + a * hw, depth + a * hw, n * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            if (mask1) CUDA_TRY(cudaMemcpyAsync(d_mask1 + a * hw, mask1 + a * hw, n * hw * fb, cudaMemcpyHostToDevice, ctx->s_in));
            uploaded = need;
        }
        return TCW_OK;
    };
    for (int i = 0; i < n_slices; ++i) {
        const int u0 = i * slice, nu = n_units - u0 < slice ? n_units - u0 : slice;
        int need = 0;
        if (flags & TCW_PER_UNIT_MAX) {
            for (int u = u0; u < u0 + nu; ++u) { int s = src_index ? src_index[u] : u; if (s + 1 > need) need = s + 1; }
        } else {
            need = n_src;      // a call-wide normaliser needs every source before the first splat
        }
        rc = upload_to(need);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(ev_in[i], ctx->s_in));
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_run, ev_in[i], 0));
        const float *sf = d_frame, *sd = d_depth, *sm = d_mask1;
        int ns = n_src;
        if (!d_si) {   // identity mapping: unit u reads source u, so offset the source pointers
            sf += (size_t)u0 * 3 * hw; sd += (size_t)u0 * hw; if (sm) sm += (size_t)u0 * hw; ns = nu;
        }
        rc = tcw_forward_warp(ctx, sf, sm, sd, d_si ? d_si + u0 : nullptr, d_M + (size_t)u0 * M_FLOATS,
                              nu, ns, h, w, kflags,
                              d_of + (size_t)u0 * 3 * hw * esz, d_om + (size_t)u0 * hw * esz,
                              d_od ? d_od + (size_t)u0 * hw : nullptr, d_ofl ? d_ofl + (size_t)u0 * 2 * hw : nullptr,
                              d_flag, base + o_ws, ws_bytes, ctx->s_run);
        if (rc) return rc;
        if (mask_u8) {
            const size_t n = (size_t)nu * hw;
            k_mask_to_u8<<<(unsigned)((n / 4 + 256) / 256), 256, 0, ctx->s_run>>>((const float*)d_om + (size_t)u0 * hw, d_om8 + (size_t)u0 * hw, n);
            LAUNCH_CHECK("k_mask_to_u8");
        }
        CUDA_TRY(cudaEventRecord(ev_run[i], ctx->s_run));
        CUDA_TRY(cudaStreamWaitEvent(ctx->s_out, ev_run[i], 0));
        CUDA_TRY(cudaMemcpyAsync((char*)out_frame + (size_t)u0 * 3 * hw * esz, d_of + (size_t)u0 * 3 * hw * esz, (size_t)nu * 3 * hw * esz, cudaMemcpyDeviceToHost, ctx->s_out));
        if (mask_u8) CUDA_TRY(cudaMemcpyAsync((unsigned char*)out_mask + (size_t)u0 * hw, d_om8 + (size_t)u0 * hw, (size_t)nu * hw, cudaMemcpyDeviceToHost, ctx->s_out));
        else CUDA_TRY(cudaMemcpyAsync((char*)out_mask + (size_t)u0 * hw * esz, d_om + (size_t)u0 * hw * esz, (size_t)nu * hw * esz, cudaMemcpyDeviceToHost, ctx->s_out));
        if (out_depth) CUDA_TRY(cudaMemcpyAsync(out_depth + (size_t)u0 * hw, d_od + (size_t)u0 * hw, (size_t)nu * hw * fb, cudaMemcpyDeviceToHost, ctx->s_out));
        if (out_flow) CUDA_TRY(cudaMemcpyAsync(out_flow + (size_t)u0 * 2 * hw, d_ofl + (size_t)u0 * 2 * hw, (size_t)nu * 2 * hw * fb, cudaMemcpyDeviceToHost, ctx->s_out));
    }
    CUDA_TRY(cudaStreamSynchronize(ctx->s_run));
    CUDA_TRY(cudaMemcpyAsync(flag_host, d_flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->s_out));
    CUDA_TRY(cudaStreamSynchronize(ctx->s_out));
    CUDA_TRY(cudaStreamSynchronize(ctx->s_in));
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
    if ((flags & TCW_HOST_MASK_U8) && (flags & TCW_CLEAN)) return fail(TCW_E_ARG, "tcw_forward_warp_host: TCW_HOST_MASK_U8 excludes TCW_CLEAN");
    if (src_index) for (int u = 0; u < n_units; ++u)
        if (src_index[u] < 0 || src_index[u] >= n_src) return fail(TCW_E_ARG, "tcw_forward_warp_host: src_index out of range");
    int prev = 0;
    CUDA_TRY(cudaGetDevice(&prev));
    CUDA_TRY(cudaSetDevice(ctx->device));
    int flag_host = 0;
    const int rc = host_body(ctx, frame, mask1, depth, src_index, M, n_units, n_src, h, w, flags, out_frame, out_mask, out_depth, out_flow, &flag_host);
    if (rc) {
        // nothing of this call may still be in flight against the caller's buffers when it returns
        char keep[sizeof(g_err)];
        memcpy(keep, g_err, sizeof(keep));
        if (ctx->s_in) { cudaStreamSynchronize(ctx->s_in); cudaStreamSynchronize(ctx->s_run); cudaStreamSynchronize(ctx->s_out); }
        cudaGetLastError();
        memcpy(g_err, keep, sizeof(keep));
    }
    if (range_flag) *range_flag = flag_host;
    cudaSetDevice(prev);
    return rc;
}

// ----------------------------------------------------------------------------------------------
// Allocation-free z-buffer render (no counterpart in the reference).
// For callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it:
// the h*w*8-byte z-buffer that tcw_pc_render_zbuffer takes as caller workspace lives in a file-scope device array instead.
// Every GPU that loads the library holds TCW_ZB_STATIC_MAX_PIXELS * 8 bytes (8 MiB) until exit.  All calls on one GPU share
// the array and it keeps the last call's contents: each call re-fills the part it uses with the "empty" key first, and two
// calls that overlap on the GPU (different streams, no ordering between them) race.
// ----------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned long long g_zb_static[TCW_ZB_STATIC_MAX_PIXELS];

// In host code `g_zb_static` is a host-side stand-in, not a device address: the kernels get what cudaGetSymbolAddress
// returns, looked up once per device with that device current.
static const int ZB_MAX_DEVICES = 64;
static void* g_zb_static_ptr[ZB_MAX_DEVICES];

static int zb_static_address(tcw_ctx* ctx, void** out) {
    int cur = -1;
    CUDA_TRY(cudaGetDevice(&cur));
    if (cur != ctx->device) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer_static: the context's device must be the current device");
    if (cur < 0 || cur >= ZB_MAX_DEVICES) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer_static: device index beyond 63");
    void* p = g_zb_static_ptr[cur];
    if (!p) {
        const cudaError_t e = cudaGetSymbolAddress(&p, g_zb_static);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(TCW_E_CUDA, "tcw_pc_render_zbuffer_static: cudaGetSymbolAddress of the static z-buffer failed");
        }
        g_zb_static_ptr[cur] = p;      // (two threads racing here store the same value)
    }
    *out = p;
    return TCW_OK;
}

TCW_API int tcw_pc_render_zbuffer_static(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                                         long long n_points, int h, int w, int point_size,
                                         float* out_frame, float* out_mask, float* out_depth, void* cuda_stream) {
    if (!ctx) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer_static: ctx is NULL");
    if (h <= 0 || w <= 0) return fail(TCW_E_ARG, "tcw_pc_render_zbuffer_static: bad image size");
    if ((long long)h * (long long)w > (long long)TCW_ZB_STATIC_MAX_PIXELS)
        return fail(TCW_E_WORKSPACE, "tcw_pc_render_zbuffer_static: h*w exceeds TCW_ZB_STATIC_MAX_PIXELS (1048576); use tcw_pc_render_zbuffer with a caller workspace");
    void* zb = nullptr;
    const int rc = zb_static_address(ctx, &zb);
    if (rc) return rc;
    // same kernels, same winner rule as tcw_pc_render_zbuffer; no cudaMalloc / cudaFree / allocating routine on this path
    return pc_zbuffer_impl(ctx, points, colors, camera, n_points, h, w, point_size, 0, out_frame, out_mask, out_depth,
                           zb, (size_t)TCW_ZB_STATIC_MAX_PIXELS * sizeof(unsigned long long), cuda_stream);
}
