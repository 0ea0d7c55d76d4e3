/*
 * tcwarp.h — C ABI of libtcwarp.so: the sm_100a forward re-rendering ("warp") path.
 *
 * The reference has no FFI layer: the boundary it exposes for this path is the Python class
 * `Warper` (reference models/utils.py:213-682).  Each entry point below names the reference
 * method it replaces.  All pointers are plain device (or, for *_host, host) pointers to
 * contiguous float32 data; no torch types cross this boundary.  Every function returns 0 on
 * success or a negative TCW_E_* code; tcw_last_error() gives the message for the calling thread.
 *
 * A "unit" is one (source frame, target pose) pair, i.e. one reference call with b=1.
 *
 * Matrix block M (per unit, 30 floats, row-major):
 *   [0..8]   K1^-1                     (torch.linalg.inv(intrinsic1),   models/utils.py:382)
 *   [9..20]  rows 0..2 of T2 * T1^-1   (torch.bmm(T2, inv(T1)),         models/utils.py:366-368)
 *   [21..29] K2                        (intrinsic2 or intrinsic1.clone(), models/utils.py:364-365)
 * The two tiny inverses are deliberately left to the host (see DESIGN.md "H2"): the reference's
 * LAPACK result feeds floor()/ceil() decisions, so the host mirror computes them with the same
 * torch CPU calls and uploads the result.
 */
#ifndef TCWARP_H
#define TCWARP_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TCW_VERSION 200

/* flags */
#define TCW_PER_UNIT_MAX   1   /* soft-z normaliser log_depth.max() taken per unit (== n_units reference calls with b=1);
                                  without it the max is over the whole call (== one reference call with b=n_units, models/utils.py:492) */
/* 2 is retired: the round-1 int64 fixed-point mode.  The DEFAULT path (no mode flag) sums every target cell in a fixed order
   (warp-private shared-memory windows, then an ordered gather) and is bitwise reproducible */
#define TCW_CLEAN          4   /* fuse clean_points (models/utils.py:585-626): 5x5 dilation of the hole mask; out_frame/out_mask are then DOUBLE (float64) buffers */
#define TCW_IS_IMAGE       8   /* tcw_splat only: holes = -1, range check, clamp to [-1,1] (models/utils.py:572-582) */
#define TCW_ATOMIC_F32    16   /* tcw_forward_warp: opt-in float reductions into L2 accumulators (the round-1 kernel); the order of the
                                  additions, and with it the last bits of colour / depth, varies from run to run */
#define TCW_HOST_MASK_U8  32   /* tcw_forward_warp_host only: out_mask is a uint8 (n,1,h,w) host buffer (0 / 1), a quarter of the download */
#define TCW_MASK_U8       64   /* tcw_forward_warp (default mode, no TCW_CLEAN): out_mask is a uint8 (n,1,h,w) DEVICE buffer (0 / 1).  With
                                  out_frame / out_mask pointing into a peer GPU's memory (tcw_ipc_open) this is the fused form of the
                                  exchange that assembles the condition video on the denoiser's rank: 13 B per pixel over NVLink,
                                  stored by the gather kernel itself (the consumer is reference demo.py:117-148) */
#define TCW_PACK         128   /* tcw_workspace_bytes only: size the workspace for tcw_forward_warp_pack (tcw_forward_warp rejects the flag) */

/* bits of *range_flag (status word) */
#define TCW_STATUS_RANGE       1   /* a valid colour left [-1.1, 1.1] before the clamp: the reference's assert (models/utils.py:580-581) */
#define TCW_STATUS_UNORDERED   2   /* a footprint did not fit the unit's slab pool and was reduced with float atomics (order not fixed) */
#define TCW_STATUS_SRC_INDEX   4
#define TCW_STATUS_BOUNDS      8   /* -DTCW_BOUNDS_CHECK=1 builds only: an access left its allocation and was skipped */   /* src_index held a value outside [0, n_src); that unit read source 0 instead */

/* error codes */
#define TCW_OK              0
#define TCW_E_ARG          -1
#define TCW_E_CUDA         -2
#define TCW_E_WORKSPACE    -3
#define TCW_E_NOMEM        -4

typedef struct tcw_ctx tcw_ctx;

/* Lifetime. Replaces Warper.__init__/get_device (models/utils.py:214-218, 668-682). */
int  tcw_create(int device, tcw_ctx** out);
void tcw_destroy(tcw_ctx* ctx);
int  tcw_version(void);
const char* tcw_last_error(void);
/* number of kernels this library has launched on behalf of ctx since creation (bench evidence) */
long long tcw_launch_count(const tcw_ctx* ctx);

/*
 * Optional per-kernel device timing (CUDA events around every launch, on the launching stream).
 * tcw_profile_read synchronises on the recorded events, sums them per kernel class
 * [0]=pre-pass (default path: project + flow + boxes; opt-in modes: depth-max), [1]=splat, [2]=gather / normalise,
 * [3]=unused, and resets the log.  ms/launches/units are arrays of 4; units = number of warp units those launches processed.
 */
int tcw_profile_enable(tcw_ctx* ctx, int enable);
int tcw_profile_read(tcw_ctx* ctx, double* ms, long long* launches, double* units);

/* Scratch size needed by tcw_forward_warp / tcw_splat for these shapes (device bytes). */
int tcw_workspace_bytes(const tcw_ctx* ctx, int n_units, int h, int w, int flags, size_t* bytes);

/*
 * Replaces Warper.forward_warp(..., twice=False)   (models/utils.py:220-293), batched over units.
 *   frame  (n_src,3,h,w)  depth (n_src,1,h,w)  mask1 (n_src,1,h,w) or NULL (== ones)
 *   src_index[n_units] -> which source frame unit u warps (NULL: identity, requires n_src==n_units)
 *   M (n_units,30) DEVICE pointer, layout above.
 *   out_frame (n_units,3,h,w)  out_mask (n_units,1,h,w)  out_depth (n_units,1,h,w) or NULL
 *   out_flow (n_units,2,h,w) or NULL
 *   range_flag: device int, OR-ed with TCW_STATUS_* bits (bit 0 = the reference's range assert, :580-581); may be NULL.
 *   flags: TCW_PER_UNIT_MAX, TCW_CLEAN, TCW_MASK_U8, and the opt-in mode flag TCW_ATOMIC_F32.
 */
int tcw_forward_warp(tcw_ctx* ctx,
                     const float* frame, const float* mask1, const float* depth,
                     const int* src_index, const float* M,
                     int n_units, int n_src, int h, int w, int flags,
                     void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                     int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * tcw_forward_warp with a destination index (multi-GPU sharding; the reference has one process and one loop,
 * demo.py:77-90): unit u of this call writes unit slot dst_index[u] of out_frame / out_mask / out_depth / out_flow, which hold
 * n_dst units.  A rank can then take ANY subset of a clip's units -- a few frames of every trajectory, so that all ranks cost the
 * same -- and still store each result where it belongs in the assembled clip, in its own memory or (tcw_ipc_open) another GPU's.
 * dst_index: n_units ints on the DEVICE, distinct, in [0, n_dst); an entry outside writes slot 0 and sets TCW_STATUS_SRC_INDEX.
 * Default mode only (no TCW_CLEAN / TCW_ATOMIC_F32); TCW_MASK_U8 allowed.  Everything else as tcw_forward_warp.
 */
int tcw_forward_warp_scatter(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                             const int* src_index, const float* M, const int* dst_index, int n_units, int n_src, int n_dst,
                             int h, int w, int flags, void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                             int* range_flag, void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * Same operation with HOST buffers (pageable or pinned): uploads inputs, runs, downloads outputs,
 * pipelined over internal streams; synchronous on return.  M and src_index are host pointers here.
 * range_flag is a host int (written).  This is the call `bench.py` times as "e2e".
 */
int tcw_forward_warp_host(tcw_ctx* ctx,
                     const float* frame, const float* mask1, const float* depth,
                     const int* src_index, const float* M,
                     int n_units, int n_src, int h, int w, int flags,
                     void* out_frame, void* out_mask, float* out_depth, float* out_flow,
                     int* range_flag);

/* Replaces Warper.compute_transformed_points (models/utils.py:350-420): out (n_units,h,w,3). */
int tcw_project(tcw_ctx* ctx, const float* depth, const int* src_index, const float* M,
                int n_units, int n_src, int h, int w, float* out_points, void* cuda_stream);

/*
 * Replaces Warper.bilinear_splatting (models/utils.py:422-583) for c in 1..4 channels.
 *   frame (b,c,h,w)  mask1 (b,1,h,w)|NULL  depth (b,h,w)  flow (b,2,h,w)  flow_mask (b,1,h,w)|NULL
 *   out (b,c,h,w)  out_mask (b,1,h,w).   flags: TCW_IS_IMAGE, TCW_PER_UNIT_MAX.  Runs the window kernels of the forward warp
 *   (fixed summation order); the soft-z normaliser is taken over the whole call unless TCW_PER_UNIT_MAX is set (:490-492).
 */
int tcw_splat(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
              const float* flow, const float* flow_mask, int b, int c, int h, int w, int flags,
              float* out, float* out_mask, int* range_flag,
              void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * Replaces Warper.clean_points (models/utils.py:585-626) for one (1,c,h,w) image, c in {1,3}:
 *   out_frame (double, c*h*w) = ((x+1)/2 in f32) * (1-m) * 2 - 1,  out_mask (double, h*w) = 1 - m,
 *   m = 5x5 dilation (constant border) of the hole mask (1-mask2 >= 0.5).
 */
int tcw_clean_points(tcw_ctx* ctx, const float* frame, const float* mask2, int c, int h, int w,
                     double* out_frame, double* out_mask, void* cuda_stream);

/*
 * The step right after the warp loop, fused into one pass (reference demo.py:91-100, 117-120 -- inline code the
 * reference repeats in every infer_* method, no function of its own):
 *   out_video (3,n,h,w) = F.interpolate((frame + 1) / 2, (h,w), 'bilinear', align_corners=False).permute(1,0,2,3)
 *   out_masks (1,n,h,w) = (1 - F.interpolate(mask, (h,w), 'nearest').permute(1,0,2,3)) * 255
 *   frame (n,3,H,W) = warped colour in [-1,1], mask (n,1,H,W) = the warp's validity mask; all float32 device pointers.
 */
int tcw_resize_pack(tcw_ctx* ctx, const float* frame, const float* mask, int n, int H, int W, int h, int w,
                    float* out_video, float* out_masks, void* cuda_stream);

/*
 * The warp loop AND the lines after it in one submission (reference demo.py:77-100, 117-120): what tcw_forward_warp followed
 * by tcw_resize_pack produces, bit for bit, without the full-resolution outputs ever reaching the caller (or HBM: the colour
 * and mask of a chunk of units stay in the workspace, L2-resident, until k_resize_pack has read them).  For callers that only
 * keep `warped, mask, _, _` to build the condition video -- every shipped pipeline; at 576x1024 -> 384x672 the results shrink
 * from 28 to 7 bytes per warped pixel, which is also what a multi-GPU exchange then has to carry.
 *   inputs as tcw_forward_warp at warp resolution H x W; flags: 0 or TCW_PER_UNIT_MAX;
 *   out_video (3, n_units, h, w), out_masks (1, n_units, h, w) float32 DEVICE; range_flag as tcw_forward_warp;
 *   workspace >= tcw_workspace_bytes(ctx, n_units, H, W, flags | TCW_PACK, ...).
 */
int tcw_forward_warp_pack(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth,
                          const int* src_index, const float* M, int n_units, int n_src, int H, int W, int h, int w,
                          int flags, float* out_video, float* out_masks, int* range_flag,
                          void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * The fork's global point-cloud renderer (reference notebooks/06_10_25_vggt/warper_point_cloud.py, GlobalPointCloudWarper).
 * tcw_pc_lift              lift_to_3d_pointcloud (:12-77): frame (b,3,h,w), mask1 (b,1,h,w)|NULL, depth (b,1,h,w),
 *                          Minv (b,21) DEVICE = K^-1 (9) | rows 0..2 of T^-1 (12) -> out_points, out_colors (b,h,w,3), both * mask1.
 *                          step > 1: the inputs are a ::step subsample and pixel (x, y) stands for (x*step, y*step) -- with the
 *                          extrinsic itself in the T slot this is VisualizationWarper.extract_3d_points_with_colors
 *                          (notebooks/28_08_25_trajectories/core.py:15-79)
 * tcw_pc_render_zbuffer    render_pointcloud_zbuffer_vectorized_point_size (:522-646): points / colors (N,3) DEVICE, camera =
 *                          21 HOST floats (rows 0..2 of the target extrinsic | intrinsic); every point's point_size footprint,
 *                          (point_size 0 = render_pointcloud_zbuffer_vectorized :445-519: one pixel, clamped into the image),
 *                          nearest entry per pixel (64-bit atomicMin on depth bits | reversed entry index: bit-reproducible);
 *                          out_frame (3,h,w) with -1 in empty pixels, out_mask (h,w), out_depth (h,w)|NULL; workspace >= h*w*8 B
 * tcw_pc_render_zbuffer_first   render_pointcloud_zbuffer_vectorized_fixed (:650-765), the sequential loop with a strict `<`:
 *                          among equal depths the FIRST entry stays (depth bits | entry index); point_size <= 1 = one pixel,
 *                          clamped into the image; same arguments and workspace
 * tcw_pc_render_bilinear   render_from_camera(use_bilinear=True) (:158-256, :351-443), one camera: depth-weighted bilinear
 *                          scatter (float reductions: tolerance parity); weights (N) DEVICE; workspace >= 256 + h*w*16 B
 * tcw_pc_render_disc       render_from_camera(use_bilinear=False) = _splat_points_to_image_with_size (:258-349), one camera:
 *                          every visible point stamped at round(pixel + offset) for the n_splats (<= 128) offsets of the disc,
 *                          weight = point weight * exp(-2 depth / (max depth + 1e-8)) * splat_weights[j]; offsets_xy
 *                          (n_splats x 2 ints: dx, dy) and splat_weights are HOST arrays built as the reference builds them;
 *                          pixels with summed weight <= 1e-6 stay 0; workspace >= 256 + h*w*16 B
 */
int tcw_pc_lift(tcw_ctx* ctx, const float* frame, const float* mask1, const float* depth, const float* Minv,
                int b, int h, int w, int step, float* out_points, float* out_colors, void* cuda_stream);
int tcw_pc_render_zbuffer(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                          long long n_points, int h, int w, int point_size,
                          float* out_frame, float* out_mask, float* out_depth,
                          void* workspace, size_t workspace_bytes, void* cuda_stream);
int tcw_pc_render_zbuffer_first(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                                long long n_points, int h, int w, int point_size,
                                float* out_frame, float* out_mask, float* out_depth,
                                void* workspace, size_t workspace_bytes, void* cuda_stream);
/*
 * tcw_pc_render_zbuffer_static   NO counterpart in the reference: added for callers that
 *                          cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 *                          forbid cudaMalloc).  Same arguments, kernels and results as tcw_pc_render_zbuffer, but without the
 *                          workspace: the h*w*8-byte z-buffer is a file-scope device array inside the library
 *                          (TCW_ZB_STATIC_MAX_PIXELS cells = 8 MiB, held by every GPU that loads the library until exit).
 *                          h*w > TCW_ZB_STATIC_MAX_PIXELS is rejected with TCW_E_WORKSPACE.  The call allocates nothing
 *                          (no cudaMalloc / cudaFree); the context's device must be current.  All calls on one GPU share the
 *                          array: each call re-initialises the part it uses, calls that overlap on the GPU race.
 */
#define TCW_ZB_STATIC_MAX_PIXELS (1 << 20)
int tcw_pc_render_zbuffer_static(tcw_ctx* ctx, const float* points, const float* colors, const float* camera,
                                 long long n_points, int h, int w, int point_size,
                                 float* out_frame, float* out_mask, float* out_depth, void* cuda_stream);
int tcw_pc_render_bilinear(tcw_ctx* ctx, const float* points, const float* colors, const float* weights, const float* camera,
                           long long n_points, int h, int w, float depth_threshold,
                           float* out_frame, float* out_mask, void* workspace, size_t workspace_bytes, void* cuda_stream);
int tcw_pc_render_disc(tcw_ctx* ctx, const float* points, const float* colors, const float* weights, const float* camera,
                       long long n_points, int h, int w, float depth_threshold,
                       const int* offsets_xy, const float* splat_weights, int n_splats,
                       float* out_frame, float* out_mask, void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * The step right before the warp, on the device (reference models/infer.py:79-92, the numpy / torch lines that end
 * DepthCrafterDemo.infer, and the radius of demo.py:539-543):
 *   res (n,h,w,c) float32 = the depth model's raw output (c = 3 channels-last as the pipeline returns it, or 1)
 *   out_depth (n,1,h,w) = clip(10000 * (1 / max(3900 * (mean_c(res) - min) / (max - min), 1e-5)), near, far), bit-identical
 *                         (torch evaluates `10000.0 / t` as t.reciprocal() * 10000.0)
 *   stats (4 floats, device): min, max, out_depth[0,0,h/2,w/2], radius = min(centre * radius_scale, 5)
 * The host reads the 16 bytes of `stats` (pose generation stays on the host, DESIGN.md "H2"); the depth map stays in HBM.
 */
int tcw_depth_post(tcw_ctx* ctx, const float* res, int n, int h, int w, int c, float near_v, float far_v,
                   float radius_scale, float* out_depth, float* stats, void* cuda_stream);

/*
 * Peer-visible device buffers for the multi-GPU exchange (the reference has no collective: every frame is warped on one
 * device, demo.py:77-90; the build shards units over ranks and assembles colour + mask on the denoiser's rank).
 * tcw_ipc_alloc: cudaMalloc on ctx's device + cudaIpcGetMemHandle; `handle` receives TCW_IPC_HANDLE_BYTES bytes to send to
 * the other ranks (any transport).  tcw_ipc_open maps another PROCESS's buffer into this one (peer access over NVLink /
 * NVSwitch); the returned pointer is valid as out_frame / out_mask of tcw_forward_warp on ctx's device.
 */
#define TCW_IPC_HANDLE_BYTES 64
int tcw_ipc_alloc(tcw_ctx* ctx, size_t bytes, void** ptr, unsigned char* handle);
int tcw_ipc_free(tcw_ctx* ctx, void* ptr);
int tcw_ipc_open(tcw_ctx* ctx, const unsigned char* handle, void** ptr);
int tcw_ipc_close(tcw_ctx* ctx, void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* TCWARP_H */
