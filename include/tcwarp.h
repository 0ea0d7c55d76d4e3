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
#define TCW_DETERMINISTIC  2   /* tcw_forward_warp: strict integer accumulation (per-cell maximum weight, then order-free int64 fixed-point sums).
                                  Slow; the DEFAULT path (no mode flag) is already reproducible run to run: it sums every cell in a fixed order
                                  (warp-private shared-memory windows, then an ordered gather) */
#define TCW_CLEAN          4   /* fuse clean_points (models/utils.py:585-626): 5x5 dilation of the hole mask; out_frame/out_mask are then DOUBLE (float64) buffers */
#define TCW_IS_IMAGE       8   /* tcw_splat only: holes = -1, range check, clamp to [-1,1] (models/utils.py:572-582) */
#define TCW_ATOMIC_F32    16   /* tcw_forward_warp: opt-in float reductions into L2 accumulators (the round-1 kernel); the order of the
                                  additions, and with it the last bits of colour / depth, varies from run to run */
#define TCW_HOST_MASK_U8  32   /* tcw_forward_warp_host only: out_mask is a uint8 (n,1,h,w) host buffer (0 / 1), a quarter of the download */

/* bits of *range_flag (status word) */
#define TCW_STATUS_RANGE       1   /* a valid colour left [-1.1, 1.1] before the clamp: the reference's assert (models/utils.py:580-581) */
#define TCW_STATUS_UNORDERED   2   /* a footprint did not fit the unit's slab pool and was reduced with float atomics (order not fixed) */
#define TCW_STATUS_SRC_INDEX   4   /* src_index held a value outside [0, n_src); that unit read source 0 instead */

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
 * [3]=max-weight pre-pass of the integer mode, and resets the log.  ms/launches/units are arrays of 4; units = number of warp units those launches processed.
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
 *   flags: TCW_PER_UNIT_MAX, TCW_CLEAN, and at most one mode flag (TCW_DETERMINISTIC | TCW_ATOMIC_F32).
 */
int tcw_forward_warp(tcw_ctx* ctx,
                     const float* frame, const float* mask1, const float* depth,
                     const int* src_index, const float* M,
                     int n_units, int n_src, int h, int w, int flags,
                     void* out_frame, void* out_mask, float* out_depth, float* out_flow,
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

#ifdef __cplusplus
}
#endif
#endif /* TCWARP_H */
