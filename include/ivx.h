/*
 * libivx -- B200 (sm_100a) kernels for ImVoxelNet's image->voxel path, C ABI.
 *
 * This is the drop-in boundary: plain pointers, sizes and a CUDA stream; no
 * torch / C++ types.  The reference (SamsungLabs/imvoxelnet, a Python/PyTorch
 * code base) has no FFI of its own for this path -- it runs ~40 ATen kernels
 * per scene.  Each entry point below therefore cites the reference *Python*
 * code it replaces; paths are relative to the reference tree.  The Python
 * mirror that binds these symbols with ctypes is imvoxelnet_b200/_lib.py;
 * INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - Every data pointer is a DEVICE pointer on the current device unless the
 *     parameter comment says "host".  The caller owns all memory; the library
 *     never allocates or frees device memory and never synchronises.
 *   - All work is enqueued on `stream` (a cudaStream_t passed as void*; NULL is
 *     the legacy default stream).
 *   - Return value: IVX_OK or a negative ivx_status; ivx_last_error() gives the
 *     message of the last failure on the calling thread.  No C++ exception
 *     crosses this boundary.
 *   - Voxel linear index n = (ix*Y + iy)*Z + iz (Z fastest), as the reference's
 *     meshgrid builds it (mmdet3d/models/detectors/imvoxelnet.py:134-138).
 *   - Strides are in ELEMENTS, not bytes.
 */
#ifndef IVX_H_
#define IVX_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* the library is built with -fvisibility=hidden: only these symbols are exported */
#if defined(__GNUC__)
#define IVX_API __attribute__((visibility("default")))
#else
#define IVX_API
#endif

typedef enum ivx_status {
    IVX_OK = 0,
    IVX_ERR_INVALID_ARGUMENT = -1,
    IVX_ERR_UNSUPPORTED = -2,
    IVX_ERR_CUDA = -3
} ivx_status;

/* element type of a volume written by the library */
typedef enum ivx_dtype {
    IVX_F32 = 0,
    IVX_BF16 = 1,
    IVX_U8 = 2,   /* masks only */
    IVX_F32X2 = 3 /* fp32 split into two tf32 values, channels (hi | lo): ivx_split_tf32_ndhwc */
} ivx_dtype;

/* what ivx_backproject_scenes writes */
typedef enum ivx_reduce {
    IVX_REDUCE_MEAN = 0,    /* sum / count, 0 where count == 0  (:70-74)          */
    IVX_REDUCE_PARTIAL = 1  /* raw fp32 sum + int32 count (view-sharded multi-GPU) */
} ivx_reduce;

#define IVX_MAX_SCENES_PER_LAUNCH 64 /* larger batches are split internally */
#define IVX_MAX_VIEWS 1024

/* Library / build identification ("libivx 0.1 sm_100a ..."). Never NULL. */
IVX_API const char *ivx_version(void);

/* Message of the last error raised on this thread ("" if none). */
IVX_API const char *ivx_last_error(void);

/*
 * Regular voxel grid, replaces get_points (detectors/imvoxelnet.py:132-141):
 * coordinate of voxel (i,j,k) on axis a is  rn(rn(idx_a * voxel_size[a]) +
 * new_origin[a]),  new_origin = origin - n_voxels/2 * voxel_size computed by
 * the caller in fp32 (:139).  The kernels rebuild the coordinates in registers
 * from this description, so no [3,N] tensor has to exist.
 */
typedef struct ivx_grid {
    int32_t n_voxels[3];  /* X, Y, Z */
    float voxel_size[3];
} ivx_grid;

/*
 * ivx_project_indices -- pixel indices and valid mask of every voxel-view.
 * Replaces the index arithmetic of backproject (detectors/imvoxelnet.py
 * :148-154): homogeneous projection with the K=4 fma chain of torch.bmm,
 * IEEE division, round-half-even, in-image & positive-depth test.
 *
 *   points      [3][N] fp32 voxel coordinates, or NULL to build them from
 *               `grid` + `new_origin` (host, 3 floats)
 *   proj        [V][12] fp32 row-major 3x4 matrices
 *   x, y        [V][N] int32 out (either may be NULL).  On lanes where valid==0
 *               the value is unspecified (the reference's own value there is
 *               platform dependent: x86 and CUDA cast inf/NaN differently).
 *   valid       [V][N] uint8 out (0/1)
 */
IVX_API int ivx_project_indices(const float *points, const ivx_grid *grid,
                        const float *new_origin /* host[3] */, int64_t n_points,
                        const float *proj, int n_views, int height, int width,
                        int32_t *x, int32_t *y, uint8_t *valid, void *stream);

/*
 * ivx_backproject_scenes -- the fused hot path: for a batch of B scenes with V
 * views each, project every voxel through every view, gather the C-channel
 * feature of the rounded pixel, and reduce over views in registers.
 * Replaces, per scene, backproject (detectors/imvoxelnet.py:145-160) and the
 * view reduction of extract_feat (:70-74); the batch replaces the Python loop
 * and torch.stack of :57-78.  The reference's [V,C,N] temporary never exists.
 *
 *   feat          fp32 feature maps, logical shape [B][V][C][Hf][Wf], addressed
 *                 through feat_strides (host, 5 element strides: b, v, c, h, w).
 *                 Channels-last maps (stride_c == 1, C % 4 == 0, 16-byte aligned
 *                 rows) take the vectorised kernel; anything else (e.g. the
 *                 reference's NCHW slice of :69) takes the strided kernel.
 *   hw            host [B][2] int32: un-padded (height, width) of each scene's
 *                 map, img_shape // stride (:67-68).  Pixels outside are invalid.
 *                 PRECONDITION: 0 < height <= Hf and 0 < width <= Wf of the
 *                 tensor behind `feat` -- the library sees only strides and
 *                 cannot check it; a larger hw reads (or, in the backward
 *                 entry point, atomically adds) out of bounds.  The Python
 *                 mirror (ops.backproject_scenes) checks it against the tensor.
 *   proj          [B][V][12] fp32
 *   points        [3][N] fp32 shared by all scenes, or NULL to use `grid` with
 *                 new_origin (host [B][3]; scenes may have different origins)
 *   volume        out, element (b, c, n) at  b*vol_strides[0] + c*vol_strides[1]
 *                 + n*vol_strides[2]  (host, 3 element strides): {C*N, N, 1} is
 *                 the reference's [B,C,X,Y,Z]; {N*C, 1, C} is channels-last
 *                 (NDHWC) which the 3D-neck kernels consume and which stores
 *                 fully coalesced.  vol_dtype IVX_BF16 is allowed with MEAN only.
 *   valid         [B][N] uint8 out (count > 0), may be NULL
 *   count         [B][N] int32 out (number of valid views), may be NULL;
 *                 required for IVX_REDUCE_PARTIAL
 */
IVX_API int ivx_backproject_scenes(const float *feat, const int64_t *feat_strides /* host[5] */,
                           int n_scenes, int n_views, int n_channels,
                           const int32_t *hw /* host[B][2] */, const float *proj,
                           const float *points, const ivx_grid *grid,
                           const float *new_origin /* host[B][3] */, int64_t n_points,
                           void *volume, const int64_t *vol_strides /* host[3] */,
                           int vol_dtype, int reduce, uint8_t *valid, int32_t *count,
                           void *stream);

/*
 * ivx_backproject_scenes_bf16 -- ivx_backproject_scenes on bf16 feature maps
 * (channels-last, C % 8 == 0, 16-byte aligned pixels): what a bf16 2D backbone emits.  The reference
 * keeps fp32 maps (its FPN output, detectors/imvoxelnet.py:50-51); with bf16
 * maps the gather moves half the bytes.  Values are widened exactly and summed
 * in fp32, so the result equals ivx_backproject_scenes on the widened maps.
 */
IVX_API int ivx_backproject_scenes_bf16(const void *feat_bf16, const int64_t *feat_strides /* host[5] */,
                                        int n_scenes, int n_views, int n_channels,
                                        const int32_t *hw /* host[B][2] */, const float *proj,
                                        const float *points, const ivx_grid *grid,
                                        const float *new_origin /* host[B][3] */, int64_t n_points,
                                        void *volume, const int64_t *vol_strides /* host[3] */,
                                        int vol_dtype, int reduce, uint8_t *valid, int32_t *count,
                                        void *stream);

/*
 * ivx_backproject_scenes_backward -- gradient of ivx_backproject_scenes with
 * respect to the features: what autograd computes in the reference through
 * the masked gather of detectors/imvoxelnet.py:157 and the view mean of
 * :70-74 (training, forward_train :82-87).  For every hit (view v, voxel n):
 *   grad_feat[b, v, :, y, x] += grad_volume[b, :, n] / count[b, n]
 * (IVX_REDUCE_PARTIAL: no division).  grad_feat must be zero-filled by the
 * caller; it and grad_volume are fp32, channels-last (stride_c == 1,
 * C % 4 == 0), addressed like feat / volume of the forward call.  Duplicated
 * pixels are accumulated with atomic vector adds (summation order not fixed).
 */
IVX_API int ivx_backproject_scenes_backward(const float *grad_volume,
                                            const int64_t *gvol_strides /* host[3] */, float *grad_feat,
                                            const int64_t *feat_strides /* host[5] */, int n_scenes,
                                            int n_views, int n_channels, const int32_t *hw /* host[B][2] */,
                                            const float *proj, const float *points, const ivx_grid *grid,
                                            const float *new_origin /* host[B][3] */, int64_t n_points,
                                            int reduce, void *stream);

/*
 * ivx_backproject_views -- strict drop-in for the reference function
 * backproject(features, points, projection) (detectors/imvoxelnet.py:145-160):
 * per-view volumes [V][C][N] fp32 (zeros where invalid) and valid [V][N].
 * volume element (v, c, n) is at v*vol_strides[0] + c*vol_strides[1] +
 * n*vol_strides[2].  feat_strides: host, 4 element strides (v, c, h, w).
 */
IVX_API int ivx_backproject_views(const float *feat, const int64_t *feat_strides /* host[4] */,
                          int n_views, int n_channels, int height, int width,
                          const float *proj, const float *points, const ivx_grid *grid,
                          const float *new_origin /* host[3] */, int64_t n_points,
                          float *volume, const int64_t *vol_strides /* host[3] */,
                          uint8_t *valid, void *stream);

/*
 * ivx_finalize_mean -- second half of the view-sharded (multi-GPU) path: after
 * the ranks' IVX_REDUCE_PARTIAL sums and counts have been all-reduced, divide
 * and zero the unseen voxels (detectors/imvoxelnet.py:72-74).
 *   sum     [B][N][C] fp32 (channels-last partial sums), count [B][N] int32
 *   volume  out, addressed like ivx_backproject_scenes' volume; valid [B][N]
 */
IVX_API int ivx_finalize_mean(const float *sum, const int32_t *count, int n_scenes,
                      int n_channels, int64_t n_points, void *volume,
                      const int64_t *vol_strides /* host[3] */, int vol_dtype,
                      uint8_t *valid, void *stream);

/*
 * ivx_finalize_mean_peers -- ivx_finalize_mean fused with its collective: the
 * partial sums / counts of all `world` ranks are read in place, either through
 * the NVSwitch multicast alias of the buffers (mc_sum / mc_count non-NULL:
 * multimem.ld_reduce, the switch does the addition) or from each rank's copy
 * over NVLink (peer_sums / peer_counts: host arrays of `world` device
 * pointers, summed in rank order), then divided and masked
 * (detectors/imvoxelnet.py:70-74).  The buffers must be peer-addressable
 * (e.g. torch symmetric memory) and every rank must have finished -- and made
 * visible -- its partial kernel before any rank calls this (a barrier over the
 * ranks on the stream); layout as for ivx_finalize_mean.  The reference has
 * no multi-GPU form of this path at all.
 */
IVX_API int ivx_finalize_mean_peers(const float *const *peer_sums /* host[world] */,
                                    const int32_t *const *peer_counts /* host[world] */, int world,
                                    const float *mc_sum, const int32_t *mc_count, int n_scenes,
                                    int n_channels, int64_t n_points, void *volume,
                                    const int64_t *vol_strides /* host[3] */, int vol_dtype,
                                    uint8_t *valid, int32_t *count_out, void *stream);

/*
 * ivx_repack_nchw_to_nhwc -- layout pass for callers that hold the reference's
 * NCHW maps: dst[i][y][x][c] = src[i][c][y][x] for i < n_images (tiled through
 * shared memory, coalesced on both sides).  src_strides: host, 4 element
 * strides (image, c, h, w); dst is dense [n_images][H][W][C].
 * Replaces nothing in the reference (it gathers 4-byte elements from NCHW at
 * detectors/imvoxelnet.py:157); it exists so the gather can be 128-bit wide.
 */
IVX_API int ivx_repack_nchw_to_nhwc(const float *src, const int64_t *src_strides /* host[4] */,
                            int n_images, int n_channels, int height, int width,
                            float *dst, void *stream);

/*
 * ivx_backproject_scenes_nchw -- ivx_backproject_scenes for callers that hold
 * the maps the way the reference does: NCHW, possibly a strided slice
 * (detectors/imvoxelnet.py:50-51, :69).  The maps are first re-laid to
 * channels-last in `workspace` (one tiled transpose over the region the
 * scenes use, max over scenes of hw), then the vectorised kernel runs on the
 * workspace; results are identical to ivx_backproject_scenes on the same
 * values.  Needs C % 4 == 0 (IVX_ERR_UNSUPPORTED otherwise: pass the NCHW
 * strides to ivx_backproject_scenes instead, which takes the strided kernel).
 *
 *   workspace        device scratch, 16-byte aligned, owned by the caller; its
 *                    contents are undefined afterwards
 *   workspace_bytes  must be >= ivx_backproject_nchw_workspace_bytes(...)
 * All other arguments as in ivx_backproject_scenes.
 */
IVX_API int64_t ivx_backproject_nchw_workspace_bytes(int n_scenes, int n_views, int n_channels,
                                                     const int32_t *hw /* host[B][2] */);

IVX_API int ivx_backproject_scenes_nchw(const float *feat, const int64_t *feat_strides /* host[5] */,
                                        int n_scenes, int n_views, int n_channels,
                                        const int32_t *hw /* host[B][2] */, const float *proj,
                                        const float *points, const ivx_grid *grid,
                                        const float *new_origin /* host[B][3] */, int64_t n_points,
                                        void *volume, const int64_t *vol_strides /* host[3] */,
                                        int vol_dtype, int reduce, uint8_t *valid, int32_t *count,
                                        void *workspace, int64_t workspace_bytes, void *stream);

/*
 * ivx_backproject_scenes_nchw_noalloc -- ivx_backproject_scenes_nchw without a
 * workspace argument.  It has no counterpart in the reference.
 *
 * Who it is for: code that cannot obtain device memory when it calls us -- a
 * plugin whose host application has given all GPU memory to its own pool
 * (cudaMalloc then fails even though that pool has room), or a serving /
 * real-time thread that may not take the driver's allocation lock nor trigger
 * the device-wide synchronisation that cudaMalloc and cudaFree bring.  The
 * function calls no cudaMalloc / cudaFree and no allocating library routine;
 * inputs and outputs are still the caller's buffers.
 *
 * The channels-last copy of the maps goes to a fixed array inside the library
 * of IVX_NOALLOC_SCRATCH_BYTES (32 MiB), which every GPU that loads the
 * library keeps until the process exits.  Requests that need more
 * (ivx_backproject_nchw_workspace_bytes(...) > IVX_NOALLOC_SCRATCH_BYTES) are
 * refused with IVX_ERR_UNSUPPORTED and nothing is launched.
 *
 * First call on each device: the address of the array on that device is
 * looked up (with that device current) and remembered; make this first call
 * before starting any stream capture on that device.  With CUDA's default lazy
 * module loading that first call is also when the driver loads this part of
 * the library onto the device and sets the 32 MiB aside, so make it at
 * start-up, BEFORE the host application hands the rest of the GPU memory to
 * its own pool -- afterwards the load itself can fail for lack of memory.
 * cudaDeviceReset() discards the array together with everything else on the
 * device and leaves the remembered address stale: do not call this function
 * again on a device that has been reset in this process.
 *
 * All calls on one device share the array and it keeps what the last call
 * left there.  Nothing needs clearing (each call overwrites every element it
 * later reads), but two calls that overlap on the GPU -- different streams, or
 * different host threads without ordering -- race on it: keep calls on one
 * device ordered, e.g. on a single stream.
 *
 * Arguments and results as ivx_backproject_scenes_nchw.
 */
#define IVX_NOALLOC_SCRATCH_BYTES (32 * 1024 * 1024)

IVX_API int ivx_backproject_scenes_nchw_noalloc(
    const float *feat, const int64_t *feat_strides /* host[5] */, int n_scenes, int n_views,
    int n_channels, const int32_t *hw /* host[B][2] */, const float *proj, const float *points,
    const ivx_grid *grid, const float *new_origin /* host[B][3] */, int64_t n_points, void *volume,
    const int64_t *vol_strides /* host[3] */, int vol_dtype, int reduce, uint8_t *valid,
    int32_t *count, void *stream);

/* ------------------------------ 3D neck ------------------------------ */

/*
 * One convolution layer of a 3D neck with everything that follows it folded
 * in.  Replaces an nn.Conv3d / nn.ConvTranspose3d + nn.BatchNorm3d (eval) +
 * residual add + nn.ReLU group of mmdet3d/models/necks/imvoxelnet.py (e.g.
 * BasicBlock3d.forward :212-230, BasicBlock3dV2.forward :248-260, _make_block
 * :48-53, _make_up_block :56-64, _get_conv :109-114, ConditionalProjection
 * :289-294 followed by the (x + y) / 2 of :367):
 *
 *   v = conv(x) * scale + shift            (BN folded; bias folded into shift)
 *   if (residual && !residual_after_act) v += residual
 *   if (relu) v = max(v, 0)
 *   if (residual &&  residual_after_act) v += residual
 *   y = v * post_scale
 *
 * Activations are channels-last: x is [batch][in_dims X,Y,Z][c_in], y and
 * residual are [batch][out_dims][c_out], dense.  (kX,kY,kZ) of the reference's
 * weight [c_out][c_in][kD][kH][kW] correspond to (X,Y,Z) of the volume.
 */
typedef struct ivx_conv3d_params {
    int32_t batch, c_in, c_out;
    int32_t in_dims[3], out_dims[3];
    int32_t kernel[3], stride[3], pad[3];
    int32_t transposed;         /* 1: ConvTranspose3d with kernel == stride, pad 0 (:57) */
    int32_t relu;
    int32_t residual_after_act;
    float post_scale;
    int32_t in_dtype, out_dtype; /* ivx_dtype of x, and of y / residual */
} ivx_conv3d_params;

/*
 * ivx_conv3d_ndhwc -- fp32-accumulating CUDA-core implicit GEMM for every layer
 * shape (the 1e-5 parity path).  weight: fp32, [kX*kY*kZ][c_in][c_out]; for
 * transposed layers [c_in][kX*kY*kZ][c_out].  scale / shift: fp32 [c_out] or
 * NULL (1 / 0).  c_in and c_out must be multiples of 4.
 */
IVX_API int ivx_conv3d_ndhwc(const ivx_conv3d_params *params, const void *x, const float *weight,
                             const float *scale, const float *shift, const void *residual, void *y,
                             void *stream);

/*
 * ivx_conv3d_ndhwc_bf16_tc -- the same layer on the 5th-generation tensor cores:
 * implicit GEMM with tcgen05.mma (bf16 operands, fp32 accumulation in TMEM),
 * activations and weights staged by TMA (the zero padding is the TMA's
 * out-of-bounds fill), fused epilogue as above, bf16 in and out (1e-2 mode).
 * weight: bf16, [c_out][kX*kY*kZ][c_in]; for transposed layers (kernel ==
 * stride, stored through a strided TMA map) [kX*kY*kZ][c_out][c_in] and no
 * residual.  Returns IVX_ERR_UNSUPPORTED for shapes it does not take (c_in or
 * c_out not a multiple of 64, non-bf16 tensors): callers then use
 * ivx_conv3d_ndhwc.
 *
 * workspace: optional device scratch for split-K.  Layers with too few output
 * tiles to fill the GPU (the 10x10x4 / 20x20x8 tails of the indoor necks)
 * spread each tile's k-steps over several CTAs, which add fp32 partials into
 * the workspace; the last CTA of a tile applies the epilogue.  It must hold
 * ivx_conv3d_tc_workspace_bytes(params) bytes (0 = this layer does not
 * split), be 16-byte aligned and ZERO-FILLED before its first use; every call
 * leaves it zero-filled again, so one buffer can serve a layer forever (but
 * not two calls that overlap on the GPU).  NULL or too small: no split.
 */
IVX_API int64_t ivx_conv3d_tc_workspace_bytes(const ivx_conv3d_params *params);

IVX_API int ivx_conv3d_ndhwc_bf16_tc(const ivx_conv3d_params *params, const void *x,
                                     const void *weight_bf16, const float *scale, const float *shift,
                                     const void *residual, void *y, void *workspace,
                                     int64_t workspace_bytes, void *stream);

/*
 * ivx_conv3d_ndhwc_tf32_tc -- the same kernel for fp32 tensors with tf32 math
 * (tcgen05.mma kind::tf32: 10-bit mantissa operands, rounded by the TMA, fp32
 * accumulation): what PyTorch / cuDNN do for fp32 convolutions by default on
 * this hardware (torch.backends.cudnn.allow_tf32).  About 1e-3 of the fp32
 * result: a third precision between the 1e-5 CUDA-core path and bf16.
 * x, y, residual: fp32 channels-last; weight: fp32 in the layout of
 * ivx_conv3d_ndhwc_bf16_tc; c_in % 32 == 0, c_out % 64 == 0; workspace as
 * above (ivx_conv3d_tc_workspace_bytes with in_dtype = out_dtype = IVX_F32).
 * Used by the "tf32" precision of the necks and by the training path.
 */
IVX_API int ivx_conv3d_ndhwc_tf32_tc(const ivx_conv3d_params *params, const float *x, const float *weight,
                                     const float *scale, const float *shift, const float *residual, float *y,
                                     void *workspace, int64_t workspace_bytes, void *stream);

/*
 * ivx_conv3d_ndhwc_f32x3_tc -- the same layer with fp32-grade results on the
 * tensor cores ("3xTF32"): what the reference computes, an fp32 nn.Conv3d
 * (necks/imvoxelnet.py:116-120 is not autocast and no config sets fp16; the
 * pinned torch 1.6 has no TF32), within 1e-5.  Every fp32 value is the exact
 * sum of a tf32 number hi and a remainder lo; the kernel accumulates
 * x_lo*w_hi + x_hi*w_lo + x_hi*w_hi in fp32 (TMEM) as three passes of the tf32
 * kernel above -- the dropped x_lo*w_lo and the rounding of lo are each about
 * 2^-22 of a product.
 *   params         in_dtype = IVX_F32X2, out_dtype = IVX_F32; c_in is the layer's
 *                  channel count (c_in % 32 == 0, c_out % 64 == 0)
 *   x_split        [batch][in_dims][2*c_in] fp32, written by ivx_split_tf32_ndhwc
 *   weight_split   fp32 [c_out][2][kX*kY*kZ][c_in]: the tf32_tc layout with every
 *                  row holding its hi taps, then its lo taps; transposed layers
 *                  [kX*kY*kZ][c_out][2][c_in].  hi = w rounded to tf32, lo = w - hi.
 *   residual, y    plain fp32 [batch][out_dims][c_out]
 *   workspace      as for ivx_conv3d_ndhwc_bf16_tc (ivx_conv3d_tc_workspace_bytes
 *                  with these params)
 *
 * ivx_split_tf32_ndhwc -- x [rows][C] fp32 -> y [rows][2*C]: hi = rna_tf32(x),
 * lo = rna_tf32(x - hi); C % 4 == 0.  One coalesced pass (4 B read, 8 B written
 * per element).
 */
IVX_API int ivx_conv3d_ndhwc_f32x3_tc(const ivx_conv3d_params *params, const float *x_split,
                                      const float *weight_split, const float *scale, const float *shift,
                                      const float *residual, float *y, void *workspace,
                                      int64_t workspace_bytes, void *stream);
IVX_API int ivx_split_tf32_ndhwc(const float *x, float *y, int64_t rows, int n_channels, void *stream);
/* ivx_conv3d_ndhwc_tf32_tc_view (below) in the 3xTF32 mode: fp32-grade data gradients of strided layers,
 * one input phase per call, written into a strided view of the gradient (training in fp32). */
IVX_API int ivx_conv3d_ndhwc_f32x3_tc_view(const ivx_conv3d_params *params, const float *x_split,
                                           const float *weight_split, const float *scale, const float *shift,
                                           float *y, const int64_t *y_strides /* host[4] */, void *workspace,
                                           int64_t workspace_bytes, void *stream);

/*
 * ivx_upsample2x_trilinear_ndhwc -- F.interpolate(x, scale_factor=2,
 * mode='trilinear', align_corners=False) of EncoderDecoder.forward
 * (necks/imvoxelnet.py:359) on a channels-last [batch][X][Y][Z][C] tensor;
 * y is [batch][2X][2Y][2Z][C], same dtype.
 */
IVX_API int ivx_upsample2x_trilinear_ndhwc(const void *x, void *y, int batch, int n_channels,
                                           const int32_t *in_dims /* host[3] */, int dtype,
                                           void *stream);

/* ----------------------- 3D neck, training mode (fp32) ----------------------- */

/*
 * In training (forward_train, detectors/imvoxelnet.py:82-87) BatchNorm3d uses
 * batch statistics and autograd runs through every layer of
 * necks/imvoxelnet.py.  A fused layer is then computed as
 *   z = conv(x) + bias                      ivx_conv3d_ndhwc (scale NULL, shift = bias, no ReLU)
 *   S1, S2 = sum z, sum z*z per channel     ivx_channel_moments_ndhwc  -> batch mean / variance
 *   y = post * (relu(z*sc + sh [+res]) [+res])   ivx_affine_act_ndhwc, sc/sh from the batch moments
 * and differentiated with
 *   g (= dL/du), g_after                    ivx_affine_act_backward_ndhwc
 *   sum g, sum g*z per channel              ivx_channel_moments_ndhwc   -> d gamma, d beta
 *   dz = A_c*g + B_c*z + C_c                ivx_channel_lincomb_ndhwc   (BatchNorm backward, closed form)
 *   dx                                      ivx_conv3d_dgrad_ndhwc (transposed layers: ivx_conv3d_ndhwc)
 *   dW                                      ivx_conv3d_wgrad_ndhwc
 * All tensors fp32, channels-last, dense: [rows][C] with rows = batch * X * Y * Z.
 */

/* acc[c] += sum_r a[r][c];  acc[C + c] += sum_r a[r][c] * b[r][c]   (acc: 2*C doubles, caller-zeroed) */
IVX_API int ivx_channel_moments_ndhwc(const float *a, const float *b, int64_t rows, int n_channels,
                                      double *acc, void *stream);

/* y = post_scale * (act(z * scale + shift [+ residual]) [+ residual]); scale / shift / residual may be NULL */
IVX_API int ivx_affine_act_ndhwc(const float *z, const float *scale, const float *shift,
                                 const float *residual, int relu, int residual_after_act,
                                 float post_scale, int64_t rows, int n_channels, float *y, void *stream);

/* grad_u = gradient w.r.t. u = z*scale + shift (also that of a residual added before the activation);
 * grad_after (may be NULL) = post_scale * grad_y, the gradient of a residual added after it */
IVX_API int ivx_affine_act_backward_ndhwc(const float *grad_y, const float *z, const float *scale,
                                          const float *shift, const float *residual, int relu,
                                          int residual_after_act, float post_scale, int64_t rows,
                                          int n_channels, float *grad_u, float *grad_after, void *stream);

/* out[r][c] = A[c] * g[r][c] + B[c] * z[r][c] + C[c] */
IVX_API int ivx_channel_lincomb_ndhwc(const float *g, const float *z, const float *A, const float *B,
                                      const float *C, int64_t rows, int n_channels, float *out,
                                      void *stream);

/* grad_x [batch][in_dims][c_in] of a (non-transposed) layer from grad_y [batch][out_dims][c_out];
 * weight_t: fp32 [kX*kY*kZ][c_out][c_in].  params describe the FORWARD layer. */
IVX_API int ivx_conv3d_dgrad_ndhwc(const ivx_conv3d_params *params, const float *grad_y,
                                   const float *weight_t, float *grad_x, void *stream);

/* grad_w += x^T * grad_z in the layout of ivx_conv3d_ndhwc's weight ([taps][c_in][c_out];
 * transposed layers [c_in][taps][c_out]); grad_w must be zero-filled by the caller (fp32 atomics) */
IVX_API int ivx_conv3d_wgrad_ndhwc(const ivx_conv3d_params *params, const float *x, const float *grad_z,
                                   float *grad_w, void *stream);

/*
 * ivx_conv3d_ndhwc_tf32_tc_view -- ivx_conv3d_ndhwc_tf32_tc writing into a strided VIEW of a larger
 * channels-last tensor: y_strides (host[4]) are the element strides of y along X, Y, Z and between
 * scenes (channels stay contiguous), out_dims the extent of the view.  out_dims need not follow from
 * in_dims (reads beyond x are zeros).  No residual.  The data gradient of a strided convolution is
 * computed with it, one call per input phase: the inputs i = s*j + r of phase r receive a stride-1
 * convolution of grad_z with the taps that reach them, stored at y = grad_x + r with strides s.
 */
IVX_API int ivx_conv3d_ndhwc_tf32_tc_view(const ivx_conv3d_params *params, const float *x, const float *weight,
                                          const float *scale, const float *shift, float *y,
                                          const int64_t *y_strides /* host[4] */, void *workspace,
                                          int64_t workspace_bytes, void *stream);

/*
 * ivx_conv3d_wgrad_tf32_tc -- the weight gradient on the tensor cores (tf32 math), for plain
 * convolutions with c_in, c_out multiples of 32 (a kernel == stride transposed layer is the plain
 * convolution from its output grid to its input grid with the roles of x and grad_z exchanged):
 *   grad_w[tap][c_out][c_in] += sum_o grad_z[o][c_out] * x[o * stride - pad + tap][c_in]
 * (taps in kX, kY, kZ order; grad_w fp32, zero-filled by the caller; fp32 vector reductions).
 * The reduction runs over voxels, so both operands are taken in the TILED CHANNEL-MAJOR layout that
 * ivx_channel_major_ndhwc writes (runs of 32 voxels along D0, all channels of a run adjacent):
 *   dz_t [batch][D2][D1][ceil(D0 / 32)][c_out][32]
 *   x_t  [k0][batch][I2 + 2 pad2][I1][ceil(D0 / 32)][c_in][32]
 * where permuted axis i is axis axes[i] of (X, Y, Z) as returned by ivx_conv3d_wgrad_tf32_axes,
 * D = out_dims and I = in_dims in that order, k0 = kernel[axes[0]], pad2 = pad[axes[2]].  x_t holds
 * one copy of x per tap position along D0, shifted and sub-sampled and indexed by OUTPUT coordinate:
 * x_t[j][..][o] = x[..][o * stride[axes[0]] - pad[axes[0]] + j], zero outside the volume (the TMA
 * can neither stride the innermost dimension nor start a swizzled box at an unaligned innermost
 * coordinate), and pad2 zero planes on either side of every volume along D2.
 * IVX_ERR_UNSUPPORTED for other layers (callers use ivx_conv3d_wgrad_ndhwc).
 */
/*
 * ivx_channel_major_ndhwc -- those operand copies, from a channels-last tensor
 * x [batch][X][Y][Z][C] (dims = {X, Y, Z}), D_i = axis axes[i]:
 *   out [n_shifts][batch][D2 + 2 halo][D1][ceil(out_len / 32)][C][32]
 *   out[j][b][i2 + halo][i1][o / 32][c][o % 32] = x[b][..][c] at position o * step + shift0 + j along D0
 * zero where that position is outside the volume, for o >= out_len and in the halo planes.
 * grad_z: n_shifts = 1, shift0 = 0, step = 1, out_len = its extent along D0, halo = 0.
 * x: n_shifts = kernel[axes[0]], shift0 = -pad[axes[0]], step = stride[axes[0]],
 * out_len = out_dims[axes[0]], halo = pad[axes[2]].
 */
IVX_API int ivx_channel_major_ndhwc(const float *x, int32_t batch, const int32_t *dims /* host[3] */,
                                    int32_t n_channels, const int32_t *axes /* host[3] */, int32_t n_shifts,
                                    int32_t shift0, int32_t step, int32_t out_len, int32_t halo, float *out,
                                    void *stream);
IVX_API int ivx_conv3d_wgrad_tf32_axes(const ivx_conv3d_params *params, int32_t *axes /* host[3] */);
IVX_API int ivx_conv3d_wgrad_tf32_tc(const ivx_conv3d_params *params, const float *x_t, const float *dz_t,
                                     float *grad_w, void *stream);

/* grad_x [batch][X][Y][Z][C] += transpose of ivx_upsample2x_trilinear_ndhwc applied to grad_y
 * [batch][2X][2Y][2Z][C]; grad_x must be zero-filled by the caller */
/* ivx_conv3d_wgrad_tf32_tc at fp32 grade (3xTF32): x_t / dz_t are the tiled channel-major copies
 * (ivx_channel_major_ndhwc with 2 * C channels) of the SPLIT tensors written by ivx_split_tf32_ndhwc;
 * params carry the layer's own c_in / c_out.  Used by fp32 training. */
IVX_API int ivx_conv3d_wgrad_f32x3_tc(const ivx_conv3d_params *params, const float *x_t_split,
                                      const float *dz_t_split, float *grad_w, void *stream);
IVX_API int ivx_upsample2x_trilinear_backward_ndhwc(const float *grad_y, float *grad_x, int batch,
                                                    int n_channels, const int32_t *in_dims /* host[3] */,
                                                    void *stream);

/* ----------------------- dense-head tail (next row f-2) ----------------------- */

/*
 * The reference's dense heads (mmdet3d/models/dense_heads/imvoxel_head_v2.py,
 * imvoxel_head.py) end with three small 3x3x3 convolutions per scale
 * (centerness 1, box 6 or 7, classes n; _init_layers, imvoxel_head_v2.py:45-49).
 * The host mirror (imvoxelnet_b200/heads.py) runs those that share an input as
 * ONE ivx_conv3d_* call with the output channels concatenated (zero-padded to
 * the kernel's granularity); the entry points below turn that channels-last
 * result into the tensors the reference returns and prepare its top-k.
 */

/*
 * out [batch][c_n][spatial] fp32 = channels [c_off, c_off + c_n) of y [batch][spatial][c_total]
 * (dtype IVX_F32 or IVX_BF16); the first n_exp channels of the slice become
 * exp(scale[0] * v) -- torch.exp(scale(reg_final[:, :6])), imvoxel_head_v2.py:306-308, 447;
 * scale: device pointer to the head's Scale parameter (NULL = 1).
 */
IVX_API int ivx_head_split_ndhwc(const void *y, int32_t dtype, int32_t batch, int64_t spatial,
                                 int32_t c_total, int32_t c_off, int32_t c_n, int32_t n_exp,
                                 const float *scale, float *out, void *stream);

/*
 * nn.Upsample(size=out_dims, mode='trilinear')(valid).round().bool()
 * (imvoxel_head_v2.py:91-93, 224-226; imvoxel_head.py:113-115): mask [batch][X][Y][Z]
 * with 0/1 entries (mask_dtype IVX_F32, or IVX_U8 for a bool tensor) ->
 * out [batch][x][y][z] bool.  align_corners = False, source index and weights as
 * ATen computes them in fp32, round-half-to-even.
 */
IVX_API int ivx_resize_mask_trilinear(const void *mask, int32_t mask_dtype, int32_t batch,
                                      const int32_t in_dims[3] /* host */,
                                      const int32_t out_dims[3] /* host */, uint8_t *out, void *stream);

/*
 * scores [batch][spatial][n_classes] = sigmoid(cls) * sigmoid(centerness) * valid and
 * best [batch][spatial] = max over classes (imvoxel_head_v2.py:262-270), from the
 * reference-layout tensors centerness [batch][1][spatial], cls [batch][n_classes][spatial],
 * valid [batch][1][spatial] bool.
 */
IVX_API int ivx_head_scores(const float *centerness, const float *cls, const uint8_t *valid, int32_t batch,
                            int64_t spatial, int32_t n_classes, float *scores, float *best, void *stream);

/* ----------------------- projection set-up on the device (next row f-4) ----------------------- */

/*
 * ImVoxelNet._compute_projection (detectors/imvoxelnet.py:114-129) for a batch:
 *   proj[b][v] = K'[b] @ E[b][v][:3],  K'[b] = K[b] with rows 0-1 divided by ratio[b]
 * with the reference's CPU rounding (fp32 division; products accumulated as
 * rn(k0*e0) then two fused multiply-adds), so proj is bit-identical to the
 * tensor the reference builds on the host and copies over (:61).
 *   intrinsic [B][9]  fp32, the [:3,:3] block of lidar2img['intrinsic'] (:117)
 *   ratio     [B]     fp32, ori_shape[0] / (img_shape[0] / stride) (:118)
 *   extrinsic [B][V][16] fp32 row-major 4x4 world->camera matrices (:126)
 *   proj      [B][V][12] out
 * All pointers are device memory.  The predicted-angles branch (get_extrinsics,
 * :121-124, SUN RGB-D Total at test time) stays on the host: its sin/cos bits
 * are libm's.
 */
IVX_API int ivx_compute_projection(const float *intrinsic, const float *ratio, const float *extrinsic,
                                   int32_t n_scenes, int32_t n_views, float *proj, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* IVX_H_ */
