// Dense-head tail of the neck outputs: what the reference's ImVoxelHead /
// ImVoxelHeadV2 do between the last fused convolution and top-k / NMS.
//   * ivx_head_split_ndhwc      : channels-last conv output -> the reference's [B,c,X,Y,Z] fp32 tensors,
//                                 with exp(scale * v) on the leading distance channels
//                                 (imvoxel_head_v2.py:305-313, 444-449; imvoxel_head.py:326-336, 454-461)
//   * ivx_resize_mask_trilinear : nn.Upsample(size, mode='trilinear')(valid).round().bool()
//                                 (imvoxel_head_v2.py:91-93, 224-226; imvoxel_head.py:113-115)
//   * ivx_head_scores           : sigmoid(cls) * sigmoid(centerness) * valid and its max over classes
//                                 (imvoxel_head_v2.py:262-270)
// All three are HBM-bound passes over small tensors (<= 80x80x32 voxels per scene): one thread per
// voxel, writes coalesced along the voxel axis.
#include <cuda_bf16.h>

#include "ivx_common.cuh"

namespace ivx {

template <typename T>
__device__ __forceinline__ float load_as_float(const T *p);
template <>
__device__ __forceinline__ float load_as_float<float>(const float *p) {
    return __ldg(p);
}
template <>
__device__ __forceinline__ float load_as_float<__nv_bfloat16>(const __nv_bfloat16 *p) {
    return __bfloat162float(*p);
}

// y [B, S, c_total] -> out [B, c_n, S]; channel c < n_exp becomes exp(scale * v)
template <typename T>
__global__ void __launch_bounds__(256) head_split_kernel(const T *__restrict__ y, long long spatial, int c_total,
                                                         int c_off, int c_n, int n_exp,
                                                         const float *__restrict__ scale_ptr,
                                                         float *__restrict__ out) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (s >= spatial) return;
    const float scale = (n_exp > 0 && scale_ptr) ? __ldg(scale_ptr) : 1.0f;
    const T *src = y + ((long long)b * spatial + s) * c_total + c_off;
    float *dst = out + (long long)b * c_n * spatial + s;
    for (int c = 0; c < c_n; ++c) {
        float v = load_as_float<T>(src + c);
        if (c < n_exp) v = expf(__fmul_rn(v, scale));
        dst[(long long)c * spatial] = v;
    }
}

// ATen's area_pixel_compute_source_index (align_corners = False, linear modes): UpSample.h
__device__ __forceinline__ void source_index(float ratio, int dst, int in_size, int &i0, int &i1, float &l0,
                                             float &l1) {
    float src = __fsub_rn(__fmul_rn(ratio, (float)dst + 0.5f), 0.5f);
    if (src < 0.0f) src = 0.0f;
    i0 = min((int)src, in_size - 1);
    i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
    l1 = fminf(fmaxf(__fsub_rn(src, (float)i0), 0.0f), 1.0f);
    l0 = __fsub_rn(1.0f, l1);
}

template <typename T>
__device__ __forceinline__ float mask_value(const T *p) {
    return (float)*p;
}

// in [B, X, Y, Z] (0/1 as T) -> out [B, x, y, z] bool = round_half_even(trilinear) != 0
template <typename T>
__global__ void __launch_bounds__(256) resize_mask_kernel(const T *__restrict__ in, int X, int Y, int Z, int ox,
                                                          int oy, int oz, float rx, float ry, float rz,
                                                          uint8_t *__restrict__ out) {
    const long long n_out = (long long)ox * oy * oz;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= n_out) return;
    const int iz = (int)(i % oz), iy = (int)((i / oz) % oy), ix = (int)(i / ((long long)oz * oy));
    int x0, x1, y0, y1, z0, z1;
    float ax0, ax1, ay0, ay1, az0, az1;
    source_index(rx, ix, X, x0, x1, ax0, ax1);
    source_index(ry, iy, Y, y0, y1, ay0, ay1);
    source_index(rz, iz, Z, z0, z1, az0, az1);
    const T *base = in + (long long)b * X * Y * Z;
    auto at = [&](int x, int y, int z) { return mask_value<T>(base + ((long long)x * Y + y) * Z + z); };
    // innermost axis first, then the middle one, then the outer one (each a two-term weighted sum)
    auto line = [&](int x, int y) { return __fadd_rn(__fmul_rn(az0, at(x, y, z0)), __fmul_rn(az1, at(x, y, z1))); };
    auto plane = [&](int x) { return __fadd_rn(__fmul_rn(ay0, line(x, y0)), __fmul_rn(ay1, line(x, y1))); };
    const float v = __fadd_rn(__fmul_rn(ax0, plane(x0)), __fmul_rn(ax1, plane(x1)));
    out[(long long)b * n_out + i] = rintf(v) != 0.0f ? 1 : 0;
}

__device__ __forceinline__ float sigmoid_f(float v) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-v))); }

// centerness [B,1,S], cls [B,n,S], valid [B,1,S] bool -> scores [B,S,n], best [B,S]
__global__ void __launch_bounds__(256) head_scores_kernel(const float *__restrict__ centerness,
                                                          const float *__restrict__ cls,
                                                          const uint8_t *__restrict__ valid, long long spatial,
                                                          int n_classes, float *__restrict__ scores,
                                                          float *__restrict__ best) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (s >= spatial) return;
    const float ctr = sigmoid_f(__ldg(centerness + (long long)b * spatial + s));
    const float ok = valid[(long long)b * spatial + s] ? 1.0f : 0.0f;
    const float *src = cls + (long long)b * n_classes * spatial + s;
    float *dst = scores + ((long long)b * spatial + s) * n_classes;
    float m = -INFINITY;
    for (int c = 0; c < n_classes; ++c) {
        const float v = __fmul_rn(__fmul_rn(sigmoid_f(__ldg(src + (long long)c * spatial)), ctr), ok);
        dst[c] = v;
        m = fmaxf(m, v);
    }
    best[(long long)b * spatial + s] = m;
}

}  // namespace ivx

using namespace ivx;

extern "C" int ivx_head_split_ndhwc(const void *y, int32_t dtype, int32_t batch, int64_t spatial, int32_t c_total,
                                    int32_t c_off, int32_t c_n, int32_t n_exp, const float *scale, float *out,
                                    void *stream) {
    IVX_REQUIRE(y && out, "NULL argument");
    IVX_REQUIRE(batch > 0 && batch <= 65535 && spatial > 0, "bad batch / spatial size");
    IVX_REQUIRE(c_total > 0 && c_off >= 0 && c_n > 0 && c_off + c_n <= c_total, "channel slice [%d, %d) outside %d",
                c_off, c_off + c_n, c_total);
    IVX_REQUIRE(n_exp >= 0 && n_exp <= c_n, "n_exp must be within the slice");
    IVX_REQUIRE(dtype == IVX_F32 || dtype == IVX_BF16, "unknown dtype");
    dim3 grid((unsigned)((spatial + 255) / 256), (unsigned)batch);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == IVX_F32)
        head_split_kernel<float><<<grid, 256, 0, s>>>((const float *)y, spatial, c_total, c_off, c_n, n_exp, scale,
                                                       out);
    else
        head_split_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16 *)y, spatial, c_total, c_off,
                                                               c_n, n_exp, scale, out);
    IVX_CHECK_LAUNCH("head split launch");
    return IVX_OK;
}

extern "C" int ivx_resize_mask_trilinear(const void *mask, int32_t mask_dtype, int32_t batch,
                                         const int32_t in_dims[3], const int32_t out_dims[3], uint8_t *out,
                                         void *stream) {
    IVX_REQUIRE(mask && out && in_dims && out_dims, "NULL argument");
    IVX_REQUIRE(batch > 0 && batch <= 65535, "bad batch");
    for (int d = 0; d < 3; ++d) IVX_REQUIRE(in_dims[d] > 0 && out_dims[d] > 0, "bad size on axis %d", d);
    IVX_REQUIRE(mask_dtype == IVX_F32 || mask_dtype == IVX_U8, "mask must be float32 or bool/uint8");
    const long long n_out = (long long)out_dims[0] * out_dims[1] * out_dims[2];
    dim3 grid((unsigned)((n_out + 255) / 256), (unsigned)batch);
    cudaStream_t s = (cudaStream_t)stream;
    // ATen: scale = float(input_size) / output_size when only `size` is given
    const float rx = (float)in_dims[0] / (float)out_dims[0], ry = (float)in_dims[1] / (float)out_dims[1],
                rz = (float)in_dims[2] / (float)out_dims[2];
    if (mask_dtype == IVX_F32)
        resize_mask_kernel<float><<<grid, 256, 0, s>>>((const float *)mask, in_dims[0], in_dims[1], in_dims[2],
                                                        out_dims[0], out_dims[1], out_dims[2], rx, ry, rz, out);
    else
        resize_mask_kernel<uint8_t><<<grid, 256, 0, s>>>((const uint8_t *)mask, in_dims[0], in_dims[1], in_dims[2],
                                                          out_dims[0], out_dims[1], out_dims[2], rx, ry, rz, out);
    IVX_CHECK_LAUNCH("mask resize launch");
    return IVX_OK;
}

extern "C" int ivx_head_scores(const float *centerness, const float *cls, const uint8_t *valid, int32_t batch,
                               int64_t spatial, int32_t n_classes, float *scores, float *best, void *stream) {
    IVX_REQUIRE(centerness && cls && valid && scores && best, "NULL argument");
    IVX_REQUIRE(batch > 0 && batch <= 65535 && spatial > 0 && n_classes > 0, "bad sizes");
    dim3 grid((unsigned)((spatial + 255) / 256), (unsigned)batch);
    head_scores_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(centerness, cls, valid, spatial, n_classes, scores,
                                                               best);
    IVX_CHECK_LAUNCH("head scores launch");
    return IVX_OK;
}
