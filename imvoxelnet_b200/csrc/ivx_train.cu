// Training-mode pieces of the 3D necks (fp32, CUDA cores), channels-last [rows][C] tensors.
//
// In training the reference's BatchNorm3d uses batch statistics and autograd
// runs through every conv / BN / ReLU / add of mmdet3d/models/necks/imvoxelnet.py
// (forward_train, detectors/imvoxelnet.py:82-87).  A fused layer
//     z = conv(x) (+ bias);  u = z * sc + sh  (sc, sh from the batch moments of z)
//     y = post * ( relu(u [+ res]) [+ res] )
// needs, besides the convolution kernels of ivx_conv3d.cu:
//   channel_moments   per-channel  S1 = sum a,  S2 = sum a*b   (batch mean / variance of z with
//                     a = b = z;  sum g and sum g*z for the BN backward), accumulated in double
//   affine_act        the element-wise tail of the layer, forward and backward
//   lincomb           dz = A_c * g + B_c * z + C_c   (the BatchNorm backward in closed form)
//   conv wgrad        dW[tap][ci][co] = sum_m x[in(m, tap)][ci] * dz[m][co]
//   upsample backward transpose of the trilinear x2 interpolation
#include "ivx_common.cuh"

namespace ivx {

// --------------------------------------------------------------------------
// per-channel moments: acc[c] += sum_r a[r][c],  acc[C + c] += sum_r a[r][c] * b[r][c]
// --------------------------------------------------------------------------
__global__ void __launch_bounds__(256) channel_moments_kernel(const float4 *__restrict__ a,
                                                              const float4 *__restrict__ b, long long rows,
                                                              int c4, int rows_per_block,
                                                              double *__restrict__ acc) {
    __shared__ float4 s1[256], s2[256];
    const int lanes = 256 / c4;              // row lanes of the block (c4 <= 256)
    const int col = threadIdx.x % c4, lane = threadIdx.x / c4;
    const long long r0 = (long long)blockIdx.x * rows_per_block;
    const long long r1 = min(rows, r0 + rows_per_block);
    float4 t1 = make_float4(0.f, 0.f, 0.f, 0.f), t2 = t1;
    if (lane < lanes) {
        for (long long r = r0 + lane; r < r1; r += lanes) {
            const float4 va = __ldg(a + r * c4 + col);
            const float4 vb = __ldg(b + r * c4 + col);
            t1.x += va.x, t1.y += va.y, t1.z += va.z, t1.w += va.w;
            t2.x = fmaf(va.x, vb.x, t2.x), t2.y = fmaf(va.y, vb.y, t2.y);
            t2.z = fmaf(va.z, vb.z, t2.z), t2.w = fmaf(va.w, vb.w, t2.w);
        }
    }
    s1[threadIdx.x] = t1, s2[threadIdx.x] = t2;
    __syncthreads();
    if (threadIdx.x < c4) {                  // one thread per float4 column finishes in double
        double d1[4] = {0, 0, 0, 0}, d2[4] = {0, 0, 0, 0};
        for (int l = 0; l < lanes; ++l) {
            const float4 u = s1[l * c4 + threadIdx.x], v = s2[l * c4 + threadIdx.x];
            d1[0] += u.x, d1[1] += u.y, d1[2] += u.z, d1[3] += u.w;
            d2[0] += v.x, d2[1] += v.y, d2[2] += v.z, d2[3] += v.w;
        }
        const int C = c4 * 4;
        for (int e = 0; e < 4; ++e) {
            atomicAdd(acc + threadIdx.x * 4 + e, d1[e]);
            atomicAdd(acc + C + threadIdx.x * 4 + e, d2[e]);
        }
    }
}

// --------------------------------------------------------------------------
// element-wise tail of a fused layer
//   u = z * sc + sh;  if (res && !after) u += res;  v = relu ? max(u, 0) : u;
//   if (res && after) v += res;  y = post * v
// --------------------------------------------------------------------------
struct TailArgs {
    const float4 *z, *res;
    const float *sc, *sh;       // per channel, or nullptr (1 / 0)
    int relu, after;
    float post;
    long long n4;               // rows * c4
    int c4;
};

__device__ __forceinline__ float4 tail_u(const TailArgs &t, long long i, float4 z) {
    const int c = (int)(i % t.c4) * 4;
    float4 u = z;
    if (t.sc) {
        const float4 s = __ldg(reinterpret_cast<const float4 *>(t.sc + c));
        u.x *= s.x, u.y *= s.y, u.z *= s.z, u.w *= s.w;
    }
    if (t.sh) {
        const float4 s = __ldg(reinterpret_cast<const float4 *>(t.sh + c));
        u.x += s.x, u.y += s.y, u.z += s.z, u.w += s.w;
    }
    if (t.res && !t.after) {
        const float4 r = __ldg(t.res + i);
        u.x += r.x, u.y += r.y, u.z += r.z, u.w += r.w;
    }
    return u;
}

__global__ void __launch_bounds__(256) affine_act_kernel(const __grid_constant__ TailArgs t,
                                                         float4 *__restrict__ y) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < t.n4;
         i += (long long)gridDim.x * blockDim.x) {
        float4 v = tail_u(t, i, __ldg(t.z + i));
        if (t.relu) v.x = fmaxf(v.x, 0.f), v.y = fmaxf(v.y, 0.f), v.z = fmaxf(v.z, 0.f), v.w = fmaxf(v.w, 0.f);
        if (t.res && t.after) {
            const float4 r = __ldg(t.res + i);
            v.x += r.x, v.y += r.y, v.z += r.z, v.w += r.w;
        }
        v.x *= t.post, v.y *= t.post, v.z *= t.post, v.w *= t.post;
        y[i] = v;
    }
}

// g = d(loss)/du  (what flows into the affine part and, if the residual was added before the
// activation, into the residual);  g_after = post * dy (into a residual added after it)
__global__ void __launch_bounds__(256) affine_act_backward_kernel(const __grid_constant__ TailArgs t,
                                                                  const float4 *__restrict__ dy,
                                                                  float4 *__restrict__ g,
                                                                  float4 *__restrict__ g_after) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < t.n4;
         i += (long long)gridDim.x * blockDim.x) {
        float4 d = __ldg(dy + i);
        d.x *= t.post, d.y *= t.post, d.z *= t.post, d.w *= t.post;
        if (g_after) g_after[i] = d;
        if (t.relu) {
            const float4 u = tail_u(t, i, __ldg(t.z + i));
            d.x = u.x > 0.f ? d.x : 0.f, d.y = u.y > 0.f ? d.y : 0.f;
            d.z = u.z > 0.f ? d.z : 0.f, d.w = u.w > 0.f ? d.w : 0.f;
        }
        g[i] = d;
    }
}

// out = A_c * g + B_c * z + C_c
__global__ void __launch_bounds__(256) lincomb_kernel(const float4 *__restrict__ g, const float4 *__restrict__ z,
                                                      const float *__restrict__ A, const float *__restrict__ B,
                                                      const float *__restrict__ Cc, long long n4, int c4,
                                                      float4 *__restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % c4) * 4;
        const float4 a = __ldg(reinterpret_cast<const float4 *>(A + c));
        const float4 b = __ldg(reinterpret_cast<const float4 *>(B + c));
        const float4 k = __ldg(reinterpret_cast<const float4 *>(Cc + c));
        const float4 vg = __ldg(g + i), vz = __ldg(z + i);
        out[i] = make_float4(fmaf(a.x, vg.x, fmaf(b.x, vz.x, k.x)), fmaf(a.y, vg.y, fmaf(b.y, vz.y, k.y)),
                             fmaf(a.z, vg.z, fmaf(b.z, vz.z, k.z)), fmaf(a.w, vg.w, fmaf(b.w, vz.w, k.w)));
    }
}

// --------------------------------------------------------------------------
// weight gradient: dW[tap][ci][co] += sum over rows m of  x[a_row(m, tap)][ci] * dz[b_row(m, tap)][co]
//   convolution: rows are output voxels, a_row = the input voxel the tap reads (or none), b_row = m
//   transposed (kernel == stride): rows are input voxels, a_row = m, b_row = the output voxel of the tap
// 64 x 64 (ci x co) tile per CTA, rows split over gridDim.z, fp32 atomics at the end.
// --------------------------------------------------------------------------
struct WgradArgs {
    ivx_conv3d_params p;
    const float *x, *dz;
    float *dw;
    long long m_total, w_tap_stride, w_ci_stride;
    int rows_per_split, co_tiles;
};

__device__ __forceinline__ void wg_decode(long long m, const int32_t d[3], int &b, int &x, int &y, int &z) {
    z = (int)(m % d[2]);
    long long t = m / d[2];
    y = (int)(t % d[1]);
    t /= d[1];
    x = (int)(t % d[0]);
    b = (int)(t / d[0]);
}

__global__ void __launch_bounds__(256) conv3d_wgrad_kernel(const __grid_constant__ WgradArgs a) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const ivx_conv3d_params &p = a.p;
    const int tid = threadIdx.x;
    const int ci0 = (blockIdx.x / a.co_tiles) * 64, co0 = (blockIdx.x % a.co_tiles) * 64;
    const int tap = blockIdx.y;
    const int kz = tap % p.kernel[2], kt = tap / p.kernel[2];
    const int ky = kt % p.kernel[1], kx = kt / p.kernel[1];
    const long long r0 = (long long)blockIdx.z * a.rows_per_split;
    const long long r1 = min(a.m_total, r0 + a.rows_per_split);
    const int lrow = tid >> 4, lc = (tid & 15) * 4;      // the (row, 4 channels) this thread stages
    const int ty = tid >> 4, tx = tid & 15;              // 4 ci x 4 co micro-tile
    const int32_t *rd = p.transposed ? p.in_dims : p.out_dims;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (long long k0 = r0; k0 < r1; k0 += 16) {
        const long long m = k0 + lrow;
        float4 av = make_float4(0.f, 0.f, 0.f, 0.f), bv = av;
        if (m < r1) {
            int b, vx, vy, vz;
            wg_decode(m, rd, b, vx, vy, vz);
            long long arow = m, brow = m;
            if (p.transposed) {
                brow = (((long long)b * p.out_dims[0] + vx * p.stride[0] + kx) * p.out_dims[1] +
                        vy * p.stride[1] + ky) * p.out_dims[2] + vz * p.stride[2] + kz;
            } else {
                const int xi = vx * p.stride[0] - p.pad[0] + kx, yi = vy * p.stride[1] - p.pad[1] + ky;
                const int zi = vz * p.stride[2] - p.pad[2] + kz;
                arow = (xi >= 0 && xi < p.in_dims[0] && yi >= 0 && yi < p.in_dims[1] && zi >= 0 && zi < p.in_dims[2])
                           ? (((long long)b * p.in_dims[0] + xi) * p.in_dims[1] + yi) * p.in_dims[2] + zi
                           : -1;
            }
            if (arow >= 0 && ci0 + lc < p.c_in)
                av = __ldg(reinterpret_cast<const float4 *>(a.x + arow * p.c_in + ci0 + lc));
            if (arow >= 0 && co0 + lc < p.c_out)
                bv = __ldg(reinterpret_cast<const float4 *>(a.dz + brow * p.c_out + co0 + lc));
        }
        __syncthreads();
        *reinterpret_cast<float4 *>(&As[lrow][lc]) = av;
        *reinterpret_cast<float4 *>(&Bs[lrow][lc]) = bv;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const float4 ar = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            const float4 br = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
            const float a4[4] = {ar.x, ar.y, ar.z, ar.w}, b4[4] = {br.x, br.y, br.z, br.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a4[i], b4[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ci = ci0 + ty * 4 + i;
        if (ci >= p.c_in) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int co = co0 + tx * 4 + j;
            if (co < p.c_out) atomicAdd(a.dw + tap * a.w_tap_stride + ci * a.w_ci_stride + co, acc[i][j]);
        }
    }
}

// transpose of upsample2x_trilinear_kernel: every fine voxel hands its gradient to the <= 8 coarse
// voxels it was interpolated from (same indices and weights as the forward kernel)
__global__ void __launch_bounds__(256) upsample2x_trilinear_backward_kernel(
    const float4 *__restrict__ dy, float *__restrict__ dx, int B, int X, int Y, int Z, int C4) {
    const long long total = (long long)B * (2 * X) * (2 * Y) * (2 * Z) * C4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c4 = (int)(i % C4);
        long long t = i / C4;
        const int zo = (int)(t % (2 * Z));
        t /= 2 * Z;
        const int yo = (int)(t % (2 * Y));
        t /= 2 * Y;
        const int xo = (int)(t % (2 * X));
        const int b = (int)(t / (2 * X));
        int i0[3], i1[3];
        float l1[3];
        const int o[3] = {xo, yo, zo}, n[3] = {X, Y, Z};
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            float s = 0.5f * ((float)o[d] + 0.5f) - 0.5f;
            s = s < 0.f ? 0.f : s;
            i0[d] = (int)s;
            i1[d] = i0[d] + (i0[d] < n[d] - 1 ? 1 : 0);
            l1[d] = s - (float)i0[d];
        }
        const float4 g = __ldg(dy + i);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int xi = (k & 4) ? i1[0] : i0[0], yi = (k & 2) ? i1[1] : i0[1], zi = (k & 1) ? i1[2] : i0[2];
            const float w = ((k & 4) ? l1[0] : 1.f - l1[0]) * ((k & 2) ? l1[1] : 1.f - l1[1]) *
                            ((k & 1) ? l1[2] : 1.f - l1[2]);
            if (w == 0.f) continue;
            float4 *dst = reinterpret_cast<float4 *>(dx) + ((((long long)b * X + xi) * Y + yi) * Z + zi) * C4 + c4;
            atomicAdd(dst, make_float4(w * g.x, w * g.y, w * g.z, w * g.w));
        }
    }
}

}  // namespace ivx

using namespace ivx;

static unsigned ew_grid(long long n) {
    long long gx = (n + 255) / 256;
    return (unsigned)(gx > kNumSMs * 32 ? kNumSMs * 32 : (gx < 1 ? 1 : gx));
}

extern "C" int ivx_channel_moments_ndhwc(const float *a, const float *b, int64_t rows, int n_channels,
                                         double *acc, void *stream) {
    IVX_REQUIRE(a && b && acc, "NULL argument");
    IVX_REQUIRE(rows > 0 && n_channels > 0 && n_channels % 4 == 0 && n_channels <= 1024,
                "channels must be a multiple of 4, at most 1024");
    const int c4 = n_channels / 4;
    long long blocks = kNumSMs * 8;
    int rpb = (int)((rows + blocks - 1) / blocks);
    if (rpb < 64) rpb = 64;
    blocks = (rows + rpb - 1) / rpb;
    channel_moments_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(a), reinterpret_cast<const float4 *>(b), rows, c4, rpb, acc);
    IVX_CHECK_LAUNCH("channel_moments launch");
    return IVX_OK;
}

static int fill_tail(TailArgs &t, const float *z, const float *scale, const float *shift, const float *residual,
                     int relu, int after, float post, int64_t rows, int n_channels) {
    IVX_REQUIRE(z != nullptr && rows > 0 && n_channels > 0 && n_channels % 4 == 0,
                "z must be given, channels a multiple of 4");
    t.z = reinterpret_cast<const float4 *>(z);
    t.res = reinterpret_cast<const float4 *>(residual);
    t.sc = scale, t.sh = shift, t.relu = relu, t.after = after, t.post = post;
    t.c4 = n_channels / 4;
    t.n4 = rows * t.c4;
    return IVX_OK;
}

extern "C" int ivx_affine_act_ndhwc(const float *z, const float *scale, const float *shift,
                                    const float *residual, int relu, int residual_after_act, float post_scale,
                                    int64_t rows, int n_channels, float *y, void *stream) {
    TailArgs t;
    int rc = fill_tail(t, z, scale, shift, residual, relu, residual_after_act, post_scale, rows, n_channels);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(y != nullptr, "NULL output");
    affine_act_kernel<<<ew_grid(t.n4), 256, 0, (cudaStream_t)stream>>>(t, reinterpret_cast<float4 *>(y));
    IVX_CHECK_LAUNCH("affine_act launch");
    return IVX_OK;
}

extern "C" int ivx_affine_act_backward_ndhwc(const float *grad_y, const float *z, const float *scale,
                                             const float *shift, const float *residual, int relu,
                                             int residual_after_act, float post_scale, int64_t rows,
                                             int n_channels, float *grad_u, float *grad_after, void *stream) {
    TailArgs t;
    int rc = fill_tail(t, z, scale, shift, residual, relu, residual_after_act, post_scale, rows, n_channels);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(grad_y && grad_u, "NULL argument");
    affine_act_backward_kernel<<<ew_grid(t.n4), 256, 0, (cudaStream_t)stream>>>(
        t, reinterpret_cast<const float4 *>(grad_y), reinterpret_cast<float4 *>(grad_u),
        reinterpret_cast<float4 *>(grad_after));
    IVX_CHECK_LAUNCH("affine_act_backward launch");
    return IVX_OK;
}

extern "C" int ivx_channel_lincomb_ndhwc(const float *g, const float *z, const float *A, const float *B,
                                         const float *C, int64_t rows, int n_channels, float *out,
                                         void *stream) {
    IVX_REQUIRE(g && z && A && B && C && out, "NULL argument");
    IVX_REQUIRE(rows > 0 && n_channels > 0 && n_channels % 4 == 0, "channels must be a multiple of 4");
    const int c4 = n_channels / 4;
    lincomb_kernel<<<ew_grid(rows * c4), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(g), reinterpret_cast<const float4 *>(z), A, B, C, rows * c4, c4,
        reinterpret_cast<float4 *>(out));
    IVX_CHECK_LAUNCH("lincomb launch");
    return IVX_OK;
}

extern "C" int ivx_conv3d_wgrad_ndhwc(const ivx_conv3d_params *p, const float *x, const float *grad_z,
                                      float *grad_w, void *stream) {
    IVX_REQUIRE(p && x && grad_z && grad_w, "NULL argument");
    IVX_REQUIRE(p->batch > 0 && p->c_in > 0 && p->c_out > 0 && p->c_in % 4 == 0 && p->c_out % 4 == 0,
                "channel counts must be multiples of 4");
    WgradArgs a;
    a.p = *p;
    a.x = x, a.dz = grad_z, a.dw = grad_w;
    const int taps = p->kernel[0] * p->kernel[1] * p->kernel[2];
    const int32_t *rd = p->transposed ? p->in_dims : p->out_dims;
    a.m_total = (long long)p->batch * rd[0] * rd[1] * rd[2];
    if (p->transposed) a.w_tap_stride = p->c_out, a.w_ci_stride = (long long)taps * p->c_out;
    else a.w_tap_stride = (long long)p->c_in * p->c_out, a.w_ci_stride = p->c_out;
    const int ci_tiles = (p->c_in + 63) / 64;
    a.co_tiles = (p->c_out + 63) / 64;
    const long long ctas = (long long)ci_tiles * a.co_tiles * taps;
    long long splits = (kNumSMs * 6 + ctas - 1) / ctas;                 // ~6 CTAs per SM in total
    if (splits > (a.m_total + 255) / 256) splits = (a.m_total + 255) / 256;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    a.rows_per_split = (int)(((a.m_total + splits - 1) / splits + 15) / 16 * 16);
    splits = (a.m_total + a.rows_per_split - 1) / a.rows_per_split;
    dim3 grid((unsigned)(ci_tiles * a.co_tiles), (unsigned)taps, (unsigned)splits);
    conv3d_wgrad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    IVX_CHECK_LAUNCH("conv3d wgrad launch");
    return IVX_OK;
}

extern "C" int ivx_upsample2x_trilinear_backward_ndhwc(const float *grad_y, float *grad_x, int batch,
                                                       int n_channels, const int32_t *in_dims, void *stream) {
    IVX_REQUIRE(grad_y && grad_x && in_dims, "NULL argument");
    IVX_REQUIRE(batch > 0 && n_channels > 0 && n_channels % 4 == 0, "channels must be a multiple of 4");
    const long long total = 8ll * batch * in_dims[0] * in_dims[1] * in_dims[2] * (n_channels / 4);
    upsample2x_trilinear_backward_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(grad_y), grad_x, batch, in_dims[0], in_dims[1], in_dims[2],
        n_channels / 4);
    IVX_CHECK_LAUNCH("upsample backward launch");
    return IVX_OK;
}
