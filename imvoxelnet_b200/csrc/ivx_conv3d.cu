// 3D-neck building blocks on CUDA cores, channels-last (NDHWC) activations.
//
// The reference's necks (mmdet3d/models/necks/imvoxelnet.py) are stacks of
// nn.Conv3d / nn.ConvTranspose3d + BatchNorm3d (+ residual) + ReLU and, in
// the Atlas encoder-decoder, trilinear x2 upsampling.  In eval mode every BN
// folds into a per-channel scale/shift, so one kernel per conv does
//     y = post_scale * ( act( conv(x) * scale + shift [+ res] ) [+ res] )
// This file holds the fp32-accumulating SIMT implicit-GEMM kernel that serves
// every layer shape (any kernel/stride/pad, 1x1x1, transposed k==s) and is the
// 1e-5 parity path; the tcgen05 tensor-core kernel for the FLOP-heavy 3x3x3
// layers lives in ivx_conv3d_tc.cu.
#include <cuda_bf16.h>

#include "ivx_common.cuh"

namespace ivx {

constexpr int BM = 64;   // output voxels per CTA
constexpr int BN = 64;   // output channels per CTA
constexpr int BK = 16;   // input channels per k-step

struct ConvArgs {
    ivx_conv3d_params p;
    const void *x;
    const float *w;      // [taps][c_in][n_cols] fp32, n_cols = c_out (or taps_t * c_out if transposed)
    const float *scale;  // [c_out] or nullptr (= 1)
    const float *shift;  // [c_out] or nullptr (= 0)
    const void *res;     // like y, or nullptr
    void *y;
    long long m_total;   // GEMM rows: output voxels (input voxels if transposed)
    int n_cols;
    int taps;
    int dgrad;           // gather for the data gradient: source voxel = (row + pad - tap) / stride when divisible
};

template <typename T>
__device__ __forceinline__ float4 load4(const T *p);

template <>
__device__ __forceinline__ float4 load4<float>(const float *p) {
    return __ldg(reinterpret_cast<const float4 *>(p));
}

template <>
__device__ __forceinline__ float4 load4<__nv_bfloat16>(const __nv_bfloat16 *p) {
    const uint2 u = __ldg(reinterpret_cast<const uint2 *>(p));
    const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162 *>(&u.x);
    const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162 *>(&u.y);
    const float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
    return make_float4(a.x, a.y, b.x, b.y);
}

template <typename T>
__device__ __forceinline__ void store4v(T *p, float4 v);

template <>
__device__ __forceinline__ void store4v<float>(float *p, float4 v) {
    *reinterpret_cast<float4 *>(p) = v;
}

template <>
__device__ __forceinline__ void store4v<__nv_bfloat16>(__nv_bfloat16 *p, float4 v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<unsigned *>(&lo);
    u.y = *reinterpret_cast<unsigned *>(&hi);
    *reinterpret_cast<uint2 *>(p) = u;
}

// GEMM row -> (batch, x, y, z) of the row's voxel grid (dims d)
__device__ __forceinline__ void decode_row(long long m, const int32_t d[3], int &b, int &x, int &y,
                                           int &z) {
    z = (int)(m % d[2]);
    long long t = m / d[2];
    y = (int)(t % d[1]);
    t /= d[1];
    x = (int)(t % d[0]);
    b = (int)(t / d[0]);
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) conv3d_simt_kernel(const __grid_constant__ ConvArgs a) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN];
    const ivx_conv3d_params &p = a.p;
    const int tid = threadIdx.x;
    const long long m0 = (long long)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;

    // the row this thread stages into As (4 consecutive input channels of it)
    const int arow = tid >> 2, ak = (tid & 3) * 4;
    const long long am = m0 + arow;
    const bool arow_ok = am < a.m_total;
    int ab = 0, ax = 0, ay = 0, az = 0;
    const int32_t *row_dims = p.transposed ? p.in_dims : p.out_dims;
    if (arow_ok) decode_row(am, row_dims, ab, ax, ay, az);
    // the weight slice this thread stages into Bs
    const int bk = tid >> 4, bc = (tid & 15) * 4;

    const int ty = tid >> 4, tx = tid & 15;  // 4x4 micro-tile: rows ty*4.., cols tx*4..
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const TIn *x = static_cast<const TIn *>(a.x);
    for (int tap = 0; tap < a.taps; ++tap) {
        long long in_off = -1;
        if (arow_ok) {
            if (p.transposed) {
                in_off = (((long long)ab * p.in_dims[0] + ax) * p.in_dims[1] + ay) * p.in_dims[2] + az;
            } else {
                const int kz = tap % p.kernel[2], t2 = tap / p.kernel[2];
                const int ky = t2 % p.kernel[1], kx = t2 / p.kernel[1];
                int xi = ax * p.stride[0] - p.pad[0] + kx;
                int yi = ay * p.stride[1] - p.pad[1] + ky;
                int zi = az * p.stride[2] - p.pad[2] + kz;
                if (a.dgrad) {   // rows are input-gradient voxels; which output voxel used them with this tap?
                    const int qx = ax + p.pad[0] - kx, qy = ay + p.pad[1] - ky, qz = az + p.pad[2] - kz;
                    const bool hit = qx >= 0 && qy >= 0 && qz >= 0 && qx % p.stride[0] == 0 &&
                                     qy % p.stride[1] == 0 && qz % p.stride[2] == 0;
                    xi = hit ? qx / p.stride[0] : -1;
                    yi = qy / p.stride[1];
                    zi = qz / p.stride[2];
                }
                if (xi >= 0 && xi < p.in_dims[0] && yi >= 0 && yi < p.in_dims[1] && zi >= 0 &&
                    zi < p.in_dims[2])
                    in_off = (((long long)ab * p.in_dims[0] + xi) * p.in_dims[1] + yi) * p.in_dims[2] + zi;
            }
        }
        const float *wt = a.w + (long long)tap * p.c_in * a.n_cols;
        for (int c0 = 0; c0 < p.c_in; c0 += BK) {
            float4 av = make_float4(0.f, 0.f, 0.f, 0.f);
            if (in_off >= 0 && c0 + ak < p.c_in) av = load4<TIn>(x + in_off * p.c_in + c0 + ak);
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c0 + bk < p.c_in && n0 + bc < a.n_cols)
                bv = __ldg(reinterpret_cast<const float4 *>(wt + (long long)(c0 + bk) * a.n_cols + n0 + bc));
            __syncthreads();  // previous k-step fully consumed
            As[ak + 0][arow] = av.x;
            As[ak + 1][arow] = av.y;
            As[ak + 2][arow] = av.z;
            As[ak + 3][arow] = av.w;
            *reinterpret_cast<float4 *>(&Bs[bk][bc]) = bv;
            __syncthreads();
#pragma unroll
            for (int k = 0; k < BK; ++k) {
                const float4 ar = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
                const float4 br = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
                const float av4[4] = {ar.x, ar.y, ar.z, ar.w};
                const float bv4[4] = {br.x, br.y, br.z, br.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av4[i], bv4[j], acc[i][j]);
            }
        }
    }

    // epilogue: folded BN, residual, ReLU, scale; 4 consecutive channels per store
    const int col = n0 + tx * 4;
    if (col >= a.n_cols) return;
    int co = col, ttap = 0;
    if (p.transposed) {
        ttap = col / p.c_out;
        co = col - ttap * p.c_out;
    }
    float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.scale) sc = __ldg(reinterpret_cast<const float4 *>(a.scale + co));
    if (a.shift) sh = __ldg(reinterpret_cast<const float4 *>(a.shift + co));
    TOut *y = static_cast<TOut *>(a.y);
    const TOut *res = static_cast<const TOut *>(a.res);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const long long m = m0 + ty * 4 + i;
        if (m >= a.m_total) break;
        long long ovox = m;
        if (p.transposed) {
            int b, xi, yi, zi;
            decode_row(m, p.in_dims, b, xi, yi, zi);
            const int kz = ttap % p.kernel[2], t2 = ttap / p.kernel[2];
            const int ky = t2 % p.kernel[1], kx = t2 / p.kernel[1];
            ovox = (((long long)b * p.out_dims[0] + xi * p.stride[0] + kx) * p.out_dims[1] +
                    yi * p.stride[1] + ky) * p.out_dims[2] + zi * p.stride[2] + kz;
        }
        float4 v = make_float4(fmaf(acc[i][0], sc.x, sh.x), fmaf(acc[i][1], sc.y, sh.y),
                               fmaf(acc[i][2], sc.z, sh.z), fmaf(acc[i][3], sc.w, sh.w));
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (res) r = load4<TOut>(res + ovox * p.c_out + co);
        if (res && !p.residual_after_act) v.x += r.x, v.y += r.y, v.z += r.z, v.w += r.w;
        if (p.relu) v.x = fmaxf(v.x, 0.f), v.y = fmaxf(v.y, 0.f), v.z = fmaxf(v.z, 0.f), v.w = fmaxf(v.w, 0.f);
        if (res && p.residual_after_act) v.x += r.x, v.y += r.y, v.z += r.z, v.w += r.w;
        if (p.post_scale != 1.f) v.x *= p.post_scale, v.y *= p.post_scale, v.z *= p.post_scale, v.w *= p.post_scale;
        store4v<TOut>(y + ovox * p.c_out + co, v);
    }
}

// F.interpolate(scale_factor=2, mode='trilinear', align_corners=False) on NDHWC
// (necks/imvoxelnet.py:359): source index (dst + 0.5) / 2 - 0.5 clamped at 0,
// i.e. weights 0.25 / 0.75 with edge replication, combined in ATen's order.
template <typename T>
__global__ void __launch_bounds__(256) upsample2x_trilinear_kernel(
    const T *__restrict__ x, T *__restrict__ y, int B, int X, int Y, int Z, int C4) {
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
        auto at = [&](int xi, int yi, int zi) {
            return load4<T>(x + ((((long long)b * X + xi) * Y + yi) * Z + zi) * (C4 * 4ll) + c4 * 4);
        };
        auto lerp = [](float4 p, float4 q, float w0, float w1) {
            return make_float4(w0 * p.x + w1 * q.x, w0 * p.y + w1 * q.y, w0 * p.z + w1 * q.z,
                               w0 * p.w + w1 * q.w);
        };
        const float x0 = 1.f - l1[0], y0 = 1.f - l1[1], z0 = 1.f - l1[2];
        const float4 a00 = lerp(at(i0[0], i0[1], i0[2]), at(i0[0], i0[1], i1[2]), z0, l1[2]);
        const float4 a01 = lerp(at(i0[0], i1[1], i0[2]), at(i0[0], i1[1], i1[2]), z0, l1[2]);
        const float4 a10 = lerp(at(i1[0], i0[1], i0[2]), at(i1[0], i0[1], i1[2]), z0, l1[2]);
        const float4 a11 = lerp(at(i1[0], i1[1], i0[2]), at(i1[0], i1[1], i1[2]), z0, l1[2]);
        const float4 v = lerp(lerp(a00, a01, y0, l1[1]), lerp(a10, a11, y0, l1[1]), x0, l1[0]);
        store4v<T>(y + i * 4, v);
    }
}


// Same operation, one thread per INPUT voxel and 16 bytes of channels (8 bf16 / 4 fp32): the eight
// outputs (2i + d) of a voxel all read its 3x3x3 neighbourhood, so the 27 vector loads and the
// z- and y-interpolations are shared between them (the per-output kernel above does 64 loads for
// the same work).  Output o = 2i + d has source index i - 0.25 (d = 0) or i + 0.25 (d = 1):
//   d = 0: 0.25 * in[i-1] + 0.75 * in[i]      (i = 0: ATen clamps the index to 0: 1 * in[0] + 0 * in[1])
//   d = 1: 0.75 * in[i]   + 0.25 * in[min(i+1, n-1)]
// evaluated with the same products and sums, in the same z, y, x order, as the kernel above.
template <int N>
struct ChanVec {
    float v[N];
};

template <typename T>
struct UpVec;
template <>
struct UpVec<float> {
    static constexpr int N = 4;
    __device__ static ChanVec<4> load(const float *p) {
        const float4 f = __ldg(reinterpret_cast<const float4 *>(p));
        return {{f.x, f.y, f.z, f.w}};
    }
    __device__ static void store(float *p, const ChanVec<4> &c) {
        __stcs(reinterpret_cast<float4 *>(p), make_float4(c.v[0], c.v[1], c.v[2], c.v[3]));
    }
};
template <>
struct UpVec<__nv_bfloat16> {
    static constexpr int N = 8;
    __device__ static ChanVec<8> load(const __nv_bfloat16 *p) {
        const uint4 u = __ldg(reinterpret_cast<const uint4 *>(p));
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u);
        ChanVec<8> c;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 f = __bfloat1622float2(h[e]);
            c.v[2 * e] = f.x, c.v[2 * e + 1] = f.y;
        }
        return c;
    }
    __device__ static void store(__nv_bfloat16 *p, const ChanVec<8> &c) {
        uint4 u;
        __nv_bfloat162 *h = reinterpret_cast<__nv_bfloat162 *>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(c.v[2 * e], c.v[2 * e + 1]);
        __stcs(reinterpret_cast<uint4 *>(p), u);
    }
};

template <typename T>
__global__ void __launch_bounds__(128) upsample2x_trilinear_voxel_kernel(
    const T *__restrict__ x, T *__restrict__ y, int B, int X, int Y, int Z, int CV) {
    constexpr int N = UpVec<T>::N;
    using V = ChanVec<N>;
    const long long total = (long long)B * X * Y * Z * CV;
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int cv = (int)(i % CV);
    long long t = i / CV;
    const int zi = (int)(t % Z);
    t /= Z;
    const int yi = (int)(t % Y);
    t /= Y;
    const int xi = (int)(t % X);
    const int b = (int)(t / X);
    // neighbourhood positions {i-1, i, i+1} (clamped) per axis and, per output parity d, which two of
    // them are combined with which weights
    const int xs[3] = {max(xi - 1, 0), xi, min(xi + 1, X - 1)};
    const int ys[3] = {max(yi - 1, 0), yi, min(yi + 1, Y - 1)};
    const int zs[3] = {max(zi - 1, 0), zi, min(zi + 1, Z - 1)};
    // lo: combine positions (0, 1), else (1, 2)
    auto taps = [](int idx, int d, bool &lo, float &wa, float &wb) {
        if (d == 1) lo = false, wa = 0.75f, wb = 0.25f;
        else if (idx > 0) lo = true, wa = 0.25f, wb = 0.75f;
        else lo = false, wa = 1.f, wb = 0.f;
    };
    // w0 * (lo ? p0 : p1) + w1 * (lo ? p1 : p2), selected element-wise so everything stays in registers
    auto lerp3 = [](const V &p0, const V &p1, const V &p2, bool lo, float w0, float w1) {
        V r;
#pragma unroll
        for (int e = 0; e < N; ++e) r.v[e] = w0 * (lo ? p0.v[e] : p1.v[e]) + w1 * (lo ? p1.v[e] : p2.v[e]);
        return r;
    };
    const long long cstride = (long long)CV * N;
    V yz[3][2][2];   // [x position][dy][dz]: interpolated along z, then y
#pragma unroll
    for (int px = 0; px < 3; ++px) {
        V zl[3][2];  // [y position][dz]
#pragma unroll
        for (int py = 0; py < 3; ++py) {
            const T *col = x + ((((long long)b * X + xs[px]) * Y + ys[py]) * Z) * cstride + (long long)cv * N;
            V in[3];
#pragma unroll
            for (int pz = 0; pz < 3; ++pz) in[pz] = UpVec<T>::load(col + zs[pz] * cstride);
#pragma unroll
            for (int dz = 0; dz < 2; ++dz) {
                bool lo;
                float wa, wb;
                taps(zi, dz, lo, wa, wb);
                zl[py][dz] = lerp3(in[0], in[1], in[2], lo, wa, wb);
            }
        }
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
            bool lo;
            float wa, wb;
            taps(yi, dy, lo, wa, wb);
#pragma unroll
            for (int dz = 0; dz < 2; ++dz) yz[px][dy][dz] = lerp3(zl[0][dz], zl[1][dz], zl[2][dz], lo, wa, wb);
        }
    }
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
        bool lo;
        float wa, wb;
        taps(xi, dx, lo, wa, wb);
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dz = 0; dz < 2; ++dz) {
                const V r = lerp3(yz[0][dy][dz], yz[1][dy][dz], yz[2][dy][dz], lo, wa, wb);
                const long long ovox = (((long long)b * (2 * X) + 2 * xi + dx) * (2 * Y) + 2 * yi + dy) * (2 * Z) +
                                       2 * zi + dz;
                UpVec<T>::store(y + ovox * cstride + (long long)cv * N, r);
            }
    }
}

}  // namespace ivx

using namespace ivx;

static int validate_conv(const ivx_conv3d_params *p) {
    IVX_REQUIRE(p != nullptr, "NULL params");
    IVX_REQUIRE(p->batch > 0 && p->c_in > 0 && p->c_out > 0, "batch/c_in/c_out must be positive");
    IVX_REQUIRE(p->c_in % 4 == 0 && p->c_out % 4 == 0, "channel counts must be multiples of 4");
    for (int d = 0; d < 3; ++d) {
        IVX_REQUIRE(p->in_dims[d] > 0 && p->out_dims[d] > 0 && p->kernel[d] > 0 && p->stride[d] > 0 &&
                        p->pad[d] >= 0,
                    "bad geometry on axis %d", d);
        if (p->transposed) {
            IVX_REQUIRE(p->kernel[d] == p->stride[d] && p->pad[d] == 0 &&
                            p->out_dims[d] == p->in_dims[d] * p->stride[d],
                        "transposed convolutions need kernel == stride, pad 0 (axis %d)", d);
        } else {
            const int expect = (p->in_dims[d] + 2 * p->pad[d] - p->kernel[d]) / p->stride[d] + 1;
            IVX_REQUIRE(expect == p->out_dims[d], "axis %d: out_dims %d, expected %d", d, p->out_dims[d],
                        expect);
        }
    }
    IVX_REQUIRE((p->in_dtype == IVX_F32 || p->in_dtype == IVX_BF16) &&
                    (p->out_dtype == IVX_F32 || p->out_dtype == IVX_BF16),
                "unknown dtype");
    return IVX_OK;
}

extern "C" int ivx_conv3d_ndhwc(const ivx_conv3d_params *params, const void *x, const float *weight,
                                const float *scale, const float *shift, const void *residual, void *y,
                                void *stream) {
    int rc = validate_conv(params);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(x && weight && y, "NULL argument");
    ConvArgs a;
    a.p = *params;
    a.x = x, a.w = weight, a.scale = scale, a.shift = shift, a.res = residual, a.y = y;
    const ivx_conv3d_params &p = a.p;
    const int ktaps = p.kernel[0] * p.kernel[1] * p.kernel[2];
    const int32_t *rd = p.transposed ? p.in_dims : p.out_dims;
    a.m_total = (long long)p.batch * rd[0] * rd[1] * rd[2];
    a.n_cols = p.transposed ? ktaps * p.c_out : p.c_out;
    a.taps = p.transposed ? 1 : ktaps;
    a.dgrad = 0;
    const long long gx = (a.m_total + BM - 1) / BM;
    IVX_REQUIRE(gx < (1ll << 31), "too many output voxels");
    dim3 grid((unsigned)gx, (unsigned)((a.n_cols + BN - 1) / BN));
    cudaStream_t s = (cudaStream_t)stream;
    if (p.in_dtype == IVX_F32 && p.out_dtype == IVX_F32)
        conv3d_simt_kernel<float, float><<<grid, 256, 0, s>>>(a);
    else if (p.in_dtype == IVX_BF16 && p.out_dtype == IVX_BF16)
        conv3d_simt_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, s>>>(a);
    else if (p.in_dtype == IVX_F32 && p.out_dtype == IVX_BF16)
        conv3d_simt_kernel<float, __nv_bfloat16><<<grid, 256, 0, s>>>(a);
    else
        conv3d_simt_kernel<__nv_bfloat16, float><<<grid, 256, 0, s>>>(a);
    IVX_CHECK_LAUNCH("conv3d launch");
    return IVX_OK;
}

extern "C" int ivx_upsample2x_trilinear_ndhwc(const void *x, void *y, int batch, int n_channels,
                                              const int32_t *in_dims, int dtype, void *stream) {
    IVX_REQUIRE(x && y && in_dims, "NULL argument");
    IVX_REQUIRE(batch > 0 && n_channels > 0 && n_channels % 4 == 0, "channels must be a multiple of 4");
    IVX_REQUIRE(in_dims[0] > 0 && in_dims[1] > 0 && in_dims[2] > 0, "bad dims");
    IVX_REQUIRE(dtype == IVX_F32 || dtype == IVX_BF16, "unknown dtype");
    cudaStream_t s = (cudaStream_t)stream;
    const long long voxels = (long long)batch * in_dims[0] * in_dims[1] * in_dims[2];
    const int per_thread = dtype == IVX_F32 ? 4 : 8;          // channels in 16 bytes
    if (n_channels % per_thread == 0 && (reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) % 16 == 0) {
        const int cv = n_channels / per_thread;
        const long long blocks = (voxels * cv + 127) / 128;
        IVX_REQUIRE(blocks < (1ll << 31), "too many voxels");
        if (dtype == IVX_F32)
            upsample2x_trilinear_voxel_kernel<float><<<(unsigned)blocks, 128, 0, s>>>(
                static_cast<const float *>(x), static_cast<float *>(y), batch, in_dims[0], in_dims[1], in_dims[2], cv);
        else
            upsample2x_trilinear_voxel_kernel<__nv_bfloat16><<<(unsigned)blocks, 128, 0, s>>>(
                static_cast<const __nv_bfloat16 *>(x), static_cast<__nv_bfloat16 *>(y), batch, in_dims[0],
                in_dims[1], in_dims[2], cv);
        IVX_CHECK_LAUNCH("upsample launch");
        return IVX_OK;
    }
    const long long total = 8ll * voxels * (n_channels / 4);
    long long gx = (total + 255) / 256;
    if (gx > kNumSMs * 32) gx = kNumSMs * 32;
    if (dtype == IVX_F32)
        upsample2x_trilinear_kernel<float><<<(unsigned)gx, 256, 0, s>>>(
            static_cast<const float *>(x), static_cast<float *>(y), batch, in_dims[0], in_dims[1],
            in_dims[2], n_channels / 4);
    else
        upsample2x_trilinear_kernel<__nv_bfloat16><<<(unsigned)gx, 256, 0, s>>>(
            static_cast<const __nv_bfloat16 *>(x), static_cast<__nv_bfloat16 *>(y), batch, in_dims[0],
            in_dims[1], in_dims[2], n_channels / 4);
    IVX_CHECK_LAUNCH("upsample launch");
    return IVX_OK;
}

// fp32 -> (hi | lo) tf32 halves for the 3xTF32 layers: hi = x rounded to tf32 (10 explicit mantissa
// bits; cvt.rna), lo = x - hi (exact in fp32) rounded to tf32 as well, so neither the TMA nor the MMA
// unit can change a bit of what is stored.  Row r of y is [hi(x[r][0..C)) | lo(x[r][0..C))].
// One thread per 16 bytes of input; reads and writes are fully coalesced 128-bit accesses.
__device__ __forceinline__ float tf32_rna(float v) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
    return __uint_as_float(u);
}

__global__ void __launch_bounds__(256) split_tf32_kernel(const float4 *__restrict__ x, float4 *__restrict__ y,
                                                          long long n_vec, int cv) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec;
         i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / cv;
        const int c = (int)(i - row * cv);
        const float4 v = __ldcs(x + i);
        float4 hi, lo;
        hi.x = tf32_rna(v.x), hi.y = tf32_rna(v.y), hi.z = tf32_rna(v.z), hi.w = tf32_rna(v.w);
        lo.x = tf32_rna(v.x - hi.x), lo.y = tf32_rna(v.y - hi.y);
        lo.z = tf32_rna(v.z - hi.z), lo.w = tf32_rna(v.w - hi.w);
        y[row * 2 * cv + c] = hi;
        y[row * 2 * cv + cv + c] = lo;
    }
}

extern "C" int ivx_split_tf32_ndhwc(const float *x, float *y, int64_t rows, int n_channels, void *stream) {
    IVX_REQUIRE(x && y, "NULL argument");
    IVX_REQUIRE(rows > 0 && n_channels > 0 && n_channels % 4 == 0, "channels must be a multiple of 4");
    IVX_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) % 16 == 0,
                "tensors must be 16-byte aligned");
    const int cv = n_channels / 4;
    const long long n_vec = rows * cv;
    long long blocks = (n_vec + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;   // grid-stride: 16 CTAs of 256 threads per SM
    split_tf32_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4 *>(x), reinterpret_cast<float4 *>(y), n_vec, cv);
    IVX_CHECK_LAUNCH("split_tf32 launch");
    return IVX_OK;
}

// Data gradient of a (non-transposed) convolution, fp32: for every input voxel, gather the output
// gradients that used it.  Same kernel as the forward pass with the channel roles exchanged.
extern "C" int ivx_conv3d_dgrad_ndhwc(const ivx_conv3d_params *params, const float *grad_y,
                                      const float *weight_t, float *grad_x, void *stream) {
    int rc = validate_conv(params);
    if (rc != IVX_OK) return rc;
    IVX_REQUIRE(grad_y && weight_t && grad_x, "NULL argument");
    IVX_REQUIRE(!params->transposed, "transposed layers: their data gradient is a plain strided convolution");
    ConvArgs a;
    a.p = *params;
    ivx_conv3d_params &q = a.p;   // the gather runs over the INPUT grid and reads the OUTPUT grid
    q.c_in = params->c_out, q.c_out = params->c_in;
    for (int d = 0; d < 3; ++d) q.in_dims[d] = params->out_dims[d], q.out_dims[d] = params->in_dims[d];
    q.relu = 0, q.residual_after_act = 0, q.post_scale = 1.f;
    q.in_dtype = q.out_dtype = IVX_F32;
    a.x = grad_y, a.w = weight_t, a.scale = nullptr, a.shift = nullptr, a.res = nullptr, a.y = grad_x;
    a.m_total = (long long)q.batch * q.out_dims[0] * q.out_dims[1] * q.out_dims[2];
    a.n_cols = q.c_out;
    a.taps = q.kernel[0] * q.kernel[1] * q.kernel[2];
    a.dgrad = 1;
    const long long gx = (a.m_total + BM - 1) / BM;
    IVX_REQUIRE(gx < (1ll << 31), "too many voxels");
    dim3 grid((unsigned)gx, (unsigned)((a.n_cols + BN - 1) / BN));
    conv3d_simt_kernel<float, float><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    IVX_CHECK_LAUNCH("conv3d dgrad launch");
    return IVX_OK;
}
